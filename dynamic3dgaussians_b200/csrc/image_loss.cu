// Fused image loss of the training loop:  w_l1 * mean|x - y| + w_ssim * (1 - mean SSIM(x, y)),  value and d/dx.
// Replaces the eager PyTorch code of REF/train.py:183 (0.8 * l1_loss_v1 + 0.2 * (1 - calc_ssim)), REF/helpers.py
// l1_loss_v1, REF/external.py:90-133 (calc_ssim / _ssim: five grouped 11x11 Gaussian convolutions with zero padding,
// sigma 1.5, c1 = 0.01^2, c2 = 0.03^2, mean over all channels and pixels) and their autograd twins.
//
// Two kernels, each one pass over the image with a 32x16 pixel tile + 5 pixel halo staged in shared memory and the
// 11x11 window applied separably (register sliding windows, two outputs per thread in either pass):
//   forward : mu_x, mu_y, E[x^2], E[y^2], E[xy] -> SSIM per pixel; sums of |x - y| and of SSIM (double atomics);
//             the three partial derivatives of SSIM w.r.t. the *convolved* maps (mu_x, E[x^2], E[xy]) are stored;
//   backward: dL/dx(p) = g * [ w_l1 sign(x - y) - w_ssim * ( conv(D_mu)(p) + 2 x(p) conv(D_xx)(p) + y(p) conv(D_xy)(p) ) ] / N
//             (the window is symmetric, so the adjoint of the convolution is the same convolution).
// The gradient feeds dL/dcolor of the rasterizer backward directly; the images are [n_img][H][W] planes (n_img = V * 3).
#include "common.cuh"

namespace gsr {

namespace {

constexpr int LTX = 32, LTY = 16;      // output tile of a CTA (256 threads, two outputs per thread in either pass)
constexpr int HALO = 5;                // window 11
constexpr int WIN = 2 * HALO + 1;
constexpr int LRX = LTX + 2 * HALO;    // staged region: 42 x 26
constexpr int LRY = LTY + 2 * HALO;
constexpr int NTH = 256;
constexpr float SSIM_C1 = 0.01f * 0.01f;
constexpr float SSIM_C2 = 0.03f * 0.03f;

struct Window { float w[WIN]; };

// stage a (LRY x LRX) region of one plane, zero outside the image (conv2d zero padding, external.py:115)
__device__ __forceinline__ void stage(const float* __restrict__ plane, int W, int H, int x0, int y0, float (*s)[LRX + 1]) {
    for (int i = threadIdx.x; i < LRY * LRX; i += NTH) {
        const int ry = i / LRX, rx = i - ry * LRX;
        const int x = x0 - HALO + rx, y = y0 - HALO + ry;
        s[ry][rx] = (x >= 0 && x < W && y >= 0 && y < H) ? plane[(size_t)y * W + x] : 0.0f;
    }
}

// Both passes of the separable window keep a sliding window in REGISTERS: a thread produces two neighbouring outputs from
// 12 staged values per map instead of 2 x 11 (the first version of these kernels was bound by shared-memory loads: 96
// per output pixel, 54 now; every output still accumulates its 11 taps in the same order, so the per-pixel values are
// unchanged -- only the order of the block sums of the loss differs).  640x360x3: forward 24.6 -> 22.2 us, backward 23.4 -> 18.1 us.
__global__ void __launch_bounds__(NTH)
image_loss_forward_kernel(const Window win, const float* __restrict__ img, const float* __restrict__ tgt, int W, int H,
                          float* __restrict__ d_mu, float* __restrict__ d_xx, float* __restrict__ d_xy,
                          double* __restrict__ sums /* [2]: sum |x-y|, sum ssim */) {
    __shared__ float sx[LRY][LRX + 1];
    __shared__ float sy[LRY][LRX + 1];
    __shared__ float sh[5][LRY][LTX + 1];
    __shared__ float s_red[2][NTH / 32];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * LTX, y0 = blockIdx.y * LTY;
    const size_t plane = (size_t)blockIdx.z * W * H;
    stage(img + plane, W, H, x0, y0, sx);
    stage(tgt + plane, W, H, x0, y0, sy);
    __syncthreads();
    // horizontal pass: LRY rows x LTX columns, two neighbouring columns per item
    for (int i = tid; i < LRY * (LTX / 2); i += NTH) {
        const int ry = i / (LTX / 2), cx = 2 * (i - ry * (LTX / 2));
        float xv[WIN + 1], yv[WIN + 1];
#pragma unroll
        for (int k = 0; k < WIN + 1; ++k) { xv[k] = sx[ry][cx + k]; yv[k] = sy[ry][cx + k]; }
#pragma unroll
        for (int o = 0; o < 2; ++o) {
            float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
            for (int k = 0; k < WIN; ++k) {
                const float w = win.w[k];
                const float x = xv[o + k], y = yv[o + k];
                a = fmaf(w, x, a);
                b = fmaf(w, y, b);
                aa = fmaf(w, x * x, aa);
                bb = fmaf(w, y * y, bb);
                ab = fmaf(w, x * y, ab);
            }
            sh[0][ry][cx + o] = a; sh[1][ry][cx + o] = b; sh[2][ry][cx + o] = aa; sh[3][ry][cx + o] = bb;
            sh[4][ry][cx + o] = ab;
        }
    }
    __syncthreads();
    // vertical pass: column tx, rows 2 typ and 2 typ + 1
    const int tx = tid & (LTX - 1), ty0 = 2 * (tid / LTX);
    float col[5][WIN + 1];
#pragma unroll
    for (int m = 0; m < 5; ++m)
#pragma unroll
        for (int k = 0; k < WIN + 1; ++k) col[m][k] = sh[m][ty0 + k][tx];
    float l1 = 0.f, ssim_sum = 0.f;
#pragma unroll
    for (int o = 0; o < 2; ++o) {
        float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
        for (int k = 0; k < WIN; ++k) {
            const float w = win.w[k];
            mu1 = fmaf(w, col[0][o + k], mu1);
            mu2 = fmaf(w, col[1][o + k], mu2);
            e11 = fmaf(w, col[2][o + k], e11);
            e22 = fmaf(w, col[3][o + k], e22);
            e12 = fmaf(w, col[4][o + k], e12);
        }
        const int ty = ty0 + o;
        const int px = x0 + tx, py = y0 + ty;
        if (px < W && py < H) {
            const float s1 = e11 - mu1 * mu1, s2 = e22 - mu2 * mu2, s12 = e12 - mu1 * mu2;      // external.py:118-120
            const float A1 = 2.f * mu1 * mu2 + SSIM_C1, A2 = 2.f * s12 + SSIM_C2;
            const float B1 = mu1 * mu1 + mu2 * mu2 + SSIM_C1, B2 = s1 + s2 + SSIM_C2;
            const float iB1 = 1.f / B1, iB2 = 1.f / B2;
            const float ssim = A1 * A2 * iB1 * iB2;                                               // external.py:125
            // partials w.r.t. the convolved maps (mu1, E[x^2], E[xy]); s1 and s12 depend on mu1 too
            const float dxx = -ssim * iB2;
            const float dxy = 2.f * A1 * iB1 * iB2;
            const float dmu = 2.f * mu2 * (A2 - A1) * iB1 * iB2 - 2.f * mu1 * ssim * iB1 + 2.f * mu1 * ssim * iB2;
            const size_t idx = plane + (size_t)py * W + px;
            d_mu[idx] = dmu;
            d_xx[idx] = dxx;
            d_xy[idx] = dxy;
            l1 += fabsf(sx[ty + HALO][tx + HALO] - sy[ty + HALO][tx + HALO]);
            ssim_sum += ssim;
        }
    }
    // block reduction of the two sums
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        ssim_sum += __shfl_xor_sync(0xffffffffu, ssim_sum, o);
    }
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = l1; s_red[1][tid >> 5] = ssim_sum; }
    __syncthreads();
    if (tid < 2) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < NTH / 32; ++w) t += s_red[tid][w];
        atomicAdd(&sums[tid], (double)t);
    }
}

__global__ void image_loss_finish_kernel(const double* __restrict__ sums, double inv_n, float w_l1, float w_ssim,
                                         float* __restrict__ loss /* [3]: total, l1, ssim */) {
    const double l1 = sums[0] * inv_n, ss = sums[1] * inv_n;
    loss[0] = (float)(w_l1 * l1 + w_ssim * (1.0 - ss));
    loss[1] = (float)l1;
    loss[2] = (float)ss;
}

__global__ void __launch_bounds__(NTH)
image_loss_backward_kernel(const Window win, const float* __restrict__ img, const float* __restrict__ tgt, int W, int H,
                           const float* __restrict__ d_mu, const float* __restrict__ d_xx,
                           const float* __restrict__ d_xy, const float* __restrict__ grad_out /* device scalar or null */,
                           float inv_n, float w_l1, float w_ssim, float* __restrict__ dL_dimg) {
    __shared__ float s0[LRY][LRX + 1];
    __shared__ float s1[LRY][LRX + 1];
    __shared__ float s2[LRY][LRX + 1];
    __shared__ float sh[3][LRY][LTX + 1];
    const int tid = threadIdx.x;
    const int x0 = blockIdx.x * LTX, y0 = blockIdx.y * LTY;
    const size_t plane = (size_t)blockIdx.z * W * H;
    stage(d_mu + plane, W, H, x0, y0, s0);
    stage(d_xx + plane, W, H, x0, y0, s1);
    stage(d_xy + plane, W, H, x0, y0, s2);
    __syncthreads();
    for (int i = tid; i < LRY * (LTX / 2); i += NTH) {
        const int ry = i / (LTX / 2), cx = 2 * (i - ry * (LTX / 2));
        float v0[WIN + 1], v1[WIN + 1], v2[WIN + 1];
#pragma unroll
        for (int k = 0; k < WIN + 1; ++k) { v0[k] = s0[ry][cx + k]; v1[k] = s1[ry][cx + k]; v2[k] = s2[ry][cx + k]; }
#pragma unroll
        for (int o = 0; o < 2; ++o) {
            float a = 0.f, b = 0.f, c = 0.f;
#pragma unroll
            for (int k = 0; k < WIN; ++k) {
                const float w = win.w[k];
                a = fmaf(w, v0[o + k], a);
                b = fmaf(w, v1[o + k], b);
                c = fmaf(w, v2[o + k], c);
            }
            sh[0][ry][cx + o] = a; sh[1][ry][cx + o] = b; sh[2][ry][cx + o] = c;
        }
    }
    __syncthreads();
    const int tx = tid & (LTX - 1), ty0 = 2 * (tid / LTX);
    float col[3][WIN + 1];
#pragma unroll
    for (int m = 0; m < 3; ++m)
#pragma unroll
        for (int k = 0; k < WIN + 1; ++k) col[m][k] = sh[m][ty0 + k][tx];
    const float g = (grad_out ? grad_out[0] : 1.0f) * inv_n;
#pragma unroll
    for (int o = 0; o < 2; ++o) {
        const int px = x0 + tx, py = y0 + ty0 + o;
        if (px >= W || py >= H) continue;
        float cm = 0.f, cxx = 0.f, cxy = 0.f;
#pragma unroll
        for (int k = 0; k < WIN; ++k) {
            const float w = win.w[k];
            cm = fmaf(w, col[0][o + k], cm);
            cxx = fmaf(w, col[1][o + k], cxx);
            cxy = fmaf(w, col[2][o + k], cxy);
        }
        const size_t idx = plane + (size_t)py * W + px;
        const float x = img[idx], y = tgt[idx];
        const float dssim = cm + 2.f * x * cxx + y * cxy;
        const float d = x - y;
        const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);            // torch.abs backward: sign(0) = 0
        dL_dimg[idx] = g * (w_l1 * sgn - w_ssim * dssim);
    }
}

// gaussian(11, 1.5) / sum exactly as REF/external.py:94-96 builds it: exp() in double, stored as float32, normalised in
// float32
Window make_window() {
    Window w;
    float sum = 0.f;
    for (int i = 0; i < WIN; ++i) {
        const double d = (double)(i - WIN / 2);
        w.w[i] = (float)exp(-(d * d) / (2.0 * 1.5 * 1.5));
        sum += w.w[i];
    }
    for (int i = 0; i < WIN; ++i) w.w[i] = w.w[i] / sum;
    return w;
}

size_t loss_layout(int64_t n_img, int W, int H, size_t off[4]) {
    size_t o = 0;
    auto take = [&](size_t b) { size_t r = o; o = align_up(o + b); return r; };
    const size_t plane = (size_t)n_img * W * H * sizeof(float);
    off[0] = take(2 * sizeof(double));
    off[1] = take(plane);
    off[2] = take(plane);
    off[3] = take(plane);
    return o;
}

}  // namespace

size_t image_loss_state_bytes(int64_t n_img, int W, int H) {
    size_t off[4];
    return loss_layout(n_img, W, H, off);
}

int launch_image_loss_forward(int64_t n_img, int W, int H, const float* img, const float* tgt, float w_l1, float w_ssim,
                              void* state, float* loss_out, cudaStream_t s) {
    size_t off[4];
    loss_layout(n_img, W, H, off);
    double* sums = at<double>(state, off[0]);
    GSR_CUDA_TRY(cudaMemsetAsync(sums, 0, 2 * sizeof(double), s));
    const dim3 grid((W + LTX - 1) / LTX, (H + LTY - 1) / LTY, (unsigned)n_img), block(NTH);
    image_loss_forward_kernel<<<grid, block, 0, s>>>(make_window(), img, tgt, W, H, at<float>(state, off[1]),
                                                     at<float>(state, off[2]), at<float>(state, off[3]), sums);
    GSR_LAUNCH_CHECK();
    image_loss_finish_kernel<<<1, 1, 0, s>>>(sums, 1.0 / ((double)n_img * W * H), w_l1, w_ssim, loss_out);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_image_loss_backward(int64_t n_img, int W, int H, const float* img, const float* tgt, float w_l1,
                               float w_ssim, const void* state, const float* grad_out, float* dL_dimg, cudaStream_t s) {
    size_t off[4];
    loss_layout(n_img, W, H, off);
    const dim3 grid((W + LTX - 1) / LTX, (H + LTY - 1) / LTY, (unsigned)n_img), block(NTH);
    image_loss_backward_kernel<<<grid, block, 0, s>>>(make_window(), img, tgt, W, H, at<float>(state, off[1]),
                                                      at<float>(state, off[2]), at<float>(state, off[3]), grad_out,
                                                      (float)(1.0 / ((double)n_img * W * H)), w_l1, w_ssim, dL_dimg);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
