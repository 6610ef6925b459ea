// Fused image loss of the training loop:  w_l1 * mean|x - y| + w_ssim * (1 - mean SSIM(x, y)),  value and d/dx.
// Replaces the eager PyTorch code of REF/train.py:183 (0.8 * l1_loss_v1 + 0.2 * (1 - calc_ssim)), REF/helpers.py
// l1_loss_v1, REF/external.py:90-133 (calc_ssim / _ssim: five grouped 11x11 Gaussian convolutions with zero padding,
// sigma 1.5, c1 = 0.01^2, c2 = 0.03^2, mean over all channels and pixels) and their autograd twins.
//
// Two kernels, each one pass over the image with a 16x16 pixel tile + 5 pixel halo staged in shared memory and the
// 11x11 window applied separably:
//   forward : mu_x, mu_y, E[x^2], E[y^2], E[xy] -> SSIM per pixel; sums of |x - y| and of SSIM (double atomics);
//             the three partial derivatives of SSIM w.r.t. the *convolved* maps (mu_x, E[x^2], E[xy]) are stored;
//   backward: dL/dx(p) = g * [ w_l1 sign(x - y) - w_ssim * ( conv(D_mu)(p) + 2 x(p) conv(D_xx)(p) + y(p) conv(D_xy)(p) ) ] / N
//             (the window is symmetric, so the adjoint of the convolution is the same convolution).
// The gradient feeds dL/dcolor of the rasterizer backward directly; the images are [n_img][H][W] planes (n_img = V * 3).
#include "common.cuh"

namespace gsr {

namespace {

constexpr int LT = 16;                 // tile edge
constexpr int HALO = 5;                // window 11
constexpr int WIN = 2 * HALO + 1;
constexpr int LR = LT + 2 * HALO;      // staged region edge (26)
constexpr float SSIM_C1 = 0.01f * 0.01f;
constexpr float SSIM_C2 = 0.03f * 0.03f;

struct Window { float w[WIN]; };

// stage a (LR x LR) region of one plane, zero outside the image (conv2d zero padding, external.py:115)
__device__ __forceinline__ void stage(const float* __restrict__ plane, int W, int H, int x0, int y0, float (*s)[LR + 1]) {
    for (int i = threadIdx.y * LT + threadIdx.x; i < LR * LR; i += LT * LT) {
        const int ry = i / LR, rx = i - ry * LR;
        const int x = x0 - HALO + rx, y = y0 - HALO + ry;
        s[ry][rx] = (x >= 0 && x < W && y >= 0 && y < H) ? plane[(size_t)y * W + x] : 0.0f;
    }
}

__global__ void __launch_bounds__(LT * LT)
image_loss_forward_kernel(const Window win, const float* __restrict__ img, const float* __restrict__ tgt, int W, int H,
                          float* __restrict__ d_mu, float* __restrict__ d_xx, float* __restrict__ d_xy,
                          double* __restrict__ sums /* [2]: sum |x-y|, sum ssim */) {
    __shared__ float sx[LR][LR + 1];
    __shared__ float sy[LR][LR + 1];
    __shared__ float sh[5][LR][LT + 1];
    __shared__ float s_red[2][LT * LT / 32];
    const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const size_t plane = (size_t)blockIdx.z * W * H;
    stage(img + plane, W, H, x0, y0, sx);
    stage(tgt + plane, W, H, x0, y0, sy);
    __syncthreads();
    // horizontal pass: LR rows x LT columns
    for (int i = threadIdx.y * LT + threadIdx.x; i < LR * LT; i += LT * LT) {
        const int ry = i / LT, cx = i - ry * LT;
        float a = 0.f, b = 0.f, aa = 0.f, bb = 0.f, ab = 0.f;
#pragma unroll
        for (int k = 0; k < WIN; ++k) {
            const float w = win.w[k];
            const float x = sx[ry][cx + k], y = sy[ry][cx + k];
            a = fmaf(w, x, a);
            b = fmaf(w, y, b);
            aa = fmaf(w, x * x, aa);
            bb = fmaf(w, y * y, bb);
            ab = fmaf(w, x * y, ab);
        }
        sh[0][ry][cx] = a; sh[1][ry][cx] = b; sh[2][ry][cx] = aa; sh[3][ry][cx] = bb; sh[4][ry][cx] = ab;
    }
    __syncthreads();
    const int tx = threadIdx.x, ty = threadIdx.y;
    float mu1 = 0.f, mu2 = 0.f, e11 = 0.f, e22 = 0.f, e12 = 0.f;
#pragma unroll
    for (int k = 0; k < WIN; ++k) {
        const float w = win.w[k];
        mu1 = fmaf(w, sh[0][ty + k][tx], mu1);
        mu2 = fmaf(w, sh[1][ty + k][tx], mu2);
        e11 = fmaf(w, sh[2][ty + k][tx], e11);
        e22 = fmaf(w, sh[3][ty + k][tx], e22);
        e12 = fmaf(w, sh[4][ty + k][tx], e12);
    }
    const int px = x0 + tx, py = y0 + ty;
    const bool inside = px < W && py < H;
    float l1 = 0.f, ssim = 0.f;
    if (inside) {
        const float s1 = e11 - mu1 * mu1, s2 = e22 - mu2 * mu2, s12 = e12 - mu1 * mu2;      // external.py:118-120
        const float A1 = 2.f * mu1 * mu2 + SSIM_C1, A2 = 2.f * s12 + SSIM_C2;
        const float B1 = mu1 * mu1 + mu2 * mu2 + SSIM_C1, B2 = s1 + s2 + SSIM_C2;
        const float iB1 = 1.f / B1, iB2 = 1.f / B2;
        ssim = A1 * A2 * iB1 * iB2;                                                           // external.py:125
        // partials w.r.t. the convolved maps (mu1, E[x^2], E[xy]); s1 and s12 depend on mu1 too
        const float dxx = -ssim * iB2;
        const float dxy = 2.f * A1 * iB1 * iB2;
        const float dmu = 2.f * mu2 * (A2 - A1) * iB1 * iB2 - 2.f * mu1 * ssim * iB1 + 2.f * mu1 * ssim * iB2;
        const size_t o = plane + (size_t)py * W + px;
        d_mu[o] = dmu;
        d_xx[o] = dxx;
        d_xy[o] = dxy;
        l1 = fabsf(sx[ty + HALO][tx + HALO] - sy[ty + HALO][tx + HALO]);
    }
    // block reduction of the two sums
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        l1 += __shfl_xor_sync(0xffffffffu, l1, o);
        ssim += __shfl_xor_sync(0xffffffffu, ssim, o);
    }
    const int tid = ty * LT + tx;
    if ((tid & 31) == 0) { s_red[0][tid >> 5] = l1; s_red[1][tid >> 5] = ssim; }
    __syncthreads();
    if (tid < 2) {
        float t = 0.f;
#pragma unroll
        for (int w = 0; w < LT * LT / 32; ++w) t += s_red[tid][w];
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

__global__ void __launch_bounds__(LT * LT)
image_loss_backward_kernel(const Window win, const float* __restrict__ img, const float* __restrict__ tgt, int W, int H,
                           const float* __restrict__ d_mu, const float* __restrict__ d_xx,
                           const float* __restrict__ d_xy, const float* __restrict__ grad_out /* device scalar or null */,
                           float inv_n, float w_l1, float w_ssim, float* __restrict__ dL_dimg) {
    __shared__ float s0[LR][LR + 1];
    __shared__ float s1[LR][LR + 1];
    __shared__ float s2[LR][LR + 1];
    __shared__ float sh[3][LR][LT + 1];
    const int x0 = blockIdx.x * LT, y0 = blockIdx.y * LT;
    const size_t plane = (size_t)blockIdx.z * W * H;
    stage(d_mu + plane, W, H, x0, y0, s0);
    stage(d_xx + plane, W, H, x0, y0, s1);
    stage(d_xy + plane, W, H, x0, y0, s2);
    __syncthreads();
    for (int i = threadIdx.y * LT + threadIdx.x; i < LR * LT; i += LT * LT) {
        const int ry = i / LT, cx = i - ry * LT;
        float a = 0.f, b = 0.f, c = 0.f;
#pragma unroll
        for (int k = 0; k < WIN; ++k) {
            const float w = win.w[k];
            a = fmaf(w, s0[ry][cx + k], a);
            b = fmaf(w, s1[ry][cx + k], b);
            c = fmaf(w, s2[ry][cx + k], c);
        }
        sh[0][ry][cx] = a; sh[1][ry][cx] = b; sh[2][ry][cx] = c;
    }
    __syncthreads();
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int px = x0 + tx, py = y0 + ty;
    if (px >= W || py >= H) return;
    float cm = 0.f, cxx = 0.f, cxy = 0.f;
#pragma unroll
    for (int k = 0; k < WIN; ++k) {
        const float w = win.w[k];
        cm = fmaf(w, sh[0][ty + k][tx], cm);
        cxx = fmaf(w, sh[1][ty + k][tx], cxx);
        cxy = fmaf(w, sh[2][ty + k][tx], cxy);
    }
    const size_t o = plane + (size_t)py * W + px;
    const float x = img[o], y = tgt[o];
    const float g = (grad_out ? grad_out[0] : 1.0f) * inv_n;
    const float dssim = cm + 2.f * x * cxx + y * cxy;
    const float d = x - y;
    const float sgn = d > 0.f ? 1.f : (d < 0.f ? -1.f : 0.f);            // torch.abs backward: sign(0) = 0
    dL_dimg[o] = g * (w_l1 * sgn - w_ssim * dssim);
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
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, (unsigned)n_img), block(LT, LT);
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
    const dim3 grid((W + LT - 1) / LT, (H + LT - 1) / LT, (unsigned)n_img), block(LT, LT);
    image_loss_backward_kernel<<<grid, block, 0, s>>>(make_window(), img, tgt, W, H, at<float>(state, off[1]),
                                                      at<float>(state, off[2]), at<float>(state, off[3]), grad_out,
                                                      (float)(1.0 / ((double)n_img * W * H)), w_l1, w_ssim, dL_dimg);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
