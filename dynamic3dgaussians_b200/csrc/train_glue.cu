// The small per-iteration ops around the rasterizer in the train.py-shaped loop, one kernel each instead of a string of
// eager elementwise / reduction kernels (at 300k Gaussians an iteration is ~1 ms, so ~40 launches of 2-6 us matter):
//   * params2rendervar (REF/helpers.py:98-107): rotations = normalize(unnorm_rotations), opacities =
//     sigmoid(logit_opacities), scales = exp(log_scales) -- forward and backward;
//   * the per-camera colour correction exp(cam_m[c]) * im + cam_c[c] (REF/train.py:181) -- forward and backward with the
//     two per-channel reductions for dL/dcam_m, dL/dcam_c;
//   * the densification statistics (REF/train.py:288-289, REF/external.py:136-140): max_2D_radius, seen,
//     means2D_gradient_accum, denom.
// Same formulas as the torch ops they replace (F.normalize's eps = 1e-12 clamp, 1 / (1 + exp(-x))).
#include "common.cuh"

namespace gsr {

namespace {

constexpr int GLUE_BLOCK = 256;

__global__ void __launch_bounds__(GLUE_BLOCK)
activations_forward_kernel(int64_t P, const float4* __restrict__ unnorm_rot, const float* __restrict__ logit_op,
                           const float* __restrict__ log_scales, float4* __restrict__ rot, float* __restrict__ op,
                           float* __restrict__ scales) {
    const int64_t i = (int64_t)blockIdx.x * GLUE_BLOCK + threadIdx.x;
    if (i >= P) return;
    const float4 q = unnorm_rot[i];
    const float n = fmaxf(sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w), 1e-12f);
    rot[i] = make_float4(q.x / n, q.y / n, q.z / n, q.w / n);
    op[i] = 1.0f / (1.0f + expf(-logit_op[i]));
#pragma unroll
    for (int k = 0; k < 3; ++k) scales[3 * i + k] = expf(log_scales[3 * i + k]);
}

__global__ void __launch_bounds__(GLUE_BLOCK)
activations_backward_kernel(int64_t P, const float4* __restrict__ unnorm_rot, const float* __restrict__ op,
                            const float* __restrict__ scales, const float4* __restrict__ d_rot,
                            const float* __restrict__ d_op, const float* __restrict__ d_scales,
                            float4* __restrict__ d_unnorm_rot, float* __restrict__ d_logit,
                            float* __restrict__ d_log_scales) {
    const int64_t i = (int64_t)blockIdx.x * GLUE_BLOCK + threadIdx.x;
    if (i >= P) return;
    if (d_unnorm_rot) {
        float4 r = make_float4(0.f, 0.f, 0.f, 0.f);
        if (d_rot) {
            const float4 q = unnorm_rot[i], g = d_rot[i];
            const float norm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
            if (norm >= 1e-12f) {          // y = q / |q|:  dq = (g - y <y, g>) / |q|
                const float inv = 1.0f / norm;
                const float4 y = make_float4(q.x * inv, q.y * inv, q.z * inv, q.w * inv);
                const float yg = y.x * g.x + y.y * g.y + y.z * g.z + y.w * g.w;
                r = make_float4((g.x - y.x * yg) * inv, (g.y - y.y * yg) * inv, (g.z - y.z * yg) * inv,
                                (g.w - y.w * yg) * inv);
            } else {                        // the clamp is active: y = q / eps
                r = make_float4(g.x * 1e12f, g.y * 1e12f, g.z * 1e12f, g.w * 1e12f);
            }
        }
        d_unnorm_rot[i] = r;
    }
    if (d_logit) {
        const float o = op[i];
        d_logit[i] = d_op ? d_op[i] * o * (1.0f - o) : 0.0f;
    }
    if (d_log_scales) {
#pragma unroll
        for (int k = 0; k < 3; ++k) d_log_scales[3 * i + k] = d_scales ? d_scales[3 * i + k] * scales[3 * i + k] : 0.0f;
    }
}

// im, out: [3][HW]; m3, c3: the camera's rows of cam_m / cam_c
__global__ void __launch_bounds__(GLUE_BLOCK)
cam_affine_forward_kernel(int64_t HW, const float* __restrict__ im, const float* __restrict__ m3,
                          const float* __restrict__ c3, float* __restrict__ out) {
    const int ch = blockIdx.y;
    const float a = expf(m3[ch]), b = c3[ch];
    const int64_t i = (int64_t)blockIdx.x * GLUE_BLOCK + threadIdx.x;
    if (i < HW) out[ch * HW + i] = a * im[ch * HW + i] + b;
}

constexpr int AFF_ITEMS = 8;
__global__ void __launch_bounds__(GLUE_BLOCK)
cam_affine_backward_kernel(int64_t HW, const float* __restrict__ im, const float* __restrict__ m3,
                           const float* __restrict__ g, float* __restrict__ d_im, float* __restrict__ d_m3,
                           float* __restrict__ d_c3) {
    __shared__ float s_m[GLUE_BLOCK / 32], s_c[GLUE_BLOCK / 32];
    const int ch = blockIdx.y;
    const float a = expf(m3[ch]);
    float sm = 0.0f, sc = 0.0f;
#pragma unroll
    for (int k = 0; k < AFF_ITEMS; ++k) {
        const int64_t i = ((int64_t)blockIdx.x * AFF_ITEMS + k) * GLUE_BLOCK + threadIdx.x;
        if (i < HW) {
            const float gi = g[ch * HW + i];
            if (d_im) d_im[ch * HW + i] = a * gi;
            sm = fmaf(gi, im[ch * HW + i], sm);
            sc += gi;
        }
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) {
        sm += __shfl_xor_sync(0xffffffffu, sm, o);
        sc += __shfl_xor_sync(0xffffffffu, sc, o);
    }
    if ((threadIdx.x & 31) == 0) { s_m[threadIdx.x >> 5] = sm; s_c[threadIdx.x >> 5] = sc; }
    __syncthreads();
    if (threadIdx.x == 0) {
        float tm = 0.0f, tc = 0.0f;
#pragma unroll
        for (int w = 0; w < GLUE_BLOCK / 32; ++w) { tm += s_m[w]; tc += s_c[w]; }
        if (d_m3) atomicAdd(d_m3 + ch, tm * a);        // d/dm exp(m) im = exp(m) im
        if (d_c3) atomicAdd(d_c3 + ch, tc);
    }
}

__global__ void __launch_bounds__(GLUE_BLOCK)
densify_stats_kernel(int64_t P, const int32_t* __restrict__ radii, const uint8_t* __restrict__ seen_in,
                     const float* __restrict__ grad2d, float* __restrict__ max_radius, float* __restrict__ accum,
                     float* __restrict__ denom, uint8_t* __restrict__ seen_out) {
    const int64_t i = (int64_t)blockIdx.x * GLUE_BLOCK + threadIdx.x;
    if (i >= P) return;
    bool seen;
    if (radii) {
        const int32_t r = radii[i];
        seen = r > 0;
        if (max_radius && seen) max_radius[i] = fmaxf((float)r, max_radius[i]);
    } else {
        seen = seen_in[i] != 0;
    }
    if (seen_out) seen_out[i] = seen ? 1 : 0;
    if (accum && seen) {
        const float gx = grad2d[3 * i], gy = grad2d[3 * i + 1];
        accum[i] += sqrtf(gx * gx + gy * gy);
    }
    if (denom && seen) denom[i] += 1.0f;
}

// L1 image loss against 8-bit targets: sum_i w_i |x_i - t_i / 255| with w_i = row_weight[view][y] (or a constant), the
// loss a training input pipeline feeds straight from uint8 frames (the reference converts its PNGs to float tensors on
// the host, REF/train.py:33-43, and takes l1_loss_v1).  One pass for the value, one for the gradient; 4 pixels per
// thread (W % 4 == 0) or 1.
constexpr int L1_ITEMS = 4;
// x - t / 255 with torch's two roundings (t * (1 / 255), then the subtraction): no FMA contraction, so that a pixel
// that equals its target gives exactly 0 and sign(0) = 0
__device__ __forceinline__ float diff255(float x, uint8_t t) { return __fsub_rn(x, __fmul_rn((float)t, 1.0f / 255.0f)); }
template <bool VEC>
__global__ void __launch_bounds__(GLUE_BLOCK)
l1_u8_forward_kernel(int64_t n, int W, int H, const float* __restrict__ img, const uint8_t* __restrict__ tgt,
                     const float* __restrict__ row_weight, float w_const, float* __restrict__ loss) {
    __shared__ float s_part[GLUE_BLOCK / 32];
    const int64_t HW = (int64_t)W * H;
    float acc = 0.0f;
    const int64_t i0 = ((int64_t)blockIdx.x * GLUE_BLOCK + threadIdx.x) * (VEC ? L1_ITEMS : 1);
    if (i0 < n) {
        float w = w_const;
        if (row_weight) {
            const int64_t plane = i0 / HW;
            w = row_weight[(plane / 3) * H + (i0 - plane * HW) / W];
        }
        if (VEC) {
            const float4 x = *reinterpret_cast<const float4*>(img + i0);
            const uchar4 t = *reinterpret_cast<const uchar4*>(tgt + i0);
            acc = fabsf(diff255(x.x, t.x)) + fabsf(diff255(x.y, t.y)) +
                  fabsf(diff255(x.z, t.z)) + fabsf(diff255(x.w, t.w));
        } else {
            acc = fabsf(diff255(img[i0], tgt[i0]));
        }
        acc *= w;
    }
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        float t = 0.0f;
#pragma unroll
        for (int k = 0; k < GLUE_BLOCK / 32; ++k) t += s_part[k];
        atomicAdd(loss, t);
    }
}

__device__ __forceinline__ float sgn(float d) { return d > 0.0f ? 1.0f : (d < 0.0f ? -1.0f : 0.0f); }

template <bool VEC>
__global__ void __launch_bounds__(GLUE_BLOCK)
l1_u8_backward_kernel(int64_t n, int W, int H, const float* __restrict__ img, const uint8_t* __restrict__ tgt,
                      const float* __restrict__ row_weight, float w_const, const float* __restrict__ g,
                      float* __restrict__ d_img) {
    const int64_t HW = (int64_t)W * H;
    const int64_t i0 = ((int64_t)blockIdx.x * GLUE_BLOCK + threadIdx.x) * (VEC ? L1_ITEMS : 1);
    if (i0 >= n) return;
    float w = w_const;
    if (row_weight) {
        const int64_t plane = i0 / HW;
        w = row_weight[(plane / 3) * H + (i0 - plane * HW) / W];
    }
    w *= g ? g[0] : 1.0f;
    if (VEC) {
        const float4 x = *reinterpret_cast<const float4*>(img + i0);
        const uchar4 t = *reinterpret_cast<const uchar4*>(tgt + i0);
        *reinterpret_cast<float4*>(d_img + i0) =
            make_float4(w * sgn(diff255(x.x, t.x)), w * sgn(diff255(x.y, t.y)),
                        w * sgn(diff255(x.z, t.z)), w * sgn(diff255(x.w, t.w)));
    } else {
        d_img[i0] = w * sgn(diff255(img[i0], tgt[i0]));
    }
}

}  // namespace

int launch_activations_forward(int64_t P, const float* unnorm_rot, const float* logit_op, const float* log_scales,
                               float* rot, float* op, float* scales, cudaStream_t s) {
    if (P == 0) return GSR_OK;
    activations_forward_kernel<<<(unsigned)((P + GLUE_BLOCK - 1) / GLUE_BLOCK), GLUE_BLOCK, 0, s>>>(
        P, (const float4*)unnorm_rot, logit_op, log_scales, (float4*)rot, op, scales);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_activations_backward(int64_t P, const float* unnorm_rot, const float* op, const float* scales,
                                const float* d_rot, const float* d_op, const float* d_scales, float* d_unnorm_rot,
                                float* d_logit, float* d_log_scales, cudaStream_t s) {
    if (P == 0) return GSR_OK;
    activations_backward_kernel<<<(unsigned)((P + GLUE_BLOCK - 1) / GLUE_BLOCK), GLUE_BLOCK, 0, s>>>(
        P, (const float4*)unnorm_rot, op, scales, (const float4*)d_rot, d_op, d_scales, (float4*)d_unnorm_rot, d_logit,
        d_log_scales);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_cam_affine_forward(int64_t HW, const float* im, const float* m3, const float* c3, float* out,
                              cudaStream_t s) {
    if (HW == 0) return GSR_OK;
    cam_affine_forward_kernel<<<dim3((unsigned)((HW + GLUE_BLOCK - 1) / GLUE_BLOCK), 3), GLUE_BLOCK, 0, s>>>(HW, im, m3,
                                                                                                             c3, out);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_cam_affine_backward(int64_t HW, const float* im, const float* m3, const float* g, float* d_im, float* d_m3,
                               float* d_c3, cudaStream_t s) {
    if (HW == 0) return GSR_OK;
    const int64_t per_block = (int64_t)GLUE_BLOCK * AFF_ITEMS;
    cam_affine_backward_kernel<<<dim3((unsigned)((HW + per_block - 1) / per_block), 3), GLUE_BLOCK, 0, s>>>(
        HW, im, m3, g, d_im, d_m3, d_c3);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_densify_stats(int64_t P, const int32_t* radii, const uint8_t* seen_in, const float* grad2d,
                         float* max_radius, float* accum, float* denom, uint8_t* seen_out, cudaStream_t s) {
    if (P == 0) return GSR_OK;
    densify_stats_kernel<<<(unsigned)((P + GLUE_BLOCK - 1) / GLUE_BLOCK), GLUE_BLOCK, 0, s>>>(
        P, radii, seen_in, grad2d, max_radius, accum, denom, seen_out);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_l1_u8_forward(int64_t n_views, int W, int H, const float* img, const uint8_t* tgt, const float* row_weight,
                         float w_const, float* loss, cudaStream_t s) {
    GSR_CUDA_TRY(cudaMemsetAsync(loss, 0, sizeof(float), s));
    const int64_t n = n_views * 3 * (int64_t)W * H;
    if (n == 0) return GSR_OK;
    if (W % 4 == 0) {
        const int64_t per = (int64_t)GLUE_BLOCK * L1_ITEMS;
        l1_u8_forward_kernel<true><<<(unsigned)((n + per - 1) / per), GLUE_BLOCK, 0, s>>>(n, W, H, img, tgt, row_weight,
                                                                                        w_const, loss);
    } else {
        l1_u8_forward_kernel<false><<<(unsigned)((n + GLUE_BLOCK - 1) / GLUE_BLOCK), GLUE_BLOCK, 0, s>>>(
            n, W, H, img, tgt, row_weight, w_const, loss);
    }
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_l1_u8_backward(int64_t n_views, int W, int H, const float* img, const uint8_t* tgt, const float* row_weight,
                          float w_const, const float* g, float* d_img, cudaStream_t s) {
    const int64_t n = n_views * 3 * (int64_t)W * H;
    if (n == 0) return GSR_OK;
    if (W % 4 == 0) {
        const int64_t per = (int64_t)GLUE_BLOCK * L1_ITEMS;
        l1_u8_backward_kernel<true><<<(unsigned)((n + per - 1) / per), GLUE_BLOCK, 0, s>>>(n, W, H, img, tgt, row_weight,
                                                                                         w_const, g, d_img);
    } else {
        l1_u8_backward_kernel<false><<<(unsigned)((n + GLUE_BLOCK - 1) / GLUE_BLOCK), GLUE_BLOCK, 0, s>>>(
            n, W, H, img, tgt, row_weight, w_const, g, d_img);
    }
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
