// (e) fused local-rigidity / rotation-similarity / isometry losses over the fixed K-NN graph of the foreground
// Gaussians, forward and backward.  Replaces ~15 eager PyTorch kernels plus their autograd twins per iteration
// (REF/train.py:253-273 == REF/cvpr_dyn.py:304-330; helpers REF/helpers.py:118-133, REF/external.py:61-78), which the
// reference's README prices at ~20 ms per training iteration (REF/README.md:90).
//
//   rel_i   = quat_mult(rot_i, prev_inv_rot_i)                         REF/train.py:259, helpers.py:124-133
//   R_i     = build_rotation(rel_i)   (normalises rel_i)               REF/external.py:61-78
//   o_ik    = pts[nbr_ik] - pts_i                                      REF/train.py:261-262
//   rigid   = mean_ik sqrt(|R_i^T o_ik - prev_offset_ik|^2 w_ik + 1e-20)   :263-267, weighted_l2_loss_v2
//   rot     = mean_ik sqrt(|rel[nbr_ik] - rel_i|^2 w_ik + 1e-20)           :269-270, weighted_l2_loss_v2
//   iso     = mean_ik sqrt((sqrt(|o_ik|^2 + 1e-20) - dist_ik)^2 w_ik + 1e-20)   :272-273, weighted_l2_loss_v1
//
// Three launches, no [F,K,3] temporaries: (1) rel_i per point, (2) one thread per point walks its K edges, keeps the
// i-side gradient sums in registers and scatters the j-side ones with float atomics, block-reduces the three loss
// sums into double atomics, (3) per point: d(rel) -> d(rot) through the (linear) quaternion product, final means.
#include "common.cuh"

namespace gsr {

namespace {

constexpr unsigned FULL = 0xffffffffu;

struct DynLayout { size_t rel, drel, sums, total; };

inline DynLayout dyn_layout(int64_t F) {
    DynLayout L;
    size_t o = 0;
    auto take = [&](size_t b) { size_t r = o; o = align_up(o + b); return r; };
    const size_t n = (size_t)(F > 0 ? F : 0);
    L.sums = take(4 * sizeof(double));
    L.rel = take(n * 16);
    L.drel = take(n * 16);
    L.total = o;
    return L;
}

__device__ __forceinline__ float4 quat_mul(const float4 a, const float4 b) {   // (w,x,y,z) stored as (x,y,z,w) fields
    const float w1 = a.x, x1 = a.y, y1 = a.z, z1 = a.w, w2 = b.x, x2 = b.y, y2 = b.z, z2 = b.w;
    return make_float4(w1 * w2 - x1 * x2 - y1 * y2 - z1 * z2, w1 * x2 + x1 * w2 + y1 * z2 - z1 * y2,
                       w1 * y2 - x1 * z2 + y1 * w2 + z1 * x2, w1 * z2 + x1 * y2 - y1 * x2 + z1 * w2);
}

__global__ void __launch_bounds__(256)
dyn_rel_rot_kernel(int F, const float4* __restrict__ rot, const float4* __restrict__ prev_inv_rot,
                   float4* __restrict__ rel, float4* __restrict__ drel, double* __restrict__ sums) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) { sums[0] = 0.0; sums[1] = 0.0; sums[2] = 0.0; }
    if (i >= F) return;
    rel[i] = quat_mul(rot[i], prev_inv_rot[i]);
    drel[i] = make_float4(0.f, 0.f, 0.f, 0.f);
}

__device__ __forceinline__ float block_sum_256(float v, float* s_red) {
#pragma unroll
    for (int o = 16; o >= 1; o >>= 1) v += __shfl_xor_sync(FULL, v, o);
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    float t = 0.0f;
    if (threadIdx.x < 8) t = s_red[threadIdx.x];
    if (threadIdx.x < 32) {
#pragma unroll
        for (int o = 4; o >= 1; o >>= 1) t += __shfl_xor_sync(FULL, t, o);
    }
    __syncthreads();
    return t;   // valid in thread 0
}

template <bool GRAD>
__global__ void __launch_bounds__(256)
dyn_edges_kernel(int F, int K, const float* __restrict__ pts, const float4* __restrict__ rel,
                 const int32_t* __restrict__ nbr_idx, const float* __restrict__ nbr_w,
                 const float* __restrict__ prev_offset, const float* __restrict__ nbr_dist,
                 const float* __restrict__ upstream /* [3] device, GRAD only */, double* __restrict__ sums,
                 float* __restrict__ dpts, float4* __restrict__ drel) {
    __shared__ float s_red[8];
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    float l_rigid = 0.0f, l_rot = 0.0f, l_iso = 0.0f;
    if (i < F) {
        const float px = pts[3 * (size_t)i], py = pts[3 * (size_t)i + 1], pz = pts[3 * (size_t)i + 2];
        const float4 q = rel[i];
        const float nrm = sqrtf(q.x * q.x + q.y * q.y + q.z * q.z + q.w * q.w);
        const float inv = 1.0f / nrm;
        const float r = q.x * inv, x = q.y * inv, y = q.z * inv, z = q.w * inv;
        const float R00 = 1.f - 2.f * (y * y + z * z), R01 = 2.f * (x * y - r * z), R02 = 2.f * (x * z + r * y);
        const float R10 = 2.f * (x * y + r * z), R11 = 1.f - 2.f * (x * x + z * z), R12 = 2.f * (y * z - r * x);
        const float R20 = 2.f * (x * z - r * y), R21 = 2.f * (y * z + r * x), R22 = 1.f - 2.f * (x * x + y * y);
        float g_rigid = 0.f, g_rot = 0.f, g_iso = 0.f;
        if (GRAD) {
            const float invN = 1.0f / ((float)F * (float)K);
            g_rigid = upstream[0] * invN; g_rot = upstream[1] * invN; g_iso = upstream[2] * invN;
        }
        float G[9] = {0, 0, 0, 0, 0, 0, 0, 0, 0};     // dL/dR (row-major)
        float dpi[3] = {0, 0, 0};
        float dri[4] = {0, 0, 0, 0};                  // dL/d rel_i from the rotation-similarity term
        for (int k = 0; k < K; ++k) {
            const size_t e = (size_t)i * K + k;
            const int j = nbr_idx[e];
            const float w = nbr_w[e];
            const float ox = pts[3 * (size_t)j] - px, oy = pts[3 * (size_t)j + 1] - py,
                        oz = pts[3 * (size_t)j + 2] - pz;
            // rigidity: c = R^T o
            const float cx = R00 * ox + R10 * oy + R20 * oz;
            const float cy = R01 * ox + R11 * oy + R21 * oz;
            const float cz = R02 * ox + R12 * oy + R22 * oz;
            const float ex = cx - prev_offset[3 * e], ey = cy - prev_offset[3 * e + 1],
                        ez = cz - prev_offset[3 * e + 2];
            const float Lr = sqrtf((ex * ex + ey * ey + ez * ez) * w + 1e-20f);
            l_rigid += Lr;
            // rotation similarity
            const float4 qj = rel[j];
            const float d0 = qj.x - q.x, d1 = qj.y - q.y, d2 = qj.z - q.z, d3 = qj.w - q.w;
            const float Lq = sqrtf((d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3) * w + 1e-20f);
            l_rot += Lq;
            // isometry
            const float m = sqrtf(ox * ox + oy * oy + oz * oz + 1e-20f);
            const float dm = m - nbr_dist[e];
            const float Li = sqrtf(dm * dm * w + 1e-20f);
            l_iso += Li;
            if (GRAD) {
                const float sr = g_rigid * w / Lr;
                const float dcx = sr * ex, dcy = sr * ey, dcz = sr * ez;
                // do = R dc ; dR[b][a] += o_b dc_a
                float dox = R00 * dcx + R01 * dcy + R02 * dcz;
                float doy = R10 * dcx + R11 * dcy + R12 * dcz;
                float doz = R20 * dcx + R21 * dcy + R22 * dcz;
                G[0] += ox * dcx; G[1] += ox * dcy; G[2] += ox * dcz;
                G[3] += oy * dcx; G[4] += oy * dcy; G[5] += oy * dcz;
                G[6] += oz * dcx; G[7] += oz * dcy; G[8] += oz * dcz;
                const float si = g_iso * w * dm / (Li * m);
                dox += si * ox; doy += si * oy; doz += si * oz;
                dpi[0] -= dox; dpi[1] -= doy; dpi[2] -= doz;
                atomicAdd(&dpts[3 * (size_t)j], dox);
                atomicAdd(&dpts[3 * (size_t)j + 1], doy);
                atomicAdd(&dpts[3 * (size_t)j + 2], doz);
                const float sq = g_rot * w / Lq;
                const float e0 = sq * d0, e1 = sq * d1, e2 = sq * d2, e3 = sq * d3;
                dri[0] -= e0; dri[1] -= e1; dri[2] -= e2; dri[3] -= e3;
                float* dj = reinterpret_cast<float*>(drel + j);
                atomicAdd(dj, e0); atomicAdd(dj + 1, e1); atomicAdd(dj + 2, e2); atomicAdd(dj + 3, e3);
            }
        }
        if (GRAD) {
            // dL/d(normalised quaternion) from dL/dR, then through q / |q|
            const float dr = 2.f * (-z * G[1] + y * G[2] + z * G[3] - x * G[5] - y * G[6] + x * G[7]);
            const float dx = 2.f * (y * G[1] + z * G[2] + y * G[3] - 2.f * x * G[4] - r * G[5] + z * G[6] + r * G[7] -
                                    2.f * x * G[8]);
            const float dy = 2.f * (-2.f * y * G[0] + x * G[1] + r * G[2] + x * G[3] + z * G[5] - r * G[6] + z * G[7] -
                                    2.f * y * G[8]);
            const float dz = 2.f * (-2.f * z * G[0] - r * G[1] + x * G[2] + r * G[3] - 2.f * z * G[4] + y * G[5] +
                                    x * G[6] + y * G[7]);
            const float dotp = r * dr + x * dx + y * dy + z * dz;
            dri[0] += (dr - r * dotp) * inv; dri[1] += (dx - x * dotp) * inv;
            dri[2] += (dy - y * dotp) * inv; dri[3] += (dz - z * dotp) * inv;
            float* di = reinterpret_cast<float*>(drel + i);
            atomicAdd(di, dri[0]); atomicAdd(di + 1, dri[1]); atomicAdd(di + 2, dri[2]); atomicAdd(di + 3, dri[3]);
            atomicAdd(&dpts[3 * (size_t)i], dpi[0]);
            atomicAdd(&dpts[3 * (size_t)i + 1], dpi[1]);
            atomicAdd(&dpts[3 * (size_t)i + 2], dpi[2]);
        }
    }
    const float t0 = block_sum_256(l_rigid, s_red);
    const float t1 = block_sum_256(l_rot, s_red);
    const float t2 = block_sum_256(l_iso, s_red);
    if (threadIdx.x == 0) {
        atomicAdd(&sums[0], (double)t0);
        atomicAdd(&sums[1], (double)t1);
        atomicAdd(&sums[2], (double)t2);
    }
}

// rel = rot (x) pinv is linear in rot:  d rot = J(pinv)^T d rel
__global__ void __launch_bounds__(256)
dyn_finish_kernel(int F, int K, const float4* __restrict__ prev_inv_rot, const float4* __restrict__ drel,
                  const double* __restrict__ sums, float* __restrict__ losses, float4* __restrict__ drot) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0 && losses) {
        const double n = (double)F * (double)K;
        losses[0] = (float)(sums[0] / n);
        losses[1] = (float)(sums[1] / n);
        losses[2] = (float)(sums[2] / n);
    }
    if (i >= F || !drot) return;
    const float4 g = drel[i], b = prev_inv_rot[i];
    const float dw = g.x, dx = g.y, dy = g.z, dz = g.w, w2 = b.x, x2 = b.y, y2 = b.z, z2 = b.w;
    drot[i] = make_float4(dw * w2 + dx * x2 + dy * y2 + dz * z2, -dw * x2 + dx * w2 - dy * z2 + dz * y2,
                          -dw * y2 + dx * z2 + dy * w2 - dz * x2, -dw * z2 - dx * y2 + dy * x2 + dz * w2);
}

}  // namespace
}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_dyn_losses_temp_bytes(int64_t F) { return dyn_layout(F).total; }

int gsr_dyn_losses(int64_t F, int K, const float* pts, const float* rot, const float* prev_inv_rot,
                   const int32_t* nbr_idx, const float* nbr_weight, const float* prev_offset, const float* nbr_dist,
                   const float* upstream_dev, float* losses, float* dL_dpts, float* dL_drot, void* tmp,
                   size_t tmp_bytes, gsr_stream_t stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (F < 0 || K < 1 || F >= (1ll << 30)) return GSR_ERR_INVALID_ARGUMENT;
    if (!losses || !tmp) return GSR_ERR_INVALID_ARGUMENT;
    const bool grad = dL_dpts != nullptr || dL_drot != nullptr;
    if (grad && (!dL_dpts || !dL_drot || !upstream_dev)) return GSR_ERR_INVALID_ARGUMENT;
    const DynLayout L = dyn_layout(F);
    if (tmp_bytes < L.total) return GSR_ERR_WORKSPACE_TOO_SMALL;
    if (F == 0) {
        GSR_CUDA_TRY(cudaMemsetAsync(losses, 0, 3 * sizeof(float), s));
        return GSR_OK;
    }
    if (!pts || !rot || !prev_inv_rot || !nbr_idx || !nbr_weight || !prev_offset || !nbr_dist)
        return GSR_ERR_INVALID_ARGUMENT;
    const int n = (int)F;
    const int blocks = (n + 255) / 256;
    double* sums = at<double>(tmp, L.sums);
    float4* rel = at<float4>(tmp, L.rel);
    float4* drel = at<float4>(tmp, L.drel);
    dyn_rel_rot_kernel<<<blocks, 256, 0, s>>>(n, (const float4*)rot, (const float4*)prev_inv_rot, rel, drel, sums);
    GSR_LAUNCH_CHECK();
    if (grad) {
        GSR_CUDA_TRY(cudaMemsetAsync(dL_dpts, 0, (size_t)F * 3 * sizeof(float), s));
        dyn_edges_kernel<true><<<blocks, 256, 0, s>>>(n, K, pts, rel, nbr_idx, nbr_weight, prev_offset, nbr_dist,
                                                     upstream_dev, sums, dL_dpts, drel);
    } else {
        dyn_edges_kernel<false><<<blocks, 256, 0, s>>>(n, K, pts, rel, nbr_idx, nbr_weight, prev_offset, nbr_dist,
                                                      nullptr, sums, nullptr, nullptr);
    }
    GSR_LAUNCH_CHECK();
    dyn_finish_kernel<<<blocks, 256, 0, s>>>(n, K, (const float4*)prev_inv_rot, drel, sums, losses,
                                             (float4*)dL_drot);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // extern "C"
