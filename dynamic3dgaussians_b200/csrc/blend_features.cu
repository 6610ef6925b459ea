// 32-channel semantic feature map (the Feature-3DGS lineage of the vendored rasterizer): forward and backward.
// Reference: NUM_SEMANTIC_CHANNELS = 32 (R/cuda_rasterizer/config.h:17); forward accumulation
// SF[ch] += feature[id][ch] * alpha * T and out = SF[ch] + T * bg[ch] (R/cuda_rasterizer/forward.cu:381-383,405-406);
// backward dL/dfeature and the feature term of dL/dalpha (backward.cu:594-611).
//
// These kernels run AFTER the colour kernels and reuse their per-tile record lists, n_contrib and final_T:
//  * forward: walks each tile's list front to back over the positions the colour pass blended (pos < n_contrib, same
//    skip tests, same T products), accumulating the 32 channels; the per-Gaussian feature rows are staged per batch in
//    shared memory (one 128-byte row per instance) instead of the reference's per-pair global gathers (forward.cu:382);
//  * backward: the gradient is linear in the image cotangents, so the feature part is a separate back-to-front walk that
//    ADDS its share of the geometry moments to the per-(view, Gaussian) gradient records the colour backward filled
//    (consumed afterwards by preprocess_backward) and writes dL/dfeature with one atomic per (tile, Gaussian, channel).
//    The reference's version of this term reads a never-written scratch buffer; this is the correct
//    gradient of the forward as written.
// Background: the reference reads bg_color[ch] for ch < 32 from a 3-float tensor (Q5); here channel ch < 3 gets bg[ch]
// (as written), channels >= 3 get 0.
#include "geom_math.cuh"
#include "packed_reduce.cuh"

namespace gsr {

namespace {

constexpr int NF = 32;
constexpr int FB = 128;                 // instances staged per batch
constexpr unsigned FULL = 0xffffffffu;

// stage one batch: records -> sA/sB, Gaussian index -> s_g, feature rows -> sF; then the per-warp work lists
__device__ __forceinline__ void stage_batch(int tid, int lane, int warp, int nb, size_t src, uint32_t view_P0,
                                            const float4* __restrict__ recA, const float4* __restrict__ recB,
                                            const float4* __restrict__ recC, const float* __restrict__ feat,
                                            float tile_x0, float tile_y0, float4* sA, float4* sB, uint32_t* s_gi,
                                            float (*sF)[NF], uint32_t (*s_wmask)[FB / 32]) {
    __syncthreads();                                   // everyone is done with the previous batch
    if (tid < FB) {
        uint32_t m8 = 0u;
        if (tid < nb) {
            const float4 a = recA[src + tid], b = recB[src + tid];
            sA[tid] = a;
            sB[tid] = b;
            s_gi[tid] = __float_as_uint(recC[src + tid].w);
            m8 = blend_patch_mask(a.x, a.y, a.z, a.w, b.x, b.w, tile_x0, tile_y0);
        }
#pragma unroll
        for (int w = 0; w < TILE_PIX / 32; ++w) {
            const uint32_t bal = __ballot_sync(FULL, (m8 >> w) & 1u);
            if (lane == 0) s_wmask[w][warp] = bal;
        }
    }
    __syncthreads();
    // 8 float4 per row, 256 threads: thread t copies float4 number t, t+256, ... of the [nb][8] block
    for (int q = tid; q < nb * (NF / 4); q += TILE_PIX) {
        const int row = q >> 3, c4 = q & 7;
        const uint32_t g = s_gi[row] - view_P0;
        reinterpret_cast<float4*>(sF[row])[c4] = reinterpret_cast<const float4*>(feat + (size_t)g * NF)[c4];
    }
    __syncthreads();
}

__global__ void __launch_bounds__(TILE_PIX, 2)
feature_forward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const float4* __restrict__ recA,
                       const float4* __restrict__ recB, const float4* __restrict__ recC,
                       const float* __restrict__ feat, int P, const CamDev* __restrict__ cams, int W, int H,
                       int tiles_x, int tiles_per_view, const uint32_t* __restrict__ n_contrib,
                       const float* __restrict__ final_T, float* __restrict__ out_feat) {
    __shared__ __align__(16) float4 sA[FB];
    __shared__ __align__(16) float4 sB[FB];
    __shared__ uint32_t s_gi[FB];
    __shared__ __align__(16) float sF[FB][NF];
    __shared__ uint32_t s_wmask[TILE_PIX / 32][FB / 32];
    __shared__ uint32_t s_max;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int vt = (int)tile_order[blockIdx.x];
    const int view = vt / tiles_per_view;
    const int tile = vt - view * tiles_per_view;
    const int tile_yi = tile / tiles_x, tile_xi = tile - tile_yi * tiles_x;
    const int px = tile_xi * TILE_X + (warp & 1) * 8 + (lane & 7);
    const int py = tile_yi * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t HW = (size_t)H * W, pix = (size_t)py * W + px;
    const uint2 range = ranges[vt];
    const uint32_t last_contributor = inside ? n_contrib[(size_t)view * HW + pix] : 0u;
    if (tid == 0) s_max = 0;
    __syncthreads();
    const uint32_t wmax = __reduce_max_sync(FULL, last_contributor);
    if (lane == 0 && wmax > 0) atomicMax(&s_max, wmax);
    __syncthreads();
    const int n_proc = (int)min(s_max, range.y - range.x);
    const float tile_x0 = (float)(tile_xi * TILE_X), tile_y0 = (float)(tile_yi * TILE_Y);

    float T = 1.0f;
    float F[NF];
#pragma unroll
    for (int k = 0; k < NF; ++k) F[k] = 0.0f;
    for (int r0 = 0; r0 < n_proc; r0 += FB) {
        const int nb = min(FB, n_proc - r0);
        stage_batch(tid, lane, warp, nb, (size_t)range.x + r0, (uint32_t)view * (uint32_t)P, recA, recB, recC, feat,
                    tile_x0, tile_y0, sA, sB, s_gi, sF, s_wmask);
#pragma unroll 1
        for (int q = 0; q < FB / 32; ++q) {
            uint32_t m = s_wmask[warp][q];
            while (m != 0u) {
                const int j = q * 32 + __ffs(m) - 1;
                m &= m - 1u;
                if ((uint32_t)(r0 + j) >= last_contributor) continue;      // this pixel stopped before position r0 + j
                const float4 a = sA[j];
                const float4 b = sB[j];
                const float dx = __fsub_rn(a.x, pxf), dy = __fsub_rn(a.y, pyf);
                const float power = blend_power(dx, dy, a.z, a.w, b.x);
                if (power > 0.0f || power < b.w) continue;
                const float alpha = fminf(0.99f, __fmul_rn(b.y, expf(power)));
                if (alpha < 1.0f / 255.0f) continue;
#pragma unroll
                for (int k4 = 0; k4 < NF / 4; ++k4) {
                    const float4 f = reinterpret_cast<const float4*>(sF[j])[k4];
                    F[4 * k4 + 0] = __fmaf_rn(T, __fmul_rn(alpha, f.x), F[4 * k4 + 0]);
                    F[4 * k4 + 1] = __fmaf_rn(T, __fmul_rn(alpha, f.y), F[4 * k4 + 1]);
                    F[4 * k4 + 2] = __fmaf_rn(T, __fmul_rn(alpha, f.z), F[4 * k4 + 2]);
                    F[4 * k4 + 3] = __fmaf_rn(T, __fmul_rn(alpha, f.w), F[4 * k4 + 3]);
                }
                T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
            }
        }
    }
    if (inside) {
        const CamDev& cam = cams[view];
        const float Tf = final_T[(size_t)view * HW + pix];
#pragma unroll
        for (int ch = 0; ch < NF; ++ch)
            out_feat[((size_t)view * NF + ch) * HW + pix] = ch < 3 ? __fmaf_rn(Tf, cam.bg[ch], F[ch]) : F[ch];
    }
}

constexpr int KA = 6 + 16;          // first packed reduction: 6 geometry moments + feature channels 0..15
constexpr int KB = 16;              // second: feature channels 16..31
constexpr int KT = 6 + NF;          // values per (warp, Gaussian)
constexpr int SUBF = 32;

__global__ void __launch_bounds__(TILE_PIX)
feature_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const float4* __restrict__ recA,
                        const float4* __restrict__ recB, const float4* __restrict__ recC,
                        const float* __restrict__ feat, int P, const CamDev* __restrict__ cams, int W, int H,
                        int tiles_x, int tiles_per_view, const uint32_t* __restrict__ n_contrib,
                        const float* __restrict__ final_T, const float* __restrict__ dL_dfeat_img,
                        float* __restrict__ grad_rec, int rec_stride, float* __restrict__ dL_dfeat) {
    __shared__ __align__(16) float4 sA[FB];
    __shared__ __align__(16) float4 sB[FB];
    __shared__ uint32_t s_gi[FB];
    __shared__ __align__(16) float sF[FB][NF];
    __shared__ uint32_t s_wmask[TILE_PIX / 32][FB / 32];
    __shared__ uint32_t s_mask[TILE_PIX / 32];
    __shared__ uint32_t s_max;
    extern __shared__ float s_part_dyn[];             // [8 warps][SUBF][KT]
    float (*s_part)[SUBF][KT] = reinterpret_cast<float (*)[SUBF][KT]>(s_part_dyn);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int vt = (int)tile_order[blockIdx.x];
    const int view = vt / tiles_per_view;
    const int tile = vt - view * tiles_per_view;
    const int tile_yi = tile / tiles_x, tile_xi = tile - tile_yi * tiles_x;
    const int px = tile_xi * TILE_X + (warp & 1) * 8 + (lane & 7);
    const int py = tile_yi * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;
    const size_t HW = (size_t)H * W, pix = (size_t)py * W + px;
    const uint2 range = ranges[vt];
    const uint32_t last_contributor = inside ? n_contrib[(size_t)view * HW + pix] : 0u;
    float T = inside ? final_T[(size_t)view * HW + pix] : 0.0f;
    if (tid == 0) s_max = 0;
    __syncthreads();
    const uint32_t warp_last = __reduce_max_sync(FULL, last_contributor);
    if (lane == 0 && warp_last > 0) atomicMax(&s_max, warp_last);
    __syncthreads();
    const int n_proc = (int)min(s_max, range.y - range.x);
    if (n_proc == 0) return;
    const float tile_x0 = (float)(tile_xi * TILE_X), tile_y0 = (float)(tile_yi * TILE_Y);

    float gF[NF];
    float acc = 0.0f;
    const CamDev& cam = cams[view];
#pragma unroll
    for (int ch = 0; ch < NF; ++ch) {
        gF[ch] = inside ? dL_dfeat_img[((size_t)view * NF + ch) * HW + pix] : 0.0f;
        if (ch < 3) acc += cam.bg[ch] * gF[ch];
    }
    float last_alpha = 0.0f, last_s = 0.0f;
    const int slotA = PackedReduce<KA>::slot(lane), slotB = PackedReduce<KB>::slot(lane);
    const bool writerA = lane == (__ffs(__match_any_sync(FULL, slotA)) - 1);
    const bool writerB = lane == (__ffs(__match_any_sync(FULL, slotB)) - 1);

    const int rounds = (n_proc + FB - 1) / FB;
    for (int c = rounds - 1; c >= 0; --c) {
        const int nb = min(FB, n_proc - c * FB);
        stage_batch(tid, lane, warp, nb, (size_t)range.x + (size_t)c * FB, (uint32_t)view * (uint32_t)P, recA, recB,
                    recC, feat, tile_x0, tile_y0, sA, sB, s_gi, sF, s_wmask);
        for (int sub = (nb - 1) / SUBF; sub >= 0; --sub) {
            const int lo = sub * SUBF;
            const int hi = min(lo + SUBF, nb);
            uint32_t hit_mask = 0;
            uint32_t m = s_wmask[warp][sub];
            {
                const uint32_t base = (uint32_t)(c * FB + lo);
                if (warp_last <= base) m = 0u;
                else if (warp_last - base < 32u) m &= (1u << (warp_last - base)) - 1u;
            }
            while (m != 0u) {
                const int jl = 31 - __clz(m);
                m &= ~(1u << jl);
                const int j = lo + jl;
                const float4 a = sA[j];
                const float4 b = sB[j];
                const uint32_t pos = (uint32_t)(c * FB + j);
                const float dx = __fsub_rn(a.x, pxf), dy = __fsub_rn(a.y, pyf);
                const float power = blend_power(dx, dy, a.z, a.w, b.x);
                const bool pre = (pos < last_contributor) && !(power > 0.0f) && !(power < b.w);
                if (!__any_sync(FULL, pre)) continue;
                const float G = expf(fminf(power, 0.0f));
                const float alpha = fminf(0.99f, __fmul_rn(b.y, G));
                const bool valid = pre && !(alpha < 1.0f / 255.0f);
                float va[KA], vb[KB];
#pragma unroll
                for (int k = 0; k < KA; ++k) va[k] = 0.0f;
#pragma unroll
                for (int k = 0; k < KB; ++k) vb[k] = 0.0f;
                if (valid) {
                    T = __fdividef(T, 1.0f - alpha);
                    const float w = alpha * T;
                    float s = 0.0f;
#pragma unroll
                    for (int k4 = 0; k4 < NF / 4; ++k4) {
                        const float4 f = reinterpret_cast<const float4*>(sF[j])[k4];
                        s = fmaf(f.x, gF[4 * k4], fmaf(f.y, gF[4 * k4 + 1], fmaf(f.z, gF[4 * k4 + 2], fmaf(f.w, gF[4 * k4 + 3], s))));
                    }
#pragma unroll
                    for (int ch = 0; ch < 16; ++ch) va[6 + ch] = w * gF[ch];
#pragma unroll
                    for (int ch = 0; ch < 16; ++ch) vb[ch] = w * gF[16 + ch];
                    acc = fmaf(last_alpha, last_s - acc, acc);
                    last_s = s;
                    last_alpha = alpha;
                    const float gq = G * ((s - acc) * T);
                    const float am = b.y * gq;
                    const float ux = am * dx, uy = am * dy;
                    va[GR_MX] = ux;
                    va[GR_MY] = uy;
                    va[GR_CA] = ux * dx;
                    va[GR_CB] = ux * dy;
                    va[GR_CC] = uy * dy;
                    va[GR_OP] = gq;
                }
                PackedReduce<KA>::run(va, lane);
                PackedReduce<KB>::run(vb, lane);
                if (writerA) s_part[warp][jl][slotA] = va[0];
                if (writerB) s_part[warp][jl][KA + slotB] = vb[0];
                hit_mask |= 1u << jl;
            }
            if (lane == 0) s_mask[warp] = hit_mask;
            __syncthreads();
            {   // combine the 8 warps; one atomic per (tile, Gaussian, value)
                const int jl = tid >> 3, kk = tid & 7;
                if (lo + jl < hi) {
                    uint32_t any = 0;
#pragma unroll
                    for (int w = 0; w < TILE_PIX / 32; ++w) any |= (s_mask[w] >> jl) & 1u;
                    if (any) {
                        const uint32_t gi = s_gi[lo + jl];
                        const uint32_t g = gi - (uint32_t)view * (uint32_t)P;
                        for (int k = kk; k < KT; k += 8) {
                            float sum = 0.0f;
#pragma unroll
                            for (int w = 0; w < TILE_PIX / 32; ++w)
                                if ((s_mask[w] >> jl) & 1u) sum += s_part[w][jl][k];
                            if (k < 6) atomicAdd(grad_rec + (size_t)gi * rec_stride + k, sum);
                            else atomicAdd(dL_dfeat + (size_t)g * NF + (k - 6), sum);
                        }
                    }
                }
            }
            __syncthreads();
        }
    }
}

}  // namespace

int launch_feature_forward(int V, int P, int W, int H, const void* gws, const void* bws, const void* iws,
                           const Layout& L, const float* features, float* out_features, cudaStream_t s) {
    const int T = L.tiles_x * L.tiles_y;
    if (V * T == 0) return GSR_OK;
    feature_forward_kernel<<<V * T, TILE_PIX, 0, s>>>(
        at<uint2>(iws, L.ranges), at<uint32_t>(iws, L.tile_order), at<float4>(bws, L.recA), at<float4>(bws, L.recB), at<float4>(bws, L.recC), features,
        P, at<CamDev>(gws, L.cams), W, H, L.tiles_x, T, at<uint32_t>(iws, L.n_contrib), at<float>(iws, L.final_T),
        out_features);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_feature_backward(int V, int P, int W, int H, int channels, void* gws, const void* bws, const void* iws, const Layout& L,
                            const float* features, const float* dL_dfeat_img, float* dL_dfeatures, cudaStream_t s) {
    const int T = L.tiles_x * L.tiles_y;
    GSR_CUDA_TRY(cudaMemsetAsync(dL_dfeatures, 0, (size_t)P * NF * sizeof(float), s));
    if (V * T == 0 || P == 0) return GSR_OK;
    const size_t dyn = (size_t)(TILE_PIX / 32) * SUBF * KT * sizeof(float);
    static bool attr_set = false;
    if (!attr_set) {
        GSR_CUDA_TRY(cudaFuncSetAttribute(feature_backward_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
        attr_set = true;
    }
    feature_backward_kernel<<<V * T, TILE_PIX, dyn, s>>>(
        at<uint2>(iws, L.ranges), at<uint32_t>(iws, L.tile_order), at<float4>(bws, L.recA), at<float4>(bws, L.recB), at<float4>(bws, L.recC), features,
        P, at<CamDev>(gws, L.cams), W, H, L.tiles_x, T, at<uint32_t>(iws, L.n_contrib), at<float>(iws, L.final_T),
        dL_dfeat_img, at<float>(gws, L.grad_rec), grad_rec_stride(channels), dL_dfeatures);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
