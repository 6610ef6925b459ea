// (a) per-Gaussian forward preprocess for all V views of a batch.
// Replaces FORWARD::preprocess / preprocessCUDA (R/cuda_rasterizer/forward.cu:166-269) and checkFrustum
// (R/cuda_rasterizer/rasterizer_impl.cu:54-66).
//
// One thread per (view, Gaussian) of a persistent grid.  Besides the reference's per-Gaussian state it emits the depth-sort pair
// (view << 32 | depth bits, Gaussian index): the (tile, depth) instance sort of the reference is split into a depth
// sort over Gaussians (here, P entries per view) followed by a stable tile sort over instances, which yields the
// identical order with ~4x less sort traffic.
#include "geom_math.cuh"

namespace gsr {

namespace {

__device__ __forceinline__ float3 eval_sh_pinned(int deg, int M, const float* __restrict__ sh_g, float3 pos,
                                                 const float* campos, bool* clamped) {
    // computeColorFromSH, forward.cu:20-71 (contraction pinned from the reference PTX)
    const float dx = __fsub_rn(pos.x, campos[0]), dy = __fsub_rn(pos.y, campos[1]), dz = __fsub_rn(pos.z, campos[2]);
    const float len = __fsqrt_rn(__fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy))));
    const float x = __fdiv_rn(dx, len), y = __fdiv_rn(dy, len), z = __fdiv_rn(dz, len);
    float res[3];
#pragma unroll
    for (int c = 0; c < 3; ++c) res[c] = __fmul_rn(GSR_SH_C0, sh_g[c]);
    auto SH = [&](int k, int c) { return sh_g[k * 3 + c]; };
    if (deg > 0) {
        const float a = __fmul_rn(y, GSR_SH_C1), b = __fmul_rn(z, GSR_SH_C1), d = __fmul_rn(x, GSR_SH_C1);
#pragma unroll
        for (int c = 0; c < 3; ++c) {
            float r = __fmaf_rn(-a, SH(1, c), res[c]);
            r = __fmaf_rn(b, SH(2, c), r);
            res[c] = __fmaf_rn(-d, SH(3, c), r);
        }
        if (deg > 1) {
            const float xx = __fmul_rn(x, x), yy = __fmul_rn(y, y), zz = __fmul_rn(z, z);
            const float xy = __fmul_rn(x, y), yz = __fmul_rn(y, z), xz = __fmul_rn(x, z);
            const float zz2 = __fadd_rn(zz, zz);
            const float xx_yy = __fsub_rn(xx, yy);
            const float k4 = __fmul_rn(xy, GSR_SH_C2_0), k5 = __fmul_rn(yz, GSR_SH_C2_1);
            const float k6 = __fmul_rn(__fsub_rn(__fsub_rn(zz2, xx), yy), GSR_SH_C2_2);
            const float k7 = __fmul_rn(xz, GSR_SH_C2_3), k8 = __fmul_rn(xx_yy, GSR_SH_C2_4);
#pragma unroll
            for (int c = 0; c < 3; ++c) {
                float r = __fmaf_rn(k4, SH(4, c), res[c]);
                r = __fmaf_rn(k5, SH(5, c), r);
                r = __fmaf_rn(k6, SH(6, c), r);
                r = __fmaf_rn(k7, SH(7, c), r);
                res[c] = __fmaf_rn(k8, SH(8, c), r);
            }
            if (deg > 2) {
                // 3xx, 3yy and 4zz only feed adds/subs, so they are fused there (single rounding)
                const float f401 = __fsub_rn(__fmaf_rn(zz, 4.0f, -xx), yy);
                const float k9 = __fmul_rn(__fmul_rn(y, GSR_SH_C3_0), __fmaf_rn(xx, 3.0f, -yy));
                const float k10 = __fmul_rn(z, __fmul_rn(xy, GSR_SH_C3_1));
                const float k11 = __fmul_rn(__fmul_rn(y, GSR_SH_C3_2), f401);
                const float k12 = __fmul_rn(__fmul_rn(z, GSR_SH_C3_3), __fmaf_rn(-yy, 3.0f, __fmaf_rn(-xx, 3.0f, zz2)));
                const float k13 = __fmul_rn(__fmul_rn(x, GSR_SH_C3_4), f401);
                const float k14 = __fmul_rn(__fmul_rn(z, GSR_SH_C3_5), xx_yy);
                const float k15 = __fmul_rn(__fmul_rn(x, GSR_SH_C3_6), __fmaf_rn(-yy, 3.0f, xx));
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    float r = __fmaf_rn(k9, SH(9, c), res[c]);
                    r = __fmaf_rn(k10, SH(10, c), r);
                    r = __fmaf_rn(k11, SH(11, c), r);
                    r = __fmaf_rn(k12, SH(12, c), r);
                    r = __fmaf_rn(k13, SH(13, c), r);
                    r = __fmaf_rn(k14, SH(14, c), r);
                    res[c] = __fmaf_rn(k15, SH(15, c), r);
                }
            }
        }
    }
    float3 out;
    float* o = &out.x;
#pragma unroll
    for (int c = 0; c < 3; ++c) {
        const float r = __fadd_rn(res[c], 0.5f);
        clamped[c] = r < 0.0f;
        o[c] = r < 0.0f ? 0.0f : r;
    }
    return out;
}

// What does not depend on the view, once per Gaussian instead of once per (view, Gaussian): the 3D covariance
// (computeCov3D, forward.cu:118-150, or the caller's cov3D_precomp), the opacity and the blend kernels' conservative skip
// threshold (a logf).  Two float4 per Gaussian, read by preprocess_kernel next to the mean.
__global__ void __launch_bounds__(256)
gaussian_static_kernel(int P, const float* __restrict__ opacities, const float* __restrict__ scales,
                       const float* __restrict__ rotations, const float* __restrict__ cov3D_precomp, float scale_modifier,
                       float4* __restrict__ gstatic) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P) return;
    float cv[6];
    if (cov3D_precomp != nullptr) {
#pragma unroll
        for (int i = 0; i < 6; ++i) cv[i] = cov3D_precomp[6 * (size_t)g + i];
    } else {
        const float3 sc = make_float3(scales[3 * g], scales[3 * g + 1], scales[3 * g + 2]);
        const float4 q = reinterpret_cast<const float4*>(rotations)[g];
        cov3d_from_scale_rot(sc, scale_modifier, q, cv);
    }
    const float op = opacities[g];
    gstatic[2 * (size_t)g] = make_float4(cv[0], cv[1], cv[2], cv[3]);
    gstatic[2 * (size_t)g + 1] = make_float4(cv[4], cv[5], op, blend_skip_threshold(op));
}

// Persistent: the (view, chunk of PP_CHUNK Gaussians) work items are dealt out as ONE contiguous run per CTA, so a CTA
// sees at most a few views.  That lets the kernel also take the depth sort's four digit histograms (what
// radix_histogram_kernel would re-read all V x P keys for): counted in shared memory, flushed to dhist[view][pass][bin]
// when the run leaves a view.  Keys of culled Gaussians (the sentinel ~0, dropped by the sort's first pass) are not counted.
#ifndef GSR_PP_CTAS
#define GSR_PP_CTAS 5
#endif
#ifndef GSR_PP_CHUNK
#define GSR_PP_CHUNK 1024
#endif
constexpr int PP_CHUNK = GSR_PP_CHUNK;
constexpr int PP_CTAS_PER_SM = GSR_PP_CTAS;
constexpr int DEPTH_PASSES = 4;            // 32-bit depth keys, 8-bit digits

__global__ void __launch_bounds__(256, PP_CTAS_PER_SM)
preprocess_kernel(const CamDev* __restrict__ cams, int V, int P, int sh_coeffs,
                  const float* __restrict__ means3D, const float4* __restrict__ gstatic,
                  const float* __restrict__ sh,
                  int tiles_x, int tiles_y,
                  int32_t* __restrict__ radii, ushort4* __restrict__ rects, float4* __restrict__ geomrec,
                  float* __restrict__ rgb, uint8_t* __restrict__ clamped,
                  uint32_t* __restrict__ dkeys, uint32_t* __restrict__ dvals, int32_t* __restrict__ radii_out,
                  uint32_t* __restrict__ counters, uint32_t* __restrict__ dhist) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ CamDev cam;
    __shared__ uint32_t s_tiles[8];
    __shared__ uint32_t s_dh[DEPTH_PASSES * RADIX];
    const int tid = threadIdx.x, lane = tid & 31;
    const int chunks = (P + PP_CHUNK - 1) / PP_CHUNK;
    const long long items = (long long)V * chunks;
    const int c_begin = (int)(items * blockIdx.x / gridDim.x), c_end = (int)(items * (blockIdx.x + 1) / gridDim.x);
    for (int i = tid; i < DEPTH_PASSES * RADIX; i += 256) s_dh[i] = 0;
    uint32_t tiles_acc = 0;
    int cur_view = -1;

    for (int it = c_begin; it < c_end; ++it) {
        const int v = it / chunks, chunk = it - v * chunks;
        if (v != cur_view) {
            __syncthreads();                                   // everyone is done with the old camera and its counts
            if (cur_view >= 0) {
                uint32_t* dst = dhist + (size_t)cur_view * DEPTH_PASSES * RADIX;
                for (int i = tid; i < DEPTH_PASSES * RADIX; i += 256) {
                    const uint32_t c = s_dh[i];
                    if (c) { atomicAdd(&dst[i], c); s_dh[i] = 0; }
                }
            }
            const uint32_t* src = reinterpret_cast<const uint32_t*>(cams + v);
            uint32_t* dstc = reinterpret_cast<uint32_t*>(&cam);
            for (int i = tid; i < (int)(sizeof(CamDev) / 4); i += 256) dstc[i] = src[i];
            cur_view = v;
            __syncthreads();
        }
#pragma unroll 1
        for (int sub = 0; sub < PP_CHUNK / 256; ++sub) {
            const int g = chunk * PP_CHUNK + sub * 256 + tid;
            const bool live = g < P;
            const size_t idx = (size_t)v * P + (live ? g : 0);

            int radius = 0;
            ushort4 rect = make_ushort4(0, 0, 0, 0);
            uint32_t depth_key = 0xffffffffu;      // culled Gaussians sort behind every visible one of their view

            const float x = live ? means3D[3 * g] : 0.0f, y = live ? means3D[3 * g + 1] : 0.0f,
                        z = live ? means3D[3 * g + 2] : 0.0f;
            const float depth = live ? xform_row(cam.view, 2, x, y, z) : 0.0f;          // p_view.z, auxiliary.h:156
            if (depth > 0.0f) {                                           // near cull: z <= 0 (auxiliary.h:158)
                const float hx = xform_row(cam.proj, 0, x, y, z);
                const float hy = xform_row(cam.proj, 1, x, y, z);
                const float hw = xform_row(cam.proj, 3, x, y, z);
                const float pw = __frcp_rn(__fadd_rn(hw, 0.0000001f));
                const float ndc_x = __fmul_rn(hx, pw), ndc_y = __fmul_rn(hy, pw);

                // view-independent part, computed once per Gaussian by gaussian_static_kernel: 3D covariance, opacity and
                // the blend's skip threshold
                const float4 gs0 = gstatic[2 * (size_t)g], gs1 = gstatic[2 * (size_t)g + 1];
                const float cv[6] = {gs0.x, gs0.y, gs0.z, gs0.w, gs1.x, gs1.y};
                Cov2DInter ci;
                cov2d(cam, x, y, z, cv, ci);
                const float det = __fmaf_rn(ci.a, ci.c, -__fmul_rn(ci.b, ci.b));   // a*c - b*b as the reference SASS has it
                if (det != 0.0f) {
                    const float det_inv = __frcp_rn(det);
                    const float conic_x = __fmul_rn(ci.c, det_inv);
                    const float conic_y = __fmul_rn(det_inv, -ci.b);
                    const float conic_z = __fmul_rn(ci.a, det_inv);
                    const float mid = __fmul_rn(__fadd_rn(ci.a, ci.c), 0.5f);
                    const float sq = __fsqrt_rn(fmaxf(__fmaf_rn(mid, mid, -det), 0.1f));
                    const float lam = fmaxf(__fadd_rn(mid, sq), __fsub_rn(mid, sq));
                    const float my_radius = ceilf(__fmul_rn(__fsqrt_rn(lam), 3.0f));
                    const float2 pix = make_float2(ndc2pix(ndc_x, cam.width), ndc2pix(ndc_y, cam.height));
                    const int r_i = (int)my_radius;
                    uint2 mn, mx;
                    tile_rect(pix, r_i, tiles_x, tiles_y, mn, mx);
                    mn.y = max(mn.y, (uint32_t)cam.row_begin);           // band of tile rows rendered by this call
                    mx.y = max(min(mx.y, (uint32_t)cam.row_end), mn.y);
                    const uint32_t area = (mx.x - mn.x) * (mx.y - mn.y);
                    if (area != 0) {
                        if (sh != nullptr) {
                            bool cl[3];
                            const float3 c = eval_sh_pinned(cam.sh_degree, sh_coeffs, sh + (size_t)g * sh_coeffs * 3,
                                                            make_float3(x, y, z), cam.campos, cl);
                            reinterpret_cast<float4*>(rgb)[idx] = make_float4(c.x, c.y, c.z, 0.0f);
                            clamped[3 * idx] = cl[0];
                            clamped[3 * idx + 1] = cl[1];
                            clamped[3 * idx + 2] = cl[2];
                        }
                        geomrec[2 * idx] = make_float4(pix.x, pix.y, conic_x, conic_y);          // the blend kernels' recA
                        geomrec[2 * idx + 1] = make_float4(conic_z, gs1.z, depth, gs1.w);        // ... and recB
                        radius = r_i;
                        rect = make_ushort4((unsigned short)mn.x, (unsigned short)mn.y, (unsigned short)mx.x,
                                            (unsigned short)mx.y);
                        depth_key = __float_as_uint(depth);
                        tiles_acc += area;
                    }
                }
            }
            else if (live && cam.prefiltered) {
                // in_frustum with prefiltered = true (auxiliary.h:160-167): the reference prints and traps; here the flag
                // comes back through gsr_read_counters and the host raises
                counters[CNT_PREFILTER] = 1u;
            }
            if (live) {
                radii[idx] = radius;
                if (radii_out) radii_out[idx] = radius;
                rects[idx] = rect;
                dkeys[idx] = depth_key;            // sorted per view (segmented sort), so the key is the depth alone; the
                dvals[idx] = (uint32_t)g;          // sort's first pass drops the culled ones (key 0xffffffff)
            }
            // digit counts of the kept keys.  The three low digits are spread (mantissa bits): plain shared atomics.  The
            // top digit (sign + 7 exponent bits) takes a handful of values per view: one atomic per distinct value per warp.
            const bool kept = depth_key != 0xffffffffu;
            if (kept) {
                atomicAdd(&s_dh[0 * RADIX + (depth_key & 0xffu)], 1u);
                atomicAdd(&s_dh[1 * RADIX + ((depth_key >> 8) & 0xffu)], 1u);
                atomicAdd(&s_dh[2 * RADIX + ((depth_key >> 16) & 0xffu)], 1u);
            }
            uint32_t todo = __ballot_sync(0xffffffffu, kept);
            const uint32_t top = depth_key >> 24;
            for (int round = 0; todo != 0u && round < 4; ++round) {
                const uint32_t lead = __shfl_sync(0xffffffffu, top, __ffs(todo) - 1);
                const uint32_t same = __ballot_sync(0xffffffffu, kept && top == lead) & todo;
                if (lane == __ffs(todo) - 1) atomicAdd(&s_dh[3 * RADIX + lead], (uint32_t)__popc(same));
                todo &= ~same;
            }
            if ((todo >> lane) & 1u) atomicAdd(&s_dh[3 * RADIX + top], 1u);
        }
    }
    __syncthreads();
    if (cur_view >= 0) {
        uint32_t* dst = dhist + (size_t)cur_view * DEPTH_PASSES * RADIX;
        for (int i = tid; i < DEPTH_PASSES * RADIX; i += 256) {
            const uint32_t c = s_dh[i];
            if (c) atomicAdd(&dst[i], c);
        }
    }
    // the batch's (Gaussian, tile) instance count -- what the reference returns as num_rendered (rasterizer_impl.cu:285-287):
    // one atomic per CTA
    const uint32_t tiles = __reduce_add_sync(0xffffffffu, tiles_acc);
    if (lane == 0) s_tiles[tid >> 5] = tiles;
    __syncthreads();
    if (tid == 0) {
        uint32_t t = 0;
#pragma unroll
        for (int w = 0; w < 8; ++w) t += s_tiles[w];
        if (t) atomicAdd(&counters[CNT_RENDERED], t);
    }
}

__global__ void mark_visible_kernel(int P, const float* __restrict__ means3D, float v2, float v6, float v10, float v14,
                                    uint8_t* __restrict__ present) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P) return;
    const float x = means3D[3 * g], y = means3D[3 * g + 1], z = means3D[3 * g + 2];
    const float d = __fadd_rn(v14, __fmaf_rn(z, v10, __fmaf_rn(x, v2, __fmul_rn(y, v6))));
    present[g] = d > 0.0f ? 1 : 0;
}

}  // namespace

__global__ void store_cams_kernel(const __grid_constant__ CamPack pack, CamDev* __restrict__ dst) {
    constexpr int WORDS = sizeof(CamDev) / 4;
    const uint32_t* src = reinterpret_cast<const uint32_t*>(pack.cam);
    uint32_t* out = reinterpret_cast<uint32_t*>(dst);
    for (int i = threadIdx.x; i < pack.n * WORDS; i += blockDim.x) out[i] = src[i];
}

int launch_store_cams(const CamPack& pack, CamDev* dst, cudaStream_t s) {
    store_cams_kernel<<<1, 256, 0, s>>>(pack, dst);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_preprocess(const CamDev* cams, int V, int P, int sh_coeffs, const float* means3D, const float* opacities,
                      const float* scales, const float* rotations, const float* cov3D_precomp, const float* sh,
                      float scale_modifier, void* ws, const Layout& L, int32_t* radii_out, cudaStream_t s) {
    if (P == 0 || V == 0) return GSR_OK;
    launch_kernel(gaussian_static_kernel, (P + 255) / 256, 256, 0, s, P, opacities, scales, rotations, cov3D_precomp, scale_modifier,
                                                           at<float4>(ws, L.cov3D));
    GSR_LAUNCH_CHECK();
    static int sms = 0;
    if (!sms) {
        int dev = 0;
        GSR_CUDA_TRY(cudaGetDevice(&dev));
        GSR_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    }
    const long long items = (long long)V * ((P + PP_CHUNK - 1) / PP_CHUNK);
    const long long full = (long long)sms * PP_CTAS_PER_SM;
    const int grid = (int)(items < full ? items : full);
    launch_kernel(preprocess_kernel, grid, 256, 0, s, 
        cams, V, P, sh_coeffs, means3D, at<float4>(ws, L.cov3D), sh, L.tiles_x, L.tiles_y,
        at<int32_t>(ws, L.radii), at<ushort4>(ws, L.rects), at<float4>(ws, L.geomrec), at<float>(ws, L.rgb),
        at<uint8_t>(ws, L.clamped), at<uint32_t>(ws, L.dkeys0), at<uint32_t>(ws, L.dvals0), radii_out,
        at<uint32_t>(ws, L.counters), at<uint32_t>(ws, L.dhist));
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_mark_visible(const float* view, int P, const float* means3D, uint8_t* present, cudaStream_t s) {
    if (P == 0) return GSR_OK;
    mark_visible_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, means3D, view[2], view[6], view[10], view[14], present);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
