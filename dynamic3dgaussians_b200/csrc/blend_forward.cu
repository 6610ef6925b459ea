// (c) 16x16-tile forward alpha blend, all views of the batch in one launch.
// Replaces FORWARD::render / renderCUDA (R/cuda_rasterizer/forward.cu:274-408).
//
// * one CTA per (view, tile); sorted per-instance records are contiguous, so each 128-instance batch is staged into
//   shared memory by 3-4 TMA bulk copies (cp.async.bulk + mbarrier) through a 3-slot ring: no per-pair global gathers
//   (the reference fetches colour/depth/semantics from global memory for every contributing pair, forward.cu:375-383);
// * colour, fg/bg segmentation colours, depth and accumulated alpha are blended in ONE pass over the list
//   (the reference renders colour and segmentation with two full rasterizer calls, REF/train.py:142,249);
// * a warp covers an 8x8 pixel patch, two pixels per lane, and walks the ring on its own (no CTA barrier per batch).
//   Per staged batch it intersects every Gaussian's opacity-aware extent (the ellipse outside of which alpha < 1/255)
//   with its patch and four ballots turn that into a 128-bit work list: a warp only iterates over Gaussians that can
//   reach its patch (2-3x fewer iterations than walking the tile's whole list as the reference does), and each lane
//   tests `power < threshold` before paying for the exp.  Both filters are conservative, so the set of blended pairs
//   is unchanged;
// * a lane leaves the loop as soon as both of its pixels are saturated, a warp stops blending when its 64 pixels are,
//   and the CTA stops fetching when all 256 are;
// * arithmetic per pair follows the reference binary exactly (see geom_math.cuh), so n_contrib and the images agree
//   with the reference bit for bit.
#include "geom_math.cuh"

#ifndef GSR_PATCH_TEST
#define GSR_PATCH_TEST blend_patch_hit_exact
#endif

namespace gsr {

namespace {

__device__ unsigned long long* g_trace_fwd = nullptr;   // diagnostic: per-CTA (start ns, end ns, SM, tile)
__device__ __forceinline__ unsigned long long trace_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned trace_smid() { unsigned r; asm volatile("mov.u32 %0, %smid;" : "=r"(r)); return r; }

#ifndef GSR_FWD_BATCH
#define GSR_FWD_BATCH 128
#endif
constexpr int BATCH = GSR_FWD_BATCH;
#ifndef GSR_FWD_RING
#define GSR_FWD_RING 3
#endif
constexpr int RING = GSR_FWD_RING;      // staged batches in flight per CTA

// CTA = 128 threads = 4 warps; a warp owns an 8x8 patch, a lane the pixels (x, y) and (x, y + 4).  Record loads, loop
// control and dx are shared by the pair and the pair's arithmetic is issued as packed fp32x2 instructions (FFMA2 /
// FMUL2 / FADD2, sm_100): the kernel is bound by instruction issue, so halving the instructions per pixel pair is what
// counts.  Each component goes through exactly the roundings of the reference binary (geom_math.cuh): a pixel that does not blend a Gaussian runs the same instructions with alpha = 0, which leaves C, D, W and
// T bit-for-bit unchanged.
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 splat(float a) { return make_float2(a, a); }

constexpr int NT2 = 128;
constexpr int NW2 = NT2 / 32;

// alpha of one Gaussian on a lane's pixel pair, and whether each pixel passes the skip tests of forward.cu:358-364
// (power in range, alpha >= 1/255).  Depends on nothing the blend loop carries.
struct PairAlpha {
    float2 alpha;
    bool ok0, ok1;
};

__device__ __forceinline__ PairAlpha pair_alpha(const float4 a, const float4 b, float npx, float2 npy) {
    // blend_power (geom_math.cuh) on the pixel pair, component-wise the same roundings
    const float dx = __fadd_rn(a.x, npx);
    const float2 dy = __fadd2_rn(splat(a.y), npy);
    const float dxca = __fmul_rn(dx, a.z), ndxcb = __fmul_rn(dx, -a.w);
    const float2 t3 = __fmul2_rn(dy, __fmul2_rn(dy, splat(b.x)));
    const float2 q2 = __ffma2_rn(splat(dx), splat(dxca), t3);
    const float2 power = __ffma2_rn(q2, splat(-0.5f), __fmul2_rn(dy, splat(ndxcb)));
    const float2 og = __fmul2_rn(splat(b.y), expf2_pinned(power));        // garbage where power is out of range: masked
    PairAlpha r;
    r.alpha = f2(fminf(0.99f, og.x), fminf(0.99f, og.y));
    r.ok0 = !(power.x > 0.0f || power.x < b.w) && !(r.alpha.x < 1.0f / 255.0f);
    r.ok1 = !(power.y > 0.0f || power.y < b.w) && !(r.alpha.y < 1.0f / 255.0f);
    return r;
}

template <int NCH, int ILP>
#ifndef GSR_FWD_MINB
#define GSR_FWD_MINB 8
#endif
#ifndef GSR_FWD_MINB_ILP
#define GSR_FWD_MINB_ILP 6
#endif
__global__ void __launch_bounds__(NT2, NCH > 3 ? 6 : (ILP > 1 ? GSR_FWD_MINB_ILP : GSR_FWD_MINB))
blend_forward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                      const float4* __restrict__ recA, const float4* __restrict__ recB,
                      const float4* __restrict__ recC, const float4* __restrict__ recD,
                      const CamDev* __restrict__ cams, int W, int H, int tiles_x, int tiles_per_view,
                      float* __restrict__ out_color, float* __restrict__ out_extra, float* __restrict__ out_depth,
                      float* __restrict__ out_alpha, uint32_t* __restrict__ n_contrib, float* __restrict__ final_T,
                      uint32_t* __restrict__ tile_cost) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    constexpr bool EXTRA = NCH > 3;
    // one block per ring slot, so that the loop addresses all of an entry's records off ONE register (base + 16 j)
    struct __align__(128) Stage {
        float4 a[BATCH], b[BATCH], c[BATCH], d[EXTRA ? BATCH : 1];
    };
    // The four warps walk the ring on their own (no CTA barrier per batch: with one, the warps of a tile waited for the
    // slowest of them every batch -- the largest stall reason in ncu): each builds the work list of its patch from the
    // staged batch, the last warp out of a ring slot (an atomic count) refills it, and a warp whose 64 pixels are all
    // finished only keeps that count going until every warp is (then nothing more is fetched).
    __shared__ Stage stage[RING];
    __shared__ __align__(8) uint64_t full[RING];
    __shared__ uint32_t s_done[RING];
    __shared__ int s_slot_round[RING];      // the last round issued into each slot (-1: none)
    __shared__ uint32_t s_fin;              // warps with nothing left to blend
    __shared__ uint32_t s_cost;             // the tile's largest n_contrib: how deep the backward will walk it

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int vt = (int)tile_order[blockIdx.x];
    unsigned long long* const trace = g_trace_fwd;
    if (trace && threadIdx.x == 0) { trace[4 * blockIdx.x] = trace_now(); trace[4 * blockIdx.x + 2] = trace_smid(); trace[4 * blockIdx.x + 3] = (unsigned long long)vt; }
    const int view = vt / tiles_per_view;
    const int tile = vt - view * tiles_per_view;
    const int tile_yi = tile / tiles_x, tile_xi = tile - tile_yi * tiles_x;
    const int px = tile_xi * TILE_X + (warp & 1) * 8 + (lane & 7);
    const int py0 = tile_yi * TILE_Y + (warp >> 1) * 8 + (lane >> 3);
    const int py1 = py0 + 4;
    const bool in0 = px < W && py0 < H, in1 = px < W && py1 < H;
    const float npx = -(float)px;
    const float2 npy = f2(-(float)py0, -(float)py1);

    const uint2 range = ranges[vt];
    const int n_total = (int)(range.y - range.x);
    const int rounds = (n_total + BATCH - 1) / BATCH;

    if (tid == 0) {
#pragma unroll
        for (int i = 0; i < RING; ++i) {
            mbar_init(&full[i], 1);
            s_done[i] = 0;
            s_slot_round[i] = -1;
        }
        s_fin = 0;
        s_cost = 0;
        mbar_fence_init();
    }
    __syncthreads();

    auto issue = [&](int r) {   // one thread
        const int buf = r % RING;
        const int nb = min(BATCH, n_total - r * BATCH);
        const uint32_t bytes = (uint32_t)nb * 16u;
        const size_t src = (size_t)range.x + (size_t)r * BATCH;
        s_slot_round[buf] = r;
        mbar_expect_tx(&full[buf], bytes * (EXTRA ? 4u : 3u));
        tma_load_1d(stage[buf].a, recA + src, bytes, &full[buf]);
        tma_load_1d(stage[buf].b, recB + src, bytes, &full[buf]);
        tma_load_1d(stage[buf].c, recC + src, bytes, &full[buf]);
        if (EXTRA) tma_load_1d(stage[buf].d, recD + src, bytes, &full[buf]);
    };
    if (tid == 0) {
        for (int r = 0; r < RING && r < rounds; ++r) issue(r);
    }

    // T > 0: the pixel's running transmittance.  T < 0: the pixel is finished (saturated, or outside the image) and |T|
    // is its final transmittance -- the flag rides in the sign bit, so the loop carries no predicates: a finished pixel
    // gives test_T < 0, which fails the "still above 0.0001" test below like a terminating Gaussian does, blends with
    // alpha = 0 and keeps its T.
    float2 T = f2(in0 ? 1.0f : -1.0f, in1 ? 1.0f : -1.0f);
    float2 C[NCH];
#pragma unroll
    for (int k = 0; k < NCH; ++k) C[k] = splat(0.0f);
    float2 D = splat(0.0f), Wsum = splat(0.0f);
    uint32_t lc0 = 0, lc1 = 0;

    const float patch_x0 = (float)(tile_xi * TILE_X + (warp & 1) * 8), patch_y0 = (float)(tile_yi * TILE_Y + (warp >> 1) * 8);
    bool fin = false;                        // warp-uniform: all 64 pixels of this warp are finished
    for (int r = 0; r < rounds; ++r) {
        const int buf = r % RING;
        const uint32_t parity = (uint32_t)((r / RING) & 1);
        if (fin) {
            // nothing left to blend here: only keep the slot count going for the warps that still do
            int st = 0;
            if (lane == 0) {
                while (st == 0) {
                    if (mbar_try_wait(&full[buf], parity)) st = 1;
                    else if (*(volatile uint32_t*)&s_fin == NW2) st = 2;
                }
            }
            st = __shfl_sync(0xffffffffu, st, 0);
            if (st == 2) break;              // every warp is finished: round r may never have been fetched
        } else {
        mbar_wait(&full[buf], parity);
        const int nb = min(BATCH, n_total - r * BATCH);
        const uint32_t contrib_base = (uint32_t)r * BATCH;
        // pre-pass: this warp's work list, 32 staged Gaussians per ballot against its 8x8 patch
        uint32_t wm[BATCH / 32];
#pragma unroll
        for (int sb = 0; sb < BATCH / 32; ++sb) {
            const int g = sb * 32 + lane;
            bool hit = false;
            if (g < nb) {
                const float4 a = stage[buf].a[g];
                const float4 b = stage[buf].b[g];
                hit = GSR_PATCH_TEST(a.x, a.y, a.z, a.w, b.x, b.w, patch_x0, patch_y0);
            }
            wm[sb] = __ballot_sync(0xffffffffu, hit);
        }
#pragma unroll
        for (int q = 0; q < BATCH / 32; ++q) {
            uint32_t m = wm[q];
            if constexpr (ILP == 1) {
                while (m != 0u && !(T.x < 0.0f && T.y < 0.0f)) {
                    const int j = q * 32 + __ffs(m) - 1;
                    m &= m - 1u;
                    const float4 a = stage[buf].a[j];
                    const float4 b = stage[buf].b[j];
                    // blend_power (geom_math.cuh) on the pixel pair, component-wise the same roundings
                    const float dx = __fadd_rn(a.x, npx);
                    const float2 dy = __fadd2_rn(splat(a.y), npy);
                    const float dxca = __fmul_rn(dx, a.z), ndxcb = __fmul_rn(dx, -a.w);
                    const float2 t3 = __fmul2_rn(dy, __fmul2_rn(dy, splat(b.x)));
                    const float2 q2 = __ffma2_rn(splat(dx), splat(dxca), t3);
                    const float2 power = __ffma2_rn(q2, splat(-0.5f), __fmul2_rn(dy, splat(ndxcb)));
                    bool ok0 = !(power.x > 0.0f || power.x < b.w);
                    bool ok1 = !(power.y > 0.0f || power.y < b.w);
                    if (!(ok0 || ok1)) continue;
                    const float2 og = __fmul2_rn(splat(b.y), expf2_pinned(power));
                    float2 alpha = f2(fminf(0.99f, og.x), fminf(0.99f, og.y));
                    ok0 = ok0 && !(alpha.x < 1.0f / 255.0f);
                    ok1 = ok1 && !(alpha.y < 1.0f / 255.0f);
                    const float2 test_T = __fmul2_rn(T, __fadd2_rn(splat(1.0f), f2(-alpha.x, -alpha.y)));
                    const bool c0 = ok0 && !(test_T.x < 0.0001f), c1 = ok1 && !(test_T.y < 0.0001f);
                    // ok && !c: this Gaussian would saturate the pixel (or the pixel already is): finished from here on
                    T.x = (ok0 && !c0) ? -fabsf(T.x) : T.x;
                    T.y = (ok1 && !c1) ? -fabsf(T.y) : T.y;
                    if (!(c0 || c1)) continue;
                    alpha.x = c0 ? alpha.x : 0.0f;
                    alpha.y = c1 ? alpha.y : 0.0f;
                    const float4 c = stage[buf].c[j];
                    C[0] = __ffma2_rn(T, __fmul2_rn(alpha, splat(c.x)), C[0]);
                    C[1] = __ffma2_rn(T, __fmul2_rn(alpha, splat(c.y)), C[1]);
                    C[2] = __ffma2_rn(T, __fmul2_rn(alpha, splat(c.z)), C[2]);
                    if (EXTRA) {
                        const float4 e = stage[buf].d[j];
                        C[NCH - 3] = __ffma2_rn(T, __fmul2_rn(alpha, splat(e.x)), C[NCH - 3]);
                        C[NCH - 2] = __ffma2_rn(T, __fmul2_rn(alpha, splat(e.y)), C[NCH - 2]);
                        C[NCH - 1] = __ffma2_rn(T, __fmul2_rn(alpha, splat(e.z)), C[NCH - 1]);
                    }
                    D = __ffma2_rn(T, __fmul2_rn(alpha, splat(b.z)), D);
                    Wsum = __ffma2_rn(alpha, T, Wsum);
                    T.x = c0 ? test_T.x : T.x;
                    T.y = c1 ? test_T.y : T.y;
                    const uint32_t pos1 = contrib_base + (uint32_t)j + 1u;
                    lc0 = c0 ? pos1 : lc0;
                    lc1 = c1 ? pos1 : lc1;
                }
            } else {
                while (m != 0u && !(T.x < 0.0f && T.y < 0.0f)) {
                    // ILP list entries per trip.  A lone heavy tile -- one warp per scheduler at the tail of a small
                    // launch -- is bound by the latency of the per-entry chain, not by issue: the alphas (and the
                    // colour loads) of the trip's entries depend on nothing the loop carries and are computed side by
                    // side; only the short T chain (multiply, compare, select) is walked entry by entry, in list order
                    // and with the same operations as the one-entry loop, so the bits are the same.
                    int jj[ILP];
                    bool has[ILP];
#pragma unroll
                    for (int e = 0; e < ILP; ++e) {
                        has[e] = m != 0u;
                        jj[e] = has[e] ? q * 32 + __ffs(m) - 1 : jj[0];
                        m &= m - 1u;                               // no-op once m is 0
                    }
                    PairAlpha pa[ILP];
                    float4 col[ILP], ext[EXTRA ? ILP : 1];
                    float dep[ILP];
#pragma unroll
                    for (int e = 0; e < ILP; ++e) {
                        const float4 a = stage[buf].a[jj[e]], b = stage[buf].b[jj[e]];
                        col[e] = stage[buf].c[jj[e]];
                        if (EXTRA) ext[e] = stage[buf].d[jj[e]];
                        dep[e] = b.z;
                        pa[e] = pair_alpha(a, b, npx, npy);
                        pa[e].ok0 = pa[e].ok0 && has[e];
                        pa[e].ok1 = pa[e].ok1 && has[e];
                    }
#pragma unroll
                    for (int e = 0; e < ILP; ++e) {
                        float2 alpha = pa[e].alpha;
                        const float2 test_T = __fmul2_rn(T, __fadd2_rn(splat(1.0f), f2(-alpha.x, -alpha.y)));
                        const bool c0 = pa[e].ok0 && !(test_T.x < 0.0001f), c1 = pa[e].ok1 && !(test_T.y < 0.0001f);
                        T.x = (pa[e].ok0 && !c0) ? -fabsf(T.x) : T.x;      // finished from here on (see the T comment)
                        T.y = (pa[e].ok1 && !c1) ? -fabsf(T.y) : T.y;
                        // a pixel that does not blend this entry runs the same instructions with alpha = 0
                        alpha.x = c0 ? alpha.x : 0.0f;
                        alpha.y = c1 ? alpha.y : 0.0f;
                        C[0] = __ffma2_rn(T, __fmul2_rn(alpha, splat(col[e].x)), C[0]);
                        C[1] = __ffma2_rn(T, __fmul2_rn(alpha, splat(col[e].y)), C[1]);
                        C[2] = __ffma2_rn(T, __fmul2_rn(alpha, splat(col[e].z)), C[2]);
                        if (EXTRA) {
                            C[NCH - 3] = __ffma2_rn(T, __fmul2_rn(alpha, splat(ext[EXTRA ? e : 0].x)), C[NCH - 3]);
                            C[NCH - 2] = __ffma2_rn(T, __fmul2_rn(alpha, splat(ext[EXTRA ? e : 0].y)), C[NCH - 2]);
                            C[NCH - 1] = __ffma2_rn(T, __fmul2_rn(alpha, splat(ext[EXTRA ? e : 0].z)), C[NCH - 1]);
                        }
                        D = __ffma2_rn(T, __fmul2_rn(alpha, splat(dep[e])), D);
                        Wsum = __ffma2_rn(alpha, T, Wsum);
                        T.x = c0 ? test_T.x : T.x;
                        T.y = c1 ? test_T.y : T.y;
                        const uint32_t pos1 = contrib_base + (uint32_t)jj[e] + 1u;
                        lc0 = c0 ? pos1 : lc0;
                        lc1 = c1 ? pos1 : lc1;
                    }
                }
            }
        }
        fin = __all_sync(0xffffffffu, T.x < 0.0f && T.y < 0.0f);
        if (fin && lane == 0) atomicAdd(&s_fin, 1u);
        }
        // done with this ring slot; the last of the four warps to say so refills it while anyone still blends
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(&s_done[buf], 1u) == NW2 - 1) {
                s_done[buf] = 0;
                __threadfence_block();
                fence_proxy_async_smem();          // the warps' generic-proxy reads of the slot before the TMA refill
                if (r + RING < rounds && *(volatile uint32_t*)&s_fin < NW2) issue(r + RING);
            }
        }
    }

    {
        const uint32_t wmax = __reduce_max_sync(0xffffffffu, max(lc0, lc1));
        if (lane == 0 && wmax > 0) atomicMax(&s_cost, wmax);
    }
    const CamDev& cam = cams[view];
    const size_t HW = (size_t)H * W;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        if (!(i == 0 ? in0 : in1)) continue;
        const size_t pix = (size_t)(i == 0 ? py0 : py1) * W + px;
        const float Ti = fabsf(i == 0 ? T.x : T.y);
        n_contrib[(size_t)view * HW + pix] = i == 0 ? lc0 : lc1;
        final_T[(size_t)view * HW + pix] = Ti;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
            out_color[((size_t)view * 3 + ch) * HW + pix] = __fmaf_rn(Ti, cam.bg[ch], i == 0 ? C[ch].x : C[ch].y);
        if (EXTRA) {
#pragma unroll
            for (int ch = 0; ch < 3; ++ch)
                out_extra[((size_t)view * 3 + ch) * HW + pix] =
                    __fmaf_rn(Ti, cam.bg[3 + ch], i == 0 ? C[NCH - 3 + ch].x : C[NCH - 3 + ch].y);
        }
        out_depth[(size_t)view * HW + pix] = i == 0 ? D.x : D.y;
        out_alpha[(size_t)view * HW + pix] = i == 0 ? Wsum.x : Wsum.y;
    }
    // the CTA must not retire under a batch that was fetched and never consumed
    __syncthreads();
    if (tid == 0) {
        tile_cost[vt] = s_cost;
#pragma unroll
        for (int i = 0; i < RING; ++i) {
            const int q = s_slot_round[i];
            if (q >= 0) mbar_wait(&full[i], (uint32_t)((q / RING) & 1));
        }
    }
    if (trace && lane == 0) atomicMax(&trace[4 * blockIdx.x + 1], trace_now());
}

}  // namespace

int set_blend_forward_trace(unsigned long long* p) {
    GSR_CUDA_TRY(cudaMemcpyToSymbol(g_trace_fwd, &p, sizeof(p)));
    return GSR_OK;
}

int blend_ilp(int blocks) {
    const int forced = options().blend_ilp;             // gsr_set_option("blend_ilp", 1 | 2): tuning and tests
    if (forced == 1 || forced == 2) return forced;
    return blocks <= BLEND_ILP_MAX_BLOCKS ? BLEND_ILP_SMALL : 1;
}

int launch_blend_forward(int V, int W, int H, int channels, const void* gws, const void* bws, void* iws,
                         const Layout& L, float* out_color, float* out_extra, float* out_depth, float* out_alpha,
                         cudaStream_t s) {
    const int T = L.tiles_x * L.tiles_y;
    const int blocks = V * T;
    if (blocks == 0) return GSR_OK;
    const uint2* ranges = at<uint2>(iws, L.ranges);
    const float4* A = at<float4>(bws, L.recA);
    const float4* B = at<float4>(bws, L.recB);
    const float4* C = at<float4>(bws, L.recC);
    const float4* D = at<float4>(bws, L.recD);
    const CamDev* cams = at<CamDev>(gws, L.cams);
    uint32_t* nc = at<uint32_t>(iws, L.n_contrib);
    float* fT = at<float>(iws, L.final_T);
    const int ilp = blend_ilp(blocks);           // see common.cuh: small launches end with lone heavy tiles
#define GSR_FWD(NCH_, ILP_)                                                                                           \
    launch_kernel(blend_forward_kernel<NCH_, ILP_>, blocks, NT2, 0, s, ranges, at<uint32_t>(iws, L.tile_order), A, B, C, D, cams, \
                                                            W, H, L.tiles_x, T, out_color, out_extra, out_depth,      \
                                                            out_alpha, nc, fT, at<uint32_t>(iws, L.tile_cost))
    if (channels == 3) {
        if (ilp > 1) GSR_FWD(3, BLEND_ILP_SMALL); else GSR_FWD(3, 1);
    } else if (channels == 6) {
        if (ilp > 1) GSR_FWD(6, BLEND_ILP_SMALL); else GSR_FWD(6, 1);
    } else {
        return GSR_ERR_UNSUPPORTED;
    }
#undef GSR_FWD
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
