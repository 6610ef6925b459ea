// (d) tile blend backward, all views of the batch in one launch.
// Replaces BACKWARD::render / renderCUDA (R/cuda_rasterizer/backward.cu:432-652).
//
// Differences in structure from the reference (the gradients computed are the same quantities):
//  * per-instance records arrive through the same ring of TMA bulk copies as the forward, walked back to front by
//    every warp on its own (no CTA barrier per batch); the walk starts at the tile's largest n_contrib instead of at
//    the end of the list;
//  * the colour / depth / alpha "colour behind me" recurrences (backward.cu:578,591,600,617) and the background term
//    (backward.cu:626-629) collapse into ONE scalar recurrence, because the pixel cotangent is fixed per thread:
//        s_j   = <c_j, dL/dC> + depth_j * dL/dD + dL/dA          (extended colour dotted with the pixel cotangent)
//        acc_j = acc_{j+1} + alpha_{j+1} * (s_{j+1} - acc_{j+1}),   acc_last = <bg, dL/dC>
//        dL/dalpha_j = T_j * (s_j - acc_j)
//    which removes ~3 registers and ~3 flops per blended channel per pair (the reference kernel needs 198 registers);
//  * what is reduced over the pixels are raw moments of a = opacity * G * dL/dalpha about d = mean - pixel (S1 = sum a d,
//    S2 = sum a d d^T, sum G dL/dalpha, and the colour / depth weights); preprocess_backward applies the chain rule
//    through the conic once per Gaussian;
//  * the 7+C partial sums of a warp (an 8x8 pixel patch) are reduced with a packed butterfly (12 shuffles for 10 values
//    instead of 50) and land in the per-(view, Gaussian) gradient record with ONE reduction per (patch, Gaussian,
//    value): the reference issues one atomic per (pixel, Gaussian, value) (backward.cu:586-649).
#include "geom_math.cuh"
#include "packed_reduce.cuh"

#ifndef GSR_PATCH_TEST
#define GSR_PATCH_TEST blend_patch_hit_exact
#endif

namespace gsr {

namespace {

__device__ unsigned long long* g_trace_bwd = nullptr;   // diagnostic: per-CTA (start ns, end ns, SM, tile)
__device__ __forceinline__ unsigned long long trace_now() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ unsigned trace_smid() { unsigned r; asm volatile("mov.u32 %0, %smid;" : "=r"(r)); return r; }

constexpr int BATCH = 128;
constexpr int SUB = 32;
#ifndef GSR_BWD_RING
#define GSR_BWD_RING 3
#endif
constexpr int RING = GSR_BWD_RING;                  // staged batches in flight per CTA
constexpr unsigned FULL = 0xffffffffu;

// CTA = 128 threads = 4 warps; a warp owns an 8x8 pixel patch, a lane the two pixels (x, y) and (x, y + 4) of it, so
// dx is shared by the pair and the per-(warp, Gaussian) costs (record loads, loop control, the packed butterfly, the
// atomics) are paid once per 64 pixels.  The pair's arithmetic is issued as packed
// fp32x2 instructions (FFMA2 / FMUL2 / FADD2, sm_100) where both pixels do the same operation; the kernel is bound by
// instruction issue, not by the FMA pipe.  Invalid pixels run the same instructions with alpha = 0 and a zeroed
// dL/dalpha, which leaves T, the recurrence and every sum unchanged (no divergent region).
__device__ __forceinline__ float2 f2(float a, float b) { return make_float2(a, b); }
__device__ __forceinline__ float2 splat(float a) { return make_float2(a, a); }
// predicated RED (no branch around the atomic)
__device__ __forceinline__ void red_add_if(bool p, float* addr, float v) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.u32 p, %2, 0;\n\t@p red.global.add.f32 [%0], %1;\n\t}" ::"l"(addr), "f"(v), "r"((unsigned)p) : "memory");
}
__device__ __forceinline__ float rcp_approx(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }

constexpr int NT2 = 128;
constexpr int NW2 = NT2 / 32;

// M2C (only with the fused extra colours, NCH == 6): additionally reduce the screen-space mean moments of the COLOUR
// render alone -- colour, depth and alpha cotangents, without the extra (segmentation) channels -- into record slots
// GR_COL + NCH and + 1.  The reference takes its densification statistic from the colour render's dL/dmeans2D only
// (REF/train.py:245, REF/external.py:136-140; REF/README.md:89 on why fusing the two renders changes t = 0 otherwise).
#ifndef GSR_BWD_MINB
#define GSR_BWD_MINB 8
#endif
template <int NCH, bool M2C, int BILP>
__global__ void __launch_bounds__(NT2, BILP > 1 ? 4 : GSR_BWD_MINB)
blend_backward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order,
                       const float4* __restrict__ recA, const float4* __restrict__ recB,
                       const float4* __restrict__ recC, const float4* __restrict__ recD,
                       const CamDev* __restrict__ cams, int W, int H, int tiles_x, int tiles_per_view,
                       const uint32_t* __restrict__ n_contrib, const float* __restrict__ final_T,
                       const float* __restrict__ dL_dcolor, const float* __restrict__ dL_dextra,
                       const float* __restrict__ dL_ddepth, const float* __restrict__ dL_dalpha,
                       float* __restrict__ grad_rec) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    constexpr bool EXTRA = NCH > 3;
    static_assert(!M2C || EXTRA, "the colour-only moments only differ from the totals when extra colours are blended");
    constexpr int K = GR_COL + NCH + (M2C ? 2 : 0);
    static_assert(K <= grad_rec_stride(NCH), "gradient record too small");
    // one block per ring slot, so that the loop addresses all of an entry's records off ONE register (base + 16 j)
    struct __align__(128) Stage {
        float4 a[BATCH], b[BATCH], c[BATCH], d[EXTRA ? BATCH : 1];
    };
    // The four warps of the CTA walk the ring on their own: each builds the work list of its patch from the staged
    // batch, and the last warp out of a ring slot (an atomic count) refills it.  No CTA barrier inside the walk --
    // with a barrier per batch the warps of a tile waited for the one with the longest work list every 128 entries
    // (ncu: barrier was the largest stall reason).
    __shared__ Stage stage[RING];
    __shared__ __align__(8) uint64_t full[RING];
    __shared__ uint32_t s_done[RING];
    __shared__ uint32_t s_max;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int vt = (int)tile_order[blockIdx.x];
    unsigned long long* const trace = g_trace_bwd;
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
    const size_t HW = (size_t)H * W;
    const size_t pix0 = (size_t)view * HW + (size_t)py0 * W + px, pix1 = pix0 + 4 * (size_t)W;

    const uint2 range = ranges[vt];
    const uint32_t lc0 = in0 ? n_contrib[pix0] : 0u, lc1 = in1 ? n_contrib[pix1] : 0u;
    float2 T = f2(in0 ? final_T[pix0] : 0.0f, in1 ? final_T[pix1] : 0.0f);

    if (tid == 0) {
        s_max = 0;
#pragma unroll
        for (int i = 0; i < RING; ++i) {
            mbar_init(&full[i], 1);
            s_done[i] = 0;
        }
        mbar_fence_init();
    }
    __syncthreads();
    const uint32_t warp_last = __reduce_max_sync(FULL, max(lc0, lc1));   // nothing at or past it matters to this warp
    if (lane == 0 && warp_last > 0) atomicMax(&s_max, warp_last);
    __syncthreads();
    const int n_proc = (int)min(s_max, range.y - range.x);
    const int rounds = (n_proc + BATCH - 1) / BATCH;
    if (rounds == 0) {
        if (trace && lane == 0) atomicMax(&trace[4 * blockIdx.x + 1], trace_now());
        return;
    }

    auto issue = [&](int r) {   // one thread; round r walks chunk (rounds-1-r)
        const int buf = r % RING;
        const int c = rounds - 1 - r;
        const int nb = min(BATCH, n_proc - c * BATCH);
        const uint32_t bytes = (uint32_t)nb * 16u;
        const size_t src = (size_t)range.x + (size_t)c * BATCH;
        mbar_expect_tx(&full[buf], bytes * (EXTRA ? 4u : 3u));
        tma_load_1d(stage[buf].a, recA + src, bytes, &full[buf]);
        tma_load_1d(stage[buf].b, recB + src, bytes, &full[buf]);
        tma_load_1d(stage[buf].c, recC + src, bytes, &full[buf]);
        if (EXTRA) tma_load_1d(stage[buf].d, recD + src, bytes, &full[buf]);
    };
    if (tid == 0) {
        for (int r = 0; r < RING && r < rounds; ++r) issue(r);
    }

    // pixel cotangents (x: first pixel, y: second)
    float2 gC[NCH];
    float2 gD = splat(0.0f), gA = splat(0.0f);
#pragma unroll
    for (int k = 0; k < NCH; ++k) gC[k] = splat(0.0f);
    const CamDev& cam = cams[view];
    float2 acc = splat(0.0f);
    float2 acc_c = splat(0.0f);                // M2C: the recurrence of the colour render alone
    {
        const size_t o0 = pix0 - (size_t)view * HW, o1 = pix1 - (size_t)view * HW;
        if (dL_dcolor) {
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float* g = dL_dcolor + ((size_t)view * 3 + ch) * HW;
                gC[ch] = f2(in0 ? g[o0] : 0.0f, in1 ? g[o1] : 0.0f);
                acc.x += cam.bg[ch] * gC[ch].x;
                acc.y += cam.bg[ch] * gC[ch].y;
            }
        }
        acc_c = acc;
        if (EXTRA && dL_dextra) {
#pragma unroll
            for (int ch = 0; ch < 3; ++ch) {
                const float* g = dL_dextra + ((size_t)view * 3 + ch) * HW;
                gC[NCH - 3 + ch] = f2(in0 ? g[o0] : 0.0f, in1 ? g[o1] : 0.0f);
                acc.x += cam.bg[3 + ch] * gC[NCH - 3 + ch].x;
                acc.y += cam.bg[3 + ch] * gC[NCH - 3 + ch].y;
            }
        }
        if (dL_ddepth) gD = f2(in0 ? dL_ddepth[pix0] : 0.0f, in1 ? dL_ddepth[pix1] : 0.0f);
        if (dL_dalpha) gA = f2(in0 ? dL_dalpha[pix0] : 0.0f, in1 ? dL_dalpha[pix1] : 0.0f);
    }

    const int my_slot = PackedReduce<K>::slot(lane);
    const unsigned slot_peers = __match_any_sync(FULL, my_slot);
    const bool writer = lane == (__ffs(slot_peers) - 1);
    const float patch_x0 = (float)(tile_xi * TILE_X + (warp & 1) * 8), patch_y0 = (float)(tile_yi * TILE_Y + (warp >> 1) * 8);

#ifdef GSR_TRACE_COUNTS
    unsigned long long n_list = 0, n_done = 0;      // work-list entries walked / entries some pixel of the patch blends
#endif
    for (int r = 0; r < rounds; ++r) {
        const int buf = r % RING;
        const int c = rounds - 1 - r;
        const int nb = min(BATCH, n_proc - c * BATCH);
        const int rel0 = (int)lc0 - c * BATCH, rel1 = (int)lc1 - c * BATCH;   // list position < n_contrib  <=>  j < rel
        mbar_wait(&full[buf], (uint32_t)((r / RING) & 1));
        // pre-pass: this warp's work list, 32 staged Gaussians per ballot against its 8x8 patch
        uint32_t wm[BATCH / SUB];
#pragma unroll
        for (int sb = 0; sb < BATCH / SUB; ++sb) {
            const int g = sb * SUB + lane;
            bool hit = false;
            if (g < nb && (uint32_t)(c * BATCH + sb * SUB) < warp_last) {
                const float4 a = stage[buf].a[g];
                const float4 b = stage[buf].b[g];
                hit = GSR_PATCH_TEST(a.x, a.y, a.z, a.w, b.x, b.w, patch_x0, patch_y0);
            }
            wm[sb] = __ballot_sync(FULL, hit);
        }
#pragma unroll
        for (int sub = BATCH / SUB - 1; sub >= 0; --sub) {
            const int lo = sub * SUB;
            uint32_t m = wm[sub];
            {
                const uint32_t base = (uint32_t)(c * BATCH + lo);
                if (warp_last <= base) m = 0u;
                else if (warp_last - base < 32u) m &= (1u << (warp_last - base)) - 1u;
            }
#ifdef GSR_TRACE_COUNTS
            n_list += __popc(m);
#endif
            if constexpr (BILP == 1) {
            while (m != 0u) {                                       // warp-uniform, back to front
                const int jl = 31 - __clz(m);
                m &= ~(1u << jl);
                const int j = lo + jl;
                const float4 a = stage[buf].a[j];
                const float4 b = stage[buf].b[j];
                // blend_power (geom_math.cuh) on the pixel pair; identical roundings per component
                const float dx = __fadd_rn(a.x, npx);
                const float2 dy = __fadd2_rn(splat(a.y), npy);
                const float dxca = __fmul_rn(dx, a.z), ndxcb = __fmul_rn(dx, -a.w);
                const float2 t3 = __fmul2_rn(dy, __fmul2_rn(dy, splat(b.x)));
                const float2 q = __ffma2_rn(splat(dx), splat(dxca), t3);
                const float2 power = __ffma2_rn(q, splat(-0.5f), __fmul2_rn(dy, splat(ndxcb)));
                const bool pre0 = (j < rel0) && !(power.x > 0.0f) && !(power.x < b.w);
                const bool pre1 = (j < rel1) && !(power.y > 0.0f) && !(power.y < b.w);
                if (!__any_sync(FULL, pre0 || pre1)) continue;      // warp-uniform
#ifdef GSR_TRACE_COUNTS
                ++n_done;
#endif
                const float2 G = expf2_pinned(power);        // garbage where !pre: masked by the selects below
                const float2 og = __fmul2_rn(splat(b.y), G);
                float2 alpha = f2(fminf(0.99f, og.x), fminf(0.99f, og.y));
                const bool valid0 = pre0 && !(alpha.x < 1.0f / 255.0f);
                const bool valid1 = pre1 && !(alpha.y < 1.0f / 255.0f);
                alpha.x = valid0 ? alpha.x : 0.0f;
                alpha.y = valid1 ? alpha.y : 0.0f;
                const float2 om = __fadd2_rn(splat(1.0f), f2(-alpha.x, -alpha.y));
                // transmittance in front of this Gaussian: T / (1 - alpha) as __fdividef computes it (MUFU.RCP, then a
                // multiply) minus its rescaling path for denormal divisors, which 1 - alpha in [0.01, 1] never takes
                T = __fmul2_rn(T, f2(rcp_approx(om.x), rcp_approx(om.y)));
                const float2 w = __fmul2_rn(alpha, T);
                const float4 col = stage[buf].c[j];
                float2 s = __ffma2_rn(splat(col.x), gC[0],
                           __ffma2_rn(splat(col.y), gC[1],
                           __ffma2_rn(splat(col.z), gC[2], __ffma2_rn(splat(b.z), gD, gA))));
                float v[K];
#pragma unroll
                for (int k = 0; k < 3; ++k) v[GR_COL + k] = fmaf(w.y, gC[k].y, w.x * gC[k].x);
                float2 s_c;
                if (M2C) s_c = s;
                if (EXTRA) {
                    const float4 e = stage[buf].d[j];
                    s = __ffma2_rn(splat(e.x), gC[NCH - 3],
                        __ffma2_rn(splat(e.y), gC[NCH - 2], __ffma2_rn(splat(e.z), gC[NCH - 1], s)));
#pragma unroll
                    for (int k = NCH - 3; k < NCH; ++k) v[GR_COL + k] = fmaf(w.y, gC[k].y, w.x * gC[k].x);
                }
                v[GR_DEPTH] = fmaf(w.y, gD.y, w.x * gD.x);
                // colour-behind recurrence (see header); alpha = 0 leaves acc as it is
                const float2 sd = __fadd2_rn(s, f2(-acc.x, -acc.y));          // s_j - acc_j
                float2 gq = __fmul2_rn(G, __fmul2_rn(sd, T));                   // G * dL/dalpha
                acc = __ffma2_rn(alpha, sd, acc);                               // what the next (nearer) entry has behind it
                gq.x = valid0 ? gq.x : 0.0f;
                gq.y = valid1 ? gq.y : 0.0f;
                const float2 am = __fmul2_rn(splat(b.y), gq);
                const float2 uy = __fmul2_rn(am, dy);
                const float sam = am.x + am.y;
                v[GR_MX] = sam * dx;
                v[GR_MY] = uy.x + uy.y;
                v[GR_CA] = v[GR_MX] * dx;
                v[GR_CB] = v[GR_MY] * dx;
                v[GR_CC] = fmaf(uy.y, dy.y, uy.x * dy.x);
                v[GR_OP] = gq.x + gq.y;
                if (M2C) {
                    const float2 sd_c = __fadd2_rn(s_c, f2(-acc_c.x, -acc_c.y));
                    float2 gq_c = __fmul2_rn(G, __fmul2_rn(sd_c, T));
                    acc_c = __ffma2_rn(alpha, sd_c, acc_c);
                    gq_c.x = valid0 ? gq_c.x : 0.0f;
                    gq_c.y = valid1 ? gq_c.y : 0.0f;
                    const float2 am_c = __fmul2_rn(splat(b.y), gq_c);
                    v[GR_COL + NCH] = (am_c.x + am_c.y) * dx;
                    v[GR_COL + NCH + 1] = fmaf(am_c.y, dy.y, am_c.x * dy.x);
                }
                PackedReduce<K>::run(v, lane);
                // one reduction (RED) per (warp patch, Gaussian, value): the K sums sit in K different lanes and land in
                // one 64-byte gradient record
                red_add_if(writer, grad_rec + (size_t)__float_as_uint(col.w) * grad_rec_stride(NCH) + my_slot, v[0]);
            }
            } else {
            // BILP list entries per trip, for launches of at most one wave of CTAs (a training view): every tile is
            // resident from the first microsecond, the kernel lasts as long as its deepest tile -- one warp per
            // scheduler for most of the time -- and that warp's time is the latency of its
            // instruction stream -- an exp, two reciprocals and five dependent shuffle stages per entry -- not issue
            // slots.  The alphas, the cotangent dot products and the two butterflies of the trip's entries depend on
            // nothing the loop carries and run side by side; only T and the colour-behind recurrence are walked entry by
            // entry, in list order and with the one-entry loop's operations, so every value has the same bits.  (On the
            // issue-bound full-size launch the extra instructions of the padding entry cost more than they hide.)
            while (m != 0u) {
                int jj[BILP];
                bool has[BILP];
#pragma unroll
                for (int e = 0; e < BILP; ++e) {
                    has[e] = m != 0u;
                    const int jl = 31 - __clz(m | 1u);
                    m &= ~(1u << jl);                              // (no-op once m is 0)
                    jj[e] = has[e] ? lo + jl : jj[0];
                }
                float dxe[BILP], opa[BILP];
                float2 dye[BILP], Ge[BILP], alpha[BILP], rom[BILP], se[BILP], sce[BILP];
                float4 cole[BILP];
                bool v0e[BILP], v1e[BILP];
                bool any_pre = false;
#pragma unroll
                for (int e = 0; e < BILP; ++e) {
                    const int j = jj[e];
                    const float4 a = stage[buf].a[j];
                    const float4 b = stage[buf].b[j];
                    cole[e] = stage[buf].c[j];
                    const float dx = __fadd_rn(a.x, npx);
                    const float2 dy = __fadd2_rn(splat(a.y), npy);
                    const float dxca = __fmul_rn(dx, a.z), ndxcb = __fmul_rn(dx, -a.w);
                    const float2 t3 = __fmul2_rn(dy, __fmul2_rn(dy, splat(b.x)));
                    const float2 q = __ffma2_rn(splat(dx), splat(dxca), t3);
                    const float2 power = __ffma2_rn(q, splat(-0.5f), __fmul2_rn(dy, splat(ndxcb)));
                    const bool pre0 = has[e] && (j < rel0) && !(power.x > 0.0f) && !(power.x < b.w);
                    const bool pre1 = has[e] && (j < rel1) && !(power.y > 0.0f) && !(power.y < b.w);
                    any_pre = any_pre || pre0 || pre1;
                    const float2 G = expf2_pinned(power);
                    const float2 og = __fmul2_rn(splat(b.y), G);
                    float2 al = f2(fminf(0.99f, og.x), fminf(0.99f, og.y));
                    v0e[e] = pre0 && !(al.x < 1.0f / 255.0f);
                    v1e[e] = pre1 && !(al.y < 1.0f / 255.0f);
                    al.x = v0e[e] ? al.x : 0.0f;
                    al.y = v1e[e] ? al.y : 0.0f;
                    const float2 om = __fadd2_rn(splat(1.0f), f2(-al.x, -al.y));
                    rom[e] = f2(rcp_approx(om.x), rcp_approx(om.y));
                    float2 s = __ffma2_rn(splat(cole[e].x), gC[0],
                               __ffma2_rn(splat(cole[e].y), gC[1],
                               __ffma2_rn(splat(cole[e].z), gC[2], __ffma2_rn(splat(b.z), gD, gA))));
                    sce[e] = s;
                    if (EXTRA) {
                        const float4 x = stage[buf].d[j];
                        s = __ffma2_rn(splat(x.x), gC[NCH - 3],
                            __ffma2_rn(splat(x.y), gC[NCH - 2], __ffma2_rn(splat(x.z), gC[NCH - 1], s)));
                    }
                    se[e] = s;
                    dxe[e] = dx; dye[e] = dy; Ge[e] = G; alpha[e] = al; opa[e] = b.y;
                }
                if (!__any_sync(FULL, any_pre)) continue;          // warp-uniform: none of the trip's entries blends here
                float ve[BILP][K];
#pragma unroll
                for (int e = 0; e < BILP; ++e) {
#ifdef GSR_TRACE_COUNTS
                    n_done += has[e] ? 1 : 0;
#endif
                    // the carried part, entry by entry: alpha = 0 (entry absent or blending nothing here) leaves T and acc
                    // exactly as they are (MUFU.RCP(1) = 1)
                    T = __fmul2_rn(T, rom[e]);
                    const float2 w = __fmul2_rn(alpha[e], T);
                    const float2 sd = __fadd2_rn(se[e], f2(-acc.x, -acc.y));
                    float2 gq = __fmul2_rn(Ge[e], __fmul2_rn(sd, T));
                    acc = __ffma2_rn(alpha[e], sd, acc);
                    gq.x = v0e[e] ? gq.x : 0.0f;
                    gq.y = v1e[e] ? gq.y : 0.0f;
                    float* v = ve[e];
#pragma unroll
                    for (int k = 0; k < NCH; ++k) v[GR_COL + k] = fmaf(w.y, gC[k].y, w.x * gC[k].x);
                    v[GR_DEPTH] = fmaf(w.y, gD.y, w.x * gD.x);
                    const float2 am = __fmul2_rn(splat(opa[e]), gq);
                    const float2 uy = __fmul2_rn(am, dye[e]);
                    const float sam = am.x + am.y;
                    v[GR_MX] = sam * dxe[e];
                    v[GR_MY] = uy.x + uy.y;
                    v[GR_CA] = v[GR_MX] * dxe[e];
                    v[GR_CB] = v[GR_MY] * dxe[e];
                    v[GR_CC] = fmaf(uy.y, dye[e].y, uy.x * dye[e].x);
                    v[GR_OP] = gq.x + gq.y;
                    if (M2C) {
                        const float2 sd_c = __fadd2_rn(sce[e], f2(-acc_c.x, -acc_c.y));
                        float2 gq_c = __fmul2_rn(Ge[e], __fmul2_rn(sd_c, T));
                        acc_c = __ffma2_rn(alpha[e], sd_c, acc_c);
                        gq_c.x = v0e[e] ? gq_c.x : 0.0f;
                        gq_c.y = v1e[e] ? gq_c.y : 0.0f;
                        const float2 am_c = __fmul2_rn(splat(opa[e]), gq_c);
                        v[GR_COL + NCH] = (am_c.x + am_c.y) * dxe[e];
                        v[GR_COL + NCH + 1] = fmaf(am_c.y, dye[e].y, am_c.x * dye[e].x);
                    }
                }
#pragma unroll
                for (int e = 0; e < BILP; ++e) PackedReduce<K>::run(ve[e], lane);
#pragma unroll
                for (int e = 0; e < BILP; ++e)
                    red_add_if(writer && has[e], grad_rec + (size_t)__float_as_uint(cole[e].w) * grad_rec_stride(NCH) + my_slot,
                               ve[e][0]);
            }
            }
        }
        // done with this ring slot; the last of the four warps to say so refills it
        __syncwarp();
        if (lane == 0) {
            __threadfence_block();
            if (atomicAdd(&s_done[buf], 1u) == NW2 - 1) {
                s_done[buf] = 0;
                __threadfence_block();
                fence_proxy_async_smem();          // the warps' generic-proxy reads of the slot before the TMA refill
                if (r + RING < rounds) issue(r + RING);
            }
        }
    }
#ifdef GSR_TRACE_COUNTS
    if (trace && lane == 0) {       // diagnostic build: counts ride in the upper bits of the SM / tile fields
        atomicAdd(&trace[4 * blockIdx.x + 2], n_list << 16);
        atomicAdd(&trace[4 * blockIdx.x + 3], n_done << 32);
    }
#endif
    if (trace && lane == 0) atomicMax(&trace[4 * blockIdx.x + 1], trace_now());
}

// Launch order of a SMALL backward launch (a few views: 3-4 waves of CTAs): tiles by the depth the forward walked them
// (their largest n_contrib -- exactly how far the backward walks), deepest first.  The list-length order of tile_scan
// starts the deepest tiles in the second wave (long lists saturate early; the deepest walks belong to mid-length lists
// of faint Gaussians), and the launch then ends on a handful of them running alone (tests/gpu_tile_trace.py -4).
// One CTA: counting sort into 1024 buckets.
constexpr int ORD_THREADS = 1024;
__global__ void __launch_bounds__(ORD_THREADS)
tile_order_by_cost_kernel(uint32_t VT, const uint32_t* __restrict__ cost, uint32_t* __restrict__ order) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ uint32_t s_bucket[ORD_THREADS];
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_maxc;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (threadIdx.x == 0) s_maxc = 0;
    s_bucket[threadIdx.x] = 0;
    __syncthreads();
    uint32_t vmax = 0;
    for (uint32_t i = threadIdx.x; i < VT; i += ORD_THREADS) vmax = max(vmax, cost[i]);
    vmax = __reduce_max_sync(FULL, vmax);
    if (lane == 0 && vmax > 0) atomicMax(&s_maxc, vmax);
    __syncthreads();
    const uint32_t q = s_maxc / (ORD_THREADS - 1) + 1;
    for (uint32_t i = threadIdx.x; i < VT; i += ORD_THREADS)
        atomicAdd(&s_bucket[(ORD_THREADS - 1) - min(cost[i] / q, (uint32_t)ORD_THREADS - 1)], 1u);   // bucket 0: deepest
    __syncthreads();
    {   // exclusive scan of the 1024 bucket counts
        const uint32_t c = s_bucket[threadIdx.x];
        uint32_t inc = c;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp[warp] = inc;
        __syncthreads();
        uint32_t w = s_warp[lane], winc = w;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULL, winc, o);
            if (lane >= o) winc += t;
        }
        const uint32_t base = __shfl_sync(FULL, winc - w, warp);
        __syncthreads();
        s_bucket[threadIdx.x] = base + inc - c;
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < VT; i += ORD_THREADS) {
        const uint32_t pos = atomicAdd(&s_bucket[(ORD_THREADS - 1) - min(cost[i] / q, (uint32_t)ORD_THREADS - 1)], 1u);
        order[pos] = i;
    }
}

}  // namespace

int set_blend_backward_trace(unsigned long long* p) {
    GSR_CUDA_TRY(cudaMemcpyToSymbol(g_trace_bwd, &p, sizeof(p)));
    return GSR_OK;
}

int launch_blend_backward(int V, int P, int W, int H, int channels, void* gws, const void* bws, void* iws,
                          const Layout& L, const float* dL_dcolor, const float* dL_dextra, const float* dL_ddepth,
                          const float* dL_dalpha, bool color_only_means2D, cudaStream_t s) {
    const int T = L.tiles_x * L.tiles_y;
    const int blocks = V * T;
    {
        StageScope sc(ST_GRAD_ZERO, s);
        GSR_CUDA_TRY(cudaMemsetAsync(at<float>(gws, L.grad_rec), 0, (size_t)V * P * grad_rec_stride(channels) * 4, s));
    }
    if (blocks == 0 || P == 0) return GSR_OK;
    StageScope sc(ST_BLEND_BWD, s);
    const uint2* ranges = at<uint2>(iws, L.ranges);
    const float4* A = at<float4>(bws, L.recA);
    const float4* B = at<float4>(bws, L.recB);
    const float4* C = at<float4>(bws, L.recC);
    const float4* D = at<float4>(bws, L.recD);
    const CamDev* cams = at<CamDev>(gws, L.cams);
    const uint32_t* nc = at<uint32_t>(iws, L.n_contrib);
    const float* fT = at<float>(iws, L.final_T);
    float* rec = at<float>(gws, L.grad_rec);
    const uint32_t* order = at<uint32_t>(iws, L.tile_order);
    if (options().bwd_cost_order && blocks <= BWD_COST_ORDER_MAX_BLOCKS) {          // small launch: deepest tiles first (see tile_order_by_cost_kernel)
        uint32_t* ord = at<uint32_t>(iws, L.tile_order_bwd);
        launch_kernel(tile_order_by_cost_kernel, 1, ORD_THREADS, 0, s, (uint32_t)blocks, at<uint32_t>(iws, L.tile_cost), ord);
        GSR_LAUNCH_CHECK();
        order = ord;
    }
    // two entries per trip for small launches (see the kernel); option "bwd_ilp": 0 = by launch size, 1 / 2 = forced
    const int forced = options().bwd_ilp;
    const bool two = forced == 2 || (forced != 1 && blocks <= BWD_ILP_MAX_BLOCKS);
#define GSR_BWD(NCH_, M2C_)                                                                                             \
    do {                                                                                                                \
        if (two)                                                                                                        \
            launch_kernel(blend_backward_kernel<NCH_, M2C_, 2>, blocks, NT2, 0, s, ranges, order, A, B, C, D, cams, W, H, \
                          L.tiles_x, T, nc, fT, dL_dcolor, dL_dextra, dL_ddepth, dL_dalpha, rec);                       \
        else                                                                                                            \
            launch_kernel(blend_backward_kernel<NCH_, M2C_, 1>, blocks, NT2, 0, s, ranges, order, A, B, C, D, cams, W, H, \
                          L.tiles_x, T, nc, fT, dL_dcolor, dL_dextra, dL_ddepth, dL_dalpha, rec);                       \
    } while (0)
    if (channels == 3)
        GSR_BWD(3, false);
    else if (channels == 6 && color_only_means2D)
        GSR_BWD(6, true);
    else if (channels == 6)
        GSR_BWD(6, false);
    else
        return GSR_ERR_UNSUPPORTED;
#undef GSR_BWD
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
