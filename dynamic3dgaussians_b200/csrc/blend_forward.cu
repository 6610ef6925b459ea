// (c) 16x16-tile forward alpha blend, all views of the batch in one launch.
// Replaces FORWARD::render / renderCUDA (R/cuda_rasterizer/forward.cu:274-408).
//
// * one CTA per (view, tile); sorted per-instance records are contiguous, so each 128-instance batch is staged into
//   shared memory by 3-4 TMA bulk copies (cp.async.bulk + mbarrier), double buffered: no per-pair global gathers
//   (the reference fetches colour/depth/semantics from global memory for every contributing pair, forward.cu:375-383);
// * colour, fg/bg segmentation colours, depth and accumulated alpha are blended in ONE pass over the list
//   (the reference renders colour and segmentation with two full rasterizer calls, REF/train.py:142,249);
// * a warp covers an 8x4 pixel patch.  Per staged batch a pre-pass (one thread per Gaussian) intersects the Gaussian's
//   opacity-aware extent (the ellipse outside of which alpha < 1/255) with the 8 patches and warp ballots turn that into
//   a 128-bit work list per warp: a warp only iterates over Gaussians that can reach its patch (2-3x fewer
//   iterations than walking the tile's whole list as the reference does), and each lane tests `power < threshold`
//   before paying for the exp.  Both filters are conservative, so the set of blended pairs is unchanged;
// * a warp leaves the loop as soon as all of its pixels are saturated and the CTA stops fetching when all 256 are;
// * arithmetic per pair follows the reference binary exactly (see geom_math.cuh), so n_contrib and the images agree
//   with the reference bit for bit.
#include "geom_math.cuh"

namespace gsr {

namespace {

constexpr int BATCH = 128;

template <int NCH>
__global__ void __launch_bounds__(TILE_PIX)
blend_forward_kernel(const uint2* __restrict__ ranges, const uint32_t* __restrict__ tile_order, const float4* __restrict__ recA,
                     const float4* __restrict__ recB, const float4* __restrict__ recC,
                     const float4* __restrict__ recD, const CamDev* __restrict__ cams, int W, int H, int tiles_x,
                     int tiles_per_view, float* __restrict__ out_color, float* __restrict__ out_extra,
                     float* __restrict__ out_depth, float* __restrict__ out_alpha, uint32_t* __restrict__ n_contrib,
                     float* __restrict__ final_T) {
    constexpr bool EXTRA = NCH > 3;
    __shared__ __align__(128) float4 sA[2][BATCH];
    __shared__ __align__(128) float4 sB[2][BATCH];
    __shared__ __align__(128) float4 sC[2][BATCH];
    __shared__ __align__(128) float4 sD[EXTRA ? 2 : 1][EXTRA ? BATCH : 1];
    __shared__ __align__(8) uint64_t full[2];
    __shared__ uint32_t s_wmask[TILE_PIX / 32][BATCH / 32];
    static_assert(BATCH <= TILE_PIX && BATCH % 32 == 0, "one pre-pass thread per staged Gaussian");

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int vt = (int)tile_order[blockIdx.x];
    const int view = vt / tiles_per_view;
    const int tile = vt - view * tiles_per_view;
    const int tile_yi = tile / tiles_x, tile_xi = tile - tile_yi * tiles_x;
    const int px = tile_xi * TILE_X + (warp & 1) * 8 + (lane & 7);
    const int py = tile_yi * TILE_Y + (warp >> 1) * 4 + (lane >> 3);
    const bool inside = px < W && py < H;
    const float pxf = (float)px, pyf = (float)py;

    const uint2 range = ranges[vt];
    const int n_total = (int)(range.y - range.x);
    const int rounds = (n_total + BATCH - 1) / BATCH;

    if (tid == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        mbar_fence_init();
    }
    __syncthreads();

    auto issue = [&](int r) {   // thread 0 only
        const int buf = r & 1;
        const int nb = min(BATCH, n_total - r * BATCH);
        const uint32_t bytes = (uint32_t)nb * 16u;
        const size_t src = (size_t)range.x + (size_t)r * BATCH;
        mbar_expect_tx(&full[buf], bytes * (EXTRA ? 4u : 3u));
        tma_load_1d(sA[buf], recA + src, bytes, &full[buf]);
        tma_load_1d(sB[buf], recB + src, bytes, &full[buf]);
        tma_load_1d(sC[buf], recC + src, bytes, &full[buf]);
        if (EXTRA) tma_load_1d(sD[buf], recD + src, bytes, &full[buf]);
    };
    if (tid == 0) {
        if (rounds > 0) issue(0);
        if (rounds > 1) issue(1);
    }

    float T = 1.0f;
    float C[NCH];
#pragma unroll
    for (int k = 0; k < NCH; ++k) C[k] = 0.0f;
    float D = 0.0f, Wsum = 0.0f;
    uint32_t last_contributor = 0;
    bool done = !inside;

    const float tile_x0 = (float)(tile_xi * TILE_X), tile_y0 = (float)(tile_yi * TILE_Y);
    for (int r = 0; r < rounds; ++r) {
        const int buf = r & 1;
        mbar_wait(&full[buf], (uint32_t)((r >> 1) & 1));
        const int nb = min(BATCH, n_total - r * BATCH);
        const uint32_t contrib_base = (uint32_t)r * BATCH;
        // ---- pre-pass: thread j decides which of the 8 warp patches Gaussian j can reach; the ballots turn that
        //      into one 128-bit "work list" per warp, so a warp never touches a Gaussian that misses its patch ----
        if (tid < BATCH) {
            uint32_t m8 = 0u;
            if (tid < nb) {
                const float4 a = sA[buf][tid];
                const float4 b = sB[buf][tid];
                m8 = blend_patch_mask(a.x, a.y, a.z, a.w, b.x, b.w, tile_x0, tile_y0);
            }
#pragma unroll
            for (int w = 0; w < TILE_PIX / 32; ++w) {
                const uint32_t bal = __ballot_sync(0xffffffffu, (m8 >> w) & 1u);
                if (lane == 0) s_wmask[w][warp] = bal;
            }
        }
        __syncthreads();
#pragma unroll 1
        for (int q = 0; q < BATCH / 32; ++q) {
            uint32_t m = s_wmask[warp][q];
            while (m != 0u && !done) {
                const int j = q * 32 + __ffs(m) - 1;
                m &= m - 1u;
                const float4 a = sA[buf][j];
                const float4 b = sB[buf][j];
                const float dx = __fsub_rn(a.x, pxf), dy = __fsub_rn(a.y, pyf);
                const float power = blend_power(dx, dy, a.z, a.w, b.x);
                if (power > 0.0f || power < b.w) continue;      // b.w: conservative skip threshold (no exp needed)
                const float alpha = fminf(0.99f, __fmul_rn(b.y, expf(power)));
                if (alpha < 1.0f / 255.0f) continue;
                const float test_T = __fmul_rn(T, __fsub_rn(1.0f, alpha));
                if (test_T < 0.0001f) {
                    done = true;
                    continue;
                }
                const float4 c = sC[buf][j];
                C[0] = __fmaf_rn(T, __fmul_rn(alpha, c.x), C[0]);
                C[1] = __fmaf_rn(T, __fmul_rn(alpha, c.y), C[1]);
                C[2] = __fmaf_rn(T, __fmul_rn(alpha, c.z), C[2]);
                if (EXTRA) {
                    const float4 e = sD[buf][j];
                    C[NCH - 3] = __fmaf_rn(T, __fmul_rn(alpha, e.x), C[NCH - 3]);
                    C[NCH - 2] = __fmaf_rn(T, __fmul_rn(alpha, e.y), C[NCH - 2]);
                    C[NCH - 1] = __fmaf_rn(T, __fmul_rn(alpha, e.z), C[NCH - 1]);
                }
                D = __fmaf_rn(T, __fmul_rn(alpha, b.z), D);
                Wsum = __fmaf_rn(alpha, T, Wsum);
                T = test_T;
                last_contributor = contrib_base + (uint32_t)j + 1u;
            }
        }
        // everyone is finished with `buf`; stop fetching once the whole tile is saturated
        const int num_done = __syncthreads_count(done);
        const bool cont = num_done < TILE_PIX;
        if (!cont) {
            // batch r+1 may still be in flight: the CTA must not retire under it
            if (tid == 0 && r + 1 < rounds) mbar_wait(&full[(r + 1) & 1], (uint32_t)(((r + 1) >> 1) & 1));
            break;
        }
        if (tid == 0 && r + 2 < rounds) issue(r + 2);
    }

    if (inside) {
        const CamDev& cam = cams[view];
        const size_t HW = (size_t)H * W;
        const size_t pix = (size_t)py * W + px;
        n_contrib[(size_t)view * HW + pix] = last_contributor;
        final_T[(size_t)view * HW + pix] = T;
#pragma unroll
        for (int ch = 0; ch < 3; ++ch)
            out_color[((size_t)view * 3 + ch) * HW + pix] = __fmaf_rn(T, cam.bg[ch], C[ch]);
        if (EXTRA) {
#pragma unroll
            for (int ch = 0; ch < 3; ++ch)
                out_extra[((size_t)view * 3 + ch) * HW + pix] = __fmaf_rn(T, cam.bg[3 + ch], C[NCH - 3 + ch]);
        }
        out_depth[(size_t)view * HW + pix] = D;
        out_alpha[(size_t)view * HW + pix] = Wsum;
    }
}

}  // namespace

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
    if (channels == 3) {
        blend_forward_kernel<3><<<blocks, TILE_PIX, 0, s>>>(ranges, at<uint32_t>(iws, L.tile_order), A, B, C, D, cams, W, H, L.tiles_x, T, out_color,
                                                           out_extra, out_depth, out_alpha, nc, fT);
    } else if (channels == 6) {
        blend_forward_kernel<6><<<blocks, TILE_PIX, 0, s>>>(ranges, at<uint32_t>(iws, L.tile_order), A, B, C, D, cams, W, H, L.tiles_x, T, out_color,
                                                           out_extra, out_depth, out_alpha, nc, fT);
    } else {
        return GSR_ERR_UNSUPPORTED;
    }
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
