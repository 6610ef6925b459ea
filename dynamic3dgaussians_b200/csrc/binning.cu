// (b) binning: instance counting, instance emission in depth order, stable tile sort, tile ranges, record gather.
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + SortPairs + identifyTileRanges
// (R/cuda_rasterizer/rasterizer_impl.cu:70-138,283-323).
//
// The reference sorts R instances on a 64-bit (tile | depth) key.  Here Gaussians are already in per-view depth order
// (sort of P entries per view), so instances are emitted in that order with a 32-bit (view*T + tile) key and a stable
// sort on the tile bits alone produces exactly the reference's order: (tile, depth bits, Gaussian index).
#include "geom_math.cuh"
#include "radix_rank.cuh"

namespace gsr {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int EMIT_HIST_PASSES = 2;      // tile sorts of up to 16 key bits get their histograms from the emit kernel

__device__ __forceinline__ uint32_t block_exclusive_scan_256(uint32_t v, uint32_t* s_warp /*[8]*/, uint32_t& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t base = 0, tot = 0;
#pragma unroll
    for (int w = 0; w < SCAN_BLOCK / 32; ++w) {
        const uint32_t c = s_warp[w];
        if (w < warp) base += c;
        tot += c;
    }
    total = tot;
    __syncthreads();
    return base + inc - v;
}

// slot i of the depth-sorted list: view = i / P, Gaussian = sorted_g[i]
__device__ __forceinline__ uint32_t slot_tiles(uint32_t i, uint32_t VP, int P, const uint32_t* __restrict__ sorted_g,
                                               const uint32_t* __restrict__ tiles_touched, uint32_t& gidx) {
    if (i >= VP) { gidx = 0; return 0; }
    const uint32_t view = i / (uint32_t)P;
    gidx = view * (uint32_t)P + sorted_g[i];
    return tiles_touched[gidx];
}

__global__ void __launch_bounds__(SCAN_BLOCK)
tile_count_reduce_kernel(uint32_t VP, int P, const uint32_t* __restrict__ sorted_g,
                         const uint32_t* __restrict__ tiles_touched, uint32_t* __restrict__ block_sums) {
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t sum = 0, gidx;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) sum += slot_tiles(base + k, VP, P, sorted_g, tiles_touched, gidx);
    uint32_t total;
    block_exclusive_scan_256(sum, s_warp, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

// single block: exclusive scan of the per-block sums in place; total -> counters[CNT_RENDERED]
__global__ void __launch_bounds__(SCAN_BLOCK)
block_sums_scan_kernel(uint32_t nblocks, uint32_t* __restrict__ block_sums, uint32_t* __restrict__ counters) {
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    uint32_t carry = 0;
    for (uint32_t start = 0; start < nblocks; start += SCAN_BLOCK) {
        const uint32_t i = start + threadIdx.x;
        const uint32_t v = i < nblocks ? block_sums[i] : 0u;
        uint32_t total;
        const uint32_t ex = block_exclusive_scan_256(v, s_warp, total);
        if (i < nblocks) block_sums[i] = carry + ex;
        carry += total;
    }
    if (threadIdx.x == 0) {
        counters[CNT_RENDERED] = carry;
        counters[CNT_OVERFLOW] = 0;
    }
}

// Emits the (view*T + tile, Gaussian) pairs of every visible (view, Gaussian) in depth order.
// Row-major tile order inside a Gaussian (y outer, x inner) as duplicateWithKeys, rasterizer_impl.cu:98-109.
// The reference (and the first version of this kernel) lets one thread write all instances of its Gaussian, so the
// lanes of a warp write to addresses `tiles_touched` apart: every store is its own 32-byte sector.  Here a CTA first
// scans its 2048 slots into shared memory, then walks its OUTPUT range with one thread per instance (binary search
// of the owning slot), so stores are fully coalesced.
__global__ void __launch_bounds__(SCAN_BLOCK)
emit_instances_kernel(uint32_t VP, int P, int tiles_x, int tiles_y, const uint32_t* __restrict__ sorted_g,
                      const uint32_t* __restrict__ tiles_touched, const uint32_t* __restrict__ block_sums,
                      const float2* __restrict__ means2D, const int32_t* __restrict__ radii, uint32_t capacity,
                      uint32_t* __restrict__ ikeys, uint32_t* __restrict__ ivals, uint32_t* __restrict__ counters,
                      int hist_passes, int tile_bits, uint32_t* __restrict__ sort_hist) {
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    __shared__ uint32_t s_hist[SCAN_BLOCK / 32][EMIT_HIST_PASSES][RADIX];   // warp-private digit histograms of the keys
    __shared__ uint32_t s_off[SCAN_TILE + 1];     // exclusive offset of every slot inside this CTA's output range
    __shared__ uint32_t s_key0[SCAN_TILE];        // key of the slot's first tile: view*T + min.y*tiles_x + min.x
    __shared__ uint32_t s_g[SCAN_TILE];           // Gaussian index
    __shared__ uint16_t s_w[SCAN_TILE];           // rect width in tiles
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t cnt[SCAN_ITEMS], gidx[SCAN_ITEMS];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        cnt[k] = slot_tiles(base + k, VP, P, sorted_g, tiles_touched, gidx[k]);
        sum += cnt[k];
    }
    uint32_t total;
    uint32_t off = block_exclusive_scan_256(sum, s_warp, total);
    const uint32_t T = (uint32_t)tiles_x * (uint32_t)tiles_y;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const int slot = threadIdx.x * SCAN_ITEMS + k;
        s_off[slot] = off;
        off += cnt[k];
        if (cnt[k] != 0) {
            const uint32_t gi = gidx[k];
            const uint32_t view = gi / (uint32_t)P;
            uint2 mn, mx;
            tile_rect(means2D[gi], radii[gi], tiles_x, tiles_y, mn, mx);
            s_key0[slot] = view * T + mn.y * (uint32_t)tiles_x + mn.x;
            s_g[slot] = gi - view * (uint32_t)P;
            s_w[slot] = (uint16_t)(mx.x - mn.x);
        }
    }
    if (threadIdx.x == 0) s_off[SCAN_TILE] = total;
    if (hist_passes > 0)
        for (int i = threadIdx.x; i < (SCAN_BLOCK / 32) * EMIT_HIST_PASSES * RADIX; i += SCAN_BLOCK) (&s_hist[0][0][0])[i] = 0;
    __syncthreads();
    const uint32_t out0 = block_sums[blockIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    bool overflow = false;
    for (uint32_t o0 = 0; o0 < total; o0 += SCAN_BLOCK) {     // warp-uniform trip count (the histogram uses ballots)
        const uint32_t o = o0 + threadIdx.x;
        bool stored = false;
        uint32_t key = 0;
        if (o < total) {
            // last slot with s_off <= o (empty slots share their successor's offset and are skipped by construction)
            uint32_t lo = 0, hi = SCAN_TILE;
#pragma unroll
            for (int step = 0; step < 11; ++step) {           // SCAN_TILE == 2^11
                const uint32_t mid = (lo + hi) >> 1;
                if (s_off[mid] <= o) lo = mid; else hi = mid;
            }
            const uint32_t l = o - s_off[lo];
            const uint32_t w = s_w[lo];
            const uint32_t ty = l / w;
            key = s_key0[lo] + ty * (uint32_t)tiles_x + (l - ty * w);
            const uint32_t dst = out0 + o;
            if (dst < capacity) {
                ikeys[dst] = key;
                ivals[dst] = s_g[lo];
                stored = true;
            } else {
                overflow = true;
            }
        }
        // the tile sort's digit histograms, taken while the keys are in registers (saves a pass over the keys)
#pragma unroll
        for (int p = 0; p < EMIT_HIST_PASSES; ++p) {
            if (p < hist_passes) {
                const int shift = p * RADIX_BITS;
                const int bits = min(RADIX_BITS, tile_bits - shift);
                uint32_t peers;
                warp_hist_add(s_hist[warp][p], (key >> shift) & ((1u << bits) - 1u), stored, lane, peers);
            }
        }
    }
    if (overflow) counters[CNT_OVERFLOW] = 1;
    if (hist_passes > 0) {
        __syncthreads();
        for (int i = threadIdx.x; i < hist_passes * RADIX; i += SCAN_BLOCK) {
            const int p = i / RADIX, d = i - p * RADIX;
            uint32_t c = 0;
#pragma unroll
            for (int w = 0; w < SCAN_BLOCK / 32; ++w) c += s_hist[w][p][d];
            if (c) atomicAdd(&sort_hist[i], c);
        }
    }
}

// identifyTileRanges, rasterizer_impl.cu:116-138, on the 32-bit (view*T + tile) keys
__global__ void __launch_bounds__(256)
tile_ranges_kernel(const uint32_t* __restrict__ counters, uint32_t capacity, const uint32_t* __restrict__ keys,
                   uint2* __restrict__ ranges) {
    const uint32_t n = min(counters[CNT_RENDERED], capacity);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const uint32_t cur = keys[i];
    if (i == 0) {
        ranges[cur].x = 0;
    } else {
        const uint32_t prev = keys[i - 1];
        if (cur != prev) {
            ranges[prev].y = i;
            ranges[cur].x = i;
        }
    }
    if (i == n - 1) ranges[cur].y = n;
}

// Gathers the per-instance records the blend kernels stream with TMA bulk copies:
//   A = (mean.x, mean.y, conic.x, conic.y)   B = (conic.z, opacity, depth, skip threshold [blend_skip_threshold])
//   C = (r, g, b, bits(view*P + g))          D = (extra colour r, g, b, 0)   [channels == 6]
// Two threads per instance: the per-(view, Gaussian) geometry is already stored as the 32-byte pair (recA | recB) by
// the preprocess kernel, so the even thread moves the first 16 bytes and the odd thread the second 16 bytes of ONE
// 32-byte sector; colours come from float4-packed per-Gaussian arrays.  A warp instruction therefore touches 16 sectors
// for 16 instances.  (The first version read five separate arrays with 4/8/12/16-byte gathers: six divergent load
// instructions = ~190 L1 tag cycles per 32 instances, which -- not DRAM -- bounded the kernel at 0.94 ms.)
// key / T is a multiply-high by a host-computed reciprocal (exact for key < 2^31).
constexpr int GATHER_ITEMS = 4;

__global__ void __launch_bounds__(256)
gather_records_kernel(const uint32_t* __restrict__ counters, uint32_t capacity, int P, uint32_t T, uint32_t T_magic,
                      int T_shift, int channels,
                      const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                      const float4* __restrict__ geomrec, const float4* __restrict__ col4,
                      const float4* __restrict__ rgb_sh4, const float4* __restrict__ ext4,
                      float4* __restrict__ recA, float4* __restrict__ recB, float4* __restrict__ recC,
                      float4* __restrict__ recD) {
    const uint32_t n = min(counters[CNT_RENDERED], capacity);
    const uint32_t half = threadIdx.x & 1u;
    const uint32_t i0 = blockIdx.x * (128 * GATHER_ITEMS) + (threadIdx.x >> 1);
    if (i0 >= n) return;
    uint32_t g[GATHER_ITEMS], gi[GATHER_ITEMS];
    bool ok[GATHER_ITEMS];
#pragma unroll
    for (int k = 0; k < GATHER_ITEMS; ++k) {
        const uint32_t i = i0 + k * 128;
        ok[k] = i < n;
        const uint32_t key = ok[k] ? keys[i] : 0u;
        g[k] = ok[k] ? vals[i] : 0u;
        const uint32_t view = T == 1u ? key : (__umulhi(key, T_magic) >> T_shift);
        gi[k] = view * (uint32_t)P + g[k];
    }
    float4 geo[GATHER_ITEMS], col[GATHER_ITEMS];
    const bool want_col = half == 0u || channels > 3;
#pragma unroll
    for (int k = 0; k < GATHER_ITEMS; ++k) {
        geo[k] = geomrec[2 * (size_t)gi[k] + half];
        if (want_col)
            col[k] = half == 0u ? (col4 ? col4[g[k]] : rgb_sh4[gi[k]]) : ext4[g[k]];
    }
    float4* __restrict__ dst_geo = half == 0u ? recA : recB;
    float4* __restrict__ dst_col = half == 0u ? recC : recD;
#pragma unroll
    for (int k = 0; k < GATHER_ITEMS; ++k) {
        if (!ok[k]) continue;
        const uint32_t i = i0 + k * 128;
        dst_geo[i] = geo[k];
        if (want_col) {
            if (half == 0u) col[k].w = __uint_as_float(gi[k]);
            dst_col[i] = col[k];
        }
    }
}

// float4 packing of the per-Gaussian colours handed to gsr_forward_render ([P][3] AoS in the reference's layout)
__global__ void __launch_bounds__(256)
pack_colors_kernel(int P, const float* __restrict__ colors, const float* __restrict__ extra,
                   float4* __restrict__ col4, float4* __restrict__ ext4) {
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P) return;
    if (colors) col4[g] = make_float4(colors[3 * g], colors[3 * g + 1], colors[3 * g + 2], 0.0f);
    if (extra) ext4[g] = make_float4(extra[3 * g], extra[3 * g + 1], extra[3 * g + 2], 0.0f);
}

// ---- export of one view in the reference's format (tests / tools) ----
__global__ void find_view_start_kernel(const uint32_t* __restrict__ counters, uint32_t capacity,
                                       const uint32_t* __restrict__ keys, uint32_t first_key,
                                       uint32_t* __restrict__ out) {
    const uint32_t n = min(counters[CNT_RENDERED], capacity);
    uint32_t lo = 0, hi = n;
    while (lo < hi) {
        const uint32_t mid = lo + (hi - lo) / 2;
        if (keys[mid] < first_key) lo = mid + 1; else hi = mid;
    }
    out[0] = lo;
}

__global__ void export_view_kernel(const uint32_t* __restrict__ counters, uint32_t capacity, int view, int P,
                                   uint32_t T, const uint32_t* __restrict__ keys, const uint32_t* __restrict__ vals,
                                   const float* __restrict__ depths, const uint32_t* __restrict__ view_start,
                                   uint64_t* __restrict__ keys_out, uint32_t* __restrict__ list_out) {
    const uint32_t n = min(counters[CNT_RENDERED], capacity);
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x + view_start[0];
    if (i >= n) return;
    const uint32_t k = keys[i];
    if (k / T != (uint32_t)view) return;
    const uint32_t local = i - view_start[0];
    const uint32_t g = vals[i];
    if (keys_out)
        keys_out[local] = ((uint64_t)(k - (uint32_t)view * T) << 32) |
                          (uint64_t)__float_as_uint(depths[(size_t)view * P + g]);
    if (list_out) list_out[local] = g;
}

__global__ void export_ranges_kernel(int view, uint32_t T, const uint2* __restrict__ ranges,
                                     const uint32_t* __restrict__ view_start, uint32_t* __restrict__ ranges_out) {
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= T) return;
    const uint2 r = ranges[(size_t)view * T + t];
    const bool empty = r.y == r.x;
    ranges_out[2 * t] = empty ? 0u : r.x - view_start[0];
    ranges_out[2 * t + 1] = empty ? 0u : r.y - view_start[0];
}

}  // namespace

int launch_tile_scan(int V, int P, void* ws, const Layout& L, cudaStream_t s) {
    const uint32_t VP = (uint32_t)((size_t)V * P);
    if (VP == 0) return GSR_OK;
    const uint32_t nblocks = (VP + SCAN_TILE - 1) / SCAN_TILE;
    const uint32_t* sorted_g = at<uint32_t>(ws, L.sorted_vals_in ? L.dvals1 : L.dvals0);
    tile_count_reduce_kernel<<<nblocks, SCAN_BLOCK, 0, s>>>(VP, P, sorted_g, at<uint32_t>(ws, L.tiles_touched),
                                                           at<uint32_t>(ws, L.block_sums));
    GSR_LAUNCH_CHECK();
    block_sums_scan_kernel<<<1, SCAN_BLOCK, 0, s>>>(nblocks, at<uint32_t>(ws, L.block_sums),
                                                    at<uint32_t>(ws, L.counters));
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_binning(int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                   const float* colors_precomp, const float* extra_colors, void* gws, void* bws, void* iws,
                   const Layout& L, cudaStream_t s) {
    const uint32_t VP = (uint32_t)((size_t)V * P);
    const uint32_t T = (uint32_t)L.tiles_x * (uint32_t)L.tiles_y;
    const size_t VT = (size_t)V * T;
    GSR_CUDA_TRY(cudaMemsetAsync(at<uint2>(iws, L.ranges), 0, VT * sizeof(uint2), s));
    if (VP == 0 || capacity <= 0) return GSR_OK;
    const uint32_t cap = (uint32_t)capacity;
    uint32_t* counters = at<uint32_t>(gws, L.counters);
    const uint32_t nblocks = (VP + SCAN_TILE - 1) / SCAN_TILE;
    const uint32_t* sorted_g = at<uint32_t>(gws, L.sorted_vals_in ? L.dvals1 : L.dvals0);
    const bool fused_hist = L.tile_passes <= EMIT_HIST_PASSES;
    stage_begin(ST_EMIT, s);
    if (fused_hist) GSR_CUDA_TRY(cudaMemsetAsync(at<uint32_t>(bws, L.ihist), 0, (size_t)L.tile_passes * RADIX * 4, s));
    emit_instances_kernel<<<nblocks, SCAN_BLOCK, 0, s>>>(
        VP, P, L.tiles_x, L.tiles_y, sorted_g, at<uint32_t>(gws, L.tiles_touched), at<uint32_t>(gws, L.block_sums),
        at<float2>(gws, L.means2D), at<int32_t>(gws, L.radii), cap, at<uint32_t>(bws, L.ikeys0),
        at<uint32_t>(bws, L.ivals0), counters, fused_hist ? L.tile_passes : 0, L.tile_bits,
        at<uint32_t>(bws, L.ihist));
    GSR_LAUNCH_CHECK();
    stage_end(ST_EMIT, s);
    stage_begin(ST_TILE_SORT, s);
    int rc = radix_sort_pairs<uint32_t>(at<uint32_t>(bws, L.ikeys0), at<uint32_t>(bws, L.ikeys1),
                                        at<uint32_t>(bws, L.ivals0), at<uint32_t>(bws, L.ivals1), capacity,
                                        counters + CNT_RENDERED, 0, L.tile_bits, at<uint32_t>(bws, L.ihist),
                                        at<uint32_t>(bws, L.ilookback), counters + CNT_INST_TILE, s, 1, fused_hist);
    stage_end(ST_TILE_SORT, s);
    if (rc != GSR_OK) return rc;
    const uint32_t* skeys = at<uint32_t>(bws, L.sorted_keys_in ? L.ikeys1 : L.ikeys0);
    const uint32_t* svals = at<uint32_t>(bws, L.sorted_keys_in ? L.ivals1 : L.ivals0);
    const uint32_t gblocks = (cap + 255) / 256;
    stage_begin(ST_RANGES, s);
    tile_ranges_kernel<<<gblocks, 256, 0, s>>>(counters, cap, skeys, at<uint2>(iws, L.ranges));
    GSR_LAUNCH_CHECK();
    stage_end(ST_RANGES, s);
    StageScope sc(ST_GATHER, s);
    // floor(key / T) == umulhi(key, magic) >> shift for every key < 2^31 (V*T < 2^31 is checked at the ABI)
    int lg = 1;                                   // T == 1 is special-cased in the kernel (magic would need 33 bits)
    while ((1u << lg) < T) ++lg;
    const int sh = 31 + lg;
    const uint32_t magic = (uint32_t)((((uint64_t)1 << sh) + T - 1) / T);
    const uint32_t rblocks = (cap + 128 * GATHER_ITEMS - 1) / (128 * GATHER_ITEMS);
    if (colors_precomp || extra_colors) {
        pack_colors_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, colors_precomp, extra_colors, at<float4>(gws, L.col4),
                                                           at<float4>(gws, L.ext4));
        GSR_LAUNCH_CHECK();
    }
    gather_records_kernel<<<rblocks, 256, 0, s>>>(
        counters, cap, P, T, magic, sh - 32, channels, skeys, svals, at<float4>(gws, L.geomrec),
        colors_precomp ? at<float4>(gws, L.col4) : nullptr, sh_coeffs > 0 ? at<float4>(gws, L.rgb) : nullptr,
        at<float4>(gws, L.ext4), at<float4>(bws, L.recA), at<float4>(bws, L.recB), at<float4>(bws, L.recC),
        channels > 3 ? at<float4>(bws, L.recD) : nullptr);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

int launch_export_view(int view, int V, int P, int W, int H, const void* gws, const void* bws, const void* iws,
                       const Layout& L, uint64_t* keys_out, uint32_t* list_out, uint32_t* ranges_out,
                       uint32_t* n_contrib_out, cudaStream_t s) {
    const uint32_t T = (uint32_t)L.tiles_x * (uint32_t)L.tiles_y;
    // scratch slot for the view's first sorted position: counters[2]
    uint32_t* counters = const_cast<uint32_t*>(at<uint32_t>(gws, L.counters));
    const uint32_t* skeys = at<uint32_t>(bws, L.sorted_keys_in ? L.ikeys1 : L.ikeys0);
    const uint32_t* svals = at<uint32_t>(bws, L.sorted_keys_in ? L.ivals1 : L.ivals0);
    // capacity is not needed beyond clamping; counters[CNT_RENDERED] <= capacity is enforced by the caller's check
    const uint32_t cap = 0xffffffffu;
    find_view_start_kernel<<<1, 1, 0, s>>>(counters, cap, skeys, (uint32_t)view * T, counters + 2);
    GSR_LAUNCH_CHECK();
    if (keys_out || list_out) {
        // upper bound on this view's instances: everything after its start; threads past the view return early
        uint32_t host_n = 0;
        GSR_CUDA_TRY(cudaMemcpyAsync(&host_n, counters + CNT_RENDERED, 4, cudaMemcpyDeviceToHost, s));
        GSR_CUDA_TRY(cudaStreamSynchronize(s));
        if (host_n > 0) {
            export_view_kernel<<<(host_n + 255) / 256, 256, 0, s>>>(counters, cap, view, P, T, skeys, svals,
                                                                    at<float>(gws, L.depths), counters + 2, keys_out,
                                                                    list_out);
            GSR_LAUNCH_CHECK();
        }
    }
    if (ranges_out) {
        export_ranges_kernel<<<(T + 255) / 256, 256, 0, s>>>(view, T, at<uint2>(iws, L.ranges), counters + 2,
                                                             ranges_out);
        GSR_LAUNCH_CHECK();
    }
    if (n_contrib_out) {
        GSR_CUDA_TRY(cudaMemcpyAsync(n_contrib_out, at<uint32_t>(iws, L.n_contrib) + (size_t)view * W * H,
                                     (size_t)W * H * 4, cudaMemcpyDeviceToDevice, s));
    }
    return GSR_OK;
}

}  // namespace gsr
