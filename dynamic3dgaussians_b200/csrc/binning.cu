// (b) binning: from depth-ordered Gaussians to per-tile, depth-ordered record lists.
// Replaces cub::DeviceScan::InclusiveSum + duplicateWithKeys + SortPairs + identifyTileRanges
// (R/cuda_rasterizer/rasterizer_impl.cu:70-138,283-323); the result -- for every tile the list of Gaussians whose
// rectangle covers it, ordered by (depth bits, Gaussian index) -- is identical to the reference's sorted list.
//
// The reference sorts R = sum(tiles touched) instances on a 64-bit (tile | depth) key: 6 radix passes over 12-byte pairs.
// Here the ordering is built hierarchically, and the only radix-sorted instance set is 3-4x smaller than R:
//   1. Gaussians are already in per-view depth order (segmented 32-bit sort, abi.cu).
//   2. In that order every Gaussian emits one instance per SUPERTILE (4x4 tiles) it touches, key = view*S + supertile;
//      a stable onesweep sort on those ~11 key bits (2 passes, histograms taken while emitting) groups them per
//      supertile, still in depth order.
//   3. Each supertile list is cut into chunks of 256 instances.  A counting kernel finds, per chunk, how many
//      instances hit each of the 16 tiles (warp ballots on a 16-bit hit mask); a per-supertile prefix over its chunks
//      and one scan over all tiles give every tile's range and every (chunk, tile) output offset.
//   4. A scatter kernel recomputes the hit masks and writes the instances' 48-byte blend records (recA|recB|recC[|recD])
//      straight into the tile lists -- order-preserving compaction, so depth order survives -- with the per-Gaussian
//      data fetched once per supertile instance instead of once per tile instance, and coalesced stores.
// No R-sized key/value arrays exist any more; `ranges` fall out of the scan (no identifyTileRanges pass).
#include "geom_math.cuh"
#include "radix_rank.cuh"

namespace gsr {

namespace {

constexpr unsigned FULL = 0xffffffffu;

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

// supertile rectangle of a tile rectangle (max exclusive); empty rect -> zero supertiles
__device__ __forceinline__ uint32_t st_rect(ushort4 r, uint32_t& sx0, uint32_t& sy0, uint32_t& sw) {
    if (r.z <= r.x || r.w <= r.y) { sx0 = sy0 = sw = 0; return 0u; }
    sx0 = r.x / ST;
    sy0 = r.y / ST;
    sw = (r.z - 1u) / ST - sx0 + 1u;
    return sw * ((r.w - 1u) / ST - sy0 + 1u);
}

// slot i of the depth-sorted list: view = i / P, Gaussian = sorted_g[i]; a view's list holds its vis_count[view] visible
// Gaussians (the rest of its P slots is unused)
__device__ __forceinline__ ushort4 slot_rect(uint32_t i, uint32_t VP, int P, const uint32_t* __restrict__ sorted_g,
                                             const ushort4* __restrict__ rects, const uint32_t* __restrict__ vis_count,
                                             uint32_t& gidx) {
    gidx = 0;
    if (i >= VP) return make_ushort4(0, 0, 0, 0);
    const uint32_t view = i / (uint32_t)P;
    if (i - view * (uint32_t)P >= vis_count[view]) return make_ushort4(0, 0, 0, 0);
    gidx = view * (uint32_t)P + sorted_g[i];
    return rects[gidx];
}

// 16-bit hit mask of a tile rectangle inside the supertile whose first tile is (ox, oy): bit ty*4+tx
__device__ __forceinline__ uint32_t st_hit_mask(ushort4 r, uint32_t ox, uint32_t oy) {
    const int xlo = max((int)r.x - (int)ox, 0), xhi = min((int)r.z - (int)ox, ST);
    const int ylo = max((int)r.y - (int)oy, 0), yhi = min((int)r.w - (int)oy, ST);
    if (xhi <= xlo || yhi <= ylo) return 0u;
    const uint32_t xb = ((1u << xhi) - 1u) & ~((1u << xlo) - 1u);
    const uint32_t yb = ((1u << yhi) - 1u) & ~((1u << ylo) - 1u);
    const uint32_t ys = (yb & 1u) | ((yb & 2u) << 3) | ((yb & 4u) << 6) | ((yb & 8u) << 9);
    return xb * ys;
}

__device__ __forceinline__ uint32_t ld_vol(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }
__device__ __forceinline__ void st_vol(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }

// decoupled look-back word of the emit kernel: 2 flag bits + 30-bit count
constexpr uint32_t LB_AGG = 1u << 30, LB_INC = 1u << 31, LB_VALUE = (1u << 30) - 1u;
constexpr int ST_KEY_BITS = 16;          // the instance key's low half is view*S + supertile, its high half the tile hit mask

// Emits the (key, view*P + Gaussian) pairs of every visible (view, Gaussian) in depth order, ONE pass over the
// depth-sorted slots (no separate counting pass: the CTAs take their slot block from a ticket and get their output offset
// by decoupled look-back over the blocks before them).
//   key = hit mask << 16 | (view*S + supertile): the 16 tiles of the supertile the Gaussian's rectangle covers ride in the
//   upper half of the sort key (the sort only looks at the low `key_bits`), so no later pass has to fetch the rectangle
//   again (the round-1 chunk_count kernel gathered 32-byte sectors for 8 useful bytes per instance: 117 us at C2).
// Every thread owns 8 consecutive slots of the depth-sorted list; each thread writes its own instances (walking its
// supertile rectangles, no search) into a shared-memory window at their output positions, and the window is copied out
// with fully coalesced stores.  On the way out the CTA counts: the digit histograms of the following sort (warp-private
// bins, radix_rank.cuh) and the instances per supertile key (shared-memory bins over the one or two views a CTA's slots
// belong to) -- the latter give st_ranges / the chunk table without a pass over the sorted keys.
#ifndef GSR_EMIT_LOOKBACK
#define GSR_EMIT_LOOKBACK 0              // 1: the emit CTAs find their output offset by decoupled look-back (one pass over the
#endif                                   // slots; measured SLOWER at C2: 311 us against 147 + 58 us for count + emit)
constexpr int EMIT_ITEMS = 4;            // depth-sorted slots per thread
constexpr int EMIT_TILE = SCAN_BLOCK * EMIT_ITEMS;
static_assert(EMIT_TILE == 1024, "compute_layout sizes the look-back words for blocks of 1024 slots");
constexpr int EMIT_WINDOW = 2048;        // staged outputs per round (a CTA of 1024 slots emits ~1450 at C2)

// per block of EMIT_TILE slots: number of supertile instances (for the emit offsets)
__global__ void __launch_bounds__(SCAN_BLOCK)
st_count_reduce_kernel(uint32_t VP, int P, const uint32_t* __restrict__ sorted_g, const ushort4* __restrict__ rects,
                       const uint32_t* __restrict__ vis_count, uint32_t* __restrict__ block_sums) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    const uint32_t base = blockIdx.x * EMIT_TILE + threadIdx.x * EMIT_ITEMS;
    uint32_t sum = 0, gidx;
#pragma unroll
    for (int k = 0; k < EMIT_ITEMS; ++k) {
        const ushort4 r = slot_rect(base + k, VP, P, sorted_g, rects, vis_count, gidx);
        uint32_t a, b, c;
        sum += st_rect(r, a, b, c);
    }
    uint32_t total;
    block_exclusive_scan_256(sum, s_warp, total);
    if (threadIdx.x == 0) block_sums[blockIdx.x] = total;
}

__device__ __forceinline__ uint32_t block_exclusive_scan_1024(uint32_t v, uint32_t* s_warp /*[32]*/, uint32_t& total) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_warp[warp] = inc;
    __syncthreads();
    uint32_t w = s_warp[lane], winc = w;                 // every warp scans the 32 warp totals
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, winc, o);
        if (lane >= o) winc += t;
    }
    total = __shfl_sync(FULL, winc, 31);
    const uint32_t base = __shfl_sync(FULL, winc - w, warp);
    __syncthreads();
    return base + inc - v;
}

// single block of 1024 threads, 4 sums per thread and step: exclusive scan of the per-block sums in place;
// total -> counters[CNT_ST_RENDERED]
__global__ void __launch_bounds__(1024)
block_sums_scan_kernel(uint32_t nblocks, uint32_t* __restrict__ block_sums, uint32_t* __restrict__ counters) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ uint32_t s_warp[32];
    uint32_t carry = 0;
    for (uint32_t start = 0; start < nblocks; start += 4096) {
        const uint32_t i0 = start + threadIdx.x * 4;
        uint32_t v[4], sum = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) { v[k] = i0 + k < nblocks ? block_sums[i0 + k] : 0u; sum += v[k]; }
        uint32_t total;
        uint32_t ex = carry + block_exclusive_scan_1024(sum, s_warp, total);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (i0 + k < nblocks) block_sums[i0 + k] = ex;
            ex += v[k];
        }
        carry += total;
    }
    if (threadIdx.x == 0) counters[CNT_ST_RENDERED] = carry;
}

__global__ void __launch_bounds__(SCAN_BLOCK, 4)
emit_st_instances_kernel(uint32_t VP, int P, int stiles_x, uint32_t S, int view_bits,
                         const uint32_t* __restrict__ sorted_g,
                         const ushort4* __restrict__ rects, const uint32_t* __restrict__ vis_count, uint32_t capacity,
                         uint32_t nblocks,
                         uint32_t* __restrict__ ikeys, uint32_t* __restrict__ ivals, uint32_t* __restrict__ counters,
                         uint32_t* __restrict__ lookback, uint32_t* __restrict__ ticket,
                         uint32_t* __restrict__ key_hist, uint32_t khist_views /* 0: global atomics */) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    extern __shared__ uint32_t s_khist[];                 // [S][khist_views]: instances per key of the block's views
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    __shared__ uint32_t s_key[EMIT_WINDOW];              // key | supertile x << 16 | supertile y << 24
    __shared__ uint16_t s_slot[EMIT_WINDOW];             // which of the block's slots the instance belongs to
    __shared__ uint32_t s_gidx[EMIT_TILE];               // per slot: view*P + Gaussian and the tile rectangle
    __shared__ ushort4 s_rect[EMIT_TILE];
    __shared__ uint32_t s_blk, s_out0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    (void)lane; (void)warp;
#if GSR_EMIT_LOOKBACK
    if (threadIdx.x == 0) s_blk = atomicAdd(ticket, 1u);  // blocks are numbered in the order they start: a block only
    __syncthreads();                                      // waits for blocks that already run
    const uint32_t blk = s_blk;
#else
    const uint32_t blk = blockIdx.x;
#endif
    const uint32_t slot0 = blk * EMIT_TILE;
    const uint32_t base = slot0 + threadIdx.x * EMIT_ITEMS;
    // per slot: the tile rectangle, view*P + Gaussian and the key of its first supertile; counts and widths are re-derived
    // from the rectangle where they are needed (registers decide how many CTAs hide the dependent loads)
    uint32_t gidx[EMIT_ITEMS], key0[EMIT_ITEMS];
    ushort4 rc[EMIT_ITEMS];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < EMIT_ITEMS; ++k) {
        rc[k] = slot_rect(base + k, VP, P, sorted_g, rects, vis_count, gidx[k]);
        uint32_t sx0, sy0, w;
        sum += st_rect(rc[k], sx0, sy0, w);
        key0[k] = ((sy0 * (uint32_t)stiles_x + sx0) << view_bits) | (gidx[k] / (uint32_t)P);
        s_gidx[threadIdx.x * EMIT_ITEMS + k] = gidx[k];
        s_rect[threadIdx.x * EMIT_ITEMS + k] = rc[k];
    }
    uint32_t total;
    const uint32_t off0 = block_exclusive_scan_256(sum, s_warp, total);
#if !GSR_EMIT_LOOKBACK
    if (threadIdx.x == 0) s_out0 = lookback[blk];         // exclusive scan of the per-block counts (block_sums_scan_kernel)
#else
    // ---- this block's first output position: decoupled look-back over the earlier blocks (warp 0, 32 words a step) ----
    if (warp == 0) {
        uint32_t excl = 0;
        if (blk == 0) {
            if (lane == 0) st_vol(&lookback[0], total | LB_INC);
        } else {
            if (lane == 0) st_vol(&lookback[blk], total | LB_AGG);
            int64_t t = (int64_t)blk - 1;
            while (true) {
                const int64_t idx = t - lane;
                uint32_t v = LB_INC;                                  // before block 0: inclusive prefix 0
                if (idx >= 0) {
                    while (((v = ld_vol(&lookback[idx])) >> 30) == 0u) __nanosleep(40);
                }
                const uint32_t inc = __ballot_sync(FULL, (v & LB_INC) != 0u);
                const int first = inc ? __ffs(inc) - 1 : 32;          // nearest block with an inclusive prefix
                uint32_t c = lane <= first ? (v & LB_VALUE) : 0u;
#pragma unroll
                for (int o = 16; o > 0; o >>= 1) c += __shfl_xor_sync(FULL, c, o);
                excl += c;
                if (inc) break;
                t -= 32;
            }
            if (lane == 0) st_vol(&lookback[blk], ((excl + total) & LB_VALUE) | LB_INC);
        }
        if (lane == 0) {
            s_out0 = excl;
            if (blk == nblocks - 1) counters[CNT_ST_RENDERED] = excl + total;
        }
    }
#endif
    // supertile keys this block can emit: those of the (khist_views) views its slots belong to
    const uint32_t view_first = slot0 / (uint32_t)P;
    const uint32_t vmask = (1u << view_bits) - 1u;
    const uint32_t khist_bins = S * khist_views;
    for (uint32_t i = threadIdx.x; i < khist_bins; i += SCAN_BLOCK) s_khist[i] = 0;
    __syncthreads();
    const uint32_t out0 = s_out0;
    bool overflow = false;
    for (uint32_t w0 = 0; w0 < total; w0 += EMIT_WINDOW) {
        if (w0) __syncthreads();                          // the previous window has been copied out
        if (off0 < w0 + EMIT_WINDOW && off0 + sum > w0) {     // some of this thread's instances fall into the window
            uint32_t o = off0;
#pragma unroll
            for (int k = 0; k < EMIT_ITEMS; ++k) {
                uint32_t sx0, sy0, wid;
                const uint32_t cnt = st_rect(rc[k], sx0, sy0, wid);
                uint32_t row = key0[k], x = 0, sy = sy0;
                for (uint32_t l = 0; l < cnt; ++l) {
                    const uint32_t rel = o - w0;           // wraps for o < w0: fails the test below
                    if (rel < (uint32_t)EMIT_WINDOW) {       // the serial walk only notes WHERE; the copy-out does the rest
                        s_key[rel] = (row + (x << view_bits)) | ((sx0 + x) << 16) | (sy << 24);
                        s_slot[rel] = (uint16_t)(threadIdx.x * EMIT_ITEMS + k);
                    }
                    ++o;
                    if (++x == wid) { x = 0; row += (uint32_t)stiles_x << view_bits; ++sy; }
                }
            }
        }
        __syncthreads();
        const uint32_t n = min((uint32_t)EMIT_WINDOW, total - w0);
        for (uint32_t o0 = 0; o0 < n; o0 += SCAN_BLOCK) {
            const uint32_t o = o0 + threadIdx.x;
            if (o < n) {
                const uint32_t packed = s_key[o];
                const uint32_t slot = s_slot[o];
                const uint32_t kk = packed & ((1u << ST_KEY_BITS) - 1u);
                const uint32_t key =
                    (st_hit_mask(s_rect[slot], ((packed >> 16) & 0xffu) * ST, (packed >> 24) * ST) << ST_KEY_BITS) | kk;
                const uint32_t dst = out0 + w0 + o;
                if (dst < capacity) {
                    ikeys[dst] = key;
                    ivals[dst] = s_gidx[slot];
                    // instances per key: everything the sort and the chunk table need to know about the key distribution
                    if (khist_views) atomicAdd(&s_khist[(kk >> view_bits) * khist_views + ((kk & vmask) - view_first)], 1u);
                    else atomicAdd(&key_hist[kk], 1u);
                } else {
                    overflow = true;
                }
            }
        }
    }
    if (overflow) counters[CNT_OVERFLOW] = 1;
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < khist_bins; i += SCAN_BLOCK) {
        const uint32_t c = s_khist[i];
        const uint32_t st = i / khist_views, vl = i - st * khist_views;
        if (c) atomicAdd(&key_hist[(st << view_bits) | (view_first + vl)], c);
    }
}

// single block: the digit histograms of the instance sort from the per-key counts (the sort digit is a bit field of the
// key's supertile part), so that no pass over the instances is needed for them
__global__ void __launch_bounds__(1024)
sort_hist_from_keys_kernel(uint32_t num_keys, int view_bits, int passes, const uint32_t* __restrict__ key_hist,
                           uint32_t* __restrict__ sort_hist) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ uint32_t s_h[MAX_PASSES][RADIX];
    for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x) (&s_h[0][0])[i] = 0;
    __syncthreads();
    for (uint32_t k = threadIdx.x; k < num_keys; k += blockDim.x) {
        const uint32_t c = key_hist[k];
        if (!c) continue;
        const uint32_t st = k >> view_bits;
        for (int p = 0; p < passes; ++p) atomicAdd(&s_h[p][(st >> (p * RADIX_BITS)) & (RADIX - 1)], c);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < passes * RADIX; i += blockDim.x) sort_hist[i] = (&s_h[0][0])[i];
}

// single block, off the critical path (it runs next to the instance sort, launch_binning): from the per-key instance counts
// the emit kernel took,
//   st_ranges[k]  = [first, last + 1) of key k in the SORTED instance list (the reference's identifyTileRanges,
//                   rasterizer_impl.cu:116-138, one level up -- but without reading the sorted keys),
//   chunk_base[k] = id of key k's first chunk (ceil(len_k / CHUNK) chunks each), counters[CNT_NCHUNKS] = number of chunks,
//   chunk_key[c]  = owning key of chunk c.
constexpr int CT_THREADS = 1024;

__global__ void __launch_bounds__(CT_THREADS)
chunk_table_kernel(uint32_t num_keys, uint32_t S, int view_bits, const uint32_t* __restrict__ key_hist,
                   uint2* __restrict__ st_ranges, uint32_t* __restrict__ chunk_base, uint32_t* __restrict__ chunk_key,
                   uint32_t max_chunks, uint32_t* __restrict__ counters) {
    __shared__ uint32_t s_warp[32];
    // (1) the sorted list is in key order (supertile-major): ranges by a scan over the keys
    uint32_t carry = 0;
    for (uint32_t start = 0; start < num_keys; start += CT_THREADS) {
        const uint32_t k = start + threadIdx.x;
        const uint32_t n = k < num_keys ? key_hist[k] : 0u;   // stored instances only: their sum never exceeds the capacity
        uint32_t tot;
        const uint32_t first = carry + block_exclusive_scan_1024(n, s_warp, tot);
        if (k < num_keys) st_ranges[k] = make_uint2(first, first + n);
        carry += tot;
    }
    // (2) chunks are NUMBERED view-major (view, supertile): the kernels that walk the chunks in id order then stay inside
    // one view's per-Gaussian arrays for a long stretch (supertile-major numbering cycled through all views' 9.6 MB
    // geometry records every few chunks and cost the record scatter 20 % more DRAM traffic)
    carry = 0;
    for (uint32_t start = 0; start < num_keys; start += CT_THREADS) {
        const uint32_t j = start + threadIdx.x;
        const uint32_t view = j / S, st = j - view * S;
        const uint32_t k = (st << view_bits) | view;
        const uint32_t n = j < num_keys ? key_hist[k] : 0u;
        const uint32_t nch = (n + CHUNK - 1) / CHUNK;
        uint32_t tot;
        const uint32_t cb = carry + block_exclusive_scan_1024(nch, s_warp, tot);
        if (j < num_keys) {
            chunk_base[k] = cb;
            for (uint32_t c = 0; c < nch; ++c)
                if (cb + c < max_chunks) chunk_key[cb + c] = k;
        }
        carry += tot;
    }
    if (threadIdx.x == 0) counters[CNT_NCHUNKS] = min(carry, max_chunks);
}

struct ChunkInfo {
    uint32_t key, first, count;      // supertile key, first sorted instance, instances in this chunk
    uint32_t view, ox, oy;           // view and first tile of the supertile
};

// per chunk: how many of its instances hit each of the supertile's 16 tiles (hit masks: upper half of the sorted keys).
// One WARP per chunk: 8 rows of 32 keys, 16 ballots a row, lane t < 16 keeps tile t's count in a register -- no shared
// memory, no CTA barrier, and eight chunks per CTA in flight (the CTA-per-chunk version was bound by the latency of its
// chunk-table look-ups: 77 us at C2).
__global__ void __launch_bounds__(256)
chunk_count_kernel(const uint32_t* __restrict__ counters, const uint32_t* __restrict__ chunk_key,
                   const uint32_t* __restrict__ chunk_base, const uint2* __restrict__ st_ranges,
                   const uint32_t* __restrict__ skeys, uint32_t* __restrict__ chunk_counts) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    const int lane = threadIdx.x & 31;
    const uint32_t wpb = blockDim.x >> 5;
    const uint32_t nch = counters[CNT_NCHUNKS];
    // persistent warps: the chunk count is only known on the device, so the grid is sized for the SMs, not the chunks
    for (uint32_t cid = blockIdx.x * wpb + (threadIdx.x >> 5); cid < nch; cid += gridDim.x * wpb) {
        const uint32_t key = chunk_key[cid];
        const uint2 r = st_ranges[key];
        const uint32_t first = r.x + (cid - chunk_base[key]) * CHUNK;
        const uint32_t count = min((uint32_t)CHUNK, r.y - first);
        uint32_t m[CHUNK / 32];
#pragma unroll
        for (int row = 0; row < CHUNK / 32; ++row) {
            const uint32_t i = row * 32 + lane;
            m[row] = i < count ? skeys[first + i] >> ST_KEY_BITS : 0u;
        }
        // warp-uniform counters (ballot -> popc -> add never leave the uniform datapath: the per-lane ALU only forms the
        // predicates; with `if (lane == t) n += ...` the kernel was ALU-bound, 86 % pipe utilisation)
        uint32_t cnt[ST_TILES];
#pragma unroll
        for (int t = 0; t < ST_TILES; ++t) cnt[t] = 0;
#pragma unroll
        for (int row = 0; row < CHUNK / 32; ++row) {
#pragma unroll
            for (int t = 0; t < ST_TILES; ++t) cnt[t] += __popc(__ballot_sync(FULL, (m[row] >> t) & 1u));
        }
        uint32_t n = 0;
#pragma unroll
        for (int t = 0; t < ST_TILES; ++t) n = lane == t ? cnt[t] : n;
        if (lane < ST_TILES) chunk_counts[(size_t)cid * ST_TILES + lane] = n;
    }
}

// One CTA per supertile key, one WARP per tile of the supertile: exclusive prefix of the tile's chunk counts over the key's
// chunks (in place, 32 chunks per step with a shuffle scan), and the tile's total -> tile_start[view*T + tile] (scanned by
// tile_scan_kernel).  (Round 1 walked the chunks serially with one lane per tile: 55 us at C2, set by the longest
// supertile list.)
__global__ void __launch_bounds__(ST_TILES * 32)
tile_offsets_kernel(int V, int view_bits, int stiles_x, int tiles_x, int tiles_y,
                    const uint32_t* __restrict__ chunk_base, const uint2* __restrict__ st_ranges, uint32_t max_chunks,
                    uint32_t* __restrict__ chunk_counts, uint32_t* __restrict__ tile_start) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    const uint32_t k = blockIdx.x;
    const uint32_t view = k & ((1u << view_bits) - 1u), s = k >> view_bits;
    if (view >= (uint32_t)V) return;                     // key slots of views the batch does not have
    const int t = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint2 kr = st_ranges[k];
    const uint32_t c0 = min(chunk_base[k], max_chunks);
    const uint32_t c1 = min(chunk_base[k] + (kr.y - kr.x + CHUNK - 1) / CHUNK, max_chunks);
    uint32_t run = 0;
    for (uint32_t cb = c0; cb < c1; cb += 32) {
        const uint32_t c = cb + lane;
        const uint32_t n = c < c1 ? chunk_counts[(size_t)c * ST_TILES + t] : 0u;
        uint32_t inc = n;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t v = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += v;
        }
        if (c < c1) chunk_counts[(size_t)c * ST_TILES + t] = run + inc - n;
        run += __shfl_sync(FULL, inc, 31);
    }
    if (lane == 0) {
        const uint32_t sy = s / (uint32_t)stiles_x, sx = s - sy * (uint32_t)stiles_x;
        const uint32_t gx = sx * ST + (t & 3), gy = sy * ST + (t >> 2);
        if (gx < (uint32_t)tiles_x && gy < (uint32_t)tiles_y)
            tile_start[(size_t)view * tiles_x * tiles_y + gy * (uint32_t)tiles_x + gx] = run;
    }
}

// single block: exclusive scan of the per-tile counts in place (tile_start), ranges clamped to the capacity
constexpr int ORDER_BUCKETS = 1024;
constexpr int TSCAN_THREADS = 1024;          // one CTA on the critical path: as many loads in flight as one CTA can have

__global__ void __launch_bounds__(TSCAN_THREADS)
tile_scan_kernel(uint32_t VT, uint32_t capacity, uint32_t* __restrict__ tile_start, uint2* __restrict__ ranges,
                 uint32_t* __restrict__ counters) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ uint32_t s_warp[32];
    uint32_t carry = 0;
    constexpr int IT = 8;
    uint32_t nxt[IT];                                    // the next step's counts are in flight while this one is scanned
    auto load8 = [&](uint32_t i, uint32_t* out) {        // tile_start is 256-byte aligned and i a multiple of 8
        if (i + IT <= VT) {
            const uint4 a = *reinterpret_cast<const uint4*>(tile_start + i), b = *reinterpret_cast<const uint4*>(tile_start + i + 4);
            out[0] = a.x; out[1] = a.y; out[2] = a.z; out[3] = a.w; out[4] = b.x; out[5] = b.y; out[6] = b.z; out[7] = b.w;
        } else {
#pragma unroll
            for (int k = 0; k < IT; ++k) out[k] = i + k < VT ? tile_start[i + k] : 0u;
        }
    };
    load8(threadIdx.x * IT, nxt);
    for (uint32_t start = 0; start < VT; start += TSCAN_THREADS * IT) {
        const uint32_t i0 = start + threadIdx.x * IT;
        uint32_t v[IT], sum = 0;
#pragma unroll
        for (int k = 0; k < IT; ++k) { v[k] = nxt[k]; sum += v[k]; }
        load8(i0 + TSCAN_THREADS * IT, nxt);
        uint32_t total;
        uint32_t ex = carry + block_exclusive_scan_1024(sum, s_warp, total);
        if (i0 + IT <= VT) {                              // whole group: 2 + 4 vector stores instead of 16 scalar ones
            uint32_t e[IT + 1];
            e[0] = ex;
#pragma unroll
            for (int k = 0; k < IT; ++k) e[k + 1] = e[k] + v[k];
            *reinterpret_cast<uint4*>(tile_start + i0) = make_uint4(e[0], e[1], e[2], e[3]);
            *reinterpret_cast<uint4*>(tile_start + i0 + 4) = make_uint4(e[4], e[5], e[6], e[7]);
            uint4* rp = reinterpret_cast<uint4*>(ranges + i0);
#pragma unroll
            for (int k = 0; k < IT; k += 2)
                rp[k / 2] = make_uint4(min(e[k], capacity), min(e[k + 1], capacity), min(e[k + 1], capacity),
                                       min(e[k + 2], capacity));
        } else {
#pragma unroll
            for (int k = 0; k < IT; ++k) {
                if (i0 + k < VT) {
                    tile_start[i0 + k] = ex;
                    ranges[i0 + k] = make_uint2(min(ex, capacity), min(ex + v[k], capacity));
                }
                ex += v[k];
            }
        }
        carry += total;
    }
    if (threadIdx.x == 0) {
        tile_start[VT] = carry;
        if (carry > capacity) counters[CNT_OVERFLOW] = 1;
    }
}

// single block, off the critical path (it runs next to the record scatter; only the blend kernels need it): counting sort of
// the (view, tile) indices by list length, longest first -> tile_order.  The blend kernels launch their CTAs in that
// order, so the longest sequential chains start first and the short tiles fill in behind them (the hardware dispatches
// CTAs in blockIdx order).
__global__ void __launch_bounds__(TSCAN_THREADS)
tile_order_kernel(uint32_t VT, const uint2* __restrict__ ranges, uint32_t* __restrict__ tile_order) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_bucket[ORDER_BUCKETS];
    __shared__ uint32_t s_maxlen;
    static_assert(ORDER_BUCKETS == TSCAN_THREADS, "one bucket per thread in the bucket scan");
    if (threadIdx.x == 0) s_maxlen = 0;
    s_bucket[threadIdx.x] = 0;
    __syncthreads();
    uint32_t vmax = 0;
    for (uint32_t i = threadIdx.x; i < VT; i += TSCAN_THREADS) {
        const uint2 r = ranges[i];
        vmax = max(vmax, r.y - r.x);
    }
    vmax = __reduce_max_sync(FULL, vmax);
    if ((threadIdx.x & 31) == 0 && vmax > 0) atomicMax(&s_maxlen, vmax);
    __syncthreads();
    // bucket 0 holds the longest lists ... ORDER_BUCKETS-1 the empty ones
    const uint32_t q = s_maxlen / (ORDER_BUCKETS - 1) + 1;
    constexpr int OT = 4;
    for (uint32_t start = 0; start < VT; start += TSCAN_THREADS * OT) {
        uint2 r[OT];
#pragma unroll
        for (int k = 0; k < OT; ++k) {
            const uint32_t i = start + k * TSCAN_THREADS + threadIdx.x;
            r[k] = i < VT ? ranges[i] : make_uint2(0u, 0u);
        }
#pragma unroll
        for (int k = 0; k < OT; ++k)
            if (start + k * TSCAN_THREADS + threadIdx.x < VT)
                atomicAdd(&s_bucket[(ORDER_BUCKETS - 1) - min((r[k].y - r[k].x) / q, (uint32_t)ORDER_BUCKETS - 1)], 1u);
    }
    __syncthreads();
    {
        const uint32_t c = s_bucket[threadIdx.x];
        uint32_t total;
        const uint32_t ex = block_exclusive_scan_1024(c, s_warp, total);
        s_bucket[threadIdx.x] = ex;
    }
    __syncthreads();
    for (uint32_t start = 0; start < VT; start += TSCAN_THREADS * OT) {
        uint2 r[OT];
#pragma unroll
        for (int k = 0; k < OT; ++k) {
            const uint32_t i = start + k * TSCAN_THREADS + threadIdx.x;
            r[k] = i < VT ? ranges[i] : make_uint2(0u, 0u);
        }
#pragma unroll
        for (int k = 0; k < OT; ++k) {
            const uint32_t i = start + k * TSCAN_THREADS + threadIdx.x;
            if (i < VT) {
                const uint32_t pos =
                    atomicAdd(&s_bucket[(ORDER_BUCKETS - 1) - min((r[k].y - r[k].x) / q, (uint32_t)ORDER_BUCKETS - 1)], 1u);
                tile_order[pos] = i;
            }
        }
    }
}

// Per chunk: order-preserving compaction of its instances into the 16 tile lists, writing the blend records.
// Thread i holds instance i's records in registers (geometry: two float4 of one 32-byte sector written by the
// preprocess kernel; colour: one float4); ballots give every hit its list position (per tile: chunk offset + hits of
// the earlier warps + hits of the lower lanes), and every thread stores its records once per tile it hits.
template <bool EXTRA>
__global__ void __launch_bounds__(CHUNK)
scatter_records_kernel(const uint32_t* __restrict__ counters, const uint32_t* __restrict__ chunk_key,
                       const uint32_t* __restrict__ chunk_base, const uint2* __restrict__ st_ranges, int view_bits,
                       int stiles_x, int tiles_x, int tiles_y, int P, uint32_t capacity,
                       const uint32_t* __restrict__ vals, const uint32_t* __restrict__ skeys,
                       const uint32_t* __restrict__ chunk_offs, const uint32_t* __restrict__ tile_start,
                       const float4* __restrict__ geomrec, const float4* __restrict__ col4,
                       const float4* __restrict__ rgb_sh4, const float4* __restrict__ ext4,
                       float4* __restrict__ recA, float4* __restrict__ recB, float4* __restrict__ recC,
                       float4* __restrict__ recD) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ uint32_t s_wcnt[ST_TILES][CHUNK / 32];
    __shared__ uint32_t s_tbase[ST_TILES];            // list position of this chunk's first hit, per tile
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t lt = (1u << lane) - 1u;
    // software pipeline over the CTA's chunks (the kernel is bound by dependent-load latency: chunk table -> instance
    // -> record): while chunk i's records are gathered and stored, the (instance, hit mask) pairs of chunk i+1 and the
    // chunk-table entries of chunk i+2 are already in flight
    const uint32_t nch = counters[CNT_NCHUNKS];
    auto info = [&](uint32_t cid, ChunkInfo& o) -> bool {
        if (cid >= nch) return false;
        o.key = chunk_key[cid];
        const uint2 r = st_ranges[o.key];
        o.first = r.x + (cid - chunk_base[o.key]) * CHUNK;
        o.count = min((uint32_t)CHUNK, r.y - o.first);
        o.view = o.key & ((1u << view_bits) - 1u);
        const uint32_t sidx = o.key >> view_bits;
        const uint32_t sy = sidx / (uint32_t)stiles_x;
        o.ox = (sidx - sy * (uint32_t)stiles_x) * ST;
        o.oy = sy * ST;
        return true;
    };
    ChunkInfo c, cn, cnn;
    bool have = info(blockIdx.x, c);
    bool have_next = info(blockIdx.x + gridDim.x, cn);
    uint32_t gi = 0, mask = 0;
    if (have && (uint32_t)tid < c.count) {
        gi = vals[c.first + tid];
        mask = skeys[c.first + tid] >> ST_KEY_BITS;
    }
    for (uint32_t cid = blockIdx.x; have; cid += gridDim.x) {
        float4 rA, rB, rC, rD;
        if (mask) {
            const uint32_t g = gi - c.view * (uint32_t)P;
            rA = geomrec[2 * (size_t)gi];
            rB = geomrec[2 * (size_t)gi + 1];
            rC = col4 ? col4[g] : rgb_sh4[gi];
            rC.w = __uint_as_float(gi);
            if (EXTRA) rD = ext4[g];
        }
        uint32_t gi_next = 0, mask_next = 0;
        if (have_next && (uint32_t)tid < cn.count) {
            gi_next = vals[cn.first + tid];
            mask_next = skeys[cn.first + tid] >> ST_KEY_BITS;
        }
        const bool have_nn = info(cid + 2 * gridDim.x, cnn);
        if (tid < ST_TILES) {
            const uint32_t gx = c.ox + (tid & 3), gy = c.oy + (tid >> 2);
            uint32_t b = 0;
            if (gx < (uint32_t)tiles_x && gy < (uint32_t)tiles_y)
                b = tile_start[(size_t)c.view * tiles_x * tiles_y + gy * (uint32_t)tiles_x + gx] +
                    chunk_offs[(size_t)cid * ST_TILES + tid];
            s_tbase[tid] = b;
        }
        uint32_t bal[ST_TILES];
#pragma unroll
        for (int t = 0; t < ST_TILES; ++t) {
            bal[t] = __ballot_sync(FULL, (mask >> t) & 1u);
            if (lane == 0) s_wcnt[t][warp] = __popc(bal[t]);
        }
        __syncthreads();
        if (tid < ST_TILES) {                             // per tile: exclusive prefix over the 8 warps
            uint32_t run = 0;
#pragma unroll
            for (int w = 0; w < CHUNK / 32; ++w) {
                const uint32_t n = s_wcnt[tid][w];
                s_wcnt[tid][w] = run;
                run += n;
            }
        }
        __syncthreads();
        // the hits of a warp on one tile occupy consecutive list slots: each store instruction writes one contiguous run
#pragma unroll
        for (int t = 0; t < ST_TILES; ++t) {
            if ((mask >> t) & 1u) {
                const uint32_t dst = s_tbase[t] + s_wcnt[t][warp] + __popc(bal[t] & lt);
                if (dst < capacity) {
                    recA[dst] = rA;
                    recB[dst] = rB;
                    recC[dst] = rC;
                    if (EXTRA) recD[dst] = rD;
                }
            }
        }
        __syncthreads();                              // shared counters are reused by the next chunk
        c = cn;
        cn = cnn;
        gi = gi_next;
        mask = mask_next;
        have = have_next;
        have_next = have_nn;
    }
}

// float4 packing of the per-Gaussian colours handed to gsr_forward_render ([P][3] AoS in the reference's layout)
__global__ void __launch_bounds__(256)
pack_colors_kernel(int P, const float* __restrict__ colors, const float* __restrict__ extra,
                   float4* __restrict__ col4, float4* __restrict__ ext4) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= P) return;
    if (colors) col4[g] = make_float4(colors[3 * g], colors[3 * g + 1], colors[3 * g + 2], 0.0f);
    if (extra) ext4[g] = make_float4(extra[3 * g], extra[3 * g + 1], extra[3 * g + 2], 0.0f);
}

// ---- export of one view in the reference's format (tests / tools): one CTA per tile ----
__global__ void __launch_bounds__(128)
export_view_kernel(int view, int P, uint32_t T, const uint2* __restrict__ ranges, const float4* __restrict__ recB,
                   const float4* __restrict__ recC, uint64_t* __restrict__ keys_out, uint32_t* __restrict__ list_out,
                   uint32_t* __restrict__ ranges_out) {
    const uint32_t tile = blockIdx.x;
    const uint32_t vs = ranges[(size_t)view * T].x;                 // first instance of the view (tile 0 may be empty:
    const uint2 r = ranges[(size_t)view * T + tile];                // its range is then [start, start))
    if (ranges_out && threadIdx.x == 0) {
        const bool empty = r.y == r.x;
        ranges_out[2 * tile] = empty ? 0u : r.x - vs;
        ranges_out[2 * tile + 1] = empty ? 0u : r.y - vs;
    }
    for (uint32_t i = r.x + threadIdx.x; i < r.y; i += blockDim.x) {
        const uint32_t gi = __float_as_uint(recC[i].w);
        if (keys_out) keys_out[i - vs] = ((uint64_t)tile << 32) | (uint64_t)__float_as_uint(recB[i].z);
        if (list_out) list_out[i - vs] = gi - (uint32_t)view * (uint32_t)P;
    }
}

}  // namespace

int launch_binning(int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                   const float* colors_precomp, const float* extra_colors, void* gws, void* bws, void* iws,
                   const Layout& L, cudaStream_t s) {
    const uint32_t VP = (uint32_t)((size_t)V * P);
    const uint32_t T = (uint32_t)L.tiles_x * (uint32_t)L.tiles_y;
    const uint32_t S = (uint32_t)L.stiles_x * (uint32_t)L.stiles_y;
    const uint32_t VT = (uint32_t)V * T, VS = (uint32_t)V * S;
    uint32_t* counters = at<uint32_t>(gws, L.counters);
    if (VP == 0 || capacity <= 0) {     // every tile list empty: zero counts through the same scan (ranges, launch order)
        if (VT == 0) return GSR_OK;
        GSR_CUDA_TRY(cudaMemsetAsync(at<uint32_t>(iws, L.tile_start), 0, ((size_t)VT + 1) * sizeof(uint32_t), s));
        launch_kernel(tile_scan_kernel, 1, TSCAN_THREADS, 0, s, VT, 0u, at<uint32_t>(iws, L.tile_start), at<uint2>(iws, L.ranges),
                                                  counters);
        GSR_LAUNCH_CHECK();
        tile_order_kernel<<<1, TSCAN_THREADS, 0, s>>>(VT, at<uint2>(iws, L.ranges), at<uint32_t>(iws, L.tile_order));
        GSR_LAUNCH_CHECK();
        return GSR_OK;
    }
    if (L.tile_bits > ST_KEY_BITS) return GSR_ERR_UNSUPPORTED;      // check_dims refuses such batches before
    (void)VS;
    const uint32_t cap = (uint32_t)capacity;
    const uint32_t nblocks = (VP + EMIT_TILE - 1) / EMIT_TILE;
    const uint32_t* sorted_g = at<uint32_t>(gws, L.sorted_vals_in ? L.dvals1 : L.dvals0);
    const ushort4* rects = at<ushort4>(gws, L.rects);
    const uint32_t NK = L.num_keys;                      // key slots: supertiles per view << view_bits
    uint32_t* key_hist = at<uint32_t>(bws, L.key_hist);
    uint2* st_ranges = at<uint2>(bws, L.st_ranges);
    uint32_t* chunk_base = at<uint32_t>(bws, L.chunk_base);
    uint32_t* chunk_key = at<uint32_t>(bws, L.chunk_key);
    uint32_t* chunk_counts = at<uint32_t>(bws, L.chunk_counts);
    uint32_t* tile_start = at<uint32_t>(iws, L.tile_start);
    const uint32_t max_chunks = (uint32_t)L.max_chunks;
    cudaStream_t side = nullptr;
    stage_begin(ST_EMIT, s);
    // instance sort histograms, tile tickets and look-back words, the emit kernel's ticket / look-back words / key counts:
    // one memset
    GSR_CUDA_TRY(cudaMemsetAsync(bws, 0, L.binning_zero_bytes, s));
    {
        // key counts in shared memory: the keys of the (at most EMIT_TILE / P + 2) views a block's slots belong to
        uint32_t nv = (uint32_t)V < (uint32_t)(EMIT_TILE / P + 2) ? (uint32_t)V : (uint32_t)(EMIT_TILE / P + 2);
        uint32_t bins = nv * S;
        constexpr uint32_t MAX_BINS = 24 * 1024;                     // 96 KB next to the kernel's 40 KB of static smem
        if (bins > MAX_BINS) bins = nv = 0;                          // (tiny P with thousands of supertiles: global atomics)
        static bool attr_set = false;
        if (!attr_set) {
            GSR_CUDA_TRY(cudaFuncSetAttribute(emit_st_instances_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              MAX_BINS * 4));
            attr_set = true;
        }
#if GSR_EMIT_LOOKBACK
        uint32_t* offs = at<uint32_t>(bws, L.emit_lookback);
#else
        uint32_t* offs = at<uint32_t>(gws, L.block_sums);
        launch_kernel(st_count_reduce_kernel, nblocks, SCAN_BLOCK, 0, s, VP, P, sorted_g, rects, at<uint32_t>(gws, L.vis_count), offs);
        GSR_LAUNCH_CHECK();
        launch_kernel(block_sums_scan_kernel, 1, 1024, 0, s, nblocks, offs, counters);
        GSR_LAUNCH_CHECK();
#endif
        launch_kernel(emit_st_instances_kernel, nblocks, SCAN_BLOCK, bins * 4, s, 
            VP, P, L.stiles_x, S, L.view_bits, sorted_g, rects, at<uint32_t>(gws, L.vis_count), cap, nblocks,
            at<uint32_t>(bws, L.ikeys0), at<uint32_t>(bws, L.ivals0), counters, offs, at<uint32_t>(bws, L.emit_ticket),
            key_hist, nv);
        GSR_LAUNCH_CHECK();
        launch_kernel(sort_hist_from_keys_kernel, 1, 1024, 0, s, NK, L.view_bits, L.tile_passes, key_hist, at<uint32_t>(bws, L.ihist));
        GSR_LAUNCH_CHECK();
    }
    stage_end(ST_EMIT, s);
    // the chunk table only needs the key counts: it runs on the side stream while the instances are sorted
    int rc = side_fork(s, &side);
    if (rc != GSR_OK) return rc;
    chunk_table_kernel<<<1, CT_THREADS, 0, side>>>(NK, S, L.view_bits, key_hist, st_ranges, chunk_base, chunk_key, max_chunks,
                                                   counters);
    GSR_LAUNCH_CHECK();
    stage_begin(ST_TILE_SORT, s);
    rc = radix_sort_pairs<uint32_t>(at<uint32_t>(bws, L.ikeys0), at<uint32_t>(bws, L.ikeys1),
                                    at<uint32_t>(bws, L.ivals0), at<uint32_t>(bws, L.ivals1), capacity,
                                    counters + CNT_ST_RENDERED, L.view_bits, L.tile_bits, at<uint32_t>(bws, L.ihist),
                                    at<uint32_t>(bws, L.ilookback), at<uint32_t>(bws, L.ihist) + MAX_PASSES * RADIX, s, 1,
                                    true, true);
    stage_end(ST_TILE_SORT, s);
    if (rc != GSR_OK) return rc;
    rc = side_join(s);
    if (rc != GSR_OK) return rc;
    const uint32_t* skeys = at<uint32_t>(bws, L.sorted_keys_in ? L.ikeys1 : L.ikeys0);
    const uint32_t* svals = at<uint32_t>(bws, L.sorted_keys_in ? L.ivals1 : L.ivals0);
    const uint32_t chunk_grid = max_chunks < 148u * 8u ? max_chunks : 148u * 8u;   // persistent CTAs, 8 per SM
    stage_begin(ST_RANGES, s);
    launch_kernel(chunk_count_kernel, chunk_grid, 256, 0, s, counters, chunk_key, chunk_base, st_ranges, skeys, chunk_counts);
    GSR_LAUNCH_CHECK();
    launch_kernel(tile_offsets_kernel, NK, ST_TILES * 32, 0, s, V, L.view_bits, L.stiles_x, L.tiles_x, L.tiles_y, chunk_base, st_ranges,
                                                    max_chunks, chunk_counts, tile_start);
    GSR_LAUNCH_CHECK();
    launch_kernel(tile_scan_kernel, 1, TSCAN_THREADS, 0, s, VT, cap, tile_start, at<uint2>(iws, L.ranges), counters);
    GSR_LAUNCH_CHECK();
    stage_end(ST_RANGES, s);
    // the launch order of the blend kernels is not needed before the records are in place: side stream again
    rc = side_fork(s, &side);
    if (rc != GSR_OK) return rc;
    tile_order_kernel<<<1, TSCAN_THREADS, 0, side>>>(VT, at<uint2>(iws, L.ranges), at<uint32_t>(iws, L.tile_order));
    GSR_LAUNCH_CHECK();
    {
        StageScope sc(ST_GATHER, s);
        if (colors_precomp || extra_colors) {
            launch_kernel(pack_colors_kernel, (P + 255) / 256, 256, 0, s, P, colors_precomp, extra_colors, at<float4>(gws, L.col4),
                                                               at<float4>(gws, L.ext4));
            GSR_LAUNCH_CHECK();
        }
        // persistent CTAs: exactly as many as are resident at once (a second, partial wave would run at low occupancy)
        static int scatter_per_sm[2] = {0, 0};
        if (scatter_per_sm[0] == 0) {
            GSR_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&scatter_per_sm[0], scatter_records_kernel<false>, CHUNK, 0));
            GSR_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&scatter_per_sm[1], scatter_records_kernel<true>, CHUNK, 0));
        }
        // ... minus one SM's worth: the tile_order CTA on the side stream takes an SM, and persistent CTAs queued behind
        // it would finish a whole tile_order late
        const uint32_t resident = 147u * (uint32_t)scatter_per_sm[channels > 3 ? 1 : 0];
        const uint32_t scatter_grid = max_chunks < resident ? max_chunks : resident;
        const float4* col4 = colors_precomp ? at<float4>(gws, L.col4) : nullptr;
        const float4* rgb4 = sh_coeffs > 0 ? at<float4>(gws, L.rgb) : nullptr;
        if (channels > 3) {
            launch_kernel(scatter_records_kernel<true>, scatter_grid, CHUNK, 0, s, 
                counters, chunk_key, chunk_base, st_ranges, L.view_bits, L.stiles_x, L.tiles_x, L.tiles_y, P, cap, svals, skeys,
                chunk_counts, tile_start, at<float4>(gws, L.geomrec), col4, rgb4, at<float4>(gws, L.ext4),
                at<float4>(bws, L.recA), at<float4>(bws, L.recB), at<float4>(bws, L.recC), at<float4>(bws, L.recD));
        } else {
            launch_kernel(scatter_records_kernel<false>, scatter_grid, CHUNK, 0, s, 
                counters, chunk_key, chunk_base, st_ranges, L.view_bits, L.stiles_x, L.tiles_x, L.tiles_y, P, cap, svals, skeys,
                chunk_counts, tile_start, at<float4>(gws, L.geomrec), col4, rgb4, nullptr, at<float4>(bws, L.recA),
                at<float4>(bws, L.recB), at<float4>(bws, L.recC), nullptr);
        }
        GSR_LAUNCH_CHECK();
    }
    return side_join(s);
}

int launch_export_view(int view, int V, int P, int W, int H, const void* gws, const void* bws, const void* iws,
                       const Layout& L, uint64_t* keys_out, uint32_t* list_out, uint32_t* ranges_out,
                       uint32_t* n_contrib_out, cudaStream_t s) {
    const uint32_t T = (uint32_t)L.tiles_x * (uint32_t)L.tiles_y;
    if (keys_out || list_out || ranges_out) {
        export_view_kernel<<<T, 128, 0, s>>>(view, P, T, at<uint2>(iws, L.ranges), at<float4>(bws, L.recB),
                                             at<float4>(bws, L.recC), keys_out, list_out, ranges_out);
        GSR_LAUNCH_CHECK();
    }
    if (n_contrib_out) {
        GSR_CUDA_TRY(cudaMemcpyAsync(n_contrib_out, at<uint32_t>(iws, L.n_contrib) + (size_t)view * W * H,
                                     (size_t)W * H * 4, cudaMemcpyDeviceToDevice, s));
    }
    return GSR_OK;
}

}  // namespace gsr
