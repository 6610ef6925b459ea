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
constexpr int EMIT_HIST_PASSES = 2;      // instance sorts of up to 16 key bits get their histograms from the emit kernel

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

// slot i of the depth-sorted list: view = i / P, Gaussian = sorted_g[i]
__device__ __forceinline__ ushort4 slot_rect(uint32_t i, uint32_t VP, int P, const uint32_t* __restrict__ sorted_g,
                                             const ushort4* __restrict__ rects, uint32_t& gidx) {
    if (i >= VP) { gidx = 0; return make_ushort4(0, 0, 0, 0); }
    const uint32_t view = i / (uint32_t)P;
    gidx = view * (uint32_t)P + sorted_g[i];
    return rects[gidx];
}

// per block of 2048 slots: number of supertile instances (for the emit offsets); the tile-instance total
// (the reference's num_rendered) is accumulated on the side
__global__ void __launch_bounds__(SCAN_BLOCK)
st_count_reduce_kernel(uint32_t VP, int P, const uint32_t* __restrict__ sorted_g, const ushort4* __restrict__ rects,
                       uint32_t* __restrict__ block_sums, uint32_t* __restrict__ counters) {
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t sum = 0, tiles = 0, gidx;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const ushort4 r = slot_rect(base + k, VP, P, sorted_g, rects, gidx);
        uint32_t a, b, c;
        sum += st_rect(r, a, b, c);
        tiles += (uint32_t)(r.z - r.x) * (uint32_t)(r.w - r.y);
    }
    uint32_t total, total_tiles;
    block_exclusive_scan_256(sum, s_warp, total);
    block_exclusive_scan_256(tiles, s_warp, total_tiles);
    if (threadIdx.x == 0) {
        block_sums[blockIdx.x] = total;
        if (total_tiles) atomicAdd(&counters[CNT_RENDERED], total_tiles);
    }
}

// single block: exclusive scan of the per-block sums in place; total -> counters[CNT_ST_RENDERED]
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
        counters[CNT_ST_RENDERED] = carry;
        counters[CNT_OVERFLOW] = 0;
    }
}

// Emits the (view*S + supertile, view*P + Gaussian) pairs of every visible (view, Gaussian) in depth order.
// Every thread owns 8 consecutive slots of the depth-sorted list; the CTA scans the slots' instance counts, each
// thread writes its own instances (walking its supertile rectangles, no search) into a shared-memory window at their
// output positions, and the window is copied out with fully coalesced stores.  The digit histograms of the following
// sort are taken on the way out (warp-private bins, radix_rank.cuh), which saves that sort's pass over the keys.
constexpr int EMIT_WINDOW = 3072;        // staged outputs per round (a CTA of 2048 slots emits ~2900 at C2)

__global__ void __launch_bounds__(SCAN_BLOCK)
emit_st_instances_kernel(uint32_t VP, int P, int stiles_x, uint32_t S, const uint32_t* __restrict__ sorted_g,
                         const ushort4* __restrict__ rects, const uint32_t* __restrict__ block_sums, uint32_t capacity,
                         uint32_t* __restrict__ ikeys, uint32_t* __restrict__ ivals, uint32_t* __restrict__ counters,
                         int hist_passes, int key_bits, uint32_t* __restrict__ sort_hist) {
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    __shared__ uint32_t s_hist[SCAN_BLOCK / 32][EMIT_HIST_PASSES][RADIX];
    __shared__ uint32_t s_key[EMIT_WINDOW];
    __shared__ uint32_t s_val[EMIT_WINDOW];
    const uint32_t base = blockIdx.x * SCAN_TILE + threadIdx.x * SCAN_ITEMS;
    uint32_t cnt[SCAN_ITEMS], gidx[SCAN_ITEMS], key0[SCAN_ITEMS], wid[SCAN_ITEMS];
    uint32_t sum = 0;
#pragma unroll
    for (int k = 0; k < SCAN_ITEMS; ++k) {
        const ushort4 r = slot_rect(base + k, VP, P, sorted_g, rects, gidx[k]);
        uint32_t sx0, sy0;
        cnt[k] = st_rect(r, sx0, sy0, wid[k]);
        key0[k] = (gidx[k] / (uint32_t)P) * S + sy0 * (uint32_t)stiles_x + sx0;
        sum += cnt[k];
    }
    uint32_t total;
    const uint32_t off0 = block_exclusive_scan_256(sum, s_warp, total);
    if (hist_passes > 0)
        for (int i = threadIdx.x; i < (SCAN_BLOCK / 32) * EMIT_HIST_PASSES * RADIX; i += SCAN_BLOCK) (&s_hist[0][0][0])[i] = 0;
    const uint32_t out0 = block_sums[blockIdx.x];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    bool overflow = false;
    for (uint32_t w0 = 0; w0 < total; w0 += EMIT_WINDOW) {
        __syncthreads();                                  // the previous window has been copied out (and s_hist zeroed)
        if (off0 < w0 + EMIT_WINDOW && off0 + sum > w0) {     // some of this thread's instances fall into the window
            uint32_t o = off0;
#pragma unroll
            for (int k = 0; k < SCAN_ITEMS; ++k) {
                uint32_t row = key0[k], x = 0;
                for (uint32_t l = 0; l < cnt[k]; ++l) {
                    const uint32_t rel = o - w0;           // wraps for o < w0: fails the test below
                    if (rel < (uint32_t)EMIT_WINDOW) {
                        s_key[rel] = row + x;
                        s_val[rel] = gidx[k];
                    }
                    ++o;
                    if (++x == wid[k]) { x = 0; row += (uint32_t)stiles_x; }
                }
            }
        }
        __syncthreads();
        const uint32_t n = min((uint32_t)EMIT_WINDOW, total - w0);
        for (uint32_t o0 = 0; o0 < n; o0 += SCAN_BLOCK) {     // warp-uniform trip count (the histogram uses ballots)
            const uint32_t o = o0 + threadIdx.x;
            bool stored = false;
            uint32_t key = 0;
            if (o < n) {
                key = s_key[o];
                const uint32_t dst = out0 + w0 + o;
                if (dst < capacity) {
                    ikeys[dst] = key;
                    ivals[dst] = s_val[o];
                    stored = true;
                } else {
                    overflow = true;
                }
            }
#pragma unroll
            for (int p = 0; p < EMIT_HIST_PASSES; ++p) {
                if (p < hist_passes) {
                    const int shift = p * RADIX_BITS;
                    const int bits = min(RADIX_BITS, key_bits - shift);
                    uint32_t peers;
                    if (bits <= 4)
                        warp_hist_add<4>(s_hist[warp][p], (key >> shift) & ((1u << bits) - 1u), stored, lane, peers);
                    else
                        warp_hist_add(s_hist[warp][p], (key >> shift) & ((1u << bits) - 1u), stored, lane, peers);
                }
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

// boundary detection on the sorted supertile keys (the reference's identifyTileRanges, rasterizer_impl.cu:116-138, one
// level up): st_ranges[key] = [first, last+1) ; st_ranges is zeroed before
__global__ void __launch_bounds__(256)
st_ranges_kernel(const uint32_t* __restrict__ counters, uint32_t capacity, const uint32_t* __restrict__ keys,
                 uint2* __restrict__ st_ranges) {
    const uint32_t n = min(counters[CNT_ST_RENDERED], capacity);
    // grid-stride: the count lives on the device, so the grid is sized for the SMs rather than for the capacity
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        const uint32_t cur = keys[i];
        if (i == 0) {
            st_ranges[cur].x = 0;
        } else {
            const uint32_t prev = keys[i - 1];
            if (cur != prev) {
                st_ranges[prev].y = i;
                st_ranges[cur].x = i;
            }
        }
        if (i == n - 1) st_ranges[cur].y = n;
    }
}

// single block: chunk_base[k] = exclusive scan of ceil(len_k / CHUNK) over the V*S supertile keys, chunk_base[VS] and
// counters[CNT_NCHUNKS] = number of chunks; chunk_key[c] = owning key of chunk c
__global__ void __launch_bounds__(SCAN_BLOCK)
chunk_table_kernel(uint32_t VS, const uint2* __restrict__ st_ranges, uint32_t* __restrict__ chunk_base,
                   uint32_t* __restrict__ chunk_key, uint32_t max_chunks, uint32_t* __restrict__ counters) {
    __shared__ uint32_t s_warp[SCAN_BLOCK / 32];
    uint32_t carry = 0;
    for (uint32_t start = 0; start < VS; start += SCAN_BLOCK) {
        const uint32_t k = start + threadIdx.x;
        uint32_t nch = 0;
        if (k < VS) {
            const uint2 r = st_ranges[k];
            nch = (r.y - r.x + CHUNK - 1) / CHUNK;
        }
        uint32_t total;
        const uint32_t cb = carry + block_exclusive_scan_256(nch, s_warp, total);
        if (k < VS) {
            chunk_base[k] = cb;
            for (uint32_t j = 0; j < nch; ++j)
                if (cb + j < max_chunks) chunk_key[cb + j] = k;
        }
        carry += total;
    }
    if (threadIdx.x == 0) {
        chunk_base[VS] = carry;
        counters[CNT_NCHUNKS] = min(carry, max_chunks);
    }
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

struct ChunkInfo {
    uint32_t key, first, count;      // supertile key, first sorted instance, instances in this chunk
    uint32_t view, ox, oy;           // view and first tile of the supertile
};

__device__ __forceinline__ bool chunk_info(uint32_t cid, const uint32_t* __restrict__ counters,
                                           const uint32_t* __restrict__ chunk_key, const uint32_t* __restrict__ chunk_base,
                                           const uint2* __restrict__ st_ranges, uint32_t S, int stiles_x, ChunkInfo& c) {
    if (cid >= counters[CNT_NCHUNKS]) return false;
    c.key = chunk_key[cid];
    const uint2 r = st_ranges[c.key];
    c.first = r.x + (cid - chunk_base[c.key]) * CHUNK;
    c.count = min((uint32_t)CHUNK, r.y - c.first);
    c.view = c.key / S;
    const uint32_t s = c.key - c.view * S;
    const uint32_t sy = s / (uint32_t)stiles_x;
    c.ox = (s - sy * (uint32_t)stiles_x) * ST;
    c.oy = sy * ST;
    return true;
}

// per chunk: how many of its instances hit each of the supertile's 16 tiles
__global__ void __launch_bounds__(CHUNK)
chunk_count_kernel(const uint32_t* __restrict__ counters, const uint32_t* __restrict__ chunk_key,
                   const uint32_t* __restrict__ chunk_base, const uint2* __restrict__ st_ranges, uint32_t S,
                   int stiles_x, const uint32_t* __restrict__ vals, const ushort4* __restrict__ rects,
                   uint16_t* __restrict__ masks, uint32_t* __restrict__ chunk_counts) {
    __shared__ uint32_t s_cnt[ST_TILES][CHUNK / 32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    ChunkInfo c;
    // persistent CTAs: the chunk count is only known on the device, so the grid is sized for the SMs, not the chunks
    for (uint32_t cid = blockIdx.x; chunk_info(cid, counters, chunk_key, chunk_base, st_ranges, S, stiles_x, c);
         cid += gridDim.x) {
        uint32_t mask = 0;
        if ((uint32_t)tid < c.count) {
            mask = st_hit_mask(rects[vals[c.first + tid]], c.ox, c.oy);
            masks[c.first + tid] = (uint16_t)mask;           // the scatter pass reads this instead of the rectangle
        }
#pragma unroll
        for (int t = 0; t < ST_TILES; ++t) {
            const uint32_t bal = __ballot_sync(FULL, (mask >> t) & 1u);
            if (lane == 0) s_cnt[t][warp] = __popc(bal);
        }
        __syncthreads();
        if (tid < ST_TILES) {
            uint32_t n = 0;
#pragma unroll
            for (int w = 0; w < CHUNK / 32; ++w) n += s_cnt[tid][w];
            chunk_counts[(size_t)cid * ST_TILES + tid] = n;
        }
        __syncthreads();
    }
}

// one warp per supertile key: per tile, exclusive prefix of the chunk counts over the key's chunks (in place), and the
// tile's total -> tile_start[view*T + tile] (scanned by tile_scan_kernel)
__global__ void __launch_bounds__(128)
tile_offsets_kernel(uint32_t VS, uint32_t S, int stiles_x, int tiles_x, int tiles_y,
                    const uint32_t* __restrict__ chunk_base, uint32_t max_chunks, uint32_t* __restrict__ chunk_counts,
                    uint32_t* __restrict__ tile_start) {
    const uint32_t k = blockIdx.x * 4 + (threadIdx.x >> 5);
    const int t = threadIdx.x & 31;
    if (k >= VS || t >= ST_TILES) return;
    const uint32_t c0 = min(chunk_base[k], max_chunks), c1 = min(chunk_base[k + 1], max_chunks);
    uint32_t run = 0;
    for (uint32_t c = c0; c < c1; ++c) {
        const uint32_t n = chunk_counts[(size_t)c * ST_TILES + t];
        chunk_counts[(size_t)c * ST_TILES + t] = run;
        run += n;
    }
    const uint32_t view = k / S, s = k - view * S;
    const uint32_t sy = s / (uint32_t)stiles_x, sx = s - sy * (uint32_t)stiles_x;
    const uint32_t gx = sx * ST + (t & 3), gy = sy * ST + (t >> 2);
    if (gx < (uint32_t)tiles_x && gy < (uint32_t)tiles_y)
        tile_start[(size_t)view * tiles_x * tiles_y + gy * (uint32_t)tiles_x + gx] = run;
}

// single block: exclusive scan of the per-tile counts in place (tile_start), ranges clamped to the capacity; then a
// counting sort of the (view, tile) indices by list length, longest first -> tile_order.  The blend kernels launch
// their CTAs in that order, so the longest sequential chains start first and the short tiles fill in behind them
// (the hardware dispatches CTAs in blockIdx order).
constexpr int ORDER_BUCKETS = 1024;
constexpr int TSCAN_THREADS = 1024;          // one CTA on the critical path: as many loads in flight as one CTA can have

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

__global__ void __launch_bounds__(TSCAN_THREADS)
tile_scan_kernel(uint32_t VT, uint32_t capacity, uint32_t* __restrict__ tile_start, uint2* __restrict__ ranges,
                 uint32_t* __restrict__ tile_order, uint32_t* __restrict__ counters) {
    __shared__ uint32_t s_warp[32];
    __shared__ uint32_t s_bucket[ORDER_BUCKETS];
    __shared__ uint32_t s_maxlen;
    static_assert(ORDER_BUCKETS == TSCAN_THREADS, "one bucket per thread in the bucket scan");
    if (threadIdx.x == 0) s_maxlen = 0;
    s_bucket[threadIdx.x] = 0;
    __syncthreads();
    uint32_t carry = 0, vmax = 0;
    constexpr int IT = 8;
    for (uint32_t start = 0; start < VT; start += TSCAN_THREADS * IT) {
        const uint32_t i0 = start + threadIdx.x * IT;
        uint32_t v[IT], sum = 0;
#pragma unroll
        for (int k = 0; k < IT; ++k) { v[k] = i0 + k < VT ? tile_start[i0 + k] : 0u; sum += v[k]; vmax = max(vmax, v[k]); }
        uint32_t total;
        uint32_t ex = carry + block_exclusive_scan_1024(sum, s_warp, total);
#pragma unroll
        for (int k = 0; k < IT; ++k) {
            if (i0 + k < VT) {
                tile_start[i0 + k] = ex;
                ranges[i0 + k] = make_uint2(min(ex, capacity), min(ex + v[k], capacity));
            }
            ex += v[k];
        }
        carry += total;
    }
    if (threadIdx.x == 0) {
        tile_start[VT] = carry;
        if (carry > capacity) counters[CNT_OVERFLOW] = 1;
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
                       const uint32_t* __restrict__ chunk_base, const uint2* __restrict__ st_ranges, uint32_t S,
                       int stiles_x, int tiles_x, int tiles_y, int P, uint32_t capacity,
                       const uint32_t* __restrict__ vals, const uint16_t* __restrict__ masks,
                       const uint32_t* __restrict__ chunk_offs, const uint32_t* __restrict__ tile_start,
                       const float4* __restrict__ geomrec, const float4* __restrict__ col4,
                       const float4* __restrict__ rgb_sh4, const float4* __restrict__ ext4,
                       float4* __restrict__ recA, float4* __restrict__ recB, float4* __restrict__ recC,
                       float4* __restrict__ recD) {
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
        o.view = o.key / S;
        const uint32_t sidx = o.key - o.view * S;
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
        mask = masks[c.first + tid];
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
            mask_next = masks[cn.first + tid];
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

int launch_tile_scan(int V, int P, void* ws, const Layout& L, cudaStream_t s) {
    const uint32_t VP = (uint32_t)((size_t)V * P);
    if (VP == 0) return GSR_OK;
    const uint32_t nblocks = (VP + SCAN_TILE - 1) / SCAN_TILE;
    const uint32_t* sorted_g = at<uint32_t>(ws, L.sorted_vals_in ? L.dvals1 : L.dvals0);
    st_count_reduce_kernel<<<nblocks, SCAN_BLOCK, 0, s>>>(VP, P, sorted_g, at<ushort4>(ws, L.rects),
                                                         at<uint32_t>(ws, L.block_sums), at<uint32_t>(ws, L.counters));
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
    const uint32_t S = (uint32_t)L.stiles_x * (uint32_t)L.stiles_y;
    const uint32_t VT = (uint32_t)V * T, VS = (uint32_t)V * S;
    if (VP == 0 || capacity <= 0) {     // every tile list empty: zero counts through the same scan (ranges, launch order)
        if (VT == 0) return GSR_OK;
        GSR_CUDA_TRY(cudaMemsetAsync(at<uint32_t>(iws, L.tile_start), 0, ((size_t)VT + 1) * sizeof(uint32_t), s));
        tile_scan_kernel<<<1, TSCAN_THREADS, 0, s>>>(VT, 0u, at<uint32_t>(iws, L.tile_start), at<uint2>(iws, L.ranges),
                                                  at<uint32_t>(iws, L.tile_order), at<uint32_t>(gws, L.counters));
        GSR_LAUNCH_CHECK();
        return GSR_OK;
    }
    const uint32_t cap = (uint32_t)capacity;
    uint32_t* counters = at<uint32_t>(gws, L.counters);
    const uint32_t nblocks = (VP + SCAN_TILE - 1) / SCAN_TILE;
    const uint32_t* sorted_g = at<uint32_t>(gws, L.sorted_vals_in ? L.dvals1 : L.dvals0);
    const ushort4* rects = at<ushort4>(gws, L.rects);
    const bool fused_hist = L.tile_passes <= EMIT_HIST_PASSES;
    stage_begin(ST_EMIT, s);
    // instance sort histograms, tile tickets and look-back words, supertile ranges: one memset
    GSR_CUDA_TRY(cudaMemsetAsync(bws, 0, L.binning_zero_bytes, s));
    emit_st_instances_kernel<<<nblocks, SCAN_BLOCK, 0, s>>>(
        VP, P, L.stiles_x, S, sorted_g, rects, at<uint32_t>(gws, L.block_sums), cap, at<uint32_t>(bws, L.ikeys0),
        at<uint32_t>(bws, L.ivals0), counters, fused_hist ? L.tile_passes : 0, L.tile_bits, at<uint32_t>(bws, L.ihist));
    GSR_LAUNCH_CHECK();
    stage_end(ST_EMIT, s);
    stage_begin(ST_TILE_SORT, s);
    int rc = radix_sort_pairs<uint32_t>(at<uint32_t>(bws, L.ikeys0), at<uint32_t>(bws, L.ikeys1),
                                        at<uint32_t>(bws, L.ivals0), at<uint32_t>(bws, L.ivals1), capacity,
                                        counters + CNT_ST_RENDERED, 0, L.tile_bits, at<uint32_t>(bws, L.ihist),
                                        at<uint32_t>(bws, L.ilookback), at<uint32_t>(bws, L.ihist) + MAX_PASSES * RADIX, s, 1,
                                        fused_hist, true);
    stage_end(ST_TILE_SORT, s);
    if (rc != GSR_OK) return rc;
    const uint32_t* skeys = at<uint32_t>(bws, L.sorted_keys_in ? L.ikeys1 : L.ikeys0);
    const uint32_t* svals = at<uint32_t>(bws, L.sorted_keys_in ? L.ivals1 : L.ivals0);
    uint2* st_ranges = at<uint2>(bws, L.st_ranges);
    uint32_t* chunk_base = at<uint32_t>(bws, L.chunk_base);
    uint32_t* chunk_key = at<uint32_t>(bws, L.chunk_key);
    uint32_t* chunk_counts = at<uint32_t>(bws, L.chunk_counts);
    uint32_t* tile_start = at<uint32_t>(iws, L.tile_start);
    const uint32_t max_chunks = (uint32_t)L.max_chunks;
    const uint32_t chunk_grid = max_chunks < 148u * 8u ? max_chunks : 148u * 8u;   // persistent CTAs, 8 per SM
    stage_begin(ST_RANGES, s);
    st_ranges_kernel<<<min((cap + 255u) / 256u, 148u * 16u), 256, 0, s>>>(counters, cap, skeys, st_ranges);
    GSR_LAUNCH_CHECK();
    chunk_table_kernel<<<1, SCAN_BLOCK, 0, s>>>(VS, st_ranges, chunk_base, chunk_key, max_chunks, counters);
    GSR_LAUNCH_CHECK();
    chunk_count_kernel<<<chunk_grid, CHUNK, 0, s>>>(counters, chunk_key, chunk_base, st_ranges, S, L.stiles_x, svals,
                                                   rects, at<uint16_t>(bws, L.imasks), chunk_counts);
    GSR_LAUNCH_CHECK();
    tile_offsets_kernel<<<(VS + 3) / 4, 128, 0, s>>>(VS, S, L.stiles_x, L.tiles_x, L.tiles_y, chunk_base, max_chunks,
                                                    chunk_counts, tile_start);
    GSR_LAUNCH_CHECK();
    tile_scan_kernel<<<1, TSCAN_THREADS, 0, s>>>(VT, cap, tile_start, at<uint2>(iws, L.ranges),
                                              at<uint32_t>(iws, L.tile_order), counters);
    GSR_LAUNCH_CHECK();
    stage_end(ST_RANGES, s);
    StageScope sc(ST_GATHER, s);
    if (colors_precomp || extra_colors) {
        pack_colors_kernel<<<(P + 255) / 256, 256, 0, s>>>(P, colors_precomp, extra_colors, at<float4>(gws, L.col4),
                                                           at<float4>(gws, L.ext4));
        GSR_LAUNCH_CHECK();
    }
    // persistent CTAs: exactly as many as are resident at once (a second, partial wave would run at low occupancy)
    static int scatter_per_sm[2] = {0, 0};
    if (scatter_per_sm[0] == 0) {
        GSR_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&scatter_per_sm[0], scatter_records_kernel<false>, CHUNK, 0));
        GSR_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&scatter_per_sm[1], scatter_records_kernel<true>, CHUNK, 0));
    }
    const uint32_t resident = 148u * (uint32_t)scatter_per_sm[channels > 3 ? 1 : 0];
    const uint32_t scatter_grid = max_chunks < resident ? max_chunks : resident;
    const float4* col4 = colors_precomp ? at<float4>(gws, L.col4) : nullptr;
    const float4* rgb4 = sh_coeffs > 0 ? at<float4>(gws, L.rgb) : nullptr;
    if (channels > 3) {
        scatter_records_kernel<true><<<scatter_grid, CHUNK, 0, s>>>(
            counters, chunk_key, chunk_base, st_ranges, S, L.stiles_x, L.tiles_x, L.tiles_y, P, cap, svals,
            at<uint16_t>(bws, L.imasks), chunk_counts, tile_start, at<float4>(gws, L.geomrec), col4, rgb4, at<float4>(gws, L.ext4),
            at<float4>(bws, L.recA), at<float4>(bws, L.recB), at<float4>(bws, L.recC), at<float4>(bws, L.recD));
    } else {
        scatter_records_kernel<false><<<scatter_grid, CHUNK, 0, s>>>(
            counters, chunk_key, chunk_base, st_ranges, S, L.stiles_x, L.tiles_x, L.tiles_y, P, cap, svals,
            at<uint16_t>(bws, L.imasks), chunk_counts, tile_start, at<float4>(gws, L.geomrec), col4, rgb4, nullptr, at<float4>(bws, L.recA),
            at<float4>(bws, L.recB), at<float4>(bws, L.recC), nullptr);
    }
    GSR_LAUNCH_CHECK();
    return GSR_OK;
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
