// Hand-written stable LSD radix sort ("onesweep": one global histogram pass, then one chained-scan
// scatter pass per 8-bit digit with decoupled look-back).  Replaces cub::DeviceRadixSort::SortPairs
// (R/cuda_rasterizer/rasterizer_impl.cu:309-314, K/simple_knn.cu:217-220).
//
// Stability is what makes "bit-exact sorted lists" well defined (equal keys keep their input order):
//   * inside a tile, keys are ranked in (warp, item, lane) order == input order (warp-striped layout);
//   * tiles are ordered by a dynamically assigned tile index, which also gives the forward-progress guarantee the
//     look-back needs (a tile only waits on tiles whose index was handed out earlier, i.e. that already run).
//
// HBM traffic per pass: read (key,value) once, write once.  The look-back state is one 32-bit word per
// (tile, digit): 2 flag bits + 30-bit count, so a single volatile load carries flag and value together.
//
// SEGMENTED mode (segments > 1): `segments` independent arrays of up to n_max elements each, laid out n_max apart, are
// sorted by one set of launches (grid.y = segment); n_dev, when given, holds one device-side element count per segment.  The rasterizer sorts every view's Gaussians by depth this way
// with 32-bit keys (4 passes) instead of one (view << 32 | depth) 64-bit sort (5 passes of 12-byte pairs).
#include "radix_rank.cuh"

namespace gsr {

namespace {

constexpr uint32_t FLAG_AGG = 1u << 30;     // tile-local count published
constexpr uint32_t FLAG_INC = 1u << 31;     // inclusive prefix (all tiles <= this one) published
constexpr uint32_t VALUE_MASK = (1u << 30) - 1u;
constexpr unsigned FULL = 0xffffffffu;

__device__ __forceinline__ uint32_t ld_volatile(const uint32_t* p) { return *reinterpret_cast<const volatile uint32_t*>(p); }
__device__ __forceinline__ void st_volatile(uint32_t* p, uint32_t v) { *reinterpret_cast<volatile uint32_t*>(p) = v; }

// Global digit histogram of one pass (blockIdx.z) of one segment (blockIdx.y).
// Every LANE owns a private byte counter per bin (8 warps x 256 bins x 32 lanes = 64 KB), so counting a key is a plain
// byte load / add / store: no atomics (shared-memory atomics cost ~2 cycles per lane on this part) and no digit matching
// (9 ballots + bit logic per row, which is what the ballot-based version spent its time on: 127 us for 8.1 M keys x 4
// digits, ncu launch list r01 v4).  A warp flushes its counters into the block's 256 totals before any of them can
// exceed 255 rows.
constexpr int HIST_ROWS_PER_BLOCK = 1024;   // rows of 32 keys per block (128 per warp)
constexpr int HIST_FLUSH_ROWS = 248;        // per warp, < 256

template <typename KeyT>
__global__ void __launch_bounds__(256)
radix_histogram_kernel(const KeyT* __restrict__ keys, uint32_t n_max, const uint32_t* __restrict__ n_dev,
                       int begin_bit, int end_bit, int passes_total, uint32_t* __restrict__ hist, bool drop_sentinel) {
    extern __shared__ __align__(16) uint8_t s_cnt[];     // [8 warps][256 bins][32 lanes]
    __shared__ uint32_t s_tot[RADIX];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int seg = blockIdx.y, pass = blockIdx.z;
    const int shift = begin_bit + pass * RADIX_BITS;
    const uint32_t mask = (1u << min(RADIX_BITS, end_bit - shift)) - 1u;
    keys += (size_t)seg * n_max;
    const uint32_t n = n_dev ? min(n_dev[seg], n_max) : n_max;      // one device-side count per segment
    s_tot[tid] = 0;
    uint4* mine4 = reinterpret_cast<uint4*>(s_cnt + warp * (RADIX * 32));
    for (int i = lane; i < RADIX * 32 / 16; i += 32) mine4[i] = make_uint4(0u, 0u, 0u, 0u);
    __syncthreads();
    uint8_t* mine = s_cnt + warp * (RADIX * 32) + lane;

    auto flush = [&]() {    // warp-local: bins k*32 + lane, 32 byte counters each
        __syncwarp();
#pragma unroll
        for (int k = 0; k < RADIX / 32; ++k) {
            const int bin = k * 32 + lane;
            uint4* p = reinterpret_cast<uint4*>(s_cnt + warp * (RADIX * 32) + bin * 32);
            const uint4 a = p[0], b = p[1];
            uint32_t c = 0;
            c = __dp4a(a.x, 0x01010101u, c); c = __dp4a(a.y, 0x01010101u, c);
            c = __dp4a(a.z, 0x01010101u, c); c = __dp4a(a.w, 0x01010101u, c);
            c = __dp4a(b.x, 0x01010101u, c); c = __dp4a(b.y, 0x01010101u, c);
            c = __dp4a(b.z, 0x01010101u, c); c = __dp4a(b.w, 0x01010101u, c);
            if (c) atomicAdd(&s_tot[bin], c);
            p[0] = make_uint4(0u, 0u, 0u, 0u);
            p[1] = make_uint4(0u, 0u, 0u, 0u);
        }
        __syncwarp();
    };

    const uint32_t rows = (n + 31u) / 32u;
    // 8 rows of keys per step, and the NEXT step's keys are already in flight while this step's byte counters are
    // updated (ncu on the round-1 loop: 3 CTAs / SM by shared memory, long-scoreboard stalls 4.6 per issue -- the loads'
    // latency was exposed)
    constexpr int UNROLL = 8;
    for (uint32_t r0 = blockIdx.x * HIST_ROWS_PER_BLOCK; r0 < rows; r0 += gridDim.x * HIST_ROWS_PER_BLOCK) {
        const uint32_t r1 = min(r0 + HIST_ROWS_PER_BLOCK, rows);
        int since_flush = 0;
        KeyT nxt[UNROLL];
        auto load = [&](uint32_t r) {
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint32_t row = r + 8 * u;
                const uint32_t e = row * 32u + lane;
                // a key outside the range reads as the sentinel and is skipped below (rows past r1 belong to other CTAs)
                nxt[u] = (row < r1 && e < n) ? keys[e] : ~(KeyT)0;
            }
        };
        load(r0 + warp);
        for (uint32_t r = r0 + warp; r < r1; r += 8 * UNROLL) {
            KeyT key[UNROLL];
            bool ok[UNROLL];
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                const uint32_t row = r + 8 * u;
                key[u] = nxt[u];
                ok[u] = row < r1 && row * 32u + lane < n && !(drop_sentinel && key[u] == ~(KeyT)0);
            }
            if (r + 8 * UNROLL < r1) load(r + 8 * UNROLL);
#pragma unroll
            for (int u = 0; u < UNROLL; ++u) {
                if (ok[u]) {
                    const uint32_t d = (uint32_t)(key[u] >> shift) & mask;
                    mine[d * 32] = (uint8_t)(mine[d * 32] + 1);
                }
            }
            since_flush += UNROLL;
            if (since_flush + UNROLL > HIST_FLUSH_ROWS) {
                flush();
                since_flush = 0;
            }
        }
        flush();
    }
    __syncthreads();
    const uint32_t c = s_tot[tid];
    if (c) atomicAdd(&hist[((size_t)seg * passes_total + pass) * RADIX + tid], c);
}

// exclusive scan of each (segment, pass)'s 256 bins, in place (one block of 256 threads each)
// n_kept (nullable): n_kept[segment] <- number of keys counted in the segment's first-pass histogram (see radix_sort_pairs)
__global__ void __launch_bounds__(RADIX) radix_scan_hist_kernel(uint32_t* __restrict__ hist, uint32_t* __restrict__ n_kept,
                                                                int passes) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    __shared__ uint32_t warp_tot[RADIX / 32];
    uint32_t* h = hist + blockIdx.x * RADIX;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t v = h[tid];
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(FULL, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t base = 0;
    for (int w = 0; w < warp; ++w) base += warp_tot[w];
    h[tid] = base + inc - v;
    if (n_kept && tid == RADIX - 1 && (int)(blockIdx.x % passes) == 0) n_kept[blockIdx.x / passes] = base + inc;
}

template <typename KeyT, int ITEMS>
__global__ void __launch_bounds__(SORT_THREADS, 4)
onesweep_pass_kernel(const KeyT* __restrict__ keys_in, KeyT* __restrict__ keys_out,
                     const uint32_t* __restrict__ vals_in, uint32_t* __restrict__ vals_out,
                     uint32_t n_max, const uint32_t* __restrict__ n_dev, int shift, uint32_t digit_mask,
                     const uint32_t* __restrict__ hist_excl, size_t hist_seg_stride, uint32_t* __restrict__ lookback,
                     size_t lookback_seg_stride, uint32_t* __restrict__ tile_counter, bool drop_sentinel) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    constexpr int TILE = SORT_THREADS * ITEMS;
    constexpr int WARPS = SORT_THREADS / 32;
    __shared__ uint32_t s_warp_hist[WARPS][RADIX];
    __shared__ uint32_t s_block_excl[RADIX];
    __shared__ uint32_t s_glob_base[RADIX];
    __shared__ uint32_t s_warp_tot[WARPS];
    __shared__ uint32_t s_tile, s_kept;
    extern __shared__ __align__(16) unsigned char s_dyn[];
    KeyT* s_keys = reinterpret_cast<KeyT*>(s_dyn);
    uint32_t* s_vals = reinterpret_cast<uint32_t*>(s_dyn + sizeof(KeyT) * TILE);

    const int seg = blockIdx.y;
    keys_in += (size_t)seg * n_max;
    keys_out += (size_t)seg * n_max;
    vals_in += (size_t)seg * n_max;
    vals_out += (size_t)seg * n_max;
    hist_excl += (size_t)seg * hist_seg_stride;
    lookback += (size_t)seg * lookback_seg_stride;
    tile_counter += seg;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) s_tile = atomicAdd(tile_counter, 1u);
#pragma unroll
    for (int w = 0; w < WARPS; ++w) s_warp_hist[w][tid] = 0;
    __syncthreads();
    const uint32_t tile = s_tile;
    const uint32_t n = n_dev ? min(n_dev[seg], n_max) : n_max;
    const uint64_t base64 = (uint64_t)tile * TILE;
    if (base64 >= n) return;                       // uniform: whole block leaves
    const uint32_t base = (uint32_t)base64;
    const uint32_t valid_count = min((uint32_t)TILE, n - base);

    // ---- load keys (warp-striped: element e = warp*32*ITEMS + i*32 + lane); values are loaded only after the keys
    //      have left the registers (two-phase scatter), which keeps the kernel at <= 64 registers = 4 CTAs / SM ----
    KeyT key[ITEMS];
    uint32_t rank[ITEMS];
    const uint32_t e0 = warp * 32 * ITEMS + lane;
    // `live` bit i: item i takes part.  Sentinel keys (~0) are dropped by the pass that is told to (the first one): they
    // get no rank and are not written, so the tile's output is shorter than its input and the later passes run on the
    // kept keys only.
    uint32_t live = 0;
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const uint32_t e = e0 + i * 32;
        const bool in = e < valid_count;
        key[i] = in ? keys_in[base + e] : (KeyT)0;
        if (in && !(drop_sentinel && key[i] == ~(KeyT)0)) live |= 1u << i;
    }

    // ---- stable in-warp ranking, one item row at a time ----
    const unsigned lt_mask = (1u << lane) - 1u;
    if (digit_mask < 16u) {                        // a narrow (<= 4 bit) last digit needs half the ballots
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const bool valid = (live >> i) & 1u;
            const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
            uint32_t peers;
            const uint32_t prev = warp_hist_add<4>(s_warp_hist[warp], d, valid, lane, peers);
            rank[i] = prev + __popc(peers & lt_mask);
        }
    } else {
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const bool valid = (live >> i) & 1u;
            const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
            uint32_t peers;
            const uint32_t prev = warp_hist_add(s_warp_hist[warp], d, valid, lane, peers);
            rank[i] = prev + __popc(peers & lt_mask);
        }
    }
    __syncthreads();

    // ---- per digit (thread == digit): warp-exclusive offsets, tile count, publish, look back ----
    {
        const int d = tid;
        uint32_t run = 0;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            const uint32_t c = s_warp_hist[w][d];
            s_warp_hist[w][d] = run;
            run += c;
        }
        const uint32_t count = run;
        uint32_t* lb = lookback + (size_t)tile * RADIX;
        st_volatile(&lb[d], count | (tile == 0 ? FLAG_INC : FLAG_AGG));

        // exclusive scan of `count` across the 256 digits
        uint32_t inc = count;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(FULL, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_warp_tot[warp] = inc;
        __syncthreads();
        uint32_t wbase = 0;
        for (int w = 0; w < warp; ++w) wbase += s_warp_tot[w];
        const uint32_t excl = wbase + inc - count;
        s_block_excl[d] = excl;
        if (d == RADIX - 1) s_kept = excl + count;       // keys of this tile that take part (all but dropped sentinels)

        uint32_t prefix = 0;
        if (tile > 0) {
            int64_t t = (int64_t)tile - 1;
            while (true) {
                const uint32_t v = ld_volatile(lookback + (size_t)t * RADIX + d);
                if ((v >> 30) == 0u) continue;           // predecessor not published yet
                prefix += v & VALUE_MASK;
                if (v & FLAG_INC) break;
                --t;                                      // tile 0 always publishes FLAG_INC, so t stays >= 0
            }
            st_volatile(&lb[d], ((prefix + count) & VALUE_MASK) | FLAG_INC);
        }
        s_glob_base[d] = hist_excl[d] + prefix - excl;   // modulo-2^32 arithmetic
    }
    __syncthreads();

    // ---- phase 1: keys through shared memory so that global writes are runs of equal digits ----
#pragma unroll
    for (int i = 0; i < ITEMS; ++i) {
        const uint32_t d = (uint32_t)(key[i] >> shift) & digit_mask;
        rank[i] += s_block_excl[d] + s_warp_hist[warp][d];      // rank[] becomes the tile-local sorted position
        if ((live >> i) & 1u) s_keys[rank[i]] = key[i];
    }
    __syncthreads();
    const uint32_t kept = s_kept;
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
        const uint32_t sidx = tid + k * SORT_THREADS;
        if (sidx < kept) {
            const KeyT kk = s_keys[sidx];
            const uint32_t d = (uint32_t)(kk >> shift) & digit_mask;
            keys_out[s_glob_base[d] + sidx] = kk;
        }
    }
    // ---- phase 2: values follow the same permutation (digits are re-read from the staged keys) ----
    {
        uint32_t val[ITEMS];
#pragma unroll
        for (int i = 0; i < ITEMS; ++i) {
            const uint32_t e = e0 + i * 32;
            val[i] = e < valid_count ? vals_in[base + e] : 0u;
        }
#pragma unroll
        for (int i = 0; i < ITEMS; ++i)
            if ((live >> i) & 1u) s_vals[rank[i]] = val[i];
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < ITEMS; ++k) {
        const uint32_t sidx = tid + k * SORT_THREADS;
        if (sidx < kept) {
            const uint32_t d = (uint32_t)(s_keys[sidx] >> shift) & digit_mask;
            vals_out[s_glob_base[d] + sidx] = s_vals[sidx];
        }
    }
}

}  // namespace

template <typename KeyT>
int radix_sort_pairs(KeyT* keys0, KeyT* keys1, uint32_t* vals0, uint32_t* vals1, int64_t n_max,
                     const uint32_t* n_dev, int begin_bit, int end_bit, uint32_t* hist, uint32_t* lookback,
                     uint32_t* tile_counters, cudaStream_t s, int segments, bool hist_ready, bool scratch_zeroed,
                     uint32_t* n_kept_dev) {
    if (n_max <= 0 || end_bit <= begin_bit || segments <= 0) return GSR_OK;
    if (n_kept_dev && n_dev) return GSR_ERR_INVALID_ARGUMENT;      // (hist_ready + n_kept_dev: the given histograms must
                                                                    //  count the non-sentinel keys only)
    if (n_max >= (int64_t)VALUE_MASK) return GSR_ERR_UNSUPPORTED;
    if (segments > 65535) return GSR_ERR_UNSUPPORTED;
    const int passes = (end_bit - begin_bit + RADIX_BITS - 1) / RADIX_BITS;
    if (passes > MAX_PASSES) return GSR_ERR_INVALID_ARGUMENT;
    constexpr int TILE_BIG = SORT_THREADS * SortCfg<KeyT>::ITEMS, TILE_SMALL = SORT_THREADS * SortCfg<KeyT>::ITEMS_SMALL;
    const bool small = ((n_max + TILE_BIG - 1) / TILE_BIG) * segments <= SORT_SMALL_GRID;     // see SortCfg
    const int TILE = small ? TILE_SMALL : TILE_BIG;
    const int tiles = (int)((n_max + TILE - 1) / TILE);
    const size_t hist_seg = (size_t)passes * RADIX;
    const size_t lb_seg = (size_t)tiles * RADIX;
    if (!scratch_zeroed) {
        if (!hist_ready) GSR_CUDA_TRY(cudaMemsetAsync(hist, 0, (size_t)segments * hist_seg * 4, s));
        GSR_CUDA_TRY(cudaMemsetAsync(lookback, 0, (size_t)passes * segments * lb_seg * 4, s));
        GSR_CUDA_TRY(cudaMemsetAsync(tile_counters, 0, (size_t)passes * segments * 4, s));
    }
    if (!hist_ready) {
        const int64_t rows = (n_max + 31) / 32;
        const int64_t hb = (rows + HIST_ROWS_PER_BLOCK - 1) / HIST_ROWS_PER_BLOCK;
        const int hist_blocks = (int)(hb < 4096 ? hb : 4096);
        constexpr int hist_smem = 8 * RADIX * 32;
        static bool attr_set = false;        // per instantiation (KeyT)
        if (!attr_set) {
            GSR_CUDA_TRY(cudaFuncSetAttribute(radix_histogram_kernel<KeyT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                              hist_smem));
            attr_set = true;
        }
        radix_histogram_kernel<KeyT><<<dim3(hist_blocks, segments, passes), 256, hist_smem, s>>>(
            keys0, (uint32_t)n_max, n_dev, begin_bit, end_bit, passes, hist, n_kept_dev != nullptr);
        GSR_LAUNCH_CHECK();
    }
    launch_kernel(radix_scan_hist_kernel, passes * segments, RADIX, 0, s, hist, n_kept_dev, passes);
    GSR_LAUNCH_CHECK();
    const size_t dyn = (size_t)TILE * (sizeof(KeyT) + 4);
    for (int p = 0; p < passes; ++p) {
        const int shift = begin_bit + p * RADIX_BITS;
        const int bits = (end_bit - shift) < RADIX_BITS ? (end_bit - shift) : RADIX_BITS;
        const KeyT* kin = (p & 1) ? keys1 : keys0;
        KeyT* kout = (p & 1) ? keys0 : keys1;
        const uint32_t* vin = (p & 1) ? vals1 : vals0;
        uint32_t* vout = (p & 1) ? vals0 : vals1;
        // with sentinel dropping the first pass reads all n_max keys and writes the kept ones; the others move those only
        const uint32_t* n_pass = (n_kept_dev && p > 0) ? n_kept_dev : n_dev;
        if (small)
            launch_kernel(onesweep_pass_kernel<KeyT, SortCfg<KeyT>::ITEMS_SMALL>, dim3(tiles, segments), SORT_THREADS, dyn, s,
                kin, kout, vin, vout, (uint32_t)n_max, n_pass, shift, (1u << bits) - 1u, hist + p * RADIX, hist_seg,
                lookback + (size_t)p * segments * lb_seg, lb_seg, tile_counters + (size_t)p * segments,
                n_kept_dev != nullptr && p == 0);
        else
            launch_kernel(onesweep_pass_kernel<KeyT, SortCfg<KeyT>::ITEMS>, dim3(tiles, segments), SORT_THREADS, dyn, s,
                kin, kout, vin, vout, (uint32_t)n_max, n_pass, shift, (1u << bits) - 1u, hist + p * RADIX, hist_seg,
                lookback + (size_t)p * segments * lb_seg, lb_seg, tile_counters + (size_t)p * segments,
                n_kept_dev != nullptr && p == 0);
        GSR_LAUNCH_CHECK();
    }
    return GSR_OK;
}

template int radix_sort_pairs<uint32_t>(uint32_t*, uint32_t*, uint32_t*, uint32_t*, int64_t, const uint32_t*, int, int,
                                        uint32_t*, uint32_t*, uint32_t*, cudaStream_t, int, bool, bool, uint32_t*);
template int radix_sort_pairs<uint64_t>(uint64_t*, uint64_t*, uint32_t*, uint32_t*, int64_t, const uint32_t*, int, int,
                                        uint32_t*, uint32_t*, uint32_t*, cudaStream_t, int, bool, bool, uint32_t*);

}  // namespace gsr
