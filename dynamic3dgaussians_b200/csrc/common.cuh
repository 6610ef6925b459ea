// Shared declarations for the sm_100a rasterizer kernels.  Internal header (the public C ABI is include/gsr_b200.h).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stddef.h>

#include "../../include/gsr_b200.h"

namespace gsr {

constexpr int TILE_X = 16;            // R/cuda_rasterizer/config.h:18-19
constexpr int TILE_Y = 16;
constexpr int TILE_PIX = TILE_X * TILE_Y;
constexpr int GRAD_REC = 16;          // upper bound on the floats per (view, Gaussian) gradient record
// floats per (view, Gaussian) gradient record written by the blend backward: 7 geometry / opacity / depth sums + one per
// blended channel, rounded up to whole float4s (12 for colour only, 16 with the fused extra colours -- the spare slots
// 13, 14 of the latter take the colour-only screen-space moments when those are asked for)
constexpr int grad_rec_stride(int channels) { return (7 + channels + 3) / 4 * 4; }
constexpr int NUM_COUNTERS = 64;

// gradient-record slots
enum : int { GR_MX = 0, GR_MY = 1, GR_CA = 2, GR_CB = 3, GR_CC = 4, GR_OP = 5, GR_DEPTH = 6, GR_COL = 7 };
// counter slots
// CNT_PREFILTER: a Gaussian was culled although the caller promised a prefiltered set (auxiliary.h:160-167 traps there)
enum : int { CNT_RENDERED = 0, CNT_OVERFLOW = 1, CNT_PREFILTER = 2, CNT_ST_RENDERED = 3, CNT_NCHUNKS = 4, CNT_INST_TILE = 24 };

// Device-side camera: GsrCamera plus the per-view constants the reference derives on the host
// (focal lengths, rasterizer_impl.cu:227-228) or in every thread (clamp limits, forward.cu:87-99).
struct CamDev {
    float view[16];
    float proj[16];
    float campos[3];
    float focal_x, focal_y;
    float tanfovx, tanfovy;
    float lim_x_pos, lim_x_neg, lim_y_pos, lim_y_neg;
    float c_x, c_y;
    float scale_modifier;
    float bg[8];
    int sh_degree;
    int prefiltered;
    int width, height;
    int row_begin, row_end;     // band of tile rows binned for this view (whole grid by default)
    int pad[1];
};

struct CamPack {            // cameras travel to the device as a kernel argument (<= 4 KB)
    static constexpr int N = 16;
    int n;
    CamDev cam[N];
};
static_assert(sizeof(CamPack) <= 4000, "kernel argument space");

// ------------------------------------------------------------------------------------------------------------
// radix sort geometry
// ------------------------------------------------------------------------------------------------------------
constexpr int SORT_THREADS = 256;
constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
template <typename KeyT> struct SortCfg;
#ifndef GSR_SORT_ITEMS32
#define GSR_SORT_ITEMS32 16
#endif
// keys per thread: ITEMS for sorts that fill the machine, ITEMS_SMALL (half the tile, twice the CTAs) for sorts whose
// big-tile grid would leave SMs idle -- one training view is 300k keys = 74 tiles of 4096 on 148 SMs
template <> struct SortCfg<uint32_t> { static constexpr int ITEMS = GSR_SORT_ITEMS32, ITEMS_SMALL = GSR_SORT_ITEMS32 / 2; };
template <> struct SortCfg<uint64_t> { static constexpr int ITEMS = 12, ITEMS_SMALL = 6; };
constexpr int SORT_SMALL_GRID = 148;     // use the small tile when the big-tile grid has at most this many CTAs
template <typename KeyT> constexpr int sort_tile() { return SORT_THREADS * SortCfg<KeyT>::ITEMS_SMALL; }   // sizing: the smaller
constexpr int MAX_PASSES = 8;

inline int sort_num_tiles(int64_t n, int key_bytes) {
    const int tile = key_bytes == 8 ? sort_tile<uint64_t>() : sort_tile<uint32_t>();
    return (int)((n + tile - 1) / tile);
}

constexpr int SCAN_BLOCK = 256;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_BLOCK * SCAN_ITEMS;

inline int ceil_log2(uint32_t n) {  // smallest b with (1<<b) >= n, at least 1
    int b = 1;
    while ((1ull << b) < n) ++b;
    return b;
}

struct Layout : GsrLayout {
    size_t geom_bytes, binning_bytes, image_bytes;
    // every array that must be zero when a forward starts sits at the head of its workspace: one memset per workspace
    size_t geom_zero_bytes, binning_zero_bytes;
    // image workspace, internal: per-tile walked depth (the forward's largest n_contrib) and the backward's launch order
    size_t tile_cost, tile_order_bwd;
    // binning workspace, internal (zeroed): instances per supertile key, the emit kernel's look-back words and ticket
    size_t key_hist, emit_lookback, emit_ticket;
    // geometry workspace, internal: per view, the number of Gaussians the depth sort kept (those that touch a tile)
    size_t vis_count;
    int view_bits, st_bits;       // instance key = supertile << view_bits | view
    uint32_t num_keys;            // supertiles per view << view_bits
};

inline size_t align_up(size_t x, size_t a = 256) { return (x + a - 1) / a * a; }

// hierarchical binning geometry: a supertile is ST x ST tiles; its sorted instance list is cut into chunks of CHUNK
constexpr int ST = 4;
constexpr int ST_TILES = ST * ST;
constexpr int CHUNK = 256;

inline void compute_layout(int P, int V, int W, int H, int channels, int sh_coeffs, int64_t capacity, Layout& L) {
    const size_t VP = (size_t)V * (size_t)P;
    const int tx = (W + TILE_X - 1) / TILE_X, ty = (H + TILE_Y - 1) / TILE_Y;
    const size_t VT = (size_t)V * tx * ty;
    L.tiles_x = tx;
    L.tiles_y = ty;
    L.stiles_x = (tx + ST - 1) / ST;
    L.stiles_y = (ty + ST - 1) / ST;
    // instance key = supertile << view_bits | view.  The instances are emitted view by view in depth order, so a STABLE
    // sort on the supertile bits alone leaves them grouped by key with every group in depth order: one radix pass for up
    // to 256 supertiles per view (a 1024x1024 image), two beyond -- the view bits are never sorted on.
    L.view_bits = ceil_log2((uint32_t)(V > 0 ? V : 1));
    L.st_bits = ceil_log2((uint32_t)(L.stiles_x * L.stiles_y));
    L.tile_bits = L.view_bits + L.st_bits;    // live key bits (the upper half of the 32-bit key carries the tile hit mask)
    const size_t VS = (size_t)(L.stiles_x * L.stiles_y) << L.view_bits;      // key slots (those of views >= V stay empty)
    L.num_keys = (uint32_t)VS;
    L.depth_passes = 32 / RADIX_BITS;         // per-view segmented sort of the 32 depth bits
    L.tile_passes = (L.st_bits + RADIX_BITS - 1) / RADIX_BITS;
    L.sorted_keys_in = (L.tile_passes & 1);   // which ping-pong buffer holds the result of the instance sort
    L.sorted_vals_in = (L.depth_passes & 1);  // ... and of the depth sort
    size_t o = 0;
    auto take = [&](size_t bytes) { size_t r = o; o = align_up(o + bytes); return r; };
    // geom
    L.counters = take(NUM_COUNTERS * sizeof(uint32_t));   // always at offset 0 (gsr_read_counters relies on it)
    L.dhist = take((size_t)V * MAX_PASSES * RADIX * 4);
    L.dlookback = take((size_t)MAX_PASSES * V * (sort_num_tiles((int64_t)P, 4) + 1) * RADIX * 4);
    L.dcounters = take((size_t)MAX_PASSES * V * 4);
    L.geom_zero_bytes = o;                // counters + the depth sort's histograms / look-back words / tile tickets
    L.vis_count = take((size_t)V * 4);    // Gaussians per view that touch a tile: what the depth sort keeps (written by it)
    L.cams = take((size_t)V * sizeof(CamDev));
    L.radii = take(VP * 4);
    L.rects = take(VP * 8);               // ushort4 tile rectangle (min.x, min.y, max.x, max.y), all zero when culled
    L.geomrec = take(VP * 32);            // (mean.x, mean.y, conic.a, conic.b | conic.c, opacity, depth, skip threshold)
    L.cov3D = take((size_t)P * 32);       // per Gaussian, view-independent: (cov3D[0..3] | cov3D[4..5], opacity, skip threshold)
    L.rgb = take(sh_coeffs > 0 ? VP * 16 : 0);     // float4 per (view, Gaussian): SH colours
    L.col4 = take((size_t)P * 16);                 // float4-packed colors_precomp / extra colours (per Gaussian)
    L.ext4 = take(channels > 3 ? (size_t)P * 16 : 0);
    L.clamped = take(sh_coeffs > 0 ? VP * 3 : 0);
    L.dkeys0 = take(VP * 4);
    L.dkeys1 = take(VP * 4);
    L.dvals0 = take(VP * 4);
    L.dvals1 = take(VP * 4);
    L.block_sums = take((VP / 1024 + 3) * 4);      // per block of 1024 depth-sorted slots (binning.cu EMIT_TILE)
    L.grad_rec = take(VP * grad_rec_stride(channels) * 4);
    L.geom_bytes = o;
    // binning
    o = 0;
    const size_t cap = (size_t)(capacity > 0 ? capacity : 0);
    const size_t max_chunks = cap / CHUNK + VS + 1;
    L.max_chunks = (int32_t)max_chunks;
    L.ihist = take(((size_t)MAX_PASSES * RADIX + MAX_PASSES) * 4);      // + the instance sort's tile tickets
    L.ilookback = take((size_t)L.tile_passes * (sort_num_tiles((int64_t)cap, 4) + 1) * RADIX * 4);
    L.key_hist = take(VS * 4);
    L.emit_lookback = take((VP / 1024 + 2) * 4);       // one word per block of EMIT_TILE (= 1024) depth-sorted slots
    L.emit_ticket = take(4);
    L.binning_zero_bytes = o;             // instance sort histograms / look-back words, emit look-back, key counts
    L.st_ranges = take(VS * 8);
    L.ikeys0 = take(cap * 4);             // supertile instances: at most one per (Gaussian, tile) instance
    L.ikeys1 = take(cap * 4);
    L.ivals0 = take(cap * 4);
    L.ivals1 = take(cap * 4);
    L.imasks = take(0);                   // (ABI v3 kept the tile hit masks here; they ride in the keys' upper half now)
    L.chunk_base = take((VS + 1) * 4);
    L.chunk_key = take(max_chunks * 4);
    L.chunk_counts = take(max_chunks * ST_TILES * 4);
    L.recA = take(cap * 16);
    L.recB = take(cap * 16);
    L.recC = take(cap * 16);
    L.recD = take(channels > 3 ? cap * 16 : 0);
    L.binning_bytes = o;
    // image
    o = 0;
    L.ranges = take(VT * 8);
    L.tile_start = take((VT + 1) * 4);
    L.tile_order = take(VT * 4);          // (view, tile) indices, longest list first: the blend kernels' launch order
    L.n_contrib = take((size_t)V * W * H * 4);
    L.final_T = take((size_t)V * W * H * 4);
    L.tile_cost = take(VT * 4);
    L.tile_order_bwd = take(VT * 4);
    L.image_bytes = o;
}

template <typename T>
__host__ __device__ inline T* at(void* base, size_t off) { return reinterpret_cast<T*>(reinterpret_cast<char*>(base) + off); }
template <typename T>
__host__ __device__ inline const T* at(const void* base, size_t off) {
    return reinterpret_cast<const T*>(reinterpret_cast<const char*>(base) + off);
}

// Process-wide switches of the library (gsr_set_option); nothing on the launch path reads the environment.
struct Options {
    int blend_ilp = 0;        // forward blend loop: 0 = by launch size (blend_ilp()), 1 / 2 = force that variant
    int bwd_cost_order = 1;   // small backward launches ordered by the depth the forward walked (tile_order_by_cost_kernel)
    int nvtx = 0;             // an NVTX range around every stage (profilers)
    int bwd_ilp = 0;          // backward blend loop: 0 = by launch size (two entries per trip for small launches), 1 / 2 = forced
    int pdl = 0;              // programmatic dependent launch between the step's kernels (launch_kernel below); measured: off
};
Options& options();

// Programmatic dependent launch (sm_90+): a kernel launched through launch_kernel() may be scheduled while its
// predecessor in the stream is still running -- its CTAs take SM slots as the predecessor's free them and park in
// pdl_wait() (griddepcontrol.wait: until the predecessor grid has completed and its writes are visible), so the launch
// latency and the ramp-up of the grid no longer sit between two kernels.  A step is ~25 short dependent kernels; at a
// 1/8 share of a timestep or one training view those gaps are a tenth of the step.  Rules: a kernel launched this way
// touches NO global memory before pdl_wait() (reads: not yet written; writes: the predecessor may still read the old
// values), and every such kernel calls pdl_trigger() (griddepcontrol.launch_dependents) at its top so that ITS successor
// may be scheduled as soon as all of its CTAs have started.  Launched without the attribute (option "pdl" = 0, or after
// a memset / event wait) both instructions are no-ops.
// Measured at C2 on a B200 (profiles/gpu_check.sh, A/B of the same library): launched one by one from the host the stages
// shorten by 2-8 % (depth sort 0.2185 -> 0.2133 ms, emit 0.184 -> 0.175, tile offsets 0.096 -> 0.091), but the step
// replayed as ONE CUDA graph -- the path the bench and the training loop use -- gets slower (4.27 -> 4.40 ms; a 1/8 share
// 0.709 -> 0.730): graph kernel nodes already follow each other within ~1 us, and the parked CTAs of the successor take
// slots from the predecessor's last wave.  Hence off by default; gsr_set_option("pdl", 1) turns it on for eager callers.
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline void launch_kernel(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t s, Args&&... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = s;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = options().pdl ? 1 : 0;
    (void)cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);     // errors: GSR_LAUNCH_CHECK() after the call
}
#endif

// optional per-stage device timing (bench.py's roofline numbers): events are recorded on the launching stream
enum Stage : int { ST_PREPROCESS = 0, ST_DEPTH_SORT, ST_TILE_SCAN, ST_EMIT, ST_TILE_SORT, ST_RANGES, ST_GATHER,
                   ST_BLEND_FWD, ST_GRAD_ZERO, ST_BLEND_BWD, ST_PREPROCESS_BWD, ST_COUNT };
void stage_begin(int stage, cudaStream_t s);
void stage_end(int stage, cudaStream_t s);
// debug mode of one call (GsrCamera.debug / GSR_FLAG_DEBUG, the reference's `debug` flag: CHECK_CUDA after every launch,
// auxiliary.h:172-179): stage_end() synchronises the stream and checks for errors; the first failing stage is recorded
// with its name (gsr_last_cuda_error) and the entry point returns GSR_ERR_CUDA.
struct DebugSyncScope {
    explicit DebugSyncScope(bool on);
    ~DebugSyncScope();
    int status() const;       // GSR_OK or GSR_ERR_CUDA
};
struct StageScope {
    int st; cudaStream_t s;
    StageScope(int stage, cudaStream_t stream) : st(stage), s(stream) { stage_begin(st, s); }
    ~StageScope() { stage_end(st, s); }
};

// error plumbing (no exceptions across the C ABI)
void set_cuda_error(cudaError_t e, const char* where);
#define GSR_CUDA_TRY(expr)                                              \
    do {                                                                \
        cudaError_t _e = (expr);                                        \
        if (_e != cudaSuccess) { ::gsr::set_cuda_error(_e, #expr); return GSR_ERR_CUDA; } \
    } while (0)
#define GSR_LAUNCH_CHECK() GSR_CUDA_TRY(cudaGetLastError())

// ------------------------------------------------------------------------------------------------------------
// stage launchers (defined in the respective .cu files)
// ------------------------------------------------------------------------------------------------------------
int launch_preprocess(const CamDev* cams, int V, int P, int sh_coeffs,
                      const float* means3D, const float* opacities, const float* scales, const float* rotations,
                      const float* cov3D_precomp, const float* sh, float scale_modifier, void* geom_ws, const Layout& L,
                      int32_t* radii_out, cudaStream_t s);   // sets counters[CNT_PREFILTER] (see preprocess.cu)
int launch_peer_allreduce(int mode, int world, int rank, const void* const* bufs_host, const void* const* flags_host,
                          void* multicast_buf, int64_t n_floats, int blocks, cudaStream_t s);
int launch_mark_visible(const float* view, int P, const float* means3D, uint8_t* present, cudaStream_t s);

template <typename KeyT>
int radix_sort_pairs(KeyT* keys0, KeyT* keys1, uint32_t* vals0, uint32_t* vals1, int64_t n_max,
                     const uint32_t* n_dev, int begin_bit, int end_bit, uint32_t* hist, uint32_t* lookback,
                     uint32_t* tile_counters, cudaStream_t s, int segments = 1, bool hist_ready = false,
                     bool scratch_zeroed = false,    // scratch_zeroed: hist / lookback / tile_counters are already 0
                     uint32_t* n_kept_dev = nullptr);
// n_kept_dev (one word per segment, written): keys equal to ~0 are SENTINELS -- counted out by the histogram, dropped by
// the first pass, and the later passes only move the n_kept_dev[segment] keys that remain (the rasterizer marks culled
// Gaussians this way: they cost one read instead of four read-write round trips).  Needs hist_ready = false, n_dev = null.

// fork / join of the library's own side stream (per host thread and device): side_fork makes the side stream wait for
// everything enqueued on `main` so far and hands it out; side_join makes `main` wait for everything enqueued on the side
// stream.  Works inside a stream capture (the side stream joins the capture).  The streams / events are created on first
// use -- outside any capture, since every graph user warms up first.
int side_fork(cudaStream_t main, cudaStream_t* side);
int side_join(cudaStream_t main);
int launch_binning(int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                   const float* colors_precomp, const float* extra_colors,
                   void* geom_ws, void* binning_ws, void* image_ws, const Layout& L, cudaStream_t s);
// Forward blend: BLEND_ILP list entries per loop trip -- the alphas (and the colour loads) of the trip's entries depend on
// nothing the loop carries and are computed side by side, only the short T chain is walked entry by entry, with a
// branch-free blend.  Round 1 used it for launches of a few views only (the issue-bound 27-view launch was 3 % faster
// with the one-entry loop and its early `continue`s); with the exact ellipse / patch test of round 2 the work lists hold
// almost only entries that blend, the early exits have little left to skip, and two entries per trip win at every size
// (C2 on a B200: 1/8 share 0.182 -> 0.162 ms, 1/2 share 0.569 -> 0.544, all 27 views 1.117 -> 1.067; C5 2.33 -> 2.20).
// More entries per trip or fewer resident CTAs change nothing (measured), and the same change makes the backward slower
// (more instructions on the same scheduler), so the backward has one variant.  gsr_set_option("blend_ilp", 1 | 2) forces
// a variant (tuning and tests; the variants produce identical bits).
constexpr int BLEND_ILP_MAX_BLOCKS = 1 << 30;
constexpr int BWD_COST_ORDER_MAX_BLOCKS = 148 * 8 * 4;
#ifndef GSR_BWD_ILP_MAX_BLOCKS
#define GSR_BWD_ILP_MAX_BLOCKS (148 * 8)
#endif
// Backward launches up to this size -- at most one wave of CTAs, i.e. a training view -- take two list entries per trip
// (blend_backward_kernel<.., .., 2>).  Measured on a B200 at C2: one 640x360 view per optimizer step 1146 -> 1241 it/s;
// larger launches are issue-bound and lose (1/8 share of a timestep 0.2385 -> 0.250 ms, all 27 views 1.63 -> 1.77 ms).
constexpr int BWD_ILP_MAX_BLOCKS = GSR_BWD_ILP_MAX_BLOCKS;   // backward launches up to this size order their CTAs by walked depth
#ifndef GSR_BLEND_ILP_N
#define GSR_BLEND_ILP_N 2
#endif
constexpr int BLEND_ILP_SMALL = GSR_BLEND_ILP_N;
int blend_ilp(int blocks);
int set_blend_forward_trace(unsigned long long* p);     // diagnostics (gsr_debug_blend_trace)
int set_blend_backward_trace(unsigned long long* p);

int launch_blend_forward(int V, int W, int H, int channels, const void* geom_ws, const void* binning_ws,
                         void* image_ws, const Layout& L, float* out_color, float* out_extra, float* out_depth,
                         float* out_alpha, cudaStream_t s);
int launch_blend_backward(int V, int P, int W, int H, int channels, void* geom_ws, const void* binning_ws,
                          void* image_ws, const Layout& L, const float* dL_dcolor, const float* dL_dextra,
                          const float* dL_ddepth, const float* dL_dalpha, bool color_only_means2D, cudaStream_t s);
size_t image_loss_state_bytes(int64_t n_img, int W, int H);
int launch_image_loss_forward(int64_t n_img, int W, int H, const float* img, const float* tgt, float w_l1, float w_ssim,
                              void* state, float* loss_out, cudaStream_t s);
int launch_image_loss_backward(int64_t n_img, int W, int H, const float* img, const float* tgt, float w_l1,
                               float w_ssim, const void* state, const float* grad_out, float* dL_dimg, cudaStream_t s);
// train_glue.cu: the small per-iteration ops around the rasterizer in the training loop
int launch_activations_forward(int64_t P, const float* unnorm_rot, const float* logit_op, const float* log_scales,
                               float* rot, float* op, float* scales, cudaStream_t s);
int launch_activations_backward(int64_t P, const float* unnorm_rot, const float* op, const float* scales,
                                const float* d_rot, const float* d_op, const float* d_scales, float* d_unnorm_rot,
                                float* d_logit, float* d_log_scales, cudaStream_t s);
int launch_cam_affine_forward(int64_t HW, const float* im, const float* m3, const float* c3, float* out, cudaStream_t s);
int launch_cam_affine_backward(int64_t HW, const float* im, const float* m3, const float* g, float* d_im, float* d_m3,
                               float* d_c3, cudaStream_t s);
int launch_l1_u8_forward(int64_t n_views, int W, int H, const float* img, const uint8_t* tgt, const float* row_weight,
                         float w_const, float* loss, cudaStream_t s);
int launch_l1_u8_backward(int64_t n_views, int W, int H, const float* img, const uint8_t* tgt, const float* row_weight,
                          float w_const, const float* g, float* d_img, cudaStream_t s);
int launch_densify_stats(int64_t P, const int32_t* radii, const uint8_t* seen_in, const float* grad2d,
                         float* max_radius, float* accum, float* denom, uint8_t* seen_out, cudaStream_t s);
int launch_adam_step(int n_tensors, const GsrAdamTensor* tensors, double beta1, double beta2, double eps, void* scratch,
                     cudaStream_t s);
int launch_store_cams(const CamPack& pack, CamDev* dst, cudaStream_t s);
int launch_feature_forward(int V, int P, int W, int H, const void* geom_ws, const void* binning_ws,
                           const void* image_ws, const Layout& L, const float* features, float* out_features,
                           cudaStream_t s);
int launch_feature_backward(int V, int P, int W, int H, int channels, void* geom_ws, const void* binning_ws, const void* image_ws,
                            const Layout& L, const float* features, const float* dL_dfeat_img, float* dL_dfeatures,
                            cudaStream_t s);
int launch_preprocess_backward(int V, int P, int W, int H, int channels, int sh_coeffs,
                               const float* means3D, const float* opacities, const float* scales,
                               const float* rotations, const float* cov3D_precomp, const float* sh,
                               const int32_t* radii, const void* geom_ws, const Layout& L, const float* label,
                               float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dmeans2D_color_only, float* dL_dcolors,
                               float* dL_dextra_colors, float* dL_dopacity, float* dL_dscales, float* dL_drotations,
                               float* dL_dcov3D, float* dL_dsh, cudaStream_t s);
int launch_export_view(int view, int V, int P, int W, int H, const void* geom_ws, const void* binning_ws,
                       const void* image_ws, const Layout& L, uint64_t* keys_out, uint32_t* list_out,
                       uint32_t* ranges_out, uint32_t* n_contrib_out, cudaStream_t s);

// ------------------------------------------------------------------------------------------------------------
// device helpers: mbarrier + 1-D TMA bulk copy (cp.async.bulk), sm_90+ PTX
// ------------------------------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_fence_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {}
}
// Orders this thread's (and, through the preceding block-scope synchronisation, the CTA's) generic-proxy accesses to
// shared memory before later async-proxy operations: a ring slot that warps have READ with ordinary loads is handed back
// to the TMA engine (cp.async.bulk WRITES it through the async proxy) only behind this fence (PTX memory model: accesses
// through different proxies need a proxy fence; an mbarrier hand-over would carry it implicitly, the ring's "last warp
// out" atomic counter does not).
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
// global -> shared bulk copy; bytes must be a multiple of 16, both addresses 16-byte aligned.
__device__ __forceinline__ void tma_load_1d(void* smem_dst, const void* gmem_src, uint32_t bytes, uint64_t* bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(smem_dst)),
        "l"(gmem_src), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}
#endif

}  // namespace gsr
