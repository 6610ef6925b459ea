/*
 * gsr_b200.h -- C ABI of the B200-native (sm_100a) differentiable Gaussian-splatting rasterizer and kNN.
 *
 * This is the drop-in boundary for the hot path of ImNotPrepared/Dynamic3DGaussians.  Paths below are relative to
 *   R/ = submodules_fsgs/diff-gaussian-rasterization-confidence/      K/ = submodules_fsgs/simple-knn/
 *
 * The reference's own boundary is C++ (std::function resize callbacks + pybind11), not a C ABI:
 *   CudaRasterizer::Rasterizer::forward      R/cuda_rasterizer/rasterizer.h:35-63   (impl rasterizer_impl.cu:198-346)
 *   CudaRasterizer::Rasterizer::backward     R/cuda_rasterizer/rasterizer.h:65-105  (impl rasterizer_impl.cu:350-467)
 *   CudaRasterizer::Rasterizer::markVisible  R/cuda_rasterizer/rasterizer.h:28-33   (impl rasterizer_impl.cu:141-153)
 *   RasterizeGaussiansCUDA / ...BackwardCUDA / markVisible (torch glue)  R/rasterize_points.cu:35-126,128-225,227-246
 *   SimpleKNN::knn / distCUDA2               K/simple_knn.cu:192-228, K/spatial.cu:15-27
 * The resize callbacks (R/rasterize_points.cu:27-33) are replaced by a query-then-run workspace protocol; the
 * caller owns every buffer and the library never allocates, frees, synchronises the device or throws.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless the name ends in _host;
 *   - all functions enqueue work on `stream` and return immediately (0 = ok, negative = GsrStatus);
 *   - V views are processed in one batch; all views of a batch share width/height;
 *   - per-view arrays are laid out [V][...]; images are planar [V][C][H][W] like the reference's [C,H,W];
 *   - matrices are the reference's transposed (column-major) 4x4 floats: element i is what the reference kernels
 *     call matrix[i] (R/cuda_rasterizer/auxiliary.h:58-77).
 */
#ifndef GSR_B200_H_INCLUDED
#define GSR_B200_H_INCLUDED

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef void* gsr_stream_t; /* a cudaStream_t */

typedef enum GsrStatus {
    GSR_OK = 0,
    GSR_ERR_INVALID_ARGUMENT = -1,
    GSR_ERR_WORKSPACE_TOO_SMALL = -2,
    GSR_ERR_CUDA = -3,           /* a CUDA runtime call failed; see gsr_last_cuda_error() */
    GSR_ERR_UNSUPPORTED = -4,
    GSR_ERR_CAPACITY = -5        /* instance capacity exceeded (reported by gsr_check_overflow) */
} GsrStatus;

/* One camera == one GaussianRasterizationSettings (R/diff_gaussian_rasterization/__init__.py:176-191). Host struct. */
typedef struct GsrCamera {
    int32_t image_width;
    int32_t image_height;
    float tanfovx;
    float tanfovy;
    float c_x;
    float c_y;
    float scale_modifier;
    int32_t sh_degree;
    int32_t prefiltered;
    float bg[8];          /* background per blended channel: [0..2] colour, [3..5] extra (segmentation) colours */
    float viewmatrix[16];
    float projmatrix[16];
    float campos[3];
    /* Optional band of TILE rows [tile_row_begin, tile_row_end) this call renders for the view (both 0 = the whole
     * image).  Gaussians are binned only into tiles of the band; pixels outside it get the background and take no part
     * in the backward.  Used to shard one view over several GPUs (dist.partition_units): the per-Gaussian gradients
     * of the bands of a view add up to the view's gradient. */
    int32_t tile_row_begin;
    int32_t tile_row_end;
    /* The reference's `debug` setting (R/diff_gaussian_rasterization/__init__.py:189; CHECK_CUDA after every launch,
     * R/cuda_rasterizer/auxiliary.h:172-179): gsr_forward_geometry synchronises `stream` after every stage and returns
     * GSR_ERR_CUDA naming the first failing stage in gsr_last_cuda_error().  Read from cams_host[0].  The other entry
     * points take the same switch as GSR_FLAG_DEBUG. */
    int32_t debug;
} GsrCamera;

/* `flags` of gsr_forward_render / gsr_backward */
#define GSR_FLAG_DEBUG 1   /* synchronise + check after every stage (see GsrCamera.debug); not inside a graph capture */

/* Byte sizes of the three caller-owned workspaces (the reference's geomBuffer / binningBuffer / imgBuffer). */
typedef struct GsrWorkspaceSizes {
    size_t geom_bytes;     /* per (view, Gaussian) state + depth-sort scratch      */
    size_t binning_bytes;  /* per (Gaussian, tile) instance state, for `capacity`  */
    size_t image_bytes;    /* per pixel / per tile state                            */
} GsrWorkspaceSizes;

/* Offsets (bytes) of the arrays inside the workspaces, for tests / tools that want to look at intermediate state. */
typedef struct GsrLayout {
    /* geom workspace (cov3D: 8 floats per Gaussian -- the 3D covariance, the opacity and the blend kernels' skip threshold,
     * i.e. everything per Gaussian that does not depend on the view) */
    size_t cams, counters, radii, rects, geomrec, cov3D, rgb, col4, ext4, clamped;
    size_t dkeys0, dkeys1, dvals0, dvals1, dhist, dlookback, dcounters, block_sums, grad_rec;
    /* binning workspace: supertile-instance sort arrays, chunk tables, per-instance records */
    size_t ikeys0, ikeys1, ivals0, ivals1, imasks, ihist, ilookback, st_ranges, chunk_base, chunk_key, chunk_counts;
    size_t recA, recB, recC, recD;
    /* image workspace */
    size_t ranges, tile_start, tile_order, n_contrib, final_T;
    /* derived integers: tiles per view, supertiles (4x4 tiles) per view, key bits / passes of the two sorts */
    int32_t tiles_x, tiles_y, stiles_x, stiles_y, tile_bits, depth_passes, tile_passes, sorted_keys_in, sorted_vals_in;
    int32_t max_chunks;
} GsrLayout;

/* Number of colour channels blended in the single pass: 3 (colour) or 6 (colour + fg/bg segmentation colours). */
#define GSR_MAX_CHANNELS 6

int gsr_abi_version(void);   /* 4 */
const char* gsr_strerror(int status);
const char* gsr_last_cuda_error(void);

/*
 * Process-wide switches.  The library never reads the environment; whoever embeds it sets these explicitly.
 *   "blend_ilp"       0 (default): forward blend loop variant picked by launch size; 1 / 2: force one / two list entries
 *                     per trip (identical bits; tuning and tests)
 *   "bwd_ilp"         0 (default): backward blend loop variant picked by launch size (two list entries per trip for small
 *                     launches); 1 / 2: force one / two entries per trip (identical bits; tuning and tests)
 *   "bwd_cost_order"  1 (default): small backward launches order their CTAs by the depth the forward walked each tile
 *   "nvtx"            1: an NVTX range around every stage (preprocess, depth_sort, ..., blend_backward), for profilers
 *   "pdl"             1: the step's kernels are chained by programmatic dependent launch (a kernel is scheduled while its
 *                     predecessor drains and waits in griddepcontrol.wait); helps host-launched steps by 1-2 %, measured
 *                     slower under CUDA-graph replay, hence 0 (plain stream order) by default
 * Unknown key or value out of range: GSR_ERR_INVALID_ARGUMENT.  Apart from these switches, the stage-timing events and the
 * blend trace pointers below, the library keeps no state: everything lives in the caller's workspaces.
 */
int gsr_set_option(const char* key, int value);
int gsr_get_option(const char* key, int* value);

/* Workspace protocol: query sizes for (P, V, W, H, channels, instance capacity). */
int gsr_workspace_sizes(int P, int V, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                        GsrWorkspaceSizes* out);
int gsr_workspace_layout(int P, int V, int W, int H, int channels, int sh_coeffs, int64_t capacity, GsrLayout* out);

/*
 * Forward, stage 1 of 2: per-Gaussian preprocess for all V views (projection, covariance, conic, radius, tile rect;
 * replaces FORWARD::preprocess, R/cuda_rasterizer/forward.cu:166-269), per-view depth ordering, and the instance count.
 *   num_rendered_dev[0]  <- total number of (Gaussian, tile) instances over all views (what the reference returns as
 *                           `num_rendered`, rasterizer_impl.cu:285-287, summed over views).
 * Exactly one of (scales+rotations | cov3D_precomp) and one of (colors_precomp | sh) must be non-null.
 * radii is [V][P] int32.
 */
int gsr_forward_geometry(const GsrCamera* cams_host, int V, int P, int channels, int sh_coeffs,
                         const float* means3D, const float* opacities, const float* scales, const float* rotations,
                         const float* cov3D_precomp, const float* sh,
                         void* geom_ws, size_t geom_bytes,
                         int32_t* radii, uint32_t* num_rendered_dev, gsr_stream_t stream);

/*
 * Forward, stage 2 of 2: supertile instance emission and sort, per-tile offsets, record scatter and the fused tile blend
 * (replaces duplicateWithKeys / SortPairs / identifyTileRanges / FORWARD::render, rasterizer_impl.cu:295-343).
 * `capacity` is the number of instances the binning workspace was sized for; if the true count exceeds it the
 * surplus instances are dropped and the overflow flag (gsr_check_overflow) is raised.
 * colors_precomp [P][3] (or null when SH colours were evaluated in stage 1); extra_colors [P][3] or null (channels==6).
 * Outputs: out_color [V][3][H][W], out_extra [V][3][H][W] (channels==6), out_depth [V][1][H][W], out_alpha [V][1][H][W].
 * Semantic features (the vendored variant's 32-channel map, R/cuda_rasterizer/config.h:17, forward.cu:381-383,405-406):
 * features [P][num_features] (nullable), num_features must be GSR_NUM_SEMANTIC_CHANNELS when features is given;
 * out_features [V][32][H][W].  Channels 0..2 are composited over bg[0..2], the others over 0.
 */
#define GSR_NUM_SEMANTIC_CHANNELS 32
int gsr_forward_render(int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                       const float* colors_precomp, const float* extra_colors,
                       const float* features, int num_features,
                       void* geom_ws, void* binning_ws, void* image_ws,
                       float* out_color, float* out_extra, float* out_depth, float* out_alpha, float* out_features,
                       int flags, gsr_stream_t stream);

/* Reads back (asynchronously, into host memory) the instance count and the two device-side flags.  Caller synchronises.
 *   [0] num_rendered   [1] overflow: the instance capacity was exceeded, surplus instances were dropped
 *   [2] prefilter violation: a camera had prefiltered = 1 and a Gaussian was culled all the same -- the reference prints
 *       "Point is filtered although prefiltered is set" and traps the kernel (R/cuda_rasterizer/auxiliary.h:160-167); here
 *       the Gaussian is skipped, the flag is raised and the binding turns it into an exception
 *   [3] supertile instances (internal) */
int gsr_read_counters(const void* geom_ws, uint32_t* counters_host /* [4] */, gsr_stream_t stream);

/*
 * Backward (replaces BACKWARD::render + BACKWARD::preprocess, R/cuda_rasterizer/backward.cu:432-652,144-429).
 * Image cotangents may be null (treated as zero).  Per-Gaussian gradients are SUMMED over the V views and written
 * (not accumulated) to the outputs; any output may be null.  dL_dmeans2D is [V][P][3] in the reference's NDC-scaled
 * units with z = 0 (backward.cu:520-521).  `label` ([P], nullable) is the per-Gaussian gradient mask the reference
 * applies in Python (R/diff_gaussian_rasterization/__init__.py:159-170); it is not applied to dL_dfeatures (:164).
 * features / dL_dfeatures_img [V][32][H][W] / dL_dfeatures [P][32]: the semantic-feature branch (backward.cu:594-611),
 * all three null when the forward had no features.
 * dL_dmeans2D_color_only [V][P][3] (nullable): dL_dmeans2D of the COLOUR render alone -- the colour, depth and alpha
 * cotangents without the extra (segmentation) channels that were blended in the same pass.  The reference renders colour
 * and segmentation with two calls and feeds densification from the colour call's means2D gradient only (REF/train.py:245,
 * REF/external.py:136-140); with channels == 3 it equals dL_dmeans2D.
 */
int gsr_backward(int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                 const float* means3D, const float* opacities, const float* scales, const float* rotations,
                 const float* cov3D_precomp, const float* sh, const float* colors_precomp, const float* extra_colors,
                 const float* features, int num_features, const int32_t* radii,
                 void* geom_ws, void* binning_ws, void* image_ws,
                 const float* dL_dcolor, const float* dL_dextra, const float* dL_ddepth, const float* dL_dalpha,
                 const float* dL_dfeatures_img, const float* label,
                 float* dL_dmeans3D, float* dL_dmeans2D, float* dL_dmeans2D_color_only, float* dL_dcolors,
                 float* dL_dextra_colors, float* dL_dopacity, float* dL_dscales, float* dL_drotations, float* dL_dcov3D,
                 float* dL_dsh, float* dL_dfeatures, int flags, gsr_stream_t stream);

/* present[i] = (view * p_i).z > 0   (checkFrustum, rasterizer_impl.cu:54-66).  present is [P] bytes. */
int gsr_mark_visible(const GsrCamera* cam_host, int P, const float* means3D, uint8_t* present, gsr_stream_t stream);

/*
 * Test / tool export of one view's binning state in the REFERENCE's format: point_list_keys (uint64:
 * tile << 32 | depth bits) and point_list (uint32 Gaussian index) in sorted order, and ranges [T][2].
 * keys_out / list_out must hold that view's instance count (ranges tell it).  Any output may be null.
 */
int gsr_export_view_binning(int view, int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                            const void* geom_ws, const void* binning_ws, const void* image_ws,
                            uint64_t* keys_out, uint32_t* list_out, uint32_t* ranges_out, uint32_t* n_contrib_out,
                            gsr_stream_t stream);

/* Stable LSD radix sort of n (key, value) pairs on key bits [begin_bit, end_bit) -- the hand-written onesweep used
 * for both sorts above, exposed for tests (replaces cub::DeviceRadixSort::SortPairs, rasterizer_impl.cu:309-314).
 * Results land in keys_out/vals_out.  tmp must hold gsr_sort_temp_bytes(n, key_bytes). key_bytes is 4 or 8. */
size_t gsr_sort_temp_bytes(int64_t n, int key_bytes);
int gsr_sort_pairs(int key_bytes, const void* keys_in, void* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                   int64_t n, int begin_bit, int end_bit, void* tmp, size_t tmp_bytes, gsr_stream_t stream);

/* The same sort with NO allocation anywhere on the call: the scratch gsr_sort_pairs takes from the caller lives in a
 * file-scope `static __device__` array inside the library (8 MiB, held by every GPU that loads the library until the
 * process exits), so the call can be made where cudaMalloc cannot (under CUDA graph capture, in a serving loop).
 * This entry point has NO counterpart in the reference.
 * Limit: n <= GSR_SORT_STATIC_MAX_PAIRS; a larger n returns GSR_ERR_WORKSPACE_TOO_SMALL with the reason in
 * gsr_last_cuda_error().  Results are those of gsr_sort_pairs, bit for bit.  Every call on a device shares the one
 * array: calls that overlap on the GPU (different streams, different threads) race -- serialise them.  The sort
 * clears the histogram / look-back / ticket words it needs at the start of each call; nothing else is read before
 * it is written. */
#define GSR_SORT_STATIC_MAX_PAIRS 262144
int64_t gsr_sort_pairs_static_max(void);   /* GSR_SORT_STATIC_MAX_PAIRS */
int gsr_sort_pairs_static(int key_bytes, const void* keys_in, void* keys_out, const uint32_t* vals_in,
                          uint32_t* vals_out, int64_t n, int begin_bit, int end_bit, gsr_stream_t stream);

/*
 * Sum of a float32 buffer over the ranks of one NVSwitch node, as ONE kernel over peer memory: the exchange step of the
 * view-sharded path (per-Gaussian gradient slab).  The reference is single-GPU (REF/train.py trains
 * on one device), so there is no reference interface behind this one; it takes the place of the ncclAllReduce call of
 * round 1.  bufs[p] / flags[p] (HOST arrays of `world` DEVICE pointers) are rank p's buffer and flag block as mapped into
 * THIS process (symmetric memory: every rank allocates the same sizes and the mappings are exchanged by the host side,
 * dynamic3dgaussians_b200/dist.py PeerSlab).  A flag block is GSR_PEER_FLAG_WORDS uint32, zeroed once before the first
 * call on any rank.  multicast_buf (nullable): the multicast (NVLS) mapping of the same buffers; when given, the
 * reduction happens in the switch (multimem.ld_reduce / multimem.st).  n_floats: a multiple of 4; buffers 16-byte aligned.
 * Every rank must make the call, with the same n_floats and blocks (<= GSR_PEER_MAX_CTAS; 0 = default), once per step and
 * in the same order; the sum lands in every rank's buffer, bit-identical on all ranks.  Stream-ordered, capturable
 * into a CUDA graph (the flag words are epoch counters kept on the device).  A peer that never arrives is given up on
 * after ~4 s: word GSR_PEER_ERROR_WORD of the local flag block is then 1.
 */
#define GSR_PEER_MAX_WORLD 8
#define GSR_PEER_MAX_CTAS 256
#define GSR_PEER_FLAG_WORDS 4608
#define GSR_PEER_ERROR_WORD (2 * GSR_PEER_MAX_CTAS * GSR_PEER_MAX_WORLD + GSR_PEER_MAX_CTAS)
#define GSR_PEER_ALLREDUCE 0       /* the sum lands in every rank's buffer */
#define GSR_PEER_REDUCE_SCATTER 1  /* rank r's buffer gets the sum of slice r only (slices: ceil(n_floats/4/world) float4s) */
#define GSR_PEER_ALLGATHER 2       /* slice r of rank r's buffer is copied into slice r of every peer's buffer */
int gsr_peer_allreduce(int mode, int world, int rank, const void* const* bufs, const void* const* flags,
                       void* multicast_buf, int64_t n_floats, int blocks, gsr_stream_t stream);

/*
 * k-nearest-neighbour search (replaces SimpleKNN::knn, K/simple_knn.cu:192-228; K = 3 is distCUDA2).
 *   mean_dist2 [P]  mean of the squared distances to the K nearest other points
 *   idx        [P][K] int32 their indices (ascending distance);  dist2 [P][K] (nullable) the squared distances
 */
size_t gsr_knn_temp_bytes(int64_t P);
int gsr_knn(int64_t P, int K, const float* points, float* mean_dist2, int32_t* idx, float* dist2,
            void* tmp, size_t tmp_bytes, gsr_stream_t stream);

/*
 * Fused image loss of the training loop (SURVEY.md 8f "next" row 3): w_l1 * mean|x - y| + w_ssim * (1 - mean SSIM(x, y))
 * and its gradient w.r.t. x.  Replaces the eager PyTorch code of REF/train.py:183 = 0.8 * l1_loss_v1 (REF/helpers.py)
 * + 0.2 * (1 - calc_ssim) (REF/external.py:90-133: 11x11 Gaussian window, sigma 1.5, zero padding, mean over all
 * channels and pixels).  img / target / dL_dimg: [n_planes][H][W] float32 (n_planes = images * channels);
 * state: gsr_image_loss_state_bytes() bytes kept from forward to backward; loss_out [3] device floats: total, L1 mean,
 * SSIM mean; grad_out: device scalar cotangent of the total (null = 1).
 */
size_t gsr_image_loss_state_bytes(int64_t n_planes, int W, int H);
int gsr_image_loss_forward(int64_t n_planes, int W, int H, const float* img, const float* target, float w_l1,
                           float w_ssim, void* state, size_t state_bytes, float* loss_out, gsr_stream_t stream);
int gsr_image_loss_backward(int64_t n_planes, int W, int H, const float* img, const float* target, float w_l1,
                            float w_ssim, const void* state, size_t state_bytes, const float* grad_out, float* dL_dimg,
                            gsr_stream_t stream);

/*
 * Fused Adam step over up to GSR_ADAM_MAX_TENSORS parameter tensors in one launch (the optimizer
 * of the train.py-shaped loop; REF/train.py:119-135 builds one torch.optim.Adam group per tensor, lr per group,
 * betas (0.9, 0.999), eps 1e-15).  Same update rule as torch.optim.Adam without weight decay / amsgrad.  All pointers
 * are device pointers; `step` is a one-float device counter per tensor (incremented by the call); tensors with a zero
 * learning rate still update their moments.  scratch: GSR_ADAM_SCRATCH_BYTES device bytes.
 */
#define GSR_ADAM_MAX_TENSORS 16
#define GSR_ADAM_SCRATCH_BYTES 256
typedef struct GsrAdamTensor {
    float* param;
    const float* grad;
    float* exp_avg;
    float* exp_avg_sq;
    float* step;
    int64_t n;
    float lr;
    int32_t pad_;
} GsrAdamTensor;
int gsr_adam_step(int n_tensors, const GsrAdamTensor* tensors_host, double beta1, double beta2, double eps,
                  void* scratch, gsr_stream_t stream);

/*
 * The small per-iteration ops around the rasterizer in the train.py-shaped loop, one kernel each.
 *   gsr_activations_*: params2rendervar, REF/helpers.py:98-107 -- rotations = F.normalize(unnorm_rotations) [P][4],
 *     opacities = sigmoid(logit_opacities) [P], scales = exp(log_scales) [P][3]; the backward takes the cotangents of the
 *     three outputs (any may be null = zero) and writes the gradients of the three raw parameters (any may be null).
 *   gsr_cam_affine_*: the per-camera colour correction exp(cam_m[c])[:, None, None] * im + cam_c[c][:, None, None] of
 *     REF/train.py:181; im / out [3][HW], cam_m3 / cam_c3 point at the camera's 3 floats.  The backward ADDS the two
 *     per-channel reductions into dL_dcam_m3 / dL_dcam_c3 (caller zeroes them) and writes dL_dim (nullable).
 *   gsr_densify_stats: REF/train.py:288-289 and accumulate_mean2d_gradient (REF/external.py:136-140).  seen = radii > 0
 *     (or seen_in when radii is null); max_2D_radius = max(radii, max_2D_radius) where seen; means2D_gradient_accum +=
 *     |means2D_grad.xy| where seen; denom += seen; every output is optional.
 */
/*
 * L1 image loss against 8-bit target frames: loss = sum_i w_i |img_i - target_i / 255| over [n_views][3][H][W], with
 * w_i = row_weight[view][y] ([n_views][H], e.g. the rows of a view a rank owns, pre-normalised) or the constant `weight`
 * (1 / N for the mean, l1_loss_v1 of REF/helpers.py) when row_weight is null.  loss_out: one device float (written).
 * Backward: dL_dimg = grad_out[0] (device scalar, null = 1) * w_i * sign(img_i - target_i / 255).
 */
int gsr_l1_u8_loss_forward(int64_t n_views, int W, int H, const float* img, const uint8_t* target,
                           const float* row_weight, float weight, float* loss_out, gsr_stream_t stream);
int gsr_l1_u8_loss_backward(int64_t n_views, int W, int H, const float* img, const uint8_t* target,
                            const float* row_weight, float weight, const float* grad_out, float* dL_dimg,
                            gsr_stream_t stream);
int gsr_activations_forward(int64_t P, const float* unnorm_rotations, const float* logit_opacities,
                            const float* log_scales, float* rotations, float* opacities, float* scales,
                            gsr_stream_t stream);
int gsr_activations_backward(int64_t P, const float* unnorm_rotations, const float* opacities, const float* scales,
                             const float* dL_drotations, const float* dL_dopacities, const float* dL_dscales,
                             float* dL_dunnorm_rotations, float* dL_dlogit_opacities, float* dL_dlog_scales,
                             gsr_stream_t stream);
int gsr_cam_affine_forward(int64_t HW, const float* im, const float* cam_m3, const float* cam_c3, float* out,
                           gsr_stream_t stream);
int gsr_cam_affine_backward(int64_t HW, const float* im, const float* cam_m3, const float* dL_dout, float* dL_dim,
                            float* dL_dcam_m3, float* dL_dcam_c3, gsr_stream_t stream);
int gsr_densify_stats(int64_t P, const int32_t* radii, const uint8_t* seen_in, const float* means2D_grad,
                      float* max_2D_radius, float* means2D_gradient_accum, float* denom, uint8_t* seen_out,
                      gsr_stream_t stream);

/*
 * Fused local-rigidity / rotation-similarity / isometry losses over a fixed K-NN graph of F foreground Gaussians,
 * forward and (optionally) backward: replaces the eager PyTorch code of REF/train.py:253-273 (== REF/cvpr_dyn.py:304-330).
 *   pts [F][3], rot [F][4] (w,x,y,z; the normalised current rotations), prev_inv_rot [F][4] (REF/train.py:303-304),
 *   nbr_idx [F][K] int32, nbr_weight [F][K], prev_offset [F][K][3] (:306), nbr_dist [F][K] (:323)
 *   losses [3] <- (rigid, rot, iso), each the mean over the F*K edges as weighted_l2_loss_v2 / _v1 (REF/helpers.py:118-123)
 * When dL_dpts [F][3] and dL_drot [F][4] are non-null they receive  sum_l upstream_dev[l] * d losses[l] / d(pts, rot);
 * upstream_dev is a DEVICE pointer to 3 floats (the loss weights or autograd's cotangents), so no host sync is needed.
 */
size_t gsr_dyn_losses_temp_bytes(int64_t F);
int gsr_dyn_losses(int64_t F, int K, const float* pts, const float* rot, const float* prev_inv_rot,
                   const int32_t* nbr_idx, const float* nbr_weight, const float* prev_offset, const float* nbr_dist,
                   const float* upstream_dev, float* losses, float* dL_dpts, float* dL_drot, void* tmp,
                   size_t tmp_bytes, gsr_stream_t stream);

/* Optional per-stage device timing (used by bench.py for the roofline numbers): when enabled, the library records
 * CUDA events on the launching stream around each stage of the forward/backward; gsr_stage_timing_read waits for the
 * most recent events and returns milliseconds per stage (-1 for a stage that did not run). Not thread-safe. */
int gsr_stage_timing_enable(int on);
int gsr_stage_timing_count(void);
const char* gsr_stage_name(int stage);
int gsr_stage_timing_read(float* ms_out, int n);

/* Utility kernel used by bench.py: dependent-free FMA loop that measures the FP32-pipe peak.  packed = 0: scalar FFMA,
 * flops = blocks * threads * iters * 256 (8 chains x 16 steps x 2);  packed = 1: fp32x2 FFMA2 (what the blend kernels
 * issue), flops = blocks * threads * iters * 512.  `out` holds blocks * threads floats. */
int gsr_fp32_peak_probe(float* out, int iters, int packed, gsr_stream_t stream);
int gsr_probe_geometry(int* blocks, int* threads);

/* Diagnostic: when non-null, every CTA of the blend forward / backward kernels writes 4 x uint64 at index blockIdx.x:
 * (first-thread start ns, last-warp end ns [atomicMax: zero the buffer first], SM id, view*T + tile).  Buffers must hold
 * 4 * V * T uint64 each; pass null to switch the trace off.  Synchronises the device (cudaMemcpyToSymbol). */
int gsr_debug_blend_trace(uint64_t* fwd, uint64_t* bwd);

#ifdef __cplusplus
}
#endif
#endif /* GSR_B200_H_INCLUDED */
