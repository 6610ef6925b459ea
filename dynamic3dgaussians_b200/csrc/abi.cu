// extern "C" entry points of libgsr_b200.so (see include/gsr_b200.h).  Plain pointers and sizes only; no torch types,
// no allocation, no device synchronisation, no exceptions.
#include <math.h>
#include <stdio.h>
#include <string.h>
#include <mutex>

#include "common.cuh"

namespace gsr {

static thread_local char g_cuda_err[512] = "";

void set_cuda_error(cudaError_t e, const char* where) {
    snprintf(g_cuda_err, sizeof(g_cuda_err), "%s: %s (%s)", where, cudaGetErrorName(e), cudaGetErrorString(e));
}

static bool g_timing = false;
static cudaEvent_t g_ev[ST_COUNT][2];
static bool g_ev_made = false, g_ev_used[ST_COUNT];
static Options g_options;
static thread_local bool g_debug_sync = false;     // this thread's current call runs in debug mode
static thread_local int g_debug_status = GSR_OK;

Options& options() { return g_options; }

static const char* const kStageNames[ST_COUNT] = {
    "preprocess", "depth_sort", "instance_count", "emit_supertile_instances", "supertile_sort", "tile_offsets",
    "scatter_records", "blend_forward", "grad_zero", "blend_backward", "preprocess_backward"};

}  // namespace gsr

#include <nvtx3/nvToolsExt.h>     // header-only; resolves the profiler's injection library lazily, a no-op without one

namespace gsr {

void stage_begin(int st, cudaStream_t s) {
    if (g_options.nvtx) nvtxRangePushA(kStageNames[st]);
    if (!g_timing) return;
    cudaEventRecord(g_ev[st][0], s);
}
void stage_end(int st, cudaStream_t s) {
    if (g_timing) {
        cudaEventRecord(g_ev[st][1], s);
        g_ev_used[st] = true;
    }
    if (g_options.nvtx) nvtxRangePop();
    if (g_debug_sync && g_debug_status == GSR_OK) {
        // the reference's debug mode: synchronise after every stage so that a faulting kernel is named
        // (CHECK_CUDA, R/cuda_rasterizer/auxiliary.h:172-179)
        cudaError_t e = cudaStreamSynchronize(s);
        if (e == cudaSuccess) e = cudaGetLastError();
        if (e != cudaSuccess) {
            char where[96];
            snprintf(where, sizeof(where), "stage '%s' (debug sync)", kStageNames[st]);
            set_cuda_error(e, where);
            g_debug_status = GSR_ERR_CUDA;
        }
    }
}

DebugSyncScope::DebugSyncScope(bool on) {
    g_debug_sync = on;
    g_debug_status = GSR_OK;
}
DebugSyncScope::~DebugSyncScope() { g_debug_sync = false; }
int DebugSyncScope::status() const { return g_debug_status; }

// ---- the library's side stream (see common.cuh) ----
namespace {
struct SideStream {
    cudaStream_t stream = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
};
constexpr int MAX_DEVICES = 64;
thread_local SideStream g_side[MAX_DEVICES];

int side_for_current_device(SideStream** out) {
    int dev = 0;
    GSR_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return GSR_ERR_UNSUPPORTED;
    SideStream& ss = g_side[dev];
    if (!ss.stream) {
        GSR_CUDA_TRY(cudaStreamCreateWithFlags(&ss.stream, cudaStreamNonBlocking));
        GSR_CUDA_TRY(cudaEventCreateWithFlags(&ss.fork, cudaEventDisableTiming));
        GSR_CUDA_TRY(cudaEventCreateWithFlags(&ss.join, cudaEventDisableTiming));
    }
    *out = &ss;
    return GSR_OK;
}
}  // namespace

int side_fork(cudaStream_t main, cudaStream_t* side) {
    SideStream* ss;
    int rc = side_for_current_device(&ss);
    if (rc) return rc;
    GSR_CUDA_TRY(cudaEventRecord(ss->fork, main));
    GSR_CUDA_TRY(cudaStreamWaitEvent(ss->stream, ss->fork, 0));
    *side = ss->stream;
    return GSR_OK;
}

int side_join(cudaStream_t main) {
    SideStream* ss;
    int rc = side_for_current_device(&ss);
    if (rc) return rc;
    GSR_CUDA_TRY(cudaEventRecord(ss->join, ss->stream));
    GSR_CUDA_TRY(cudaStreamWaitEvent(main, ss->join, 0));
    return GSR_OK;
}

static void fill_cam(const GsrCamera& c, CamDev& d) {
    memset(&d, 0, sizeof(d));
    memcpy(d.view, c.viewmatrix, sizeof(d.view));
    memcpy(d.proj, c.projmatrix, sizeof(d.proj));
    memcpy(d.campos, c.campos, sizeof(d.campos));
    memcpy(d.bg, c.bg, sizeof(d.bg));
    const int W = c.image_width, H = c.image_height;
    // host float arithmetic exactly as rasterizer_impl.cu:227-228 and forward.cu:87-99 (IEEE single ops)
    d.focal_y = (float)H / (2.0f * c.tanfovy);
    d.focal_x = (float)W / (2.0f * c.tanfovx);
    d.tanfovx = c.tanfovx;
    d.tanfovy = c.tanfovy;
    d.c_x = c.c_x;
    d.c_y = c.c_y;
    d.lim_x_pos = fmaf(c.tanfovx, 0.3f, ((float)W - c.c_x) / d.focal_x);
    d.lim_x_neg = fmaf(c.tanfovx, 0.3f, c.c_x / d.focal_x);
    d.lim_y_pos = fmaf(c.tanfovy, 0.3f, ((float)H - c.c_y) / d.focal_y);
    d.lim_y_neg = fmaf(c.tanfovy, 0.3f, c.c_y / d.focal_y);
    d.scale_modifier = c.scale_modifier;
    d.sh_degree = c.sh_degree;
    d.prefiltered = c.prefiltered;
    d.width = W;
    d.height = H;
    const int tiles_y = (H + TILE_Y - 1) / TILE_Y;
    const bool whole = c.tile_row_begin == 0 && c.tile_row_end == 0;
    d.row_begin = whole ? 0 : (c.tile_row_begin < 0 ? 0 : c.tile_row_begin);
    d.row_end = whole ? tiles_y : (c.tile_row_end > tiles_y ? tiles_y : c.tile_row_end);
    if (d.row_end < d.row_begin) d.row_end = d.row_begin;
}

// FP32-pipe peak probes: 8 independent chains per thread, 16 steps per chain per trip (128 FFMA or 128 FFMA2 between
// two loop branches, so the loop overhead is < 2 % of the issued instructions; the round-1 probe had 8 per trip and
// read 14 % under nominal).  PACKED issues fp32x2 FFMA2 -- what the blend kernels issue.
template <bool PACKED>
__global__ void __launch_bounds__(256) fp32_peak_kernel(float* out, int iters) {
    const float m = 0.9999f, c = 1e-4f;
    if constexpr (!PACKED) {
        float a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = threadIdx.x * 1e-3f + (float)k;
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int u = 0; u < 16; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = __fmaf_rn(a[k], m, c);
        }
        float sum = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) sum += a[k];
        out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
    } else {
        float2 a[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) a[k] = make_float2(threadIdx.x * 1e-3f + (float)k, (float)k - threadIdx.x * 1e-3f);
        const float2 m2 = make_float2(m, m), c2 = make_float2(c, c);
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int u = 0; u < 16; ++u)
#pragma unroll
                for (int k = 0; k < 8; ++k) a[k] = __ffma2_rn(a[k], m2, c2);
        }
        float sum = 0.f;
#pragma unroll
        for (int k = 0; k < 8; ++k) sum += a[k].x + a[k].y;
        out[blockIdx.x * blockDim.x + threadIdx.x] = sum;
    }
}

}  // namespace gsr

using namespace gsr;

extern "C" {

int gsr_abi_version(void) { return 4; }

const char* gsr_strerror(int status) {
    switch (status) {
        case GSR_OK: return "ok";
        case GSR_ERR_INVALID_ARGUMENT: return "invalid argument";
        case GSR_ERR_WORKSPACE_TOO_SMALL: return "workspace too small";
        case GSR_ERR_CUDA: return "CUDA error";
        case GSR_ERR_UNSUPPORTED: return "unsupported configuration";
        case GSR_ERR_CAPACITY: return "instance capacity exceeded";
        default: return "unknown status";
    }
}

const char* gsr_last_cuda_error(void) { return g_cuda_err; }

static int* option_slot(const char* key) {
    if (!key) return nullptr;
    if (!strcmp(key, "blend_ilp")) return &g_options.blend_ilp;
    if (!strcmp(key, "bwd_cost_order")) return &g_options.bwd_cost_order;
    if (!strcmp(key, "nvtx")) return &g_options.nvtx;
    if (!strcmp(key, "pdl")) return &g_options.pdl;
    if (!strcmp(key, "bwd_ilp")) return &g_options.bwd_ilp;
    return nullptr;
}

int gsr_set_option(const char* key, int value) {
    int* slot = option_slot(key);
    if (!slot) return GSR_ERR_INVALID_ARGUMENT;
    if ((slot == &g_options.blend_ilp || slot == &g_options.bwd_ilp) && (value < 0 || value > 2)) return GSR_ERR_INVALID_ARGUMENT;
    *slot = value;
    return GSR_OK;
}

int gsr_get_option(const char* key, int* value) {
    const int* slot = option_slot(key);
    if (!slot || !value) return GSR_ERR_INVALID_ARGUMENT;
    *value = *slot;
    return GSR_OK;
}

static int check_dims(int P, int V, int W, int H, int channels, int sh_coeffs, int64_t capacity) {
    if (P < 0 || V < 0 || W <= 0 || H <= 0 || capacity < 0 || sh_coeffs < 0 || sh_coeffs > 16)
        return GSR_ERR_INVALID_ARGUMENT;
    if (channels != 3 && channels != 6) return GSR_ERR_UNSUPPORTED;
    if ((int64_t)V * P >= (1ll << 30) || capacity >= (1ll << 30)) return GSR_ERR_UNSUPPORTED;
    const int64_t tx = (W + TILE_X - 1) / TILE_X, ty = (H + TILE_Y - 1) / TILE_Y;
    const int64_t VT = (int64_t)V * tx * ty;
    if (VT >= (1ll << 28) || tx > 1024 || ty > 1024) return GSR_ERR_UNSUPPORTED;   // supertile coordinates are bytes (16384 px)
    // the instance key keeps supertile << view_bits | view in 16 bits (its upper half carries the tile hit mask):
    // ceil(log2 views) + ceil(log2 supertiles per view) <= 16 -- e.g. 128 views of 1920x1080; split larger batches
    if (ceil_log2((uint32_t)(V > 0 ? V : 1)) + ceil_log2((uint32_t)(((tx + ST - 1) / ST) * ((ty + ST - 1) / ST))) > 16)
        return GSR_ERR_UNSUPPORTED;
    return GSR_OK;
}

int gsr_workspace_layout(int P, int V, int W, int H, int channels, int sh_coeffs, int64_t capacity, GsrLayout* out) {
    if (!out) return GSR_ERR_INVALID_ARGUMENT;
    int rc = check_dims(P, V, W, H, channels, sh_coeffs, capacity);
    if (rc) return rc;
    Layout L;
    compute_layout(P, V, W, H, channels, sh_coeffs, capacity, L);
    *out = static_cast<GsrLayout&>(L);
    return GSR_OK;
}

int gsr_workspace_sizes(int P, int V, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                        GsrWorkspaceSizes* out) {
    if (!out) return GSR_ERR_INVALID_ARGUMENT;
    int rc = check_dims(P, V, W, H, channels, sh_coeffs, capacity);
    if (rc) return rc;
    Layout L;
    compute_layout(P, V, W, H, channels, sh_coeffs, capacity, L);
    out->geom_bytes = L.geom_bytes;
    out->binning_bytes = L.binning_bytes;
    out->image_bytes = L.image_bytes;
    return GSR_OK;
}

int gsr_forward_geometry(const GsrCamera* cams_host, int V, int P, int channels, int sh_coeffs, const float* means3D,
                         const float* opacities, const float* scales, const float* rotations,
                         const float* cov3D_precomp, const float* sh, void* geom_ws, size_t geom_bytes,
                         int32_t* radii, uint32_t* num_rendered_dev, gsr_stream_t stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (!cams_host || V <= 0 || !geom_ws) return GSR_ERR_INVALID_ARGUMENT;
    if (P > 0 && (!means3D || !opacities)) return GSR_ERR_INVALID_ARGUMENT;
    const bool has_sr = scales != nullptr && rotations != nullptr;
    if (P > 0 && (has_sr == (cov3D_precomp != nullptr))) return GSR_ERR_INVALID_ARGUMENT;
    if ((sh != nullptr) != (sh_coeffs > 0)) return GSR_ERR_INVALID_ARGUMENT;
    const int W = cams_host[0].image_width, H = cams_host[0].image_height;
    for (int v = 1; v < V; ++v) {
        if (cams_host[v].image_width != W || cams_host[v].image_height != H) return GSR_ERR_UNSUPPORTED;
        if (cams_host[v].scale_modifier != cams_host[0].scale_modifier) return GSR_ERR_UNSUPPORTED;
    }
    int rc = check_dims(P, V, W, H, channels, sh_coeffs, 0);
    if (rc) return rc;
    Layout L;
    compute_layout(P, V, W, H, channels, sh_coeffs, 0, L);
    if (geom_bytes < L.geom_bytes) return GSR_ERR_WORKSPACE_TOO_SMALL;
    DebugSyncScope dbg(cams_host[0].debug != 0);

    // cameras -> device as kernel arguments (a few hundred bytes each): unlike a copy from pageable host memory this
    // never makes the driver wait for the stream, and it can be captured into a CUDA graph
    {
        CamDev* cd = at<CamDev>(geom_ws, L.cams);
        for (int v0 = 0; v0 < V; v0 += CamPack::N) {
            CamPack pack;
            pack.n = V - v0 < CamPack::N ? V - v0 : CamPack::N;
            for (int i = 0; i < pack.n; ++i) fill_cam(cams_host[v0 + i], pack.cam[i]);
            rc = launch_store_cams(pack, cd + v0, s);
            if (rc) return rc;
        }
    }
    uint32_t* counters = at<uint32_t>(geom_ws, L.counters);
    // counters, the depth sort's histograms, look-back words and tile tickets: one memset
    GSR_CUDA_TRY(cudaMemsetAsync(geom_ws, 0, L.geom_zero_bytes, s));
    {
        StageScope sc(ST_PREPROCESS, s);
        rc = launch_preprocess(at<CamDev>(geom_ws, L.cams), V, P, sh_coeffs, means3D, opacities, scales, rotations,
                               cov3D_precomp, sh, cams_host[0].scale_modifier, geom_ws, L, radii, s);
    }
    if (rc) return rc;
    {
        // every view's Gaussians by depth: V segments of P (depth bits, Gaussian index) pairs, 4 passes of 32-bit keys
        StageScope sc(ST_DEPTH_SORT, s);
        // Gaussians that touch no tile carry the sentinel key ~0: the first pass drops them and leaves vis_count[v] per view
        rc = radix_sort_pairs<uint32_t>(at<uint32_t>(geom_ws, L.dkeys0), at<uint32_t>(geom_ws, L.dkeys1),
                                        at<uint32_t>(geom_ws, L.dvals0), at<uint32_t>(geom_ws, L.dvals1), P, nullptr,
                                        0, 32, at<uint32_t>(geom_ws, L.dhist), at<uint32_t>(geom_ws, L.dlookback),
                                        at<uint32_t>(geom_ws, L.dcounters), s, V, /*hist_ready=*/true, true,
                                        at<uint32_t>(geom_ws, L.vis_count));   // (histograms: the preprocess kernel)
    }
    if (rc) return rc;
    // (the instance count -- the reference's num_rendered -- is summed by the preprocess kernel itself)
    if (num_rendered_dev)
        GSR_CUDA_TRY(cudaMemcpyAsync(num_rendered_dev, counters + CNT_RENDERED, 4, cudaMemcpyDeviceToDevice, s));
    return dbg.status();
}

int gsr_forward_render(int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                       const float* colors_precomp, const float* extra_colors, const float* features,
                       int num_features, void* geom_ws, void* binning_ws, void* image_ws, float* out_color,
                       float* out_extra, float* out_depth, float* out_alpha, float* out_features, int flags,
                       gsr_stream_t stream) {
    cudaStream_t s = (cudaStream_t)stream;
    int rc = check_dims(P, V, W, H, channels, sh_coeffs, capacity);
    if (rc) return rc;
    DebugSyncScope dbg((flags & GSR_FLAG_DEBUG) != 0);
    if (!geom_ws || !image_ws || !out_color || !out_depth || !out_alpha) return GSR_ERR_INVALID_ARGUMENT;
    if (capacity > 0 && !binning_ws) return GSR_ERR_INVALID_ARGUMENT;
    if (channels == 6 && (!extra_colors || !out_extra)) return GSR_ERR_INVALID_ARGUMENT;
    if ((sh_coeffs > 0) == (colors_precomp != nullptr) && P > 0) return GSR_ERR_INVALID_ARGUMENT;
    if (features && (num_features != GSR_NUM_SEMANTIC_CHANNELS || !out_features)) return GSR_ERR_INVALID_ARGUMENT;
    Layout L;
    compute_layout(P, V, W, H, channels, sh_coeffs, capacity, L);
    rc = launch_binning(V, P, W, H, channels, sh_coeffs, capacity, colors_precomp, extra_colors, geom_ws, binning_ws,
                        image_ws, L, s);
    if (rc) return rc;
    {
        StageScope sc(ST_BLEND_FWD, s);
        rc = launch_blend_forward(V, W, H, channels, geom_ws, binning_ws, image_ws, L, out_color, out_extra, out_depth,
                                  out_alpha, s);
    }
    if (rc) return rc;
    if (features) {
        rc = launch_feature_forward(V, P, W, H, geom_ws, binning_ws, image_ws, L, features, out_features, s);
        if (rc) return rc;
        if (flags & GSR_FLAG_DEBUG) GSR_CUDA_TRY(cudaStreamSynchronize(s));
    }
    return dbg.status();
}

int gsr_read_counters(const void* geom_ws, uint32_t* counters_host, gsr_stream_t stream) {
    if (!geom_ws || !counters_host) return GSR_ERR_INVALID_ARGUMENT;
    // the counter block sits at offset 0 of the geometry workspace for every problem size (compute_layout):
    // [CNT_RENDERED, CNT_OVERFLOW, CNT_PREFILTER, supertile instances]
    GSR_CUDA_TRY(cudaMemcpyAsync(counters_host, geom_ws, 4 * sizeof(uint32_t), cudaMemcpyDeviceToHost,
                                 (cudaStream_t)stream));
    return GSR_OK;
}

int gsr_backward(int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity, const float* means3D,
                 const float* opacities, const float* scales, const float* rotations, const float* cov3D_precomp,
                 const float* sh, const float* colors_precomp, const float* extra_colors, const float* features,
                 int num_features, const int32_t* radii, void* geom_ws, void* binning_ws, void* image_ws,
                 const float* dL_dcolor, const float* dL_dextra, const float* dL_ddepth, const float* dL_dalpha,
                 const float* dL_dfeatures_img, const float* label, float* dL_dmeans3D, float* dL_dmeans2D,
                 float* dL_dmeans2D_color_only, float* dL_dcolors, float* dL_dextra_colors, float* dL_dopacity,
                 float* dL_dscales, float* dL_drotations, float* dL_dcov3D, float* dL_dsh, float* dL_dfeatures,
                 int flags, gsr_stream_t stream) {
    cudaStream_t s = (cudaStream_t)stream;
    int rc = check_dims(P, V, W, H, channels, sh_coeffs, capacity);
    if (rc) return rc;
    DebugSyncScope dbg((flags & GSR_FLAG_DEBUG) != 0);
    if (!geom_ws || !image_ws) return GSR_ERR_INVALID_ARGUMENT;
    if (capacity > 0 && !binning_ws) return GSR_ERR_INVALID_ARGUMENT;
    if (P > 0 && !means3D) return GSR_ERR_INVALID_ARGUMENT;
    (void)opacities; (void)colors_precomp; (void)extra_colors;
    if (features && num_features != GSR_NUM_SEMANTIC_CHANNELS) return GSR_ERR_INVALID_ARGUMENT;
    Layout L;
    compute_layout(P, V, W, H, channels, sh_coeffs, capacity, L);
    rc = launch_blend_backward(V, P, W, H, channels, geom_ws, binning_ws, image_ws, L, dL_dcolor, dL_dextra, dL_ddepth,
                               dL_dalpha, dL_dmeans2D_color_only != nullptr && channels == 6, s);
    if (rc) return rc;
    if (features && dL_dfeatures) {
        if (dL_dfeatures_img)
            rc = launch_feature_backward(V, P, W, H, channels, geom_ws, binning_ws, image_ws, L, features, dL_dfeatures_img,
                                         dL_dfeatures, s);
        else
            rc = cudaMemsetAsync(dL_dfeatures, 0, (size_t)P * GSR_NUM_SEMANTIC_CHANNELS * sizeof(float), s) == cudaSuccess
                     ? GSR_OK : GSR_ERR_CUDA;
        if (rc) return rc;
    }
    {
        StageScope sc(ST_PREPROCESS_BWD, s);
        rc = launch_preprocess_backward(V, P, W, H, channels, sh_coeffs, means3D, opacities, scales, rotations,
                                        cov3D_precomp, sh, radii, geom_ws, L, label, dL_dmeans3D, dL_dmeans2D,
                                        dL_dmeans2D_color_only, dL_dcolors, dL_dextra_colors, dL_dopacity, dL_dscales,
                                        dL_drotations, dL_dcov3D, dL_dsh, s);
    }
    return rc ? rc : dbg.status();
}

int gsr_mark_visible(const GsrCamera* cam_host, int P, const float* means3D, uint8_t* present, gsr_stream_t stream) {
    if (!cam_host || P < 0 || (P > 0 && (!means3D || !present))) return GSR_ERR_INVALID_ARGUMENT;
    return launch_mark_visible(cam_host->viewmatrix, P, means3D, present, (cudaStream_t)stream);
}

int gsr_export_view_binning(int view, int V, int P, int W, int H, int channels, int sh_coeffs, int64_t capacity,
                            const void* geom_ws, const void* binning_ws, const void* image_ws, uint64_t* keys_out,
                            uint32_t* list_out, uint32_t* ranges_out, uint32_t* n_contrib_out, gsr_stream_t stream) {
    int rc = check_dims(P, V, W, H, channels, sh_coeffs, capacity);
    if (rc) return rc;
    if (view < 0 || view >= V || !geom_ws || !binning_ws || !image_ws) return GSR_ERR_INVALID_ARGUMENT;
    Layout L;
    compute_layout(P, V, W, H, channels, sh_coeffs, capacity, L);
    return launch_export_view(view, V, P, W, H, geom_ws, binning_ws, image_ws, L, keys_out, list_out, ranges_out,
                              n_contrib_out, (cudaStream_t)stream);
}

// ---- fused L1 + SSIM image loss ----
size_t gsr_image_loss_state_bytes(int64_t n_planes, int W, int H) {
    if (n_planes <= 0 || W <= 0 || H <= 0) return 0;
    return image_loss_state_bytes(n_planes, W, H);
}

static int check_loss_args(int64_t n_planes, int W, int H, const void* a, const void* b, const void* state,
                           size_t state_bytes) {
    if (n_planes <= 0 || n_planes > 65535 || W <= 0 || H <= 0 || W >= (1 << 20) || H >= (1 << 20))
        return GSR_ERR_INVALID_ARGUMENT;
    if (!a || !b || !state) return GSR_ERR_INVALID_ARGUMENT;
    if (state_bytes < image_loss_state_bytes(n_planes, W, H)) return GSR_ERR_WORKSPACE_TOO_SMALL;
    return GSR_OK;
}

int gsr_image_loss_forward(int64_t n_planes, int W, int H, const float* img, const float* target, float w_l1,
                           float w_ssim, void* state, size_t state_bytes, float* loss_out, gsr_stream_t stream) {
    int rc = check_loss_args(n_planes, W, H, img, target, state, state_bytes);
    if (rc) return rc;
    if (!loss_out) return GSR_ERR_INVALID_ARGUMENT;
    return launch_image_loss_forward(n_planes, W, H, img, target, w_l1, w_ssim, state, loss_out, (cudaStream_t)stream);
}

int gsr_image_loss_backward(int64_t n_planes, int W, int H, const float* img, const float* target, float w_l1,
                            float w_ssim, const void* state, size_t state_bytes, const float* grad_out, float* dL_dimg,
                            gsr_stream_t stream) {
    int rc = check_loss_args(n_planes, W, H, img, target, state, state_bytes);
    if (rc) return rc;
    if (!dL_dimg) return GSR_ERR_INVALID_ARGUMENT;
    return launch_image_loss_backward(n_planes, W, H, img, target, w_l1, w_ssim, state, grad_out, dL_dimg,
                                      (cudaStream_t)stream);
}

// ---- fused Adam ----
int gsr_adam_step(int n_tensors, const GsrAdamTensor* tensors_host, double beta1, double beta2, double eps,
                  void* scratch, gsr_stream_t stream) {
    if (n_tensors < 0 || n_tensors > GSR_ADAM_MAX_TENSORS || (n_tensors > 0 && !tensors_host) || !scratch)
        return GSR_ERR_INVALID_ARGUMENT;
    for (int i = 0; i < n_tensors; ++i) {
        const GsrAdamTensor& t = tensors_host[i];
        if (t.n < 0 || !t.step || (t.n > 0 && (!t.param || !t.grad || !t.exp_avg || !t.exp_avg_sq)))
            return GSR_ERR_INVALID_ARGUMENT;
    }
    if (n_tensors == 0) return GSR_OK;
    return launch_adam_step(n_tensors, tensors_host, beta1, beta2, eps, scratch, (cudaStream_t)stream);
}

// ---- training-loop glue ----
int gsr_activations_forward(int64_t P, const float* unnorm_rotations, const float* logit_opacities,
                            const float* log_scales, float* rotations, float* opacities, float* scales,
                            gsr_stream_t stream) {
    if (P < 0 || (P > 0 && (!unnorm_rotations || !logit_opacities || !log_scales || !rotations || !opacities || !scales)))
        return GSR_ERR_INVALID_ARGUMENT;
    return launch_activations_forward(P, unnorm_rotations, logit_opacities, log_scales, rotations, opacities, scales,
                                      (cudaStream_t)stream);
}

int gsr_activations_backward(int64_t P, const float* unnorm_rotations, const float* opacities, const float* scales,
                             const float* dL_drotations, const float* dL_dopacities, const float* dL_dscales,
                             float* dL_dunnorm_rotations, float* dL_dlogit_opacities, float* dL_dlog_scales,
                             gsr_stream_t stream) {
    if (P < 0 || (P > 0 && (!unnorm_rotations || !opacities || !scales))) return GSR_ERR_INVALID_ARGUMENT;
    return launch_activations_backward(P, unnorm_rotations, opacities, scales, dL_drotations, dL_dopacities, dL_dscales,
                                       dL_dunnorm_rotations, dL_dlogit_opacities, dL_dlog_scales, (cudaStream_t)stream);
}

int gsr_cam_affine_forward(int64_t HW, const float* im, const float* cam_m3, const float* cam_c3, float* out,
                           gsr_stream_t stream) {
    if (HW < 0 || (HW > 0 && (!im || !cam_m3 || !cam_c3 || !out))) return GSR_ERR_INVALID_ARGUMENT;
    return launch_cam_affine_forward(HW, im, cam_m3, cam_c3, out, (cudaStream_t)stream);
}

int gsr_cam_affine_backward(int64_t HW, const float* im, const float* cam_m3, const float* dL_dout, float* dL_dim,
                            float* dL_dcam_m3, float* dL_dcam_c3, gsr_stream_t stream) {
    if (HW < 0 || (HW > 0 && (!im || !cam_m3 || !dL_dout))) return GSR_ERR_INVALID_ARGUMENT;
    return launch_cam_affine_backward(HW, im, cam_m3, dL_dout, dL_dim, dL_dcam_m3, dL_dcam_c3, (cudaStream_t)stream);
}

int gsr_densify_stats(int64_t P, const int32_t* radii, const uint8_t* seen_in, const float* means2D_grad,
                      float* max_2D_radius, float* means2D_gradient_accum, float* denom, uint8_t* seen_out,
                      gsr_stream_t stream) {
    if (P < 0 || (P > 0 && ((!radii && !seen_in) || (means2D_gradient_accum && !means2D_grad))))
        return GSR_ERR_INVALID_ARGUMENT;
    return launch_densify_stats(P, radii, seen_in, means2D_grad, max_2D_radius, means2D_gradient_accum, denom, seen_out,
                                (cudaStream_t)stream);
}

int gsr_l1_u8_loss_forward(int64_t n_views, int W, int H, const float* img, const uint8_t* target,
                           const float* row_weight, float weight, float* loss_out, gsr_stream_t stream) {
    if (n_views < 0 || W <= 0 || H <= 0 || !loss_out || (n_views > 0 && (!img || !target)))
        return GSR_ERR_INVALID_ARGUMENT;
    return launch_l1_u8_forward(n_views, W, H, img, target, row_weight, weight, loss_out, (cudaStream_t)stream);
}

int gsr_l1_u8_loss_backward(int64_t n_views, int W, int H, const float* img, const uint8_t* target,
                            const float* row_weight, float weight, const float* grad_out, float* dL_dimg,
                            gsr_stream_t stream) {
    if (n_views < 0 || W <= 0 || H <= 0 || (n_views > 0 && (!img || !target || !dL_dimg)))
        return GSR_ERR_INVALID_ARGUMENT;
    return launch_l1_u8_backward(n_views, W, H, img, target, row_weight, weight, grad_out, dL_dimg,
                                 (cudaStream_t)stream);
}

// ---- stand-alone sort (tests) ----
static size_t sort_layout(int64_t n, int key_bytes, size_t off[7]) {
    size_t o = 0;
    auto take = [&](size_t b) { size_t r = o; o = align_up(o + b); return r; };
    const size_t nn = (size_t)(n > 0 ? n : 0);
    off[0] = take(nn * key_bytes);                 // keys A
    off[1] = take(nn * key_bytes);                 // keys B
    off[2] = take(nn * 4);                         // vals A
    off[3] = take(nn * 4);                         // vals B
    off[4] = take((size_t)MAX_PASSES * RADIX * 4); // hist
    off[5] = take((size_t)MAX_PASSES * (sort_num_tiles(n, key_bytes) + 1) * RADIX * 4);  // lookback
    off[6] = take(MAX_PASSES * 4);                 // tile counters
    return o;
}

size_t gsr_sort_temp_bytes(int64_t n, int key_bytes) {
    size_t off[7];
    return sort_layout(n, key_bytes, off);
}

int gsr_sort_pairs(int key_bytes, const void* keys_in, void* keys_out, const uint32_t* vals_in, uint32_t* vals_out,
                   int64_t n, int begin_bit, int end_bit, void* tmp, size_t tmp_bytes, gsr_stream_t stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (key_bytes != 4 && key_bytes != 8) return GSR_ERR_INVALID_ARGUMENT;
    if (n < 0 || begin_bit < 0 || end_bit > key_bytes * 8 || end_bit < begin_bit) return GSR_ERR_INVALID_ARGUMENT;
    if (n == 0) return GSR_OK;
    if (!keys_in || !keys_out || !vals_in || !vals_out || !tmp) return GSR_ERR_INVALID_ARGUMENT;
    size_t off[7];
    if (tmp_bytes < sort_layout(n, key_bytes, off)) return GSR_ERR_WORKSPACE_TOO_SMALL;
    char* t = (char*)tmp;
    GSR_CUDA_TRY(cudaMemcpyAsync(t + off[0], keys_in, (size_t)n * key_bytes, cudaMemcpyDeviceToDevice, s));
    GSR_CUDA_TRY(cudaMemcpyAsync(t + off[2], vals_in, (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
    const int passes = (end_bit - begin_bit + RADIX_BITS - 1) / RADIX_BITS;
    int rc;
    if (key_bytes == 4)
        rc = radix_sort_pairs<uint32_t>((uint32_t*)(t + off[0]), (uint32_t*)(t + off[1]), (uint32_t*)(t + off[2]),
                                        (uint32_t*)(t + off[3]), n, nullptr, begin_bit, end_bit,
                                        (uint32_t*)(t + off[4]), (uint32_t*)(t + off[5]), (uint32_t*)(t + off[6]), s);
    else
        rc = radix_sort_pairs<uint64_t>((uint64_t*)(t + off[0]), (uint64_t*)(t + off[1]), (uint32_t*)(t + off[2]),
                                        (uint32_t*)(t + off[3]), n, nullptr, begin_bit, end_bit,
                                        (uint32_t*)(t + off[4]), (uint32_t*)(t + off[5]), (uint32_t*)(t + off[6]), s);
    if (rc) return rc;
    const int res = passes & 1;
    GSR_CUDA_TRY(cudaMemcpyAsync(keys_out, t + off[res ? 1 : 0], (size_t)n * key_bytes, cudaMemcpyDeviceToDevice, s));
    GSR_CUDA_TRY(cudaMemcpyAsync(vals_out, t + off[res ? 3 : 2], (size_t)n * 4, cudaMemcpyDeviceToDevice, s));
    return GSR_OK;
}

// ---- the same sort with its scratch in a library-owned device array: no allocation anywhere on the call ----
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (CUDA graph capture, serving loops).  GSR_SORT_STATIC_MAX_PAIRS pairs of 8-byte keys need
// sort_layout(GSR_SORT_STATIC_MAX_PAIRS, 8) = 7.4 MiB; every GPU that loads the library holds the array until exit.
constexpr size_t SORT_STATIC_BYTES = 8u << 20;
static __device__ __align__(256) unsigned char g_sort_static_scratch[SORT_STATIC_BYTES];

// In host code `g_sort_static_scratch` is only a handle for the runtime: the device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
static int sort_static_scratch(char** out) {
    constexpr int MAX_DEVICES = 64;
    static std::mutex mu;
    static char* addr[MAX_DEVICES] = {};
    int dev = -1;
    GSR_CUDA_TRY(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES) return GSR_ERR_INVALID_ARGUMENT;
    std::lock_guard<std::mutex> lock(mu);
    if (!addr[dev]) {
        void* p = nullptr;
        GSR_CUDA_TRY(cudaGetSymbolAddress(&p, g_sort_static_scratch));
        addr[dev] = (char*)p;
    }
    *out = addr[dev];
    return GSR_OK;
}

int64_t gsr_sort_pairs_static_max(void) { return GSR_SORT_STATIC_MAX_PAIRS; }

int gsr_sort_pairs_static(int key_bytes, const void* keys_in, void* keys_out, const uint32_t* vals_in,
                          uint32_t* vals_out, int64_t n, int begin_bit, int end_bit, gsr_stream_t stream) {
    if (key_bytes != 4 && key_bytes != 8) return GSR_ERR_INVALID_ARGUMENT;
    if (n < 0 || begin_bit < 0 || end_bit > key_bytes * 8 || end_bit < begin_bit) return GSR_ERR_INVALID_ARGUMENT;
    if (n > GSR_SORT_STATIC_MAX_PAIRS) {
        snprintf(g_cuda_err, sizeof(g_cuda_err), "gsr_sort_pairs_static: n = %lld exceeds GSR_SORT_STATIC_MAX_PAIRS = %d; "
                 "use gsr_sort_pairs with caller-owned scratch", (long long)n, GSR_SORT_STATIC_MAX_PAIRS);
        return GSR_ERR_WORKSPACE_TOO_SMALL;
    }
    size_t off[7];
    if (sort_layout(n, key_bytes, off) > SORT_STATIC_BYTES) return GSR_ERR_WORKSPACE_TOO_SMALL;
    if (n == 0) return GSR_OK;
    char* scratch = nullptr;
    int rc = sort_static_scratch(&scratch);
    if (rc != GSR_OK) return rc;
    // radix_sort_pairs clears the histogram, look-back and ticket words it uses at the start of every call (the
    // array keeps the previous call's contents); the key/value ping-pong buffers are fully overwritten.
    return gsr_sort_pairs(key_bytes, keys_in, keys_out, vals_in, vals_out, n, begin_bit, end_bit, scratch,
                          SORT_STATIC_BYTES, stream);
}

int gsr_peer_allreduce(int mode, int world, int rank, const void* const* bufs, const void* const* flags,
                       void* multicast_buf, int64_t n_floats, int blocks, gsr_stream_t stream) {
    if (!bufs || !flags) return GSR_ERR_INVALID_ARGUMENT;
    if (n_floats == 0) return GSR_OK;
    return launch_peer_allreduce(mode, world, rank, bufs, flags, multicast_buf, n_floats, blocks, (cudaStream_t)stream);
}

int gsr_stage_timing_enable(int on) {
    if (on && !g_ev_made) {
        for (int i = 0; i < ST_COUNT; ++i)
            for (int j = 0; j < 2; ++j) GSR_CUDA_TRY(cudaEventCreate(&g_ev[i][j]));
        g_ev_made = true;
    }
    for (int i = 0; i < ST_COUNT; ++i) g_ev_used[i] = false;
    g_timing = on != 0;
    return GSR_OK;
}

int gsr_stage_timing_count(void) { return ST_COUNT; }

const char* gsr_stage_name(int stage) { return stage >= 0 && stage < ST_COUNT ? kStageNames[stage] : "?"; }

int gsr_stage_timing_read(float* ms_out, int n) {
    if (!ms_out || !g_ev_made) return GSR_ERR_INVALID_ARGUMENT;
    for (int i = 0; i < n && i < ST_COUNT; ++i) {
        ms_out[i] = -1.0f;
        if (!g_ev_used[i]) continue;
        GSR_CUDA_TRY(cudaEventSynchronize(g_ev[i][1]));
        GSR_CUDA_TRY(cudaEventElapsedTime(&ms_out[i], g_ev[i][0], g_ev[i][1]));
    }
    return GSR_OK;
}

int gsr_debug_blend_trace(uint64_t* fwd, uint64_t* bwd) {
    int rc = gsr::set_blend_forward_trace((unsigned long long*)fwd);
    if (rc != GSR_OK) return rc;
    return gsr::set_blend_backward_trace((unsigned long long*)bwd);
}

int gsr_probe_geometry(int* blocks, int* threads) {
    if (blocks) *blocks = 148 * 8;
    if (threads) *threads = 256;
    return GSR_OK;
}

int gsr_fp32_peak_probe(float* out, int iters, int packed, gsr_stream_t stream) {
    if (!out || iters <= 0) return GSR_ERR_INVALID_ARGUMENT;
    if (packed)
        fp32_peak_kernel<true><<<148 * 8, 256, 0, (cudaStream_t)stream>>>(out, iters);
    else
        fp32_peak_kernel<false><<<148 * 8, 256, 0, (cudaStream_t)stream>>>(out, iters);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // extern "C"
