// TEST INFRASTRUCTURE ONLY.  Thin extern "C" shim around the UNMODIFIED reference CUDA rasterizer
// (R/cuda_rasterizer/*.cu compiled from /root/reference where they lie, see oracle/build_ref.py), so that tests and
// the benchmark's reference arm can call CudaRasterizer::Rasterizer::forward/backward (R/cuda_rasterizer/rasterizer.h)
// without the reference's torch extension.  Nothing here is part of the product.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>

#include <functional>

#include "cuda_rasterizer/rasterizer_impl.h"   // resolved through -I <reference dir>

namespace {

struct GrowBuf {
    char* ptr = nullptr;
    size_t cap = 0;
    char* get(size_t n) {
        if (n > cap) {
            if (ptr) cudaFree(ptr);
            size_t want = n + n / 4 + 4096;
            if (cudaMalloc(&ptr, want) != cudaSuccess) { ptr = nullptr; cap = 0; return nullptr; }
            cap = want;
        }
        return ptr;
    }
};

GrowBuf g_geom, g_bin, g_img;
int g_P = 0, g_R = 0, g_W = 0, g_H = 0;

}  // namespace

extern "C" {

// Returns num_rendered (>= 0) or -1 on CUDA error.  All pointers are device pointers.
int ref_forward(int P, int D, int M, const float* background, int W, int H, const float* means3D, const float* shs,
                const float* colors_precomp, const float* semantic_feature, const float* opacities,
                const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                const float* viewmatrix, const float* projmatrix, const float* campos, float c_x, float c_y,
                float tan_fovx, float tan_fovy, float* out_color, float* out_feature_map, float* out_depth,
                float* out_alpha, int* radii) {
    std::function<char*(size_t)> fg = [](size_t n) { return g_geom.get(n); };
    std::function<char*(size_t)> fb = [](size_t n) { return g_bin.get(n); };
    std::function<char*(size_t)> fi = [](size_t n) { return g_img.get(n); };
    int R = CudaRasterizer::Rasterizer::forward(fg, fb, fi, P, D, M, background, W, H, means3D, shs, colors_precomp,
                                                semantic_feature, opacities, scales, scale_modifier, rotations,
                                                cov3D_precomp, viewmatrix, projmatrix, campos, c_x, c_y, tan_fovx,
                                                tan_fovy, false, out_color, out_feature_map, out_depth, out_alpha,
                                                radii, false);
    g_P = P; g_R = R; g_W = W; g_H = H;
    return cudaGetLastError() == cudaSuccess ? R : -1;
}

// Backward of the most recent ref_forward (uses its geometry/binning/image buffers), as Rasterizer::backward.
int ref_backward(int P, int D, int M, const float* background, int W, int H, const float* means3D, const float* shs,
                 const float* colors_precomp, const float* semantic_feature, const float* alphas,
                 const float* scales, float scale_modifier, const float* rotations, const float* cov3D_precomp,
                 const float* viewmatrix, const float* projmatrix, const float* campos, float c_x, float c_y,
                 float tan_fovx, float tan_fovy, const int* radii, const float* dL_dpix, const float* dL_dfeaturepix,
                 const float* dL_dpix_depth, const float* dL_dalphas, float* dL_dmean2D, float* dL_dconic,
                 float* dL_dopacity, float* dL_dcolor, float* dL_dsemantic_feature, float* dL_ddepth,
                 float* dL_dmean3D, float* dL_dcov3D, float* dL_dsh, float* dL_dscale, float* dL_drot) {
    CudaRasterizer::Rasterizer::backward(P, D, M, g_R, background, W, H, means3D, shs, colors_precomp,
                                         semantic_feature, alphas, scales, scale_modifier, rotations, cov3D_precomp,
                                         viewmatrix, projmatrix, campos, c_x, c_y, tan_fovx, tan_fovy, radii,
                                         g_geom.ptr, g_bin.ptr, g_img.ptr, dL_dpix, dL_dfeaturepix, dL_dpix_depth,
                                         dL_dalphas, dL_dmean2D, dL_dconic, dL_dopacity, dL_dcolor,
                                         dL_dsemantic_feature, dL_ddepth, dL_dmean3D, dL_dcov3D, dL_dsh, dL_dscale,
                                         dL_drot, false);
    return cudaGetLastError() == cudaSuccess ? 0 : -1;
}

// Copies the internal state of the most recent ref_forward into caller buffers (device pointers, any may be null):
// sorted keys [R] u64, sorted point list [R] u32, ranges [tiles][2] u32, n_contrib [W*H] u32, tiles_touched [P] u32,
// means2D [P][2] f32, depths [P] f32, conic_opacity [P][4] f32.
int ref_export_state(uint64_t* keys, uint32_t* point_list, uint32_t* ranges, uint32_t* n_contrib,
                     uint32_t* tiles_touched, float* means2D, float* depths, float* conic_opacity) {
    using namespace CudaRasterizer;
    char* gp = g_geom.ptr;
    char* bp = g_bin.ptr;
    char* ip = g_img.ptr;
    GeometryState geom = GeometryState::fromChunk(gp, g_P);
    ImageState img = ImageState::fromChunk(ip, (size_t)g_W * g_H);
    const int tiles = ((g_W + 15) / 16) * ((g_H + 15) / 16);
    auto cp = [](void* dst, const void* src, size_t n) {
        return dst == nullptr || n == 0 || cudaMemcpy(dst, src, n, cudaMemcpyDeviceToDevice) == cudaSuccess;
    };
    bool ok = true;
    if (g_R > 0 && bp) {
        BinningState bin = BinningState::fromChunk(bp, g_R);
        ok &= cp(keys, bin.point_list_keys, (size_t)g_R * 8);
        ok &= cp(point_list, bin.point_list, (size_t)g_R * 4);
    }
    ok &= cp(ranges, img.ranges, (size_t)tiles * 8);
    ok &= cp(n_contrib, img.n_contrib, (size_t)g_W * g_H * 4);
    ok &= cp(tiles_touched, geom.tiles_touched, (size_t)g_P * 4);
    ok &= cp(means2D, geom.means2D, (size_t)g_P * 8);
    ok &= cp(depths, geom.depths, (size_t)g_P * 4);
    ok &= cp(conic_opacity, geom.conic_opacity, (size_t)g_P * 16);
    return ok ? 0 : -1;
}

void ref_release(void) {
    if (g_geom.ptr) cudaFree(g_geom.ptr);
    if (g_bin.ptr) cudaFree(g_bin.ptr);
    if (g_img.ptr) cudaFree(g_img.ptr);
    g_geom = GrowBuf(); g_bin = GrowBuf(); g_img = GrowBuf();
}

}  // extern "C"
