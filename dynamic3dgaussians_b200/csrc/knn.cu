// (e) exact k-nearest neighbours over a point cloud: replaces SimpleKNN::knn (K/simple_knn.cu:192-228), K = 3 being
// distCUDA2 (K/spatial.cu:15-27), K = 20 the rigidity graph the reference builds on the CPU with Open3D
// (REF/helpers.py:136-146, REF/train.py:321).
//
// Same idea as the reference (Morton order + box-pruned exhaustive search, exact result), different structure:
//   * scene bounds by one grid-stride kernel with warp-shuffle reduction and ordered-integer atomics -- no
//     cub::DeviceReduce, no device->host round trips (K/simple_knn.cu:200-207), no allocation (thrust vectors :201-223);
//   * 30-bit Morton codes sorted by the hand-written onesweep sort on 30 live bits (the reference sorts all 32);
//   * points are re-ordered into Morton order as float4 (x, y, z, original index) so the search streams them;
//   * a TWO-level box hierarchy, 32-point leaves inside 1024-point super boxes (the reference has only the
//     1024-point level, K/simple_knn.cu:19, so a surviving box costs 1024 distance evaluations; here it costs 32 box
//     tests + 32 evaluations per surviving leaf);
//   * one thread per query in Morton order, so the lanes of a warp walk nearly the same boxes and box / point loads
//     are warp-uniform broadcasts.
// The squared distance is evaluated with the reference binary's contraction, fma(dz,dz, fma(dx,dx, dy*dy))
// (read off its SASS), the running list is updated with the same strict `>` insertion (K/simple_knn.cu:131-148), and
// self is excluded by position, not by distance (:163,182) -- duplicates of a point are neighbours at distance 0.
#include <float.h>

#include "common.cuh"

namespace gsr {

namespace {

constexpr unsigned FULL = 0xffffffffu;
constexpr int LEAF = 32;              // points per leaf box
constexpr int SUPER = 32;             // leaves per super box
constexpr int MORTON_BITS = 30;
constexpr int KNN_MAX_K = 32;

struct KnnLayout {
    size_t bounds, keys0, keys1, vals0, vals1, hist, lookback, counters, pts, leaf, super, total;
    int nleaf, nsuper;
};

inline KnnLayout knn_layout(int64_t P) {
    KnnLayout L;
    size_t o = 0;
    auto take = [&](size_t b) { size_t r = o; o = align_up(o + b); return r; };
    const size_t n = (size_t)(P > 0 ? P : 0);
    L.nleaf = (int)((n + LEAF - 1) / LEAF);
    L.nsuper = (L.nleaf + SUPER - 1) / SUPER;
    L.bounds = take(8 * 4);
    L.keys0 = take(n * 4);
    L.keys1 = take(n * 4);
    L.vals0 = take(n * 4);
    L.vals1 = take(n * 4);
    L.hist = take((size_t)MAX_PASSES * RADIX * 4);
    L.lookback = take((size_t)MAX_PASSES * (sort_num_tiles((int64_t)n, 4) + 1) * RADIX * 4);
    L.counters = take(MAX_PASSES * 4);
    L.pts = take(n * 16);
    L.leaf = take((size_t)L.nleaf * 32);
    L.super = take((size_t)L.nsuper * 32);
    L.total = o;
    return L;
}

// order-preserving float <-> uint32 map, so atomicMin / atomicMax on integers order floats
__device__ __forceinline__ uint32_t f2ord(float f) {
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float ord2f(uint32_t o) {
    return __uint_as_float((o & 0x80000000u) ? (o & 0x7fffffffu) : ~o);
}

__global__ void knn_bounds_init_kernel(uint32_t* __restrict__ bounds) {
    if (threadIdx.x < 3) bounds[threadIdx.x] = 0xffffffffu;          // min x,y,z
    else if (threadIdx.x < 6) bounds[threadIdx.x] = 0u;               // max x,y,z
}

__global__ void __launch_bounds__(256)
knn_bounds_kernel(uint32_t P, const float* __restrict__ pts, uint32_t* __restrict__ bounds) {
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (uint32_t i = blockIdx.x * blockDim.x + threadIdx.x; i < P; i += gridDim.x * blockDim.x) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            const float v = pts[3 * (size_t)i + a];
            mn[a] = fminf(mn[a], v);
            mx[a] = fmaxf(mx[a], v);
        }
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(FULL, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(FULL, mx[a], o));
        }
    }
    if ((threadIdx.x & 31) == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            atomicMin(&bounds[a], f2ord(mn[a]));
            atomicMax(&bounds[3 + a], f2ord(mx[a]));
        }
    }
}

// K/simple_knn.cu:45-62
__device__ __forceinline__ uint32_t spread10(uint32_t x) {
    x = (x | (x << 16)) & 0x030000FFu;
    x = (x | (x << 8)) & 0x0300F00Fu;
    x = (x | (x << 4)) & 0x030C30C3u;
    x = (x | (x << 2)) & 0x09249249u;
    return x;
}

__global__ void __launch_bounds__(256)
knn_morton_kernel(uint32_t P, const float* __restrict__ pts, const uint32_t* __restrict__ bounds,
                  uint32_t* __restrict__ codes, uint32_t* __restrict__ ids) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    uint32_t code = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
        const float lo = ord2f(bounds[a]), hi = ord2f(bounds[3 + a]);
        const float ext = hi - lo;
        float t = ext > 0.0f ? (pts[3 * (size_t)i + a] - lo) / ext : 0.0f;
        t = fminf(fmaxf(t, 0.0f), 1.0f);
        code |= spread10((uint32_t)(t * 1023.0f)) << a;
    }
    codes[i] = code;
    ids[i] = i;
}

// One CTA of 1024 threads == one super box: re-orders its points into Morton order, reduces the 32 leaf boxes with
// warp shuffles (the reference's boxMinMax is a 10-step shared-memory tree over 24 KiB, K/simple_knn.cu:78-117).
__global__ void __launch_bounds__(LEAF * SUPER)
knn_reorder_boxes_kernel(uint32_t P, const float* __restrict__ pts, const uint32_t* __restrict__ sorted_ids,
                         float4* __restrict__ sorted_pts, float4* __restrict__ leaf_boxes,
                         float4* __restrict__ super_boxes) {
    __shared__ float s_mn[SUPER][3], s_mx[SUPER][3];
    const uint32_t i = blockIdx.x * (LEAF * SUPER) + threadIdx.x;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    if (i < P) {
        const uint32_t id = sorted_ids[i];
        const float x = pts[3 * (size_t)id], y = pts[3 * (size_t)id + 1], z = pts[3 * (size_t)id + 2];
        sorted_pts[i] = make_float4(x, y, z, __uint_as_float(id));
        mn[0] = mx[0] = x; mn[1] = mx[1] = y; mn[2] = mx[2] = z;
    }
#pragma unroll
    for (int a = 0; a < 3; ++a) {
#pragma unroll
        for (int o = 16; o >= 1; o >>= 1) {
            mn[a] = fminf(mn[a], __shfl_xor_sync(FULL, mn[a], o));
            mx[a] = fmaxf(mx[a], __shfl_xor_sync(FULL, mx[a], o));
        }
    }
    const uint32_t leaf = blockIdx.x * SUPER + warp;
    if (lane == 0) {
        if ((uint64_t)leaf * LEAF < P) {
            leaf_boxes[2 * (size_t)leaf] = make_float4(mn[0], mn[1], mn[2], 0.0f);
            leaf_boxes[2 * (size_t)leaf + 1] = make_float4(mx[0], mx[1], mx[2], 0.0f);
        }
#pragma unroll
        for (int a = 0; a < 3; ++a) { s_mn[warp][a] = mn[a]; s_mx[warp][a] = mx[a]; }
    }
    __syncthreads();
    if (warp == 0) {
#pragma unroll
        for (int a = 0; a < 3; ++a) {
            float m0 = s_mn[lane][a], m1 = s_mx[lane][a];
#pragma unroll
            for (int o = 16; o >= 1; o >>= 1) {
                m0 = fminf(m0, __shfl_xor_sync(FULL, m0, o));
                m1 = fmaxf(m1, __shfl_xor_sync(FULL, m1, o));
            }
            mn[a] = m0; mx[a] = m1;
        }
        if (lane == 0) {
            super_boxes[2 * (size_t)blockIdx.x] = make_float4(mn[0], mn[1], mn[2], 0.0f);
            super_boxes[2 * (size_t)blockIdx.x + 1] = make_float4(mx[0], mx[1], mx[2], 0.0f);
        }
    }
}

// distBoxPoint, K/simple_knn.cu:119-129 (squared distance from p to the box; 0 inside)
__device__ __forceinline__ float box_dist2(const float4 bmin, const float4 bmax, float px, float py, float pz) {
    float dx = 0.0f, dy = 0.0f, dz = 0.0f;
    if (px < bmin.x || px > bmax.x) dx = fminf(fabsf(px - bmin.x), fabsf(px - bmax.x));
    if (py < bmin.y || py > bmax.y) dy = fminf(fabsf(py - bmin.y), fabsf(py - bmax.y));
    if (pz < bmin.z || pz > bmax.z) dz = fminf(fabsf(pz - bmin.z), fabsf(pz - bmax.z));
    return dx * dx + dy * dy + dz * dz;
}

template <int KM>
struct Best {
    float d[KM];
    uint32_t id[KM];
    __device__ __forceinline__ void reset() {
#pragma unroll
        for (int j = 0; j < KM; ++j) { d[j] = FLT_MAX; id[j] = 0u; }
    }
    // updateKBest, K/simple_knn.cu:131-148: strict '>' keeps the earlier candidate on ties
    __device__ __forceinline__ void offer(float dist, uint32_t pid, int k) {
        if (d[KM - 1] > dist) {      // entries past k stay FLT_MAX, so this also guards the k < KM case
#pragma unroll
            for (int j = 0; j < KM; ++j) {
                const bool sw = d[j] > dist;
                const float t = d[j]; const uint32_t ti = id[j];
                d[j] = sw ? dist : t; id[j] = sw ? pid : ti;
                dist = sw ? t : dist; pid = sw ? ti : pid;
            }
        }
    }
};

// EXACT: k == KM at compile time (k = 3, 20: the list lives in registers); otherwise k <= KM at run time.
template <int KM, bool EXACT>
__global__ void __launch_bounds__(128)
knn_search_kernel(uint32_t P, int k_rt, const float4* __restrict__ spts, const float4* __restrict__ leaf_boxes,
                  const float4* __restrict__ super_boxes, int nleaf, int nsuper, float* __restrict__ mean_dist2,
                  int32_t* __restrict__ idx_out, float* __restrict__ dist2_out) {
    const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= P) return;
    const int k = EXACT ? KM : k_rt;
    const float4 me = spts[i];
    Best<KM> best;
    best.reset();
    auto dist2 = [&](const float4 q) {
        const float dx = __fsub_rn(q.x, me.x), dy = __fsub_rn(q.y, me.y), dz = __fsub_rn(q.z, me.z);
        return __fmaf_rn(dz, dz, __fmaf_rn(dx, dx, __fmul_rn(dy, dy)));
    };
    // seed: the +-k neighbours in Morton order give an upper bound on the k-th distance (K/simple_knn.cu:160-172)
    {
        const uint32_t lo = i >= (uint32_t)k ? i - k : 0u;
        const uint32_t hi = min(P - 1, i + (uint32_t)k);
        for (uint32_t j = lo; j <= hi; ++j) {
            if (j == i) continue;
            const float4 q = spts[j];
            best.offer(dist2(q), __float_as_uint(q.w), k);
        }
    }
    float reject = FLT_MAX;
#pragma unroll
    for (int j = 0; j < KM; ++j)
        if (j == k - 1) reject = best.d[j];
    best.reset();
    float kth = FLT_MAX;
    for (int s = 0; s < nsuper; ++s) {
        const float ds = box_dist2(super_boxes[2 * s], super_boxes[2 * s + 1], me.x, me.y, me.z);
        if (ds > reject || ds > kth) continue;
        const int l1 = min(nleaf, (s + 1) * SUPER);
        for (int l = s * SUPER; l < l1; ++l) {
            const float dl = box_dist2(leaf_boxes[2 * l], leaf_boxes[2 * l + 1], me.x, me.y, me.z);
            if (dl > reject || dl > kth) continue;
            const uint32_t p0 = (uint32_t)l * LEAF;
            const uint32_t p1 = min(P, p0 + LEAF);
            for (uint32_t j = p0; j < p1; ++j) {
                if (j == i) continue;
                const float4 q = spts[j];
                best.offer(dist2(q), __float_as_uint(q.w), k);
            }
#pragma unroll
            for (int j = 0; j < KM; ++j)
                if (j == k - 1) kth = best.d[j];
        }
    }
    const uint32_t self = __float_as_uint(me.w);
    float sum = 0.0f;
#pragma unroll
    for (int j = 0; j < KM; ++j) {
        if (j < k) {
            sum = __fadd_rn(sum, best.d[j]);                 // ((d0 + d1) + d2) as K/simple_knn.cu:186
            idx_out[(size_t)self * k + j] = (int32_t)best.id[j];
            if (dist2_out) dist2_out[(size_t)self * k + j] = best.d[j];
        }
    }
    mean_dist2[self] = __fdiv_rn(sum, (float)k);
}

}  // namespace
}  // namespace gsr

using namespace gsr;

extern "C" {

size_t gsr_knn_temp_bytes(int64_t P) { return knn_layout(P).total; }

int gsr_knn(int64_t P, int K, const float* points, float* mean_dist2, int32_t* idx, float* dist2, void* tmp,
            size_t tmp_bytes, gsr_stream_t stream) {
    cudaStream_t s = (cudaStream_t)stream;
    if (P < 0 || K < 1 || K > KNN_MAX_K) return GSR_ERR_INVALID_ARGUMENT;
    if (P == 0) return GSR_OK;
    if (P >= (1ll << 30)) return GSR_ERR_UNSUPPORTED;
    if (!points || !mean_dist2 || !idx || !tmp) return GSR_ERR_INVALID_ARGUMENT;
    const KnnLayout L = knn_layout(P);
    if (tmp_bytes < L.total) return GSR_ERR_WORKSPACE_TOO_SMALL;
    const uint32_t n = (uint32_t)P;
    uint32_t* bounds = at<uint32_t>(tmp, L.bounds);
    knn_bounds_init_kernel<<<1, 32, 0, s>>>(bounds);
    GSR_LAUNCH_CHECK();
    const int rb = (int)((n + 255) / 256 < 148 * 4 ? (n + 255) / 256 : 148 * 4);
    knn_bounds_kernel<<<rb, 256, 0, s>>>(n, points, bounds);
    GSR_LAUNCH_CHECK();
    knn_morton_kernel<<<(n + 255) / 256, 256, 0, s>>>(n, points, bounds, at<uint32_t>(tmp, L.keys0),
                                                      at<uint32_t>(tmp, L.vals0));
    GSR_LAUNCH_CHECK();
    int rc = radix_sort_pairs<uint32_t>(at<uint32_t>(tmp, L.keys0), at<uint32_t>(tmp, L.keys1),
                                        at<uint32_t>(tmp, L.vals0), at<uint32_t>(tmp, L.vals1), P, nullptr, 0,
                                        MORTON_BITS, at<uint32_t>(tmp, L.hist), at<uint32_t>(tmp, L.lookback),
                                        at<uint32_t>(tmp, L.counters), s);
    if (rc) return rc;
    constexpr int passes = (MORTON_BITS + RADIX_BITS - 1) / RADIX_BITS;
    const uint32_t* sorted_ids = at<uint32_t>(tmp, (passes & 1) ? L.vals1 : L.vals0);
    float4* spts = at<float4>(tmp, L.pts);
    float4* leaf = at<float4>(tmp, L.leaf);
    float4* super = at<float4>(tmp, L.super);
    knn_reorder_boxes_kernel<<<L.nsuper, LEAF * SUPER, 0, s>>>(n, points, sorted_ids, spts, leaf, super);
    GSR_LAUNCH_CHECK();
    const int sb = (int)((n + 127) / 128);
    if (K == 3)
        knn_search_kernel<3, true><<<sb, 128, 0, s>>>(n, K, spts, leaf, super, L.nleaf, L.nsuper, mean_dist2, idx, dist2);
    else if (K == 20)
        knn_search_kernel<20, true><<<sb, 128, 0, s>>>(n, K, spts, leaf, super, L.nleaf, L.nsuper, mean_dist2, idx, dist2);
    else if (K <= 8)
        knn_search_kernel<8, false><<<sb, 128, 0, s>>>(n, K, spts, leaf, super, L.nleaf, L.nsuper, mean_dist2, idx, dist2);
    else
        knn_search_kernel<KNN_MAX_K, false><<<sb, 128, 0, s>>>(n, K, spts, leaf, super, L.nleaf, L.nsuper, mean_dist2, idx,
                                                        dist2);
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // extern "C"
