// Packed butterfly reduction of K per-lane values over a warp (shared by the blend backward kernels).
#pragma once
#include "common.cuh"

namespace gsr {

// Packed butterfly: after run(), v[0] of lane l holds the warp-wide sum of the original v[slot(l)].
template <int K>
struct PackedReduce {
    template <int KC, int OFF>
    __device__ __forceinline__ static void step(float (&v)[K], int lane) {
        if constexpr (OFF >= 1) {
            constexpr int H = (KC + 1) / 2;
            constexpr int PAIRS = KC / 2;
            const bool up = (lane & OFF) != 0;
#pragma unroll
            for (int i = 0; i < PAIRS; ++i) {
                const float send = up ? v[i] : v[i + H];
                const float keep = up ? v[i + H] : v[i];
                v[i] = keep + __shfl_xor_sync(0xffffffffu, send, OFF);
            }
            if constexpr ((KC & 1) != 0) v[H - 1] += __shfl_xor_sync(0xffffffffu, v[H - 1], OFF);
            step<H, OFF / 2>(v, lane);
        }
    }
    __device__ __forceinline__ static void run(float (&v)[K], int lane) { step<K, 16>(v, lane); }

    template <int KC, int OFF>
    __device__ __forceinline__ static void slot_step(int (&idx)[K], int lane) {
        if constexpr (OFF >= 1) {
            constexpr int H = (KC + 1) / 2;
            constexpr int PAIRS = KC / 2;
            const bool up = (lane & OFF) != 0;
#pragma unroll
            for (int i = 0; i < PAIRS; ++i) idx[i] = up ? idx[i + H] : idx[i];
            slot_step<H, OFF / 2>(idx, lane);
        }
    }
    __device__ __forceinline__ static int slot(int lane) {
        int idx[K];
#pragma unroll
        for (int i = 0; i < K; ++i) idx[i] = i;
        slot_step<K, 16>(idx, lane);
        return idx[0];
    }
};

}  // namespace gsr
