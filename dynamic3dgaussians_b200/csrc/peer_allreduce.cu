// (e) Sum of the per-Gaussian gradient slab over the ranks of one NVSwitch node, as ONE kernel over peer memory.
// Replaces nothing in the reference (it is single-GPU; REF/train.py trains one scene on one device): this is the exchange
// step of the view-sharded path, which round 1 handed to ncclAllReduce.
//
// Every rank's slab lives in symmetric memory (the same allocation mapped into every process; the host side exchanges
// the mappings, dist.PeerSlab), so a kernel on rank r can load from and store to the slabs of all ranks through NVLink:
//
//   barrier A   CTA c of every rank tells CTA c of every peer "my rank's backward kernels are done" (stream order put
//               them before this kernel) and waits for the same word from every peer;
//   reduce      rank r owns slice r of the slab: it loads that slice from all `world` slabs (world - 1 of them remote),
//               adds them in rank order 0, 1, ... -- the same order on every rank, so every rank ends up with the same
//               bits -- and stores the sum into slice r of ALL slabs (the all-gather as a push);
//   barrier B   "my pushes are visible": after it, every CTA c of every peer has finished pushing, and since a kernel ends
//               when all its CTAs have, the whole slab is complete on every rank at kernel end.
//
// Per rank and step: (world - 1) / world of the slab in over NVLink and as much out, two flag round trips, no staging
// copy, no second kernel, nothing on the host -- the kernel sits in the step's CUDA graph.  The flag words are epoch
// counters kept in the symmetric flag block itself, so a captured launch can be replayed forever.
//
// With multicast (NVLS) mappings the reduce is one multimem.ld_reduce per 16 bytes (the switch adds the `world` copies)
// and the push one multimem.st.
#include "common.cuh"

namespace gsr {

namespace {

constexpr int PEER_MAX_WORLD = GSR_PEER_MAX_WORLD;        // 8
constexpr int PEER_MAX_CTAS = GSR_PEER_MAX_CTAS;          // 256
// flag block of one rank, uint32 words: [2 barriers][PEER_MAX_CTAS][PEER_MAX_WORLD] arrival words written by the peers,
// then [PEER_MAX_CTAS] epochs (local), then one error word (local)
constexpr int PEER_EPOCH_OFF = 2 * PEER_MAX_CTAS * PEER_MAX_WORLD;
constexpr int PEER_ERR_OFF = PEER_EPOCH_OFF + PEER_MAX_CTAS;
static_assert(PEER_ERR_OFF + 1 <= GSR_PEER_FLAG_WORDS, "flag block layout");

struct PeerPtrs {
    float* buf[PEER_MAX_WORLD];
    uint32_t* flags[PEER_MAX_WORLD];
};

__device__ __forceinline__ void st_relaxed_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_relaxed_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.relaxed.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ float4 ld_peer(const float4* p) {      // not through L1: the line may be from the last step
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ float4 multimem_ld_reduce(const float4* mc) {
    float4 v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
                 : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void multimem_st(float4* mc, float4 v) {
    asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "f"(v.x), "f"(v.y), "f"(v.z),
                 "f"(v.w) : "memory");
}

// One cross-rank barrier of CTA `cta`: threads 0..world-1 each handle one peer.  The arrival words are relaxed
// system-scope stores / loads; ordering comes from ONE system-scope fence where it is needed:
//   A (RELEASE = false): nothing this kernel wrote has to be published -- the data the peers are about to load was written
//     by EARLIER kernels of this rank's stream, complete and in this GPU's L2 (where peer loads are served from) before
//     this kernel started; the peer loads that follow the wait are volatile (no stale L1 line) and not issued before the
//     wait's last load returns;
//   B (RELEASE = true): this CTA's pushes must have landed in the peers' memory before its arrival word does: bar.sync
//     orders every thread's stores before the signalling thread's fence.sys (cumulativity), the fence before the word.
//     Nothing is loaded after the wait; the NEXT kernel reads what the peers pushed, and their words arrived after their
//     own fences.
// A peer that never arrives (a rank that died) must not hang the GPU: after ~4 s the wait gives up and leaves a mark the
// host can read (word GSR_PEER_ERROR_WORD).
template <bool RELEASE>
__device__ __forceinline__ void peer_barrier(const PeerPtrs& pp, int world, int rank, int which, int cta, uint32_t epoch) {
    const int tid = threadIdx.x;
    __syncthreads();
    if (tid < world) {
        const int slot = (which * PEER_MAX_CTAS + cta) * PEER_MAX_WORLD;
        if (RELEASE) __threadfence_system();
        st_relaxed_sys(pp.flags[tid] + slot + rank, epoch);
        const uint32_t* mine = pp.flags[rank] + slot + tid;
        const long long t0 = clock64();
        while ((int32_t)(ld_relaxed_sys(mine) - epoch) < 0) {
            if (clock64() - t0 > 8000000000ll) {
                pp.flags[rank][PEER_ERR_OFF] = 1u;
                break;
            }
        }
    }
    __syncthreads();
}

// MODE: GSR_PEER_ALLREDUCE (sum everywhere) | GSR_PEER_REDUCE_SCATTER (rank r keeps the sum of slice r only; nothing is
// pushed) | GSR_PEER_ALLGATHER (rank r's slice r goes to every peer as it is) | MODE_MULTIMEM (all-reduce in the switch).
constexpr int MODE_MULTIMEM = 3;

template <int WORLD, int MODE>
__global__ void __launch_bounds__(512)
peer_allreduce_kernel(const __grid_constant__ PeerPtrs pp, float* __restrict__ mc_buf, int world_rt, int rank, size_t n4) {
    pdl_wait();          // (common.cuh) nothing global is touched before the predecessor kernel has completed
    pdl_trigger();
    constexpr int WMAX = WORLD > 0 ? WORLD : PEER_MAX_WORLD;
    const int world = WORLD > 0 ? WORLD : world_rt;
    const int cta = blockIdx.x, tid = threadIdx.x;
    __shared__ uint32_t s_epoch;
    uint32_t* my_flags = pp.flags[rank];
    if (tid == 0) s_epoch = my_flags[PEER_EPOCH_OFF + cta] + 1u;
    __syncthreads();
    const uint32_t epoch = s_epoch;
    // A: every peer's stream has reached its call -- its buffer holds this step's data and is no longer read by
    // anything it enqueued earlier
    peer_barrier<false>(pp, world, rank, 0, cta, epoch);

    const size_t per = (n4 + world - 1) / world;
    const size_t lo = per * rank < n4 ? per * rank : n4;
    const size_t hi = lo + per < n4 ? lo + per : n4;
    const size_t stride = (size_t)gridDim.x * blockDim.x;
    // Remote loads take microseconds: a thread keeps U x world of them in flight (at world = 8 a thread has only one or
    // two 16-byte elements of its rank's slice to do, at world = 2 it has seven)
    constexpr int U = MODE == MODE_MULTIMEM ? 4 : (WORLD == 2 ? 4 : (WORLD == 4 ? 2 : 1));
    for (size_t i0 = lo + (size_t)cta * blockDim.x + tid; i0 < hi; i0 += stride * U) {
        if (MODE == MODE_MULTIMEM) {
            float4* mc = reinterpret_cast<float4*>(mc_buf);
            float4 v[U];
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (i0 + u * stride < hi) v[u] = multimem_ld_reduce(mc + i0 + u * stride);
#pragma unroll
            for (int u = 0; u < U; ++u)
                if (i0 + u * stride < hi) multimem_st(mc + i0 + u * stride, v[u]);
        } else if (MODE == GSR_PEER_ALLGATHER) {
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
                if (i < hi) {
                    const float4 v = reinterpret_cast<const float4*>(pp.buf[rank])[i];
#pragma unroll
                    for (int p = 0; p < WMAX; ++p)
                        if (p < world && p != rank) reinterpret_cast<float4*>(pp.buf[p])[i] = v;
                }
            }
        } else {
            float4 v[U][WMAX];
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
#pragma unroll
                for (int p = 0; p < WMAX; ++p)
                    if (p < world && i < hi) v[u][p] = ld_peer(reinterpret_cast<const float4*>(pp.buf[p]) + i);
            }
#pragma unroll
            for (int u = 0; u < U; ++u) {
                const size_t i = i0 + u * stride;
                if (i >= hi) continue;
                float4 acc = v[u][0];
#pragma unroll
                for (int p = 1; p < WMAX; ++p)
                    if (p < world) {
                        acc.x += v[u][p].x; acc.y += v[u][p].y; acc.z += v[u][p].z; acc.w += v[u][p].w;
                    }
                if (MODE == GSR_PEER_REDUCE_SCATTER) {
                    reinterpret_cast<float4*>(pp.buf[rank])[i] = acc;
                } else {
#pragma unroll
                    for (int p = 0; p < WMAX; ++p)
                        if (p < world) reinterpret_cast<float4*>(pp.buf[p])[i] = acc;
                }
            }
        }
    }
    // B: every peer is done reading this rank's buffer and its pushes into it have landed
    peer_barrier<true>(pp, world, rank, 1, cta, epoch);
    if (tid == 0) my_flags[PEER_EPOCH_OFF + cta] = epoch;
}

}  // namespace

int launch_peer_allreduce(int mode, int world, int rank, const void* const* bufs_host, const void* const* flags_host,
                          void* multicast_buf, int64_t n_floats, int blocks, cudaStream_t s) {
    if (world < 1 || world > PEER_MAX_WORLD || rank < 0 || rank >= world) return GSR_ERR_INVALID_ARGUMENT;
    if (mode != GSR_PEER_ALLREDUCE && mode != GSR_PEER_REDUCE_SCATTER && mode != GSR_PEER_ALLGATHER)
        return GSR_ERR_INVALID_ARGUMENT;
    if (n_floats < 0 || (n_floats & 3) != 0) return GSR_ERR_INVALID_ARGUMENT;        // whole float4s
    if (multicast_buf && mode != GSR_PEER_ALLREDUCE) return GSR_ERR_INVALID_ARGUMENT;
    // measured on 8 B200s, 16.8 MB: P2P 67 us at 148 CTAs (77 at 32, 82 at 256); multimem 57 us at 32 CTAs (60 at 148)
    if (blocks <= 0) blocks = multicast_buf ? 32 : 148;
    if (blocks > PEER_MAX_CTAS) blocks = PEER_MAX_CTAS;
    PeerPtrs pp;
    memset(&pp, 0, sizeof(pp));
    for (int p = 0; p < world; ++p) {
        if (!bufs_host[p] || !flags_host[p] || ((uintptr_t)bufs_host[p] & 15u)) return GSR_ERR_INVALID_ARGUMENT;
        pp.buf[p] = (float*)bufs_host[p];
        pp.flags[p] = (uint32_t*)flags_host[p];
    }
    const size_t n4 = (size_t)n_floats / 4;
    float* mc = (float*)multicast_buf;
#define GSR_PEER(W_, M_) launch_kernel(peer_allreduce_kernel<W_, M_>, blocks, 512, 0, s, pp, mc, world, rank, n4)
#define GSR_PEER_W(M_)                                                                                                 \
    switch (world) {                                                                                                   \
        case 2: GSR_PEER(2, M_); break;                                                                                \
        case 4: GSR_PEER(4, M_); break;                                                                                \
        case 8: GSR_PEER(8, M_); break;                                                                                \
        default: GSR_PEER(0, M_); break;                                                                               \
    }
    if (mc) {
        GSR_PEER(0, MODE_MULTIMEM);
    } else if (mode == GSR_PEER_ALLREDUCE) {
        GSR_PEER_W(GSR_PEER_ALLREDUCE)
    } else if (mode == GSR_PEER_REDUCE_SCATTER) {
        GSR_PEER_W(GSR_PEER_REDUCE_SCATTER)
    } else {
        GSR_PEER_W(GSR_PEER_ALLGATHER)
    }
#undef GSR_PEER_W
#undef GSR_PEER
    GSR_LAUNCH_CHECK();
    return GSR_OK;
}

}  // namespace gsr
