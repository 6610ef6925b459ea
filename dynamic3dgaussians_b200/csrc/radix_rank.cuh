// Warp-level digit matching for the radix sort and its histograms.
//
// match_digit() returns, for every valid lane, the mask of valid lanes holding the same 8-bit digit.  It is eight
// ballots + bit logic rather than one MATCH.ANY: on B200 the two-pass tile sort spent its time in MATCH.ANY (16 per
// thread per pass; ncu launch list r01: histogram 390 us, scatter pass 360 us for 42.6 M keys), while ballots issue at
// full rate.  Digits narrower than 8 bits just have constant-zero upper bits.
#pragma once
#include "common.cuh"

namespace gsr {

template <int NBITS = RADIX_BITS>
__device__ __forceinline__ uint32_t match_digit(uint32_t d, bool valid) {
    uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < NBITS; ++b) {
        const bool bit = (d >> b) & 1u;
        const uint32_t bal = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? bal : ~bal;
    }
    return peers;
}

// Adds one row of 32 digits to a WARP-PRIVATE 256-bin histogram in shared memory without atomics (the first lane of
// every group of equal digits does a plain read-modify-write; groups have distinct digits, hence distinct addresses).
// Shared-memory atomics would cost ~2 cycles per lane on this part (B300_MICROARCH.md), i.e. more than the whole pass.
// Returns the histogram value BEFORE this row in the group's leader lane broadcast to the group (the rank base).
// NBITS: live bits of the digit (the last pass of a sort is often narrower than 8 bits: fewer ballots).
template <int NBITS = RADIX_BITS>
__device__ __forceinline__ uint32_t warp_hist_add(uint32_t* __restrict__ hist_warp, uint32_t d, bool valid, int lane,
                                                  uint32_t& peers_out) {
    const uint32_t peers = match_digit<NBITS>(d, valid);
    peers_out = peers;
    const int leader = __ffs(peers) - 1;
    uint32_t prev = 0;
    if (valid && lane == leader) {
        prev = hist_warp[d];
        hist_warp[d] = prev + __popc(peers);
    }
    __syncwarp();
    return __shfl_sync(0xffffffffu, prev, valid ? leader : 0);
}

}  // namespace gsr
