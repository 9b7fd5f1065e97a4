// Block-wide radix select and key helpers shared by rank.cu and sparse.cu.
#pragma once

#include "common.cuh"

namespace mcb {

__device__ __forceinline__ uint32_t lanemask_lt() {
    uint32_t mask;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(mask));
    return mask;
}

// ascending order of `sort_key`  ==  requested order of the values
__device__ __forceinline__ uint32_t sort_key(float value, bool ascending) {
    uint32_t key = float_to_ordered(canonical_zero(value));
    return ascending ? key : ~key;
}

// The lanes of the warp whose (valid) 8-bit digit equals this lane's: eight ballots.  MATCH.ANY gives
// the same mask in one instruction, but its latency grows with the number of distinct values in the
// warp (~230 cycles for random digits, measured with ncu on B200), which made it the bottleneck.
__device__ __forceinline__ uint32_t warp_peers_8bit(uint32_t digit, bool valid) {
    uint32_t peers = __ballot_sync(0xffffffffu, valid);
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        const bool bit = (digit >> b) & 1u;
        const uint32_t ballot = __ballot_sync(0xffffffffu, bit);
        peers &= bit ? ballot : ~ballot;
    }
    return peers;
}

// Shared state of one block-wide radix select.
struct SelectState {
    int hist[256];
    uint32_t prefix;
    int rank;
};

// The key of 0-based rank `rank` in ascending key order among key_at(0..n-1).  On return
// *residual is the rank of that key inside its group of equal keys (so `rank - *residual` keys
// are strictly smaller).  All threads of the block must call; all get the same result.
template <typename KeyAt>
__device__ uint32_t block_radix_select(KeyAt key_at, int n, int rank, SelectState* state, int* residual) {
    const int tid = threadIdx.x, lane = tid & 31;
    const uint32_t lower_lanes = lanemask_lt();
    if (tid == 0) {
        state->prefix = 0;
        state->rank = rank;
    }
    uint32_t known_mask = 0;
    for (int shift = 24; shift >= 0; shift -= 8) {
        for (int i = tid; i < 256; i += blockDim.x) state->hist[i] = 0;
        __syncthreads();
        const uint32_t prefix = state->prefix;
        const int rounds = (n + (int)blockDim.x - 1) / (int)blockDim.x;
        for (int round = 0; round < rounds; ++round) {
            const int i = round * (int)blockDim.x + tid;
            uint32_t digit = 0;
            bool valid = false;
            if (i < n) {
                const uint32_t key = key_at(i);
                valid = (key & known_mask) == prefix;
                digit = (key >> shift) & 255u;
            }
            const uint32_t peers = warp_peers_8bit(digit, valid);
            if (valid && (peers & lower_lanes) == 0)
                atomicAdd(&state->hist[digit], __popc(peers));
        }
        __syncthreads();
        if (tid < 32) {
            // lane owns 8 consecutive bins
            int counts[8], sum = 0;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                counts[b] = state->hist[lane * 8 + b];
                sum += counts[b];
            }
            const int inclusive = warp_inclusive_scan(sum, lane);
            int before = inclusive - sum;
            const int wanted = state->rank;
            __syncwarp();
            if (wanted >= before && wanted < inclusive) {
#pragma unroll
                for (int b = 0; b < 8; ++b) {
                    if (wanted >= before && wanted < before + counts[b]) {
                        state->prefix = prefix | ((uint32_t)(lane * 8 + b) << shift);
                        state->rank = wanted - before;
                    }
                    before += counts[b];
                }
            }
        }
        known_mask |= 255u << shift;
        __syncthreads();
    }
    *residual = state->rank;
    const uint32_t result = state->prefix;
    __syncthreads();
    return result;
}

}  // namespace mcb
