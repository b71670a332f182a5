// Decoupled look-back tile prefix (single-pass chained scan) used by every
// compaction kernel on the path. One 64-bit status word per tile:
//   bits 63..62  flag (0 invalid, 1 tile aggregate, 2 inclusive prefix)
//   bits 61..0   value
// Tiles must be numbered in launch order (atomic ticket) so that every
// predecessor of a running tile is itself running or finished.
#pragma once
#include "mtg_common.cuh"

namespace mtg {

constexpr unsigned long long SCAN_FLAG_AGG = 1ull << 62;
constexpr unsigned long long SCAN_FLAG_INCL = 2ull << 62;
constexpr unsigned long long SCAN_VALUE_MASK = (1ull << 62) - 1;

__device__ __forceinline__ uint64_t warp_sum_u64(uint64_t v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    return __shfl_sync(0xffffffffu, v, 0);
}

// Called by ALL threads of the block (contains __syncthreads). Thread 0 supplies
// `aggregate`; returns the exclusive prefix of tile `tile` to every thread.
// `slot` is a block-shared uint64_t.
__device__ __forceinline__ uint64_t lookback_exclusive(unsigned long long *status, uint32_t tile,
                                                       uint64_t aggregate, uint64_t *slot) {
    if (threadIdx.x < 32) {
        const unsigned lane = threadIdx.x;
        aggregate = __shfl_sync(0xffffffffu, aggregate, 0);
        uint64_t excl = 0;
        if (tile == 0) {
            if (lane == 0) st_relaxed_u64(status, SCAN_FLAG_INCL | aggregate);
        } else {
            if (lane == 0) st_relaxed_u64(status + tile, SCAN_FLAG_AGG | aggregate);
            // LBW x 32 predecessors are requested per round, nearest window first. Measured on B200
            // (scripts/unique_bench.py, 1.2 G keys): LBW 1 -> 3.84 ms, 4 -> 4.21 ms; without any
            // look-back the same kernel takes 1.66 ms, i.e. the wait is for the predecessors' tile
            // loads (a tile cannot finish before its 32 predecessors have counted), not for the
            // window width.
            constexpr int LBW = 1;
            long long base = (long long)tile - 1;
            bool done = false;
            while (!done) {
                unsigned long long s[LBW];
#pragma unroll
                for (int k = 0; k < LBW; ++k) {
                    long long idx = base - 32 * k - (long long)lane;
                    s[k] = idx >= 0 ? ld_relaxed_u64(status + idx) : SCAN_FLAG_INCL;
                }
#pragma unroll
                for (int k = 0; k < LBW; ++k) {
                    if (done) continue;
                    long long idx = base - 32 * k - (long long)lane;
                    while (__any_sync(0xffffffffu, (s[k] >> 62) == 0)) {
                        if ((s[k] >> 62) == 0) s[k] = ld_relaxed_u64(status + idx);
                    }
                    unsigned incl = __ballot_sync(0xffffffffu, (s[k] >> 62) == 2);
                    uint64_t v = s[k] & SCAN_VALUE_MASK;
                    if (incl) {
                        unsigned first = __ffs(incl) - 1;
                        excl += warp_sum_u64(lane <= first ? v : 0);
                        done = true;
                    } else {
                        excl += warp_sum_u64(v);
                    }
                }
                base -= 32 * LBW;
            }
            if (lane == 0) st_relaxed_u64(status + tile, SCAN_FLAG_INCL | (excl + aggregate));
        }
        if (lane == 0) *slot = excl;
    }
    __syncthreads();
    return *slot;
}

// Block-wide exclusive scan of one uint32 per thread (BLOCK threads, multiple of 32).
// Returns the exclusive prefix; *total (shared) receives the block sum.
template <int BLOCK>
__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t *warp_tot /*[BLOCK/32+1]*/) {
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int w = 0; w < BLOCK / 32; ++w) {
            uint32_t t = warp_tot[w];
            warp_tot[w] = run;
            run += t;
        }
        warp_tot[BLOCK / 32] = run;
    }
    __syncthreads();
    return warp_tot[warp] + inc - v;
}

}  // namespace mtg
