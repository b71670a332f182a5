// Onesweep LSD radix sort kernels. See mtg_sort.cuh.
#include <cstdlib>

#include "mtg_sort.cuh"

namespace mtg {

namespace {

__device__ __forceinline__ uint32_t digit_of(uint64_t key, uint32_t val, const DigitPass &d) {
    uint32_t m = (1u << d.bits) - 1;
    return d.src == 0 ? (uint32_t)(key >> d.shift) & m : (val >> d.shift) & m;
}
__device__ __forceinline__ uint32_t digit_of(u128 key, uint32_t val, const DigitPass &d) {
    uint32_t m = (1u << d.bits) - 1;
    if (d.src != 0) return (val >> d.shift) & m;
    uint64_t x;
    if (d.shift >= 64) x = key.hi >> (d.shift - 64);
    else if (d.shift + d.bits <= 64) x = key.lo >> d.shift;
    else x = (key.lo >> d.shift) | (key.hi << (64 - d.shift));
    return (uint32_t)x & m;
}

// ---------------------------------------------------------------------------
// Kernel S1: all digit histograms in one read of the input.
// hist[p * RADIX + d] += #keys with digit d in pass p.
// ---------------------------------------------------------------------------
// NP > 0: the number of passes is a compile-time constant and the digit positions live in
// registers (4 instructions per digit instead of ~20 with the plan read per key); NP = 0: any plan.
template <typename KeyT, bool HAS_VAL, int NP>
__global__ void __launch_bounds__(512)
sort_histogram_kernel(const KeyT *__restrict__ keys, const uint32_t *__restrict__ vals, uint64_t n,
                      SortPlan plan, unsigned long long *__restrict__ hist) {
    extern __shared__ unsigned int sh_hist[];   // [plan.n * RADIX]
    const int np = NP ? NP : plan.n;
    for (int i = threadIdx.x; i < np * RADIX; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    DigitPass dp[NP ? NP : 1];
#pragma unroll
    for (int p = 0; p < NP; ++p) dp[p] = plan.p[p];
    // 4 independent coalesced loads per thread and iteration (64 KB in flight per SM)
    constexpr int HU = 4;
    const uint64_t chunk = (uint64_t)blockDim.x * HU;
    for (uint64_t base = (uint64_t)blockIdx.x * chunk; base < n; base += (uint64_t)gridDim.x * chunk) {
        KeyT k[HU];
        uint32_t v[HU];
#pragma unroll
        for (int u = 0; u < HU; ++u) {
            const uint64_t i = base + (uint64_t)u * blockDim.x + threadIdx.x;
            if (i < n) {
                k[u] = keys[i];
                v[u] = HAS_VAL ? vals[i] : 0;
            }
        }
#pragma unroll
        for (int u = 0; u < HU; ++u) {
            const uint64_t i = base + (uint64_t)u * blockDim.x + threadIdx.x;
            if (i < n) {
                if (NP) {
#pragma unroll
                    for (int p = 0; p < NP; ++p) atomicAdd(&sh_hist[p * RADIX + digit_of(k[u], v[u], dp[p])], 1u);
                } else {
                    for (int p = 0; p < np; ++p) atomicAdd(&sh_hist[p * RADIX + digit_of(k[u], v[u], plan.p[p])], 1u);
                }
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < np * RADIX; i += blockDim.x) {
        unsigned int c = sh_hist[i];
        if (c) atomicAdd(&hist[i], (unsigned long long)c);
    }
}

template <typename KeyT, bool HAS_VAL, int NP>
void launch_histogram(unsigned grid, size_t sh, const KeyT *keys, const uint32_t *vals, uint64_t n,
                      const SortPlan &plan, unsigned long long *hist, cudaStream_t stream) {
    auto kern = sort_histogram_kernel<KeyT, HAS_VAL, NP>;
    MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
    kern<<<grid, 512, sh, stream>>>(keys, vals, n, plan, hist);
}

// Kernel S2: exclusive scan of every pass' 256 bins (one block per pass).
__global__ void sort_scan_hist_kernel(unsigned long long *hist) {
    __shared__ unsigned long long sh[RADIX];
    unsigned long long *h = hist + (size_t)blockIdx.x * RADIX;
    sh[threadIdx.x] = h[threadIdx.x];
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long run = 0;
        for (int i = 0; i < RADIX; ++i) {
            unsigned long long t = sh[i];
            sh[i] = run;
            run += t;
        }
    }
    __syncthreads();
    h[threadIdx.x] = sh[threadIdx.x];
}

// L2 prefetch of a whole tile (TMA bulk prefetch, one instruction): the tile that will be loaded
// `ahead` tickets later is pulled from DRAM into L2 while this block works
__device__ __forceinline__ void prefetch_tile_l2(const void *p, uint32_t bytes) {
    asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(p), "r"(bytes) : "memory");
}

constexpr unsigned long long OS_VALUE_MASK = (1ull << 56) - 1;
__device__ __forceinline__ unsigned long long os_pack(unsigned flag, unsigned epoch,
                                                      unsigned long long v) {
    return ((unsigned long long)flag << 62) | ((unsigned long long)epoch << 56) | v;
}

// Decoupled look-back of one digit column: sum of the aggregates of the predecessor tiles back
// to the nearest published inclusive prefix. LB status words are requested per step so that
// LB L2 round trips overlap instead of one dependent load per predecessor.
template <int LB>
__device__ __forceinline__ unsigned long long lookback_digit(const unsigned long long *status,
                                                             unsigned tile, unsigned tid, unsigned epoch) {
    unsigned long long excl = 0;
    long long t = (long long)tile - 1;
    while (t >= 0) {
        unsigned long long s[LB];
#pragma unroll
        for (int j = 0; j < LB; ++j)
            s[j] = t - j >= 0 ? ld_relaxed_u64(status + (size_t)(t - j) * RADIX + tid)
                              : os_pack(2, epoch, 0);
        bool done = false;
#pragma unroll
        for (int j = 0; j < LB; ++j) {
            if (done) continue;
            unsigned flag = (unsigned)(s[j] >> 62);
            while (flag == 0 || ((s[j] >> 56) & 63) != epoch) {   // not published yet
                s[j] = ld_relaxed_u64(status + (size_t)(t - j) * RADIX + tid);
                flag = (unsigned)(s[j] >> 62);
            }
            excl += s[j] & OS_VALUE_MASK;
            if (flag == 2) done = true;
        }
        if (done) break;
        t -= LB;
    }
    return excl;
}

// ---------------------------------------------------------------------------
// Kernel S3 (v3): MATCH-free onesweep pass.
// Measured on B200 (scripts/micro/match_bench.cu): MATCH.ANY sustains one warp-instruction per
// ~62 cycles per SM, 8 ballots ~28 cycles, a shared-memory atomic (random bank, with or without
// return) ~2.5 cycles. One MATCH per key made v1 MATCH-throughput-bound (128 x 62 cycles per
// 4096-key tile = the measured tile rate). Here
//   * the tile histogram ("early counts") comes from shared-memory atomicAdd -> the aggregate is
//     published before ranking;
//   * the peer mask of a round is built by atomicOr(1 << lane) into a per-warp, per-digit mask
//     word and read back -- the same value MATCH.ANY returns, at shared-memory-atomic cost;
//   * rank = running per-warp digit offset (already holding the tile position) + popc(peers
//     below my lane): stable, deterministic, no dependence on atomic ordering.
// ---------------------------------------------------------------------------
// ROUTE: the keys of digit d go to the array at address route[d] (local or peer-mapped memory of
// another GPU) instead of out_keys + digit_offsets[d]: the scatter of the pass IS the exchange
// (grouping by routing bucket + all-to-all in one kernel; its stores leave over NVLink).
template <typename KeyT, bool HAS_VAL, int THREADS, int IPT, int MINB, int LB, bool ROUTE = false>
__global__ void __launch_bounds__(THREADS, MINB)
onesweep_pass_v3_kernel(const KeyT *__restrict__ in_keys, KeyT *__restrict__ out_keys,
                        const uint32_t *__restrict__ in_vals, uint32_t *__restrict__ out_vals,
                        uint64_t n, DigitPass dp, const unsigned long long *__restrict__ digit_offsets,
                        unsigned long long *__restrict__ status, unsigned int *tile_counter,
                        unsigned epoch, const unsigned char *__restrict__ route_owner = nullptr,
                        unsigned prefetch_ahead = 0) {
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * IPT;
    static_assert(THREADS >= RADIX, "one digit per thread in the scan phases");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    KeyT *sh_keys = reinterpret_cast<KeyT *>(smem_raw);                       // [TILE]
    uint32_t *sh_vals = reinterpret_cast<uint32_t *>(sh_keys + TILE);        // [HAS_VAL ? TILE : 0]
    uint32_t *whist = sh_vals + (HAS_VAL ? TILE : 0);                        // [WARPS][RADIX]
    uint32_t *wmask = whist + WARPS * RADIX;                                 // [WARPS][RADIX]
    long long *goff = reinterpret_cast<long long *>(wmask + WARPS * RADIX);  // [RADIX]
    uint32_t *warp_tot = reinterpret_cast<uint32_t *>(goff + RADIX);         // [WARPS + 1]
    __shared__ unsigned int sh_tile;
    // ROUTE: the scatter digit is the OWNER of the key's routing bucket (a handful of values), so a
    // tile leaves in a few long contiguous runs, which is what the NVLink stores need
    __shared__ unsigned char sh_owner[ROUTE ? RADIX : 1];
    auto dig = [&](KeyT k, uint32_t v) -> uint32_t {
        const uint32_t d = digit_of(k, v, dp);
        return ROUTE ? (uint32_t)sh_owner[d] : d;
    };

    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (ROUTE && tid < RADIX) sh_owner[tid] = route_owner[tid];
    if (tid == 0) {
        const unsigned t = atomicAdd(tile_counter, 1u);
        sh_tile = t;
        if (prefetch_ahead && ((uint64_t)t + prefetch_ahead + 1) * TILE <= n && (reinterpret_cast<uintptr_t>(in_keys) & 15) == 0) {
            prefetch_tile_l2(in_keys + ((uint64_t)t + prefetch_ahead) * TILE, (uint32_t)(sizeof(KeyT) * TILE));
            if (HAS_VAL && (reinterpret_cast<uintptr_t>(in_vals) & 15) == 0) prefetch_tile_l2(in_vals + ((uint64_t)t + prefetch_ahead) * TILE, (uint32_t)(4 * TILE));
        }
    }
    for (int i = tid; i < 2 * WARPS * RADIX; i += THREADS) whist[i] = 0;   // whist and wmask
    __syncthreads();
    const unsigned tile = sh_tile;
    const uint64_t tile_base = (uint64_t)tile * TILE;
    const uint32_t tile_valid = (uint32_t)(n - tile_base < (uint64_t)TILE ? n - tile_base : TILE);
    const bool full = tile_valid == TILE;

    // ---- load (coalesced, warp-striped) + early per-warp digit counts ----
    KeyT key[IPT];
    uint32_t val[IPT];
    uint32_t *my_hist = whist + warp * RADIX;
    uint32_t *my_mask = wmask + warp * RADIX;
    const uint32_t warp_base = warp * (IPT * 32);
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        uint32_t idx = warp_base + r * 32 + lane;
        if (full || idx < tile_valid) {
            key[r] = in_keys[tile_base + idx];
            val[r] = HAS_VAL ? in_vals[tile_base + idx] : 0;
        } else {
            key[r] = WordTraits<KeyT>::zero();
            val[r] = 0;
        }
    }
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        uint32_t idx = warp_base + r * 32 + lane;
        if (full || idx < tile_valid) atomicAdd(&my_hist[dig(key[r], val[r])], 1u);
    }
    __syncthreads();

    // ---- per digit: tile count, early publish, tile-local digit start ----
    uint32_t count = 0;
    if (tid < RADIX) {
#pragma unroll
        for (int w = 0; w < WARPS; ++w) count += whist[w * RADIX + tid];
        if (tile > 0)
            st_relaxed_u64(status + (size_t)tile * RADIX + tid, os_pack(1, epoch, count));
    }
    uint32_t inc = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t local_base = 0;
    if (tid < RADIX) {
        for (unsigned w = 0; w < warp; ++w) local_base += warp_tot[w];
        local_base += inc - count;
        // whist[w][d] := tile position of warp w's first key with digit d
        uint32_t run = local_base;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            uint32_t t = whist[w * RADIX + tid];
            whist[w * RADIX + tid] = run;
            run += t;
        }
    }
    __syncthreads();

    // ---- rank (peer mask through shared memory) + scatter into tile-sorted order ----
    const unsigned lt = lanemask_lt();
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        const uint32_t idx = warp_base + r * 32 + lane;
        const bool valid = full || idx < tile_valid;
        const uint32_t d = dig(key[r], val[r]);
        if (valid) atomicOr(&my_mask[d], 1u << lane);
        __syncwarp();
        unsigned peers = 0;
        uint32_t before = 0;
        if (valid) { peers = my_mask[d]; before = my_hist[d]; }
        __syncwarp();
        if (valid) {
            if ((peers & lt) == 0) {   // lowest lane of the group
                my_mask[d] = 0;
                my_hist[d] = before + __popc(peers);
            }
            uint32_t pos = before + __popc(peers & lt);
            sh_keys[pos] = key[r];
            if (HAS_VAL) sh_vals[pos] = val[r];
        }
        __syncwarp();
    }

    // ---- decoupled look-back per digit ----
    if (tid < RADIX) {
        unsigned long long excl = 0;
        if (LB == 0) {
            if (tile > 0) {
                long long t = (long long)tile - 1;
                while (true) {
                    unsigned long long s = ld_relaxed_u64(status + (size_t)t * RADIX + tid);
                    unsigned flag = (unsigned)(s >> 62);
                    if (flag == 0 || ((s >> 56) & 63) != epoch) continue;   // not published yet
                    excl += s & OS_VALUE_MASK;
                    if (flag == 2) break;
                    --t;
                }
            }
        } else {
            excl = lookback_digit<(LB > 0 ? LB : 1)>(status, tile, tid, epoch);
        }
        st_relaxed_u64(status + (size_t)tile * RADIX + tid, os_pack(2, epoch, excl + count));
        if (ROUTE)   // digit_offsets = the 2^bits destination addresses; goff = byte address of tile slot 0
            goff[tid] = (long long)digit_offsets[tid]
                        + ((long long)excl - (long long)local_base) * (long long)sizeof(KeyT);
        else
            goff[tid] = (long long)(digit_offsets[tid] + excl) - (long long)local_base;
    }
    __syncthreads();

    // ---- coalesced write-out ----
#pragma unroll 4
    for (uint32_t p = tid; p < tile_valid; p += THREADS) {
        KeyT k = sh_keys[p];
        uint32_t v = HAS_VAL ? sh_vals[p] : 0;
        uint32_t d = dig(k, v);
        if (ROUTE) {
            *reinterpret_cast<KeyT *>(goff[d] + (long long)p * (long long)sizeof(KeyT)) = k;
        } else {
            size_t dst = (size_t)(goff[d] + (long long)p);
            out_keys[dst] = k;
            if (HAS_VAL) out_vals[dst] = v;
        }
    }
}

// ---------------------------------------------------------------------------
// Kernel S3 (v5): the default pass. Same tile structure as v3 (early per-warp digit counts by shared
// atomics, early publish, tile-sorted staging, decoupled look-back, coalesced write-out); what changes is
// the ranking, which is where v3 spends its shared-memory bandwidth (ncu, profiles/r01_ncu_full_v6_82ms.txt:
// L1TEX 81 %, 57 % of the wavefronts bank conflicts): per key v3 issues atomicOr(mask), a mask read-back,
// a counter read and two leader stores. Here a key takes its slot with ONE atomicAdd (with return) on the
// running per-warp digit offset:
//   UNSTABLE: that is all -- lanes of one warp-row that share a digit get consecutive slots in
//     arbitrary order. Legal for the FIRST pass of an LSD sort (no earlier order to preserve) and for the
//     k-mer routing scatter.
//   stable: the lanes read the offset before the atomics of the row (`before`); a lane whose slot differs
//     from `before` shares its digit with another lane of the row (12 % of the keys of a random 8-bit
//     digit). Only those groups are repaired: per group one shuffle + one ballot give the group mask and
//     slot = before + popc(group & lanes below). ~1.9 groups per row on random digits; rows without
//     a collision pay one ballot.
// ---------------------------------------------------------------------------
template <typename KeyT, bool HAS_VAL, int THREADS, int IPT, int MINB, bool UNSTABLE, bool ROUTE = false>
__global__ void __launch_bounds__(THREADS, MINB)
onesweep_pass_v5_kernel(const KeyT *__restrict__ in_keys, KeyT *__restrict__ out_keys,
                        const uint32_t *__restrict__ in_vals, uint32_t *__restrict__ out_vals,
                        uint64_t n, DigitPass dp, const unsigned long long *__restrict__ digit_offsets,
                        unsigned long long *__restrict__ status, unsigned int *tile_counter,
                        unsigned epoch, const unsigned char *__restrict__ route_owner = nullptr,
                        unsigned prefetch_ahead = 0) {
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * IPT;
    static_assert(THREADS >= RADIX, "one digit per thread in the scan phases");
    extern __shared__ __align__(16) unsigned char smem_raw[];
    KeyT *sh_keys = reinterpret_cast<KeyT *>(smem_raw);                       // [TILE]
    uint32_t *sh_vals = reinterpret_cast<uint32_t *>(sh_keys + TILE);        // [HAS_VAL ? TILE : 0]
    uint32_t *whist = sh_vals + (HAS_VAL ? TILE : 0);                        // [WARPS][RADIX]
    long long *goff = reinterpret_cast<long long *>(whist + WARPS * RADIX);  // [RADIX]
    uint32_t *warp_tot = reinterpret_cast<uint32_t *>(goff + RADIX);         // [WARPS + 1]
    __shared__ unsigned int sh_tile;
    __shared__ unsigned char sh_owner[ROUTE ? RADIX : 1];
    auto dig = [&](KeyT k, uint32_t v) -> uint32_t {
        const uint32_t d = digit_of(k, v, dp);
        return ROUTE ? (uint32_t)sh_owner[d] : d;
    };

    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (ROUTE && tid < RADIX) sh_owner[tid] = route_owner[tid];
    if (tid == 0) {
        const unsigned t = atomicAdd(tile_counter, 1u);
        sh_tile = t;
        if (prefetch_ahead && ((uint64_t)t + prefetch_ahead + 1) * TILE <= n && (reinterpret_cast<uintptr_t>(in_keys) & 15) == 0) {
            prefetch_tile_l2(in_keys + ((uint64_t)t + prefetch_ahead) * TILE, (uint32_t)(sizeof(KeyT) * TILE));
            if (HAS_VAL && (reinterpret_cast<uintptr_t>(in_vals) & 15) == 0) prefetch_tile_l2(in_vals + ((uint64_t)t + prefetch_ahead) * TILE, (uint32_t)(4 * TILE));
        }
    }
    for (int i = tid; i < WARPS * RADIX; i += THREADS) whist[i] = 0;
    __syncthreads();
    const unsigned tile = sh_tile;
    const uint64_t tile_base = (uint64_t)tile * TILE;
    const uint32_t tile_valid = (uint32_t)(n - tile_base < (uint64_t)TILE ? n - tile_base : TILE);
    const bool full = tile_valid == TILE;

    // ---- load (coalesced, warp-striped) + early per-warp digit counts ----
    KeyT key[IPT];
    uint32_t val[IPT];
    uint32_t *my_hist = whist + warp * RADIX;
    const uint32_t warp_base = warp * (IPT * 32);
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        uint32_t idx = warp_base + r * 32 + lane;
        if (full || idx < tile_valid) {
            key[r] = in_keys[tile_base + idx];
            val[r] = HAS_VAL ? in_vals[tile_base + idx] : 0;
        } else {
            key[r] = WordTraits<KeyT>::zero();
            val[r] = 0;
        }
    }
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        uint32_t idx = warp_base + r * 32 + lane;
        if (full || idx < tile_valid) atomicAdd(&my_hist[dig(key[r], val[r])], 1u);
    }
    __syncthreads();

    // ---- per digit: tile count, early publish, tile-local digit start ----
    uint32_t count = 0;
    if (tid < RADIX) {
#pragma unroll
        for (int w = 0; w < WARPS; ++w) count += whist[w * RADIX + tid];
        if (tile > 0)
            st_relaxed_u64(status + (size_t)tile * RADIX + tid, os_pack(1, epoch, count));
    }
    uint32_t inc = count;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= (unsigned)o) inc += t;
    }
    if (lane == 31) warp_tot[warp] = inc;
    __syncthreads();
    uint32_t local_base = 0;
    if (tid < RADIX) {
        for (unsigned w = 0; w < warp; ++w) local_base += warp_tot[w];
        local_base += inc - count;
        // whist[w][d] := tile position of warp w's first key with digit d
        uint32_t run = local_base;
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            uint32_t t = whist[w * RADIX + tid];
            whist[w * RADIX + tid] = run;
            run += t;
        }
    }
    __syncthreads();

    // ---- rank + scatter into tile-sorted order ----
    const unsigned lt = lanemask_lt();
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        const uint32_t idx = warp_base + r * 32 + lane;
        const bool valid = full || idx < tile_valid;
        const uint32_t d = dig(key[r], val[r]);
        uint32_t slot = 0;
        if (UNSTABLE) {
            if (valid) slot = atomicAdd(&my_hist[d], 1u);
        } else {
            uint32_t before = 0;
            if (valid) before = my_hist[d];
            __syncwarp();                       // every lane has read the offsets of the row
            if (valid) slot = atomicAdd(&my_hist[d], 1u);
            unsigned todo = __ballot_sync(0xffffffffu, valid && slot != before);
            while (todo) {                      // warp-uniform: one iteration per group of equal digits
                const int leader = __ffs(todo) - 1;
                const uint32_t dl = __shfl_sync(0xffffffffu, d, leader);
                const bool mine = valid && d == dl;
                const unsigned grp = __ballot_sync(0xffffffffu, mine);
                if (mine) slot = before + __popc(grp & lt);
                todo &= ~grp;
            }
            __syncwarp();                       // the atomics are visible to the next row's reads
        }
        if (valid) {
            sh_keys[slot] = key[r];
            if (HAS_VAL) sh_vals[slot] = val[r];
        }
    }

    // ---- decoupled look-back per digit ----
    if (tid < RADIX) {
        unsigned long long excl = 0;
        if (tile > 0) {
            long long t = (long long)tile - 1;
            while (true) {
                unsigned long long s = ld_relaxed_u64(status + (size_t)t * RADIX + tid);
                unsigned flag = (unsigned)(s >> 62);
                if (flag == 0 || ((s >> 56) & 63) != epoch) continue;   // not published yet
                excl += s & OS_VALUE_MASK;
                if (flag == 2) break;
                --t;
            }
        }
        st_relaxed_u64(status + (size_t)tile * RADIX + tid, os_pack(2, epoch, excl + count));
        if (ROUTE)   // digit_offsets = the 2^bits destination addresses; goff = byte address of tile slot 0
            goff[tid] = (long long)digit_offsets[tid]
                        + ((long long)excl - (long long)local_base) * (long long)sizeof(KeyT);
        else
            goff[tid] = (long long)(digit_offsets[tid] + excl) - (long long)local_base;
    }
    __syncthreads();

    // ---- coalesced write-out ----
#pragma unroll 4
    for (uint32_t p = tid; p < tile_valid; p += THREADS) {
        KeyT k = sh_keys[p];
        uint32_t v = HAS_VAL ? sh_vals[p] : 0;
        uint32_t d = dig(k, v);
        if (ROUTE) {
            *reinterpret_cast<KeyT *>(goff[d] + (long long)p * (long long)sizeof(KeyT)) = k;
        } else {
            size_t dst = (size_t)(goff[d] + (long long)p);
            out_keys[dst] = k;
            if (HAS_VAL) out_vals[dst] = v;
        }
    }
}

template <typename KeyT, bool HAS_VAL, int THREADS, int IPT, int MINB, bool UNSTABLE, bool ROUTE = false>
void launch_v5(const KeyT *in_keys, KeyT *out_keys, const uint32_t *in_vals, uint32_t *out_vals,
               uint64_t n, DigitPass dp, const unsigned long long *offs, unsigned long long *status,
               unsigned int *ctr, unsigned epoch, cudaStream_t stream, const unsigned char *route_owner = nullptr,
               unsigned prefetch = 0) {
    constexpr int TILE = THREADS * IPT;
    size_t smem = sizeof(KeyT) * TILE + (HAS_VAL ? sizeof(uint32_t) * TILE : 0)
                  + sizeof(uint32_t) * (THREADS / 32) * RADIX + sizeof(long long) * RADIX
                  + sizeof(uint32_t) * (THREADS / 32 + 1);
    auto kern = onesweep_pass_v5_kernel<KeyT, HAS_VAL, THREADS, IPT, MINB, UNSTABLE, ROUTE>;
    static bool attr_set = false;
    if (!attr_set) {
        MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    kern<<<(unsigned)div_up(n, (size_t)TILE), THREADS, smem, stream>>>(in_keys, out_keys, in_vals, out_vals, n,
                                                                      dp, offs, status, ctr, epoch, route_owner,
                                                                      prefetch);
}

template <typename KeyT, bool HAS_VAL, int THREADS, int IPT, int MINB, int LB, bool ROUTE = false>
void launch_v3(const KeyT *in_keys, KeyT *out_keys, const uint32_t *in_vals, uint32_t *out_vals,
               uint64_t n, DigitPass dp, const unsigned long long *offs, unsigned long long *status,
               unsigned int *ctr, unsigned epoch, cudaStream_t stream, const unsigned char *route_owner = nullptr,
               unsigned prefetch = 0) {
    constexpr int TILE = THREADS * IPT;
    size_t smem = sizeof(KeyT) * TILE + (HAS_VAL ? sizeof(uint32_t) * TILE : 0)
                  + 2 * sizeof(uint32_t) * (THREADS / 32) * RADIX + sizeof(long long) * RADIX
                  + sizeof(uint32_t) * (THREADS / 32 + 1);
    auto kern = onesweep_pass_v3_kernel<KeyT, HAS_VAL, THREADS, IPT, MINB, LB, ROUTE>;
    static bool attr_set = false;
    if (!attr_set) {
        MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    kern<<<(unsigned)div_up(n, (size_t)TILE), THREADS, smem, stream>>>(in_keys, out_keys, in_vals, out_vals, n,
                                                                      dp, offs, status, ctr, epoch, route_owner,
                                                                      prefetch);
}

// Stable passes: v3 (peer masks through shared memory). MTG_SORT_VARIANT=5 selects v5's stable ranking
// (atomic slot + collision repair) instead -- measured on B200 (profiles/r02_sort_variants.jsonl): 4 %
// faster than v3 on random 64-bit keys, 11 % on 128-bit keys, but slower on clustered digits (the
// reverse-complement sort, the passes after the fold): 85.4 vs 81.3 ms on the whole cfg #2 build, so it
// is not the default. The UNSTABLE v5 ranking (first pass of a sort that allows it) is 17 % faster
// than a stable pass on random digits and is used; MTG_SORT_UNSTABLE_FIRST=0 switches it off.
static int g_sort_variant = -1, g_unstable_first = -1;
// launch shape of a pass: 0 = 256 threads x 16 keys, 1 = 512 x 8 (same 4096-key tile, twice the warps);
// L2 prefetch distance in tiles (0 = off)
static int g_launch_shape = -1, g_prefetch = -1;
static int launch_shape() {
    if (g_launch_shape < 0) {
        const char *e = getenv("MTG_SORT_SHAPE");
        g_launch_shape = e ? atoi(e) : 0;
    }
    return g_launch_shape;
}
// Default 600 tiles ~ the blocks resident on 148 SMs (4 per SM): a block prefetches the tile that a block
// of the NEXT wave will load. Measured (profiles/r02_pass_configs.log, 400 M keys): stable pass 2.43 ->
// 2.26 ms with 600, no gain with 1200 (the first use of a loaded key carries ~30 % of the warp stall
// samples, profiles/r02_ncu_full_sort_pass_kernels.txt).
static unsigned prefetch_tiles() {
    if (g_prefetch < 0) {
        const char *e = getenv("MTG_SORT_PREFETCH");
        g_prefetch = e ? atoi(e) : 600;
    }
    return (unsigned)g_prefetch;
}
static int sort_variant() {
    if (g_sort_variant < 0) {
        const char *e = getenv("MTG_SORT_VARIANT");
        g_sort_variant = e ? atoi(e) : 3;
    }
    return g_sort_variant;
}
static bool unstable_first_allowed() {
    if (g_unstable_first < 0) {
        const char *e = getenv("MTG_SORT_UNSTABLE_FIRST");
        g_unstable_first = e ? atoi(e) : 1;
    }
    return g_unstable_first != 0;
}

template <typename KeyT>
static int variant_tile(int) {
    return sizeof(KeyT) == 8 ? 4096 : 2048;   // u128 tiles hold half the keys
}

template <typename KeyT, bool HAS_VAL>
int radix_sort_impl(KeyT *keys[2], uint32_t *vals[2], uint64_t n, const SortPlan &plan,
                    SortWorkspace &ws, cudaStream_t stream, unsigned *launches) {
    constexpr int S = sizeof(KeyT) == 8 ? 1 : 2;
    if (n == 0 || plan.n == 0) return 0;
    const int variant = sort_variant();
    const size_t TILE = variant_tile<KeyT>(variant);
    const size_t tiles = div_up(n, (size_t)TILE);
    MTG_CUDA(cudaMemsetAsync(ws.hist, 0, sizeof(unsigned long long) * MAX_SORT_PASSES * RADIX, stream));
    MTG_CUDA(cudaMemsetAsync(ws.tile_counters, 0, sizeof(unsigned int) * MAX_SORT_PASSES, stream));
    MTG_CUDA(cudaMemsetAsync(ws.status, 0, sizeof(unsigned long long) * tiles * RADIX, stream));

    int dev, sms;
    MTG_CUDA(cudaGetDevice(&dev));
    MTG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    {
        size_t want = div_up(n, (size_t)512 * 8);
        unsigned grid = (unsigned)(want < (size_t)sms * 4 ? (want ? want : 1) : (size_t)sms * 4);
        size_t sh = sizeof(unsigned int) * plan.n * RADIX;
        if (ws.events) cudaEventRecord(ws.events[2 * MAX_SORT_PASSES], stream);
        const uint32_t *v0 = vals ? vals[0] : nullptr;
        const KeyT *k0 = ws.first_src ? static_cast<const KeyT *>(ws.first_src) : keys[0];
#define HIST(NP) launch_histogram<KeyT, HAS_VAL, NP>(grid, sh, k0, v0, n, plan, ws.hist, stream)
        switch (plan.n) {
            case 1: HIST(1); break;
            case 2: HIST(2); break;
            case 3: HIST(3); break;
            case 4: HIST(4); break;
            case 5: HIST(5); break;
            case 6: HIST(6); break;
            case 7: HIST(7); break;
            case 8: HIST(8); break;
            case 10: HIST(10); break;
            default: HIST(0); break;
        }
#undef HIST
        if (ws.events) cudaEventRecord(ws.events[2 * MAX_SORT_PASSES + 1], stream);
        sort_scan_hist_kernel<<<plan.n, RADIX, 0, stream>>>(ws.hist);
        *launches += 2;
    }
    int cur = 0;
    for (int p = 0; p < plan.n; ++p) {
        if (ws.events) cudaEventRecord(ws.events[2 * p], stream);
        const KeyT *ik = p == 0 && ws.first_src ? static_cast<const KeyT *>(ws.first_src) : keys[cur];
        KeyT *ok = keys[cur ^ 1];
        const uint32_t *iv = HAS_VAL ? vals[cur] : nullptr;
        uint32_t *ov = HAS_VAL ? vals[cur ^ 1] : nullptr;
        const unsigned long long *offs = ws.hist + (size_t)p * RADIX;
        unsigned int *ctr = ws.tile_counters + p;
        const unsigned ep = (unsigned)(p + 1);
        // the first pass of a sort has no earlier order to preserve: unstable ranking (one shared atomic
        // per key) when the caller allows it (equal keys may then leave the sort in any order)
        const bool unstable = p == 0 && plan.unstable_first && unstable_first_allowed();
        const unsigned pf = prefetch_tiles();
        const bool wide_block = launch_shape() == 1;
#define PASS_ARGS ik, ok, iv, ov, n, plan.p[p], offs, ws.status, ctr, ep, stream, nullptr, pf
        if (unstable) {
            if (wide_block) launch_v5<KeyT, HAS_VAL, 512, 8 / S, 3, true>(PASS_ARGS);
            else launch_v5<KeyT, HAS_VAL, 256, 16 / S, 4, true>(PASS_ARGS);
        } else if (variant == 5) {
            if (wide_block) launch_v5<KeyT, HAS_VAL, 512, 8 / S, 3, false>(PASS_ARGS);
            else launch_v5<KeyT, HAS_VAL, 256, 16 / S, 4, false>(PASS_ARGS);
        } else {
            if (wide_block) launch_v3<KeyT, HAS_VAL, 512, 8 / S, 3, 0>(PASS_ARGS);
            else launch_v3<KeyT, HAS_VAL, 256, 16 / S, 3, 0>(PASS_ARGS);
        }
#undef PASS_ARGS
        if (ws.events) cudaEventRecord(ws.events[2 * p + 1], stream);
        *launches += 1;
        cur ^= 1;
    }
    MTG_CUDA(cudaGetLastError());
    return cur;
}

}  // namespace

template <typename KeyT>
int radix_sort(KeyT *keys[2], uint32_t *vals[2], uint64_t n, const SortPlan &plan,
               SortWorkspace &ws, cudaStream_t stream, unsigned *launches) {
    if (vals && vals[0]) return radix_sort_impl<KeyT, true>(keys, vals, n, plan, ws, stream, launches);
    return radix_sort_impl<KeyT, false>(keys, nullptr, n, plan, ws, stream, launches);
}

// Counts of one digit (raw, not scanned): hist[0 .. 2^bits)
template <typename KeyT>
void digit_histogram(const KeyT *keys, uint64_t n, DigitPass dp, unsigned long long *hist, cudaStream_t stream,
                     unsigned *launches) {
    MTG_CUDA(cudaMemsetAsync(hist, 0, sizeof(unsigned long long) * RADIX, stream));
    if (n == 0) return;
    int dev, sms;
    MTG_CUDA(cudaGetDevice(&dev));
    MTG_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    SortPlan plan;
    plan.n = 1;
    plan.p[0] = dp;
    size_t want = div_up(n, (size_t)512 * 8);
    unsigned grid = (unsigned)(want < (size_t)sms * 4 ? (want ? want : 1) : (size_t)sms * 4);
    launch_histogram<KeyT, false, 1>(grid, sizeof(unsigned int) * RADIX, keys, nullptr, n, plan, hist, stream);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}
template void digit_histogram<uint64_t>(const uint64_t *, uint64_t, DigitPass, unsigned long long *,
                                        cudaStream_t, unsigned *);
template void digit_histogram<u128>(const u128 *, uint64_t, DigitPass, unsigned long long *, cudaStream_t,
                                    unsigned *);

// One stable scatter pass: a key whose digit dp is b goes to owner[b] (device array of RADIX bytes)
// and is written to the array at address route[owner] (device array of RADIX addresses; local or
// peer-mapped memory). status: tiles * RADIX words, tile_counter: one word; both are cleared here.
template <typename KeyT>
void route_scatter(const KeyT *keys, uint64_t n, DigitPass dp, const unsigned char *owner,
                   const unsigned long long *route, unsigned long long *status, unsigned int *tile_counter,
                   cudaStream_t stream, unsigned *launches, bool stable) {
    if (n == 0) return;
    constexpr int S = sizeof(KeyT) == 8 ? 1 : 2;
    const size_t tiles = div_up(n, (size_t)4096 / S);
    MTG_CUDA(cudaMemsetAsync(tile_counter, 0, sizeof(unsigned int), stream));
    MTG_CUDA(cudaMemsetAsync(status, 0, sizeof(unsigned long long) * tiles * RADIX, stream));
    if (!stable && unstable_first_allowed())
        launch_v5<KeyT, false, 256, 16 / S, 4, true, true>(keys, nullptr, nullptr, nullptr, n, dp, route, status,
                                                           tile_counter, 1u, stream, owner);
    else
        launch_v3<KeyT, false, 256, 16 / S, 3, 0, true>(keys, nullptr, nullptr, nullptr, n, dp, route, status,
                                                        tile_counter, 1u, stream, owner);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}
template void route_scatter<uint64_t>(const uint64_t *, uint64_t, DigitPass, const unsigned char *,
                                      const unsigned long long *, unsigned long long *, unsigned int *,
                                      cudaStream_t, unsigned *, bool);
template void route_scatter<u128>(const u128 *, uint64_t, DigitPass, const unsigned char *,
                                  const unsigned long long *, unsigned long long *, unsigned int *, cudaStream_t,
                                  unsigned *, bool);

namespace {
__global__ void fill_random_kernel(uint64_t *words, uint64_t nwords, uint64_t seed, uint64_t hi_mask,
                                   int wide) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= nwords) return;
    uint64_t x = i + seed * 0x9E3779B97F4A7C15ull;
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    x ^= x >> 31;
    if (!wide || (i & 1)) x &= hi_mask;
    words[i] = x;
}
}  // namespace

namespace {
__global__ void fill_iota_div_kernel(uint64_t *words, uint64_t n, uint32_t div) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) words[i] = i / div;
}
}  // namespace
void fill_iota_div(uint64_t *words, uint64_t n, uint32_t div, cudaStream_t stream) {
    if (!n) return;
    fill_iota_div_kernel<<<(unsigned)div_up(n, (size_t)256), 256, 0, stream>>>(words, n, div);
}

void fill_random_words(uint64_t *words, uint64_t nwords, uint64_t seed, uint64_t hi_mask, bool wide,
                       cudaStream_t stream) {
    if (!nwords) return;
    fill_random_kernel<<<(unsigned)div_up(nwords, (size_t)256), 256, 0, stream>>>(words, nwords, seed,
                                                                                 hi_mask, wide ? 1 : 0);
}

// tuning hook (mtg_debug_sort_config): -1 keeps a setting
void set_sort_config(int variant, int unstable_first, int shape, int prefetch) {
    if (variant >= 0) g_sort_variant = variant;
    if (unstable_first >= 0) g_unstable_first = unstable_first;
    if (shape >= 0) g_launch_shape = shape;
    if (prefetch >= 0) g_prefetch = prefetch;
}

size_t sort_tile_keys(bool wide) {
    return wide ? variant_tile<u128>(sort_variant()) : variant_tile<uint64_t>(sort_variant());
}

template int radix_sort<uint64_t>(uint64_t *[2], uint32_t *[2], uint64_t, const SortPlan &,
                                  SortWorkspace &, cudaStream_t, unsigned *);
template int radix_sort<u128>(u128 *[2], uint32_t *[2], uint64_t, const SortPlan &, SortWorkspace &,
                              cudaStream_t, unsigned *);

}  // namespace mtg
