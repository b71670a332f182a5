// Onesweep LSD radix sort kernels. See mtg_sort.cuh.
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
template <typename KeyT, bool HAS_VAL>
__global__ void __launch_bounds__(512)
sort_histogram_kernel(const KeyT *__restrict__ keys, const uint32_t *__restrict__ vals, uint64_t n,
                      SortPlan plan, unsigned long long *__restrict__ hist) {
    extern __shared__ unsigned int sh_hist[];   // [plan.n * RADIX]
    const int np = plan.n;
    for (int i = threadIdx.x; i < np * RADIX; i += blockDim.x) sh_hist[i] = 0;
    __syncthreads();
    const uint64_t stride = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
        KeyT k = keys[i];
        uint32_t v = HAS_VAL ? vals[i] : 0;
        for (int p = 0; p < np; ++p) atomicAdd(&sh_hist[p * RADIX + digit_of(k, v, plan.p[p])], 1u);
    }
    __syncthreads();
    for (int i = threadIdx.x; i < np * RADIX; i += blockDim.x) {
        unsigned int c = sh_hist[i];
        if (c) atomicAdd(&hist[i], (unsigned long long)c);
    }
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

constexpr unsigned long long OS_VALUE_MASK = (1ull << 56) - 1;
__device__ __forceinline__ unsigned long long os_pack(unsigned flag, unsigned epoch,
                                                      unsigned long long v) {
    return ((unsigned long long)flag << 62) | ((unsigned long long)epoch << 56) | v;
}

// ---------------------------------------------------------------------------
// Kernel S3: one onesweep pass. Tile = THREADS*IPT keys, warp-striped so that
// (warp, item, lane) order == memory order (stability).
// ---------------------------------------------------------------------------
template <typename KeyT, bool HAS_VAL, int THREADS, int IPT>
__global__ void __launch_bounds__(THREADS)
onesweep_pass_kernel(const KeyT *__restrict__ in_keys, KeyT *__restrict__ out_keys,
                     const uint32_t *__restrict__ in_vals, uint32_t *__restrict__ out_vals,
                     uint64_t n, DigitPass dp, const unsigned long long *__restrict__ digit_offsets,
                     unsigned long long *__restrict__ status, unsigned int *tile_counter,
                     unsigned epoch) {
    constexpr int WARPS = THREADS / 32;
    constexpr int TILE = THREADS * IPT;
    extern __shared__ __align__(16) unsigned char smem_raw[];
    KeyT *sh_keys = reinterpret_cast<KeyT *>(smem_raw);                       // [TILE]
    uint32_t *sh_vals = reinterpret_cast<uint32_t *>(sh_keys + TILE);        // [HAS_VAL ? TILE : 0]
    uint32_t *whist = sh_vals + (HAS_VAL ? TILE : 0);                        // [WARPS][RADIX]
    long long *goff = reinterpret_cast<long long *>(whist + WARPS * RADIX);  // [RADIX]
    uint32_t *warp_tot = reinterpret_cast<uint32_t *>(goff + RADIX);         // [WARPS + 1]
    __shared__ unsigned int sh_tile;

    const unsigned tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) sh_tile = atomicAdd(tile_counter, 1u);
    for (int i = tid; i < WARPS * RADIX; i += THREADS) whist[i] = 0;
    __syncthreads();
    const unsigned tile = sh_tile;
    const uint64_t tile_base = (uint64_t)tile * TILE;
    const uint32_t tile_valid = (uint32_t)(n - tile_base < (uint64_t)TILE ? n - tile_base : TILE);

    // ---- load (coalesced, warp-striped) ----
    KeyT key[IPT];
    uint32_t val[IPT];
    uint32_t dig[IPT];
    const uint32_t warp_base = warp * (IPT * 32);
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        uint32_t idx = warp_base + r * 32 + lane;
        if (idx < tile_valid) {
            key[r] = in_keys[tile_base + idx];
            val[r] = HAS_VAL ? in_vals[tile_base + idx] : 0;
            dig[r] = digit_of(key[r], val[r], dp);
        } else {
            key[r] = WordTraits<KeyT>::low_mask(WordTraits<KeyT>::BITS);
            val[r] = 0xffffffffu;
            dig[r] = RADIX - 1;   // padding sorts to the very end of the tile
        }
    }

    // ---- rank inside the warp, running per-warp digit counters ----
    uint32_t rank[IPT];
    uint32_t *my_hist = whist + warp * RADIX;
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        unsigned peers = __match_any_sync(0xffffffffu, dig[r]);
        uint32_t before = my_hist[dig[r]];
        __syncwarp();
        if ((peers & lanemask_lt()) == 0) my_hist[dig[r]] = before + __popc(peers);
        __syncwarp();
        rank[r] = before + __popc(peers & lanemask_lt());
    }
    __syncthreads();

    // ---- per digit: exclusive prefix over warps, tile count ----
    uint32_t count = 0;
    if (tid < RADIX) {
#pragma unroll
        for (int w = 0; w < WARPS; ++w) {
            uint32_t t = whist[w * RADIX + tid];
            whist[w * RADIX + tid] = count;
            count += t;
        }
        if (tid == RADIX - 1) count -= (TILE - tile_valid);   // drop the padding
        // publish the tile aggregate early so successors can start looking back
        if (tile > 0)
            st_relaxed_u64(status + (size_t)tile * RADIX + tid, os_pack(1, epoch, count));
    }
    // exclusive scan of the digit counts inside the tile (first RADIX threads carry data)
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
        goff[tid] = local_base;   // temporarily: local start of the digit
    }
    __syncthreads();

    // ---- scatter into shared memory in tile-sorted order ----
#pragma unroll
    for (int r = 0; r < IPT; ++r) {
        uint32_t pos = (uint32_t)goff[dig[r]] + my_hist[dig[r]] + rank[r];
        sh_keys[pos] = key[r];
        if (HAS_VAL) sh_vals[pos] = val[r];
    }
    __syncthreads();

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
        goff[tid] = (long long)(digit_offsets[tid] + excl) - (long long)local_base;
    }
    __syncthreads();

    // ---- coalesced write-out ----
    for (uint32_t p = tid; p < tile_valid; p += THREADS) {
        KeyT k = sh_keys[p];
        uint32_t v = HAS_VAL ? sh_vals[p] : 0;
        uint32_t d = digit_of(k, v, dp);
        size_t dst = (size_t)(goff[d] + (long long)p);
        out_keys[dst] = k;
        if (HAS_VAL) out_vals[dst] = v;
    }
}

template <typename KeyT, bool HAS_VAL>
int radix_sort_impl(KeyT *keys[2], uint32_t *vals[2], uint64_t n, const SortPlan &plan,
                    SortWorkspace &ws, cudaStream_t stream, unsigned *launches) {
    constexpr int THREADS = SortTile<KeyT>::THREADS, IPT = SortTile<KeyT>::IPT;
    constexpr int TILE = THREADS * IPT;
    if (n == 0 || plan.n == 0) return 0;
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
        auto kern = sort_histogram_kernel<KeyT, HAS_VAL>;
        MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sh));
        if (ws.events) cudaEventRecord(ws.events[2 * MAX_SORT_PASSES], stream);
        kern<<<grid, 512, sh, stream>>>(keys[0], vals ? vals[0] : nullptr, n, plan, ws.hist);
        if (ws.events) cudaEventRecord(ws.events[2 * MAX_SORT_PASSES + 1], stream);
        sort_scan_hist_kernel<<<plan.n, RADIX, 0, stream>>>(ws.hist);
        *launches += 2;
    }
    size_t smem = sizeof(KeyT) * TILE + (HAS_VAL ? sizeof(uint32_t) * TILE : 0)
                  + sizeof(uint32_t) * (THREADS / 32) * RADIX + sizeof(long long) * RADIX
                  + sizeof(uint32_t) * (THREADS / 32 + 1);
    auto kern = onesweep_pass_kernel<KeyT, HAS_VAL, THREADS, IPT>;
    MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int cur = 0;
    for (int p = 0; p < plan.n; ++p) {
        if (ws.events) cudaEventRecord(ws.events[2 * p], stream);
        kern<<<(unsigned)tiles, THREADS, smem, stream>>>(
            keys[cur], keys[cur ^ 1], HAS_VAL ? vals[cur] : nullptr, HAS_VAL ? vals[cur ^ 1] : nullptr,
            n, plan.p[p], ws.hist + (size_t)p * RADIX, ws.status, ws.tile_counters + p,
            (unsigned)(p + 1));
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

template int radix_sort<uint64_t>(uint64_t *[2], uint32_t *[2], uint64_t, const SortPlan &,
                                  SortWorkspace &, cudaStream_t, unsigned *);
template int radix_sort<u128>(u128 *[2], uint32_t *[2], uint64_t, const SortPlan &, SortWorkspace &,
                              cudaStream_t, unsigned *);

}  // namespace mtg
