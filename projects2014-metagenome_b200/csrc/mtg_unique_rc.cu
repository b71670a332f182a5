// K3 (segmented unique / count-reduce) and K1b (reverse-complement completion).
// Replaces std::unique / the saturating count merge of SortedSet / SortedMultiset
// (common/sorted_sets/sorted_set.cpp:41-48, sorted_multiset.cpp:54-84) and
// add_reverse_complements (boss_chunk_construct.cpp:179-222).
#include <cstdlib>

#include "mtg_scan.cuh"
#include "mtg_stages.cuh"

namespace mtg {

namespace {

constexpr int UQ_THREADS = 256;
constexpr int UQ_IPT = 8;
constexpr int UQ_TILE = UQ_THREADS * UQ_IPT;

// Single-pass compaction of run heads with decoupled look-back. Warp-striped (coalesced)
// loads; the previous key comes from the neighbouring lane, heads are compacted with one ballot
// per row so that consecutive heads are stored to consecutive addresses.
template <typename KeyT>
__global__ void __launch_bounds__(UQ_THREADS)
unique_kernel(const KeyT *__restrict__ keys, uint64_t n, KeyT *__restrict__ out_keys,
              uint64_t *__restrict__ out_starts, unsigned long long *d_num_out,
              unsigned long long *status, unsigned int *tile_counter) {
    constexpr int WARPS = UQ_THREADS / 32;
    __shared__ unsigned int sh_tile;
    __shared__ uint32_t warp_cnt[WARPS + 1];
    __shared__ uint64_t slot;
    if (threadIdx.x == 0) sh_tile = atomicAdd(tile_counter, 1u);
    __syncthreads();
    const uint32_t tile = sh_tile;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint64_t wbase = (uint64_t)tile * UQ_TILE + (uint64_t)warp * (32 * UQ_IPT);
    KeyT k[UQ_IPT];
#pragma unroll
    for (int r = 0; r < UQ_IPT; ++r) {
        uint64_t idx = wbase + r * 32 + lane;
        k[r] = idx < n ? keys[idx] : WordTraits<KeyT>::zero();
    }
    KeyT carry = WordTraits<KeyT>::zero();   // key just before this warp's chunk
    if (lane == 0 && wbase > 0 && wbase <= n) carry = keys[wbase - 1];
    unsigned bal[UQ_IPT];
    uint32_t wcount = 0;
#pragma unroll
    for (int r = 0; r < UQ_IPT; ++r) {
        uint64_t idx = wbase + r * 32 + lane;
        KeyT prev;
        if constexpr (sizeof(KeyT) == 8) {
            uint64_t up = __shfl_up_sync(0xffffffffu, w_low64(k[r]), 1);
            uint64_t last = __shfl_sync(0xffffffffu, r ? w_low64(k[r ? r - 1 : 0]) : 0ull, 31);
            uint64_t pv = lane ? up : (r ? last : w_low64(carry));
            prev = WordTraits<KeyT>::from_u64(pv);
        } else {
            const u128 cur = *reinterpret_cast<const u128 *>(&k[r]);
            const u128 prv = *reinterpret_cast<const u128 *>(&k[r ? r - 1 : 0]);
            const u128 cr = *reinterpret_cast<const u128 *>(&carry);
            u128 up = make_u128(__shfl_up_sync(0xffffffffu, cur.lo, 1), __shfl_up_sync(0xffffffffu, cur.hi, 1));
            u128 last = make_u128(__shfl_sync(0xffffffffu, prv.lo, 31), __shfl_sync(0xffffffffu, prv.hi, 31));
            u128 pv = lane ? up : (r ? last : cr);
            prev = *reinterpret_cast<const KeyT *>(&pv);
        }
        bool head = idx < n && (idx == 0 || !w_eq(k[r], prev));
        bal[r] = __ballot_sync(0xffffffffu, head);
        wcount += __popc(bal[r]);
    }
    if (lane == 0) warp_cnt[warp] = wcount;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int w = 0; w < WARPS; ++w) { uint32_t t = warp_cnt[w]; warp_cnt[w] = run; run += t; }
        warp_cnt[WARPS] = run;
    }
    __syncthreads();
    const uint32_t tile_total = warp_cnt[WARPS];
    const uint64_t tile_excl = lookback_exclusive(status, tile, tile_total, &slot);
    uint64_t pos = tile_excl + warp_cnt[warp];
    const unsigned lt = lanemask_lt();
#pragma unroll
    for (int r = 0; r < UQ_IPT; ++r) {
        if (bal[r] & (1u << lane)) {
            uint64_t dst = pos + __popc(bal[r] & lt);
            out_keys[dst] = k[r];
            if (out_starts) out_starts[dst] = wbase + r * 32 + lane;
        }
        pos += __popc(bal[r]);
    }
    // the tile holding the last element publishes the total and the end marker
    if (threadIdx.x == 0 && (uint64_t)(tile + 1) * UQ_TILE >= n) {
        uint64_t m = tile_excl + tile_total;
        *d_num_out = m;
        if (out_starts) out_starts[m] = n;
    }
}

// Same contract as unique_kernel, tile staged in shared memory by one TMA bulk copy (keys are
// never held in registers while in flight): 64 KB tiles, 3 resident blocks per SM = 192 KB of
// loads outstanding per SM, which hides the look-back latency that bound unique_kernel (61 % of
// its stall samples at the barrier after the look-back, profiles/r01_ncu_full_v2_101ms.txt).
// The 16-byte left halo holds keys[tile_base - 1], so a head is simply sk[i] != sk[i - 1].
constexpr int UB_THREADS = 256;

// Tiles are numbered by an atomic ticket (like the onesweep pass): the look-back spins on the
// predecessor tiles, so a tile may only wait for tiles whose blocks have already started -- CUDA
// does not promise that blocks are dispatched in blockIdx order.
template <typename KeyT, int UB_TILE_BYTES>
__global__ void __launch_bounds__(UB_THREADS)
unique_bulk_kernel(const KeyT *__restrict__ keys, uint64_t n, KeyT *__restrict__ out_keys,
                   uint64_t *__restrict__ out_starts, unsigned long long *d_num_out,
                   unsigned long long *status, unsigned int *tile_counter) {
    constexpr int TILE = UB_TILE_BYTES / sizeof(KeyT);
    constexpr int HALO = 16 / sizeof(KeyT);
    constexpr int WARPS = UB_THREADS / 32;
    constexpr int ROWS = TILE / UB_THREADS;          // rows of 32 keys per warp
    extern __shared__ __align__(128) unsigned char ub_smem[];
    KeyT *sk = reinterpret_cast<KeyT *>(ub_smem) + HALO;   // sk[-1] = key before the tile
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint32_t warp_cnt[WARPS + 1];
    __shared__ uint64_t slot;
    __shared__ uint32_t sh_tile;
    if (threadIdx.x == 0) {
        const uint32_t t = atomicAdd(tile_counter, 1u);
        sh_tile = t;
        const uint64_t tb = (uint64_t)t * TILE;
        const uint32_t v = (uint32_t)(n - tb < (uint64_t)TILE ? n - tb : TILE);
        mbar_init(&bar, 1);
        const uint32_t halo = t ? 16u : 0u;
        const uint32_t bytes = ((v * (uint32_t)sizeof(KeyT) + 15u) & ~15u) + halo;
        bulk_load_g2s(reinterpret_cast<unsigned char *>(sk) - halo,
                      reinterpret_cast<const unsigned char *>(keys + tb) - halo, bytes, &bar);
    }
    __syncthreads();   // the barrier is initialised (and the ticket drawn) before anyone waits on it
    const uint32_t tile = sh_tile;
    const uint64_t tile_base = (uint64_t)tile * TILE;
    const uint32_t valid = (uint32_t)(n - tile_base < (uint64_t)TILE ? n - tile_base : TILE);
    mbar_wait(&bar, 0);

    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t wbase = warp * (32 * ROWS);
    uint32_t wcount = 0;
#pragma unroll 8
    for (int r = 0; r < ROWS; ++r) {
        const uint32_t idx = wbase + r * 32 + lane;
        bool head = idx < valid && ((tile_base + idx) == 0 || !w_eq(sk[idx], sk[(int)idx - 1]));
        wcount += __popc(__ballot_sync(0xffffffffu, head));
    }
    if (lane == 0) warp_cnt[warp] = wcount;
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t run = 0;
        for (int w = 0; w < WARPS; ++w) { uint32_t t = warp_cnt[w]; warp_cnt[w] = run; run += t; }
        warp_cnt[WARPS] = run;
    }
    __syncthreads();
    const uint32_t tile_total = warp_cnt[WARPS];
    const uint64_t tile_excl = lookback_exclusive(status, tile, tile_total, &slot);
    uint64_t pos = tile_excl + warp_cnt[warp];
    const unsigned lt = lanemask_lt();
#pragma unroll 8
    for (int r = 0; r < ROWS; ++r) {
        const uint32_t idx = wbase + r * 32 + lane;
        KeyT k = sk[idx];
        bool head = idx < valid && ((tile_base + idx) == 0 || !w_eq(k, sk[(int)idx - 1]));
        const unsigned bal = __ballot_sync(0xffffffffu, head);
        if (head) {
            uint64_t dst = pos + __popc(bal & lt);
            out_keys[dst] = k;
            if (out_starts) out_starts[dst] = tile_base + idx;
        }
        pos += __popc(bal);
    }
    if (threadIdx.x == 0 && tile_base + TILE >= n) {
        uint64_t m = tile_excl + tile_total;
        *d_num_out = m;
        if (out_starts) out_starts[m] = n;
    }
}

__global__ void run_counts_kernel(const uint64_t *__restrict__ starts, const uint32_t *__restrict__ vals,
                                  uint64_t m, uint32_t cmax, uint32_t *__restrict__ counts) {
    // saturating sum of the counts of run j (sorted_multiset.cpp:70-82). A thread sums a short run on its
    // own; a long one (a low-complexity k-mer with 10^6 copies) is summed by the whole warp, strided and
    // coalesced, and stops once it saturates.
    constexpr uint64_t LONG_RUN = 64;
    const uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = j < m;
    const unsigned lane = threadIdx.x & 31;
    uint64_t a = 0, b = 0, c = 0;
    if (valid) { a = starts[j]; b = starts[j + 1]; }
    const bool is_long = valid && vals && b - a > LONG_RUN;
    if (valid && !is_long) {
        if (vals) {
            for (uint64_t i = a; i < b && c < cmax; ++i) c += vals[i];
        } else {
            c = b - a;
        }
    }
    unsigned todo = __ballot_sync(0xffffffffu, is_long);
    while (todo) {
        const int l = __ffs(todo) - 1;
        todo &= todo - 1;
        const uint64_t la = __shfl_sync(0xffffffffu, a, l), lb = __shfl_sync(0xffffffffu, b, l);
        uint64_t sum = 0;
        for (uint64_t i = la; i < lb; i += 32 * 8) {
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const uint64_t q = i + (uint64_t)u * 32 + lane;
                if (q < lb) sum += vals[q];
            }
            if (__any_sync(0xffffffffu, sum >= cmax)) break;   // saturated
        }
#pragma unroll
        for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if ((int)lane == l) c = sum;
    }
    if (valid) counts[j] = c > cmax ? cmax : (uint32_t)c;
}

constexpr int SC_THREADS = 256;
constexpr int SC_IPT = 8;
constexpr int SC_TILE = SC_THREADS * SC_IPT;

__global__ void __launch_bounds__(SC_THREADS)
scan_u32_kernel(const uint32_t *__restrict__ in, uint64_t n, uint64_t *__restrict__ out,
                unsigned long long *status, unsigned int *tile_counter) {
    __shared__ unsigned int sh_tile;
    __shared__ uint32_t warp_tot[SC_THREADS / 32 + 1];
    __shared__ uint64_t slot;
    if (threadIdx.x == 0) sh_tile = atomicAdd(tile_counter, 1u);
    __syncthreads();
    const uint32_t tile = sh_tile;
    const uint64_t base = (uint64_t)tile * SC_TILE + (uint64_t)threadIdx.x * SC_IPT;
    uint32_t v[SC_IPT];
    uint32_t sum = 0;
#pragma unroll
    for (int i = 0; i < SC_IPT; ++i) {
        v[i] = base + i < n ? in[base + i] : 0;
        sum += v[i];
    }
    uint32_t excl = block_exclusive_scan<SC_THREADS>(sum, warp_tot);
    uint32_t tile_total = warp_tot[SC_THREADS / 32];
    uint64_t run = lookback_exclusive(status, tile, tile_total, &slot) + excl;
#pragma unroll
    for (int i = 0; i < SC_IPT; ++i) {
        if (base + i < n) out[base + i] = run;
        run += v[i];
    }
    if (threadIdx.x == 0 && (uint64_t)(tile + 1) * SC_TILE >= n)
        out[n] = slot + tile_total;
}

__global__ void fill_u32_kernel(uint32_t *p, uint32_t v, uint64_t n) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}
__global__ void flip_info_kernel(uint32_t *p, uint64_t n) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = (127u - (p[i] & 127u)) | (p[i] & INFO_W_DOLLAR);
}

__device__ __forceinline__ uint64_t rev2_64(uint64_t x) {
    x = __brevll(x);
    return ((x >> 1) & 0x5555555555555555ull) | ((x & 0x5555555555555555ull) << 1);
}
__device__ __forceinline__ uint64_t rev2w(uint64_t x) { return rev2_64(x); }
__device__ __forceinline__ u128 rev2w(u128 x) { return make_u128(rev2_64(x.hi), rev2_64(x.lo)); }

// kmer::reverse_complement (kmer/kmer_transform.hpp:14-35) for the tight DNA
// alphabet (complement = code ^ 3) on the KMerBOSS layout.
template <typename KeyT>
__device__ __forceinline__ KeyT rc_boss_dna(KeyT x, unsigned K) {
    const unsigned L = 2, KL = 2 * K;
    // undo the rotation: v = sum s_i << L(i-1)
    KeyT v = w_or(w_shr(x, L), w_shl(w_and(x, WordTraits<KeyT>::low_mask(L)), L * (K - 1)));
    KeyT r = w_shr(rev2w(w_not(v)), WordTraits<KeyT>::BITS - KL);
    KeyT hi = w_and(w_shl(r, L), WordTraits<KeyT>::low_mask(KL));
    return w_or(hi, w_shr(r, L * (K - 1)));
}

constexpr int RC_THREADS = 256;
constexpr int RC_IPT = 8;
constexpr int RC_TILE = RC_THREADS * RC_IPT;

// One global atomic per block (a per-warp atomic on the single counter serialises in L2:
// 5.8 ms for 231 M keys); the block's reverse complements are staged in shared memory and
// written out coalesced. Their order is irrelevant: they are sorted afterwards.
template <typename KeyT>
__global__ void __launch_bounds__(RC_THREADS)
append_rc_kernel(const KeyT *__restrict__ keys, KeyT *__restrict__ rc_out, uint32_t *counts,
                 uint32_t *__restrict__ rc_counts, uint64_t m, uint32_t K, uint32_t cmax,
                 unsigned long long *d_appended) {
    __shared__ KeyT s_keys[RC_TILE];
    __shared__ uint32_t s_cnt[RC_TILE];
    __shared__ uint32_t warp_tot[RC_THREADS / 32 + 1];
    __shared__ unsigned long long s_base;
    const uint64_t base = (uint64_t)blockIdx.x * RC_TILE;
    KeyT rc[RC_IPT];
    uint32_t c[RC_IPT];
    uint32_t flags = 0, n_emit = 0;
#pragma unroll
    for (int it = 0; it < RC_IPT; ++it) {
        uint64_t i = base + it * RC_THREADS + threadIdx.x;
        c[it] = 0;
        rc[it] = WordTraits<KeyT>::zero();
        if (i < m) {
            KeyT x = keys[i];
            rc[it] = rc_boss_dna(x, K);
            bool emit = !w_eq(rc[it], x);
            if (counts) {
                c[it] = counts[i];
                if (!emit) counts[i] = c[it] > cmax / 2 ? cmax : 2 * c[it];   // palindrome (:205-211)
            }
            if (emit) { flags |= 1u << it; ++n_emit; }
        }
    }
    uint32_t excl = block_exclusive_scan<RC_THREADS>(n_emit, warp_tot);
    const uint32_t total = warp_tot[RC_THREADS / 32];
    if (threadIdx.x == 0) s_base = total ? atomicAdd(d_appended, (unsigned long long)total) : 0ull;
#pragma unroll
    for (int it = 0; it < RC_IPT; ++it)
        if (flags & (1u << it)) {
            s_keys[excl] = rc[it];
            if (counts) s_cnt[excl] = c[it];
            ++excl;
        }
    __syncthreads();
    const unsigned long long gbase = s_base;
    for (uint32_t j = threadIdx.x; j < total; j += RC_THREADS) {
        rc_out[gbase + j] = s_keys[j];
        if (counts) rc_counts[gbase + j] = s_cnt[j];
    }
}

}  // namespace

size_t unique_status_words(uint64_t n) { return div_up(n, (size_t)UQ_TILE) + 1; }   // >= bulk tiles

template <typename KeyT>
void unique_keys(const KeyT *keys, uint64_t n, KeyT *out_keys, uint64_t *out_starts,
                 unsigned long long *d_num_out, unsigned long long *scan_status,
                 unsigned int *tile_counter, cudaStream_t stream, unsigned *launches) {
    size_t tiles = div_up(n, (size_t)UQ_TILE);
    MTG_CUDA(cudaMemsetAsync(scan_status, 0, sizeof(unsigned long long) * (tiles + 1), stream));
    MTG_CUDA(cudaMemsetAsync(tile_counter, 0, sizeof(unsigned int), stream));
    MTG_CUDA(cudaMemsetAsync(d_num_out, 0, sizeof(unsigned long long), stream));
    if (n == 0) return;
    if (((uintptr_t)keys & 15) == 0 && !getenv("MTG_UNIQUE_NO_BULK")) {
        // TMA-staged tiles (the input buffer is an arena block: readable up to the next 16 bytes)
        static int tile_kb = -1;
        if (tile_kb < 0) {
            const char *e = getenv("MTG_UB_TILE_KB");
            tile_kb = e ? atoi(e) : 32;
        }
        auto launch = [&](auto kern, size_t tile_bytes) {
            const size_t smem = tile_bytes + 16;
            MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            kern<<<(unsigned)div_up(n, tile_bytes / sizeof(KeyT)), UB_THREADS, smem, stream>>>(
                keys, n, out_keys, out_starts, d_num_out, scan_status, tile_counter);
        };
        if (tile_kb >= 64) launch(unique_bulk_kernel<KeyT, 64 * 1024>, 64 * 1024);
        else if (tile_kb >= 32) launch(unique_bulk_kernel<KeyT, 32 * 1024>, 32 * 1024);
        else launch(unique_bulk_kernel<KeyT, 16 * 1024>, 16 * 1024);
        MTG_CUDA(cudaGetLastError());
        *launches += 1;
        return;
    }
    unique_kernel<KeyT><<<(unsigned)tiles, UQ_THREADS, 0, stream>>>(keys, n, out_keys, out_starts,
                                                                   d_num_out, scan_status, tile_counter);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}
template void unique_keys<uint64_t>(const uint64_t *, uint64_t, uint64_t *, uint64_t *,
                                    unsigned long long *, unsigned long long *, unsigned int *,
                                    cudaStream_t, unsigned *);
template void unique_keys<u128>(const u128 *, uint64_t, u128 *, uint64_t *, unsigned long long *,
                                unsigned long long *, unsigned int *, cudaStream_t, unsigned *);

void run_counts(const uint64_t *starts, const uint32_t *vals, uint64_t m, uint32_t cmax,
                uint32_t *counts, cudaStream_t stream, unsigned *launches) {
    if (m == 0) return;
    run_counts_kernel<<<(unsigned)div_up(m, (size_t)256), 256, 0, stream>>>(starts, vals, m, cmax, counts);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

void exclusive_scan_u32(const uint32_t *in, uint64_t n, uint64_t *out, unsigned long long *scan_status,
                        unsigned int *tile_counter, cudaStream_t stream, unsigned *launches) {
    size_t tiles = div_up(n, (size_t)SC_TILE);
    MTG_CUDA(cudaMemsetAsync(scan_status, 0, sizeof(unsigned long long) * (tiles + 1), stream));
    MTG_CUDA(cudaMemsetAsync(tile_counter, 0, sizeof(unsigned int), stream));
    if (n == 0) {
        MTG_CUDA(cudaMemsetAsync(out, 0, sizeof(uint64_t), stream));
        return;
    }
    scan_u32_kernel<<<(unsigned)tiles, SC_THREADS, 0, stream>>>(in, n, out, scan_status, tile_counter);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

void fill_u32(uint32_t *p, uint32_t v, uint64_t n, cudaStream_t stream, unsigned *launches) {
    if (n == 0) return;
    fill_u32_kernel<<<(unsigned)div_up(n, (size_t)256), 256, 0, stream>>>(p, v, n);
    *launches += 1;
}
void flip_info(uint32_t *info, uint64_t n, cudaStream_t stream, unsigned *launches) {
    if (n == 0) return;
    flip_info_kernel<<<(unsigned)div_up(n, (size_t)256), 256, 0, stream>>>(info, n);
    *launches += 1;
}

template <typename KeyT>
void append_reverse_complements(const KeyT *keys, KeyT *rc_out, uint32_t *counts, uint32_t *rc_counts,
                                uint64_t m, Alpha al, uint32_t cmax, unsigned long long *d_appended,
                                cudaStream_t stream, unsigned *launches) {
    MTG_CUDA(cudaMemsetAsync(d_appended, 0, sizeof(unsigned long long), stream));
    if (m == 0) return;
    append_rc_kernel<KeyT><<<(unsigned)div_up(m, (size_t)RC_TILE), RC_THREADS, 0, stream>>>(keys, rc_out, counts, rc_counts,
                                                                               m, al.K, cmax, d_appended);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}
template void append_reverse_complements<uint64_t>(const uint64_t *, uint64_t *, uint32_t *, uint32_t *,
                                                   uint64_t, Alpha, uint32_t, unsigned long long *,
                                                   cudaStream_t, unsigned *);
template void append_reverse_complements<u128>(const u128 *, u128 *, uint32_t *, uint32_t *, uint64_t, Alpha,
                                               uint32_t, unsigned long long *, cudaStream_t, unsigned *);

}  // namespace mtg
