// K4: dummy sink / source k-mer recovery, K5: merge of real and dummy rows.
// Set-level restatement of the reference's serial merge-join loops:
//   add_dummy_sink_kmers   (boss_chunk_construct.cpp:54-98)
//   add_dummy_source_kmers (boss_chunk_construct.cpp:123-168)
//   dummy-prefix loop      (boss_chunk_construct.cpp:280-306)
//   merge real + dummy     (boss_chunk_construct.cpp:308-348)
// With T = {target nodes x[2..K]} and S = {source nodes x[1..K-1]} of the real k-mers:
//   sinks      = T \ S  (as k-mers  t$)
//   sources-1  = S \ T  (as k-mers  $ n[1..K-2] -> n[K-1])
//   sources-c  = distinct prefixes of the sources-1 nodes, '$'-padded (c = 2..K-1).
// Membership tests are binary searches in the sorted real array (probes are
// sorted per leading character, so neighbouring threads walk the same cache lines).
#include "mtg_scan.cuh"
#include "mtg_stages.cuh"

namespace mtg {

namespace {

constexpr int MTG_MAX_LEVELS = 128;   // == MTG_MAX_K (include/mtg_gpu.h)

template <typename KeyT>
__device__ __forceinline__ uint64_t lower_bound_dev(const KeyT *a, uint64_t n, KeyT x) {
    uint64_t lo = 0, hi = n;
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (w_lt(a[mid], x)) lo = mid + 1; else hi = mid;
    }
    return lo;
}

// Bucket index over the sorted reals: idx[b] = first position whose top B key bits are >= b.
// A membership probe then searches ~m / 2^B keys instead of m (the probes of neighbouring
// threads are sorted per leading character, so the index and the bucket stay cache resident).
struct BucketIndex {
    const uint64_t *idx;   // [2^B + 1]
    uint32_t B, shift;     // bucket of x = x >> shift
};
template <typename KeyT>
__device__ __forceinline__ uint64_t lower_bound_idx(const KeyT *a, uint64_t n, BucketIndex bi, KeyT x) {
    if (!bi.idx) return lower_bound_dev(a, n, x);
    uint64_t b = w_low64(w_shr(x, bi.shift));
    uint64_t lo = bi.idx[b], hi = bi.idx[b + 1];
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (w_lt(a[mid], x)) lo = mid + 1; else hi = mid;
    }
    return lo;
}
template <typename KeyT>
__global__ void bucket_index_kernel(const KeyT *__restrict__ a, uint64_t n, uint32_t B, uint32_t shift,
                                    uint64_t *__restrict__ idx) {
    uint64_t b = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (b > (1ull << B)) return;
    if (b == (1ull << B)) { idx[b] = n; return; }
    idx[b] = lower_bound_dev(a, n, w_shl(WordTraits<KeyT>::from_u64(b), shift));
}

// append one element per emitting lane through a warp-aggregated atomic
template <typename KeyT>
__device__ __forceinline__ void warp_append(bool emit, KeyT v, KeyT *out, uint64_t cap,
                                            unsigned long long *counter) {
    unsigned bal = __ballot_sync(0xffffffffu, emit);
    if (bal == 0) return;
    unsigned long long base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(counter, (unsigned long long)__popc(bal));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (emit) {
        uint64_t dst = base + __popc(bal & lanemask_lt());
        if (dst < cap) out[dst] = v;
    }
}

// ---- windowed merge-join version of the candidate search (single-GPU path) ------------------
// The sink probe of a real x with edge label c is  sink(x, c) = [c, s_{K-1}..s_2 | 0]: for a
// fixed c it is monotone in the position of x, so the probes of DC_TILE consecutive reals with
// label c fall into one contiguous window of the reals (inside the range whose last node
// character is c) -- the cursor it[c] of add_dummy_sink_kmers (boss_chunk_construct.cpp:66-93),
// tile by tile. dc_partition_kernel finds the window bounds, dc_sinks_kernel stages the windows
// in shared memory and searches there. A hit at position p (the head of node t) also records
// "node p has an incoming real edge" in a bitmap, which turns the redundancy test of
// add_dummy_source_kmers (:151-163) into one bit lookup per node head (dc_sources_kernel).
constexpr int DC_THREADS = 256;
constexpr int DC_IPT = 8;
constexpr int DC_TILE = DC_THREADS * DC_IPT;
constexpr int DC_POOL = DC_TILE + DC_TILE / 2 + 64;    // window keys staged per tile (all labels together)
constexpr int DC_MAX_CH = 32;

template <typename KeyT>
__device__ __forceinline__ KeyT sink_probe(KeyT x, KeyT c, uint32_t L, uint32_t K) {
    return w_and(w_or(w_shr(x, L), w_shl(c, L * (K - 1))), w_not(WordTraits<KeyT>::low_mask(L)));
}

// part[j * nch + c] = lower_bound(reals, sink(reals[min(j * DC_TILE, m - 1)], c)), j = 0..tiles
template <typename KeyT>
__global__ void dc_partition_kernel(const KeyT *__restrict__ reals, uint64_t m, Alpha al, uint32_t nch,
                                    BucketIndex bi, uint64_t tiles, uint64_t *__restrict__ part) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (tiles + 1) * nch) return;
    uint64_t j = t / nch;
    uint32_t c = (uint32_t)(t % nch);
    uint64_t pos = j * DC_TILE < m ? j * DC_TILE : m - 1;
    KeyT probe = sink_probe(reals[pos], WordTraits<KeyT>::from_u64(c), al.L, al.K);
    part[t] = lower_bound_idx(reals, m, bi, probe);
}

// Tile and windows are staged by TMA bulk copies (one elected thread, one mbarrier): no registers
// hold data in flight and the class / node-head tests read neighbours from shared memory.
constexpr int DC_HALO = 16;   // reals staged before the tile (class test looks back <= sigma^2 rows)

template <typename KeyT>
__global__ void __launch_bounds__(DC_THREADS)
dc_sinks_kernel(const KeyT *__restrict__ reals, uint64_t m, Alpha al, uint32_t nch, BucketIndex bi,
                const uint64_t *__restrict__ part, uint32_t *__restrict__ incoming,
                uint32_t *__restrict__ heads, KeyT *sink_keys, uint64_t sink_cap,
                unsigned long long *d_num_sink) {
    using WT = WordTraits<KeyT>;
    constexpr uint32_t AL = 16 / sizeof(KeyT);            // keys per 16-byte TMA granule
    extern __shared__ __align__(128) unsigned char dc_smem[];
    KeyT *tk = reinterpret_cast<KeyT *>(dc_smem) + DC_HALO;   // tk[-DC_HALO .. DC_TILE)
    KeyT *win = tk + DC_TILE;                                 // [DC_POOL]
    __shared__ __align__(8) unsigned long long bar;
    __shared__ uint64_t w_lo[DC_MAX_CH];
    __shared__ uint32_t w_off[DC_MAX_CH];                 // pool offset of window[0], ~0u = not staged
    __shared__ uint32_t w_len[DC_MAX_CH];
    const uint32_t L = al.L, K = al.K;
    const KeyT maskL = WT::low_mask(L);
    const uint64_t j = blockIdx.x;
    const uint64_t i0 = j * DC_TILE;
    const uint32_t valid = (uint32_t)(m - i0 < (uint64_t)DC_TILE ? m - i0 : DC_TILE);
    const uint32_t halo = j ? DC_HALO : 0;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        uint32_t off = 0, total = 0;
        uint32_t c_off[DC_MAX_CH], c_bytes[DC_MAX_CH];
        for (uint32_t c = 0; c < nch; ++c) {
            uint64_t lo = part[j * nch + c];
            uint64_t hi = part[(j + 1) * nch + c] + 1;
            if (hi > m) hi = m;
            const uint64_t len = hi > lo ? hi - lo : 0;
            const uint32_t lead = (uint32_t)(lo & (AL - 1));           // keys before lo in its granule
            const uint64_t span = (lead + len + AL - 1) & ~(uint64_t)(AL - 1);
            w_lo[c] = lo;
            w_len[c] = (uint32_t)(len < 0xffffffffull ? len : 0xffffffffull);
            c_bytes[c] = 0;
            if (span <= (uint64_t)DC_POOL - off) {
                w_off[c] = off + lead;
                c_off[c] = off;
                c_bytes[c] = (uint32_t)(span * sizeof(KeyT));
                off += (uint32_t)span;
                total += c_bytes[c];
            } else {
                w_off[c] = ~0u;
            }
        }
        const uint32_t tile_bytes = (((halo + valid) * (uint32_t)sizeof(KeyT)) + 15u) & ~15u;
        mbar_expect(&bar, total + tile_bytes);
        bulk_copy_g2s(tk - halo, reals + i0 - halo, tile_bytes, &bar);
        for (uint32_t c = 0; c < nch; ++c)
            if (c_bytes[c])
                bulk_copy_g2s(win + c_off[c], reals + (w_lo[c] & ~(uint64_t)(AL - 1)), c_bytes[c], &bar);
    }
    __syncthreads();
    mbar_wait(&bar, 0);

#pragma unroll 2
    for (int r = 0; r < DC_IPT; ++r) {
        const uint32_t t = (uint32_t)r * DC_THREADS + threadIdx.x;
        const uint64_t i = i0 + t;
        bool emit_sink = false, node_head = false;
        KeyT sink = WT::zero();
        if (t < valid) {
            const KeyT x = tk[t];
            const KeyT W = w_and(x, maskL);
            const KeyT grp = w_shr(x, 2 * L);
            bool first = true;   // first element of its (x >> 2L, W) class = distinct target node
            node_head = true;    // first k-mer of its node
            for (uint64_t q = i; q-- > 0;) {
                const long long s = (long long)t - (long long)(i - q);
                const KeyT y = s >= -(long long)halo ? tk[s] : reals[q];
                if (q + 1 == i) node_head = !w_eq(w_shr(y, L), w_shr(x, L));
                if (!w_eq(w_shr(y, 2 * L), grp)) break;
                if (w_eq(w_and(y, maskL), W)) { first = false; break; }
            }
            if (first) {
                sink = sink_probe(x, W, L, K);
                const uint32_t c = (uint32_t)w_low64(W);
                uint64_t p;
                KeyT at = WT::zero();
                bool have = false;
                if (w_off[c] != ~0u) {
                    const KeyT *wk = win + w_off[c];
                    const uint32_t len = w_len[c];
                    uint32_t base = 0, n = len;   // branch-free lower bound
                    while (n > 1) {
                        const uint32_t half = n >> 1;
                        if (w_lt(wk[base + half - 1], sink)) base += half;
                        n -= half;
                    }
                    if (n == 1 && w_lt(wk[base], sink)) ++base;
                    p = w_lo[c] + base;
                    if (base < len) { at = wk[base]; have = true; }
                    else if (p < m) { at = reals[p]; have = true; }
                } else {
                    p = lower_bound_idx(reals, m, bi, sink);
                    if (p < m) { at = reals[p]; have = true; }
                }
                if (have && w_eq(w_shr(at, L), w_shr(sink, L)))
                    atomicOr(&incoming[p >> 5], 1u << (p & 31));   // node at p has an incoming edge
                else
                    emit_sink = true;
            }
        }
        // i is a multiple of 32 at lane 0: one bitmap word per warp row
        const unsigned hb = __ballot_sync(0xffffffffu, node_head);
        if ((threadIdx.x & 31) == 0 && i < m) heads[i >> 5] = hb;
        warp_append(emit_sink, sink, sink_keys, sink_cap, d_num_sink);
    }
}

// node heads without an incoming real edge -> source-1 candidates (forward-lex packing)
template <typename KeyT>
__global__ void __launch_bounds__(256)
dc_sources_kernel(const KeyT *__restrict__ reals, uint64_t m, Alpha al,
                  const uint32_t *__restrict__ incoming, const uint32_t *__restrict__ heads,
                  KeyT *src_fwd, uint64_t src_cap, unsigned long long *d_num_src) {
    // one bitmap word (32 reals) per thread; only the rare candidates touch the k-mers
    using WT = WordTraits<KeyT>;
    const uint32_t L = al.L, K = al.K;
    const KeyT maskL = WT::low_mask(L);
    const uint64_t wi = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t nwords = (m + 31) / 32;
    uint32_t cand = wi < nwords ? heads[wi] & ~incoming[wi] : 0u;
    while (__any_sync(0xffffffffu, cand != 0)) {
        bool emit_src = cand != 0;
        KeyT fwd = WT::zero();
        if (emit_src) {
            const uint32_t b = __ffs(cand) - 1;
            cand &= cand - 1;
            KeyT x = reals[wi * 32 + b];
            for (uint32_t c = 1; c < K; ++c) {
                KeyT ch = w_and(w_shr(x, L * c), maskL);
                fwd = w_or(fwd, w_shl(ch, L * (K - 1 - c)));
            }
        }
        warp_append(emit_src, fwd, src_fwd, src_cap, d_num_src);
    }
}

template <typename KeyT>
__device__ __forceinline__ int w_clz(KeyT x);
template <> __device__ __forceinline__ int w_clz<uint64_t>(uint64_t x) { return __clzll(x); }
template <> __device__ __forceinline__ int w_clz<u128>(u128 x) {
    return x.hi ? __clzll(x.hi) : 64 + __clzll(x.lo);
}

// Key range of the chunk being built: dummies outside [lo, hi) belong to another range
// (multi-GPU) and are neither counted for the layout nor written.
template <typename KeyT>
struct KeyRange {
    KeyT lo, hi;
    int has_lo, has_hi;
    __device__ __forceinline__ bool holds(KeyT k) const {
        return !(has_lo && w_lt(k, lo)) && !(has_hi && !w_lt(k, hi));
    }
};

// Node i (forward-lex packing f, sorted) contributes the prefixes of length j = lcp+1 .. K-1, i.e.
// the dummy sources with c = K - j leading '$': tight word x_{c+t} = s_t (t = 1..j-1), W = s_j.
// Calls fn(key, c) for each of them, built incrementally: from j to j+1 every node char moves one
// slot down and s_j enters at slot K-1.
template <typename KeyT, typename F>
__device__ __forceinline__ void for_each_new_prefix(KeyT f, uint32_t lcp, Alpha al, F fn) {
    using WT = WordTraits<KeyT>;
    const uint32_t L = al.L, K = al.K, nchars = K - 1;
    const KeyT maskL = WT::low_mask(L);
    const uint32_t j0 = lcp + 1;
    KeyT node = WT::zero();
    for (uint32_t t = 1; t < j0; ++t) {
        KeyT ch = w_and(w_shr(f, L * (nchars - t)), maskL);
        node = w_or(node, w_shl(ch, L * (K - j0 + t)));
    }
    for (uint32_t j = j0; j <= nchars; ++j) {
        const uint32_t c = K - j;
        KeyT sj = w_and(w_shr(f, L * (nchars - j)), maskL);   // W = s_j
        fn(w_or(node, sj), c);
        node = w_or(w_shr(node, L), w_shl(sj, L * (K - 1)));
        node = w_and(node, w_not(maskL));
    }
}

template <typename KeyT>
__device__ __forceinline__ uint32_t fwd_lcp(const KeyT *fwd, uint64_t i, Alpha al) {
    const uint32_t nchars = al.K - 1, bits = nchars * al.L;
    if (i == 0) return 0;
    KeyT d = w_xor(fwd[i], fwd[i - 1]);
    if (w_eq(d, WordTraits<KeyT>::zero())) return nchars;   // cannot happen for distinct nodes
    int lead = w_clz<KeyT>(d) - (WordTraits<KeyT>::BITS - (int)bits);   // equal leading bits
    return (uint32_t)lead / al.L;
}

template <typename KeyT>
__global__ void source_prefix_counts_kernel(const KeyT *__restrict__ fwd, uint64_t d1, Alpha al,
                                            KeyRange<KeyT> range, uint32_t *__restrict__ per_entry) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= d1) return;
    const uint32_t lcp = fwd_lcp(fwd, i, al);
    uint32_t cnt = al.K - 1 - lcp;
    if (range.has_lo || range.has_hi) {
        cnt = 0;
        for_each_new_prefix(fwd[i], lcp, al, [&](KeyT key, uint32_t) { cnt += range.holds(key) ? 1u : 0u; });
    }
    per_entry[i] = cnt;
}

template <typename KeyT>
__global__ void generate_source_dummies_kernel(const KeyT *__restrict__ fwd,
                                               const uint64_t *__restrict__ offsets, uint64_t d1,
                                               Alpha al, KeyRange<KeyT> range, KeyT *__restrict__ out_keys,
                                               uint32_t *__restrict__ out_info,
                                               unsigned long long *level_hist) {
    __shared__ unsigned int sh_levels[MTG_MAX_LEVELS];   // per-block level histogram (ALL ranges)
    for (int c = threadIdx.x; c < MTG_MAX_LEVELS; c += blockDim.x) sh_levels[c] = 0;
    __syncthreads();
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < d1) {
        uint64_t dst = offsets[i];
        for_each_new_prefix(fwd[i], fwd_lcp(fwd, i, al), al, [&](KeyT key, uint32_t c) {
            atomicAdd(&sh_levels[c - 1], 1u);
            if (range.holds(key)) {
                out_keys[dst] = key;
                out_info[dst] = c;
                ++dst;
            }
        });
    }
    __syncthreads();
    for (int c = threadIdx.x; c < MTG_MAX_LEVELS; c += blockDim.x)
        if (sh_levels[c]) atomicAdd(&level_hist[c], (unsigned long long)sh_levels[c]);
}

template <typename KeyT>
__global__ void dummy_ranks_kernel(const KeyT *__restrict__ reals, uint64_t m,
                                   const KeyT *__restrict__ dummy_keys, uint64_t d, Alpha al,
                                   BucketIndex bi, uint64_t *__restrict__ rank) {
    uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= d) return;
    // '$' sorts before every real character: the dummy precedes all reals that agree
    // with it on the non-'$' leading characters -> lower bound of the zero-filled word
    KeyT probe = w_and(dummy_keys[j], w_not(WordTraits<KeyT>::low_mask(al.L)));
    rank[j] = lower_bound_idx(reals, m, bi, probe);
}

template <typename KeyT>
__global__ void merge_dummies_kernel(const KeyT *__restrict__ dummy_keys,
                                     const uint32_t *__restrict__ dummy_info,
                                     const uint64_t *__restrict__ rank, uint64_t d, uint32_t K,
                                     uint32_t row_off, KeyT *__restrict__ row_keys,
                                     uint8_t *__restrict__ row_info, uint32_t *__restrict__ row_counts) {
    uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j == 0 && row_off == 1) {   // not the first key range: only the placeholder row
        row_keys[0] = WordTraits<KeyT>::zero();
        row_info[0] = (uint8_t)((K - 1) | INFO_W_DOLLAR);
        if (row_counts) row_counts[0] = 0;
    }
    if (j == 0 && row_off == 2) {
        // row 0: placeholder; row 1: the all-'$' k-mer (boss_chunk_construct.cpp:310-315)
        row_keys[0] = WordTraits<KeyT>::zero();
        row_keys[1] = WordTraits<KeyT>::zero();
        row_info[0] = (uint8_t)((K - 1) | INFO_W_DOLLAR);
        row_info[1] = (uint8_t)((K - 1) | INFO_W_DOLLAR);
        if (row_counts) { row_counts[0] = 0; row_counts[1] = 0; }
    }
    if (j >= d) return;
    uint64_t pos = rank[j] + j + row_off;
    row_keys[pos] = dummy_keys[j];
    row_info[pos] = (uint8_t)dummy_info[j];
    if (row_counts) row_counts[pos] = 0;
}

constexpr int MR_THREADS = 256;
constexpr int MR_TILE = 2048;
constexpr int MR_CAP = 1024;

// Each block places MR_TILE consecutive reals. The dummies that interleave with them are
// rank[jlo..jhi): two block-wide searches find the range, the per-real count is a search in
// shared memory (the range is short: dummies are a few percent of the rows).
template <typename KeyT>
__global__ void __launch_bounds__(MR_THREADS)
merge_reals_kernel(const KeyT *__restrict__ reals, const uint32_t *__restrict__ real_counts, uint64_t m,
                   const uint64_t *__restrict__ rank, uint64_t d, uint32_t row_off,
                   KeyT *__restrict__ row_keys, uint8_t *__restrict__ row_info,
                   uint32_t *__restrict__ row_counts) {
    __shared__ uint64_t s_rank[MR_CAP];
    __shared__ uint64_t s_range[2];
    const uint64_t i0 = (uint64_t)blockIdx.x * MR_TILE;
    const uint64_t i1 = i0 + MR_TILE < m ? i0 + MR_TILE : m;
    if (threadIdx.x < 2) {
        // cnt(i) = #{j : rank[j] <= i}; [0] = cnt(i0 - 1), [1] = cnt(i1 - 1)
        uint64_t lo = 0, hi = d;
        if (threadIdx.x == 0 && i0 == 0) hi = 0;
        const uint64_t bound = threadIdx.x == 0 ? i0 - 1 : i1 - 1;
        while (lo < hi) {
            uint64_t mid = (lo + hi) >> 1;
            if (rank[mid] <= bound) lo = mid + 1; else hi = mid;
        }
        s_range[threadIdx.x] = lo;
    }
    __syncthreads();
    const uint64_t jlo = s_range[0], jhi = s_range[1];
    const uint32_t nd = (uint32_t)(jhi - jlo < (uint64_t)MR_CAP + 1 ? jhi - jlo : MR_CAP + 1);
    const bool in_smem = nd <= MR_CAP;
    if (in_smem)
        for (uint32_t t = threadIdx.x; t < nd; t += MR_THREADS) s_rank[t] = rank[jlo + t];
    __syncthreads();
    for (uint64_t i = i0 + threadIdx.x; i < i1; i += MR_THREADS) {
        uint64_t cnt;
        if (in_smem) {
            uint32_t lo = 0, hi = nd;
            while (lo < hi) {
                uint32_t mid = (lo + hi) >> 1;
                if (s_rank[mid] <= i) lo = mid + 1; else hi = mid;
            }
            cnt = jlo + lo;
        } else {
            uint64_t lo = jlo, hi = jhi;
            while (lo < hi) {
                uint64_t mid = (lo + hi) >> 1;
                if (rank[mid] <= i) lo = mid + 1; else hi = mid;
            }
            cnt = lo;
        }
        uint64_t pos = i + cnt + row_off;
        row_keys[pos] = reals[i];
        row_info[pos] = 0;
        if (row_counts) row_counts[pos] = real_counts[i];
    }
}

// ---- range-partitioned (multi-GPU) variants, SURVEY.md 8e ----------------------------------
// Target node of x = s_1..s_K is t = s_2..s_K, as a (K-1)-char node word in the layout of
// (key >> L): t = (W << L(K-2)) | (x >> 2L). One entry per real (duplicates are folded later).
template <typename KeyT>
__global__ void target_nodes_kernel(const KeyT *__restrict__ reals, uint64_t m, Alpha al,
                                    KeyT *__restrict__ out) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint32_t L = al.L, K = al.K;
    KeyT x = reals[i];
    KeyT W = w_and(x, WordTraits<KeyT>::low_mask(L));
    out[i] = w_or(w_shl(W, L * (K - 2)), w_shr(x, 2 * L));
}

// reals: the sorted real k-mers of THIS key range; targets: sorted distinct target nodes that
// fall into this range (collected from all ranges). sinks = targets \ source nodes,
// sources-1 = source nodes \ targets (forward-lex packing). Both arrays are sorted by node, so
// the probes of RJ_TILE consecutive reals fall into one contiguous window of the targets: the
// window is staged in shared memory by a TMA bulk copy and searched there; a hit sets the
// target's bit, the sinks are the targets whose bit stays clear (rj_sinks_kernel).
constexpr int RJ_THREADS = 256;
constexpr int RJ_IPT = 8;
constexpr int RJ_TILE = RJ_THREADS * RJ_IPT;
constexpr int RJ_POOL = 3 * RJ_TILE;

// part[j] = lower_bound(targets, node(reals[min(j * RJ_TILE, m - 1)])), j = 0..tiles
template <typename KeyT>
__global__ void rj_partition_kernel(const KeyT *__restrict__ reals, uint64_t m, const KeyT *__restrict__ targets,
                                    uint64_t nt, uint32_t L, uint64_t tiles, uint64_t *__restrict__ part) {
    uint64_t j = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j > tiles) return;
    uint64_t pos = j * RJ_TILE < m ? j * RJ_TILE : m - 1;
    part[j] = lower_bound_dev(targets, nt, w_shr(reals[pos], L));
}

template <typename KeyT>
__global__ void __launch_bounds__(RJ_THREADS)
rj_join_kernel(const KeyT *__restrict__ reals, uint64_t m, const KeyT *__restrict__ targets, uint64_t nt,
               Alpha al, const uint64_t *__restrict__ part, uint32_t *__restrict__ hits, KeyT *src_fwd,
               uint64_t src_cap, unsigned long long *d_num_src) {
    using WT = WordTraits<KeyT>;
    constexpr uint32_t AL = 16 / sizeof(KeyT);
    extern __shared__ __align__(128) unsigned char rj_smem[];
    KeyT *win = reinterpret_cast<KeyT *>(rj_smem);   // [RJ_POOL + AL]
    __shared__ __align__(8) unsigned long long bar;
    const uint32_t L = al.L, K = al.K;
    const KeyT maskL = WT::low_mask(L);
    const uint64_t j = blockIdx.x, i0 = j * RJ_TILE;
    const uint64_t lo = part[j];
    uint64_t hi = part[j + 1] + 1;   // the node of the next tile's first real may start in this tile
    if (hi > nt) hi = nt;
    const uint64_t len = hi > lo ? hi - lo : 0;
    const uint32_t lead = (uint32_t)(lo & (AL - 1));
    const uint64_t span = (lead + len + AL - 1) & ~(uint64_t)(AL - 1);
    const bool staged = span <= (uint64_t)RJ_POOL && len > 0;
    if (threadIdx.x == 0) {
        mbar_init(&bar, 1);
        if (staged) {
            mbar_expect(&bar, (uint32_t)(span * sizeof(KeyT)));
            bulk_copy_g2s(win, targets + (lo - lead), (uint32_t)(span * sizeof(KeyT)), &bar);
        }
    }
    __syncthreads();
    if (staged) mbar_wait(&bar, 0);
    const KeyT *wk = win + lead;
    const uint32_t wlen = (uint32_t)(staged ? len : 0);

#pragma unroll 2
    for (int r = 0; r < RJ_IPT; ++r) {
        const uint64_t i = i0 + (uint64_t)r * RJ_THREADS + threadIdx.x;
        bool emit_src = false;
        KeyT fwd = WT::zero();
        if (i < m) {
            const KeyT x = reals[i];
            const KeyT node = w_shr(x, L);
            const bool node_head = i == 0 || !w_eq(w_shr(reals[i - 1], L), node);
            if (node_head) {
                uint64_t p;
                bool found;
                if (staged) {
                    uint32_t base = 0, n = wlen;   // branch-free lower bound
                    while (n > 1) {
                        const uint32_t half = n >> 1;
                        if (w_lt(wk[base + half - 1], node)) base += half;
                        n -= half;
                    }
                    if (n == 1 && w_lt(wk[base], node)) ++base;
                    p = lo + base;
                    found = base < wlen ? w_eq(wk[base], node) : (p < nt && w_eq(targets[p], node));
                } else {
                    p = lower_bound_dev(targets, nt, node);
                    found = p < nt && w_eq(targets[p], node);
                }
                if (found) {
                    atomicOr(&hits[p >> 5], 1u << (p & 31));   // target p is a source node: no sink
                } else {
                    emit_src = true;
                    for (uint32_t c = 1; c < K; ++c) {
                        KeyT ch = w_and(w_shr(x, L * c), maskL);
                        fwd = w_or(fwd, w_shl(ch, L * (K - 1 - c)));
                    }
                }
            }
        }
        warp_append(emit_src, fwd, src_fwd, src_cap, d_num_src);
    }
}

// targets without a hit -> dummy sinks  t$  (node t, W = '$')
template <typename KeyT>
__global__ void __launch_bounds__(256)
rj_sinks_kernel(const KeyT *__restrict__ targets, uint64_t nt, uint32_t L, const uint32_t *__restrict__ hits,
                KeyT *sink_keys, uint64_t sink_cap, unsigned long long *d_num_sink) {
    const uint64_t wi = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const uint64_t nwords = (nt + 31) / 32;
    uint32_t cand = 0;
    if (wi < nwords) {
        cand = ~hits[wi];
        const uint64_t left = nt - wi * 32;
        if (left < 32) cand &= (1u << left) - 1;
    }
    while (__any_sync(0xffffffffu, cand != 0)) {
        const bool emit = cand != 0;
        KeyT sink = WordTraits<KeyT>::zero();
        if (emit) {
            const uint32_t b = __ffs(cand) - 1;
            cand &= cand - 1;
            sink = w_shl(targets[wi * 32 + b], L);
        }
        warp_append(emit, sink, sink_keys, sink_cap, d_num_sink);
    }
}

}  // namespace

size_t bucket_index_words(Alpha al) {
    uint32_t KL = al.K * al.L;
    uint32_t B = KL < 22 ? KL : 22;
    return ((size_t)1 << B) + 1;
}
static BucketIndex make_bi(Alpha al, const uint64_t *idx) {
    uint32_t KL = al.K * al.L;
    uint32_t B = KL < 22 ? KL : 22;
    return BucketIndex{idx, B, KL - B};
}
template <typename KeyT>
void build_bucket_index(const KeyT *reals, uint64_t m, Alpha al, uint64_t *idx, cudaStream_t stream,
                        unsigned *launches) {
    BucketIndex bi = make_bi(al, idx);
    size_t nthreads = ((size_t)1 << bi.B) + 1;
    bucket_index_kernel<KeyT><<<(unsigned)div_up(nthreads, (size_t)256), 256, 0, stream>>>(reals, m, bi.B, bi.shift,
                                                                                        idx);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

size_t dummy_scratch_bytes(uint64_t m, Alpha al) {
    const uint64_t tiles = div_up(m, (size_t)DC_TILE);
    const uint32_t nch = al.sigma - 1;
    return 8 * (tiles + 1) * nch + 2 * 4 * (m / 32 + 2);
}

template <typename KeyT>
void find_dummy_candidates(const KeyT *reals, uint64_t m, Alpha al, const uint64_t *bucket_idx,
                           void *scratch, KeyT *sink_keys, uint64_t sink_cap,
                           unsigned long long *d_num_sink, KeyT *src_fwd, uint64_t src_cap,
                           unsigned long long *d_num_src, cudaStream_t stream, unsigned *launches) {
    MTG_CUDA(cudaMemsetAsync(d_num_sink, 0, sizeof(unsigned long long), stream));
    MTG_CUDA(cudaMemsetAsync(d_num_src, 0, sizeof(unsigned long long), stream));
    if (m == 0) return;
    const uint64_t tiles = div_up(m, (size_t)DC_TILE);
    const uint32_t nch = al.sigma - 1;
    uint64_t *part = reinterpret_cast<uint64_t *>(scratch);
    uint32_t *incoming = reinterpret_cast<uint32_t *>(part + (tiles + 1) * nch);
    uint32_t *heads = incoming + (m / 32 + 2);
    MTG_CUDA(cudaMemsetAsync(incoming, 0, 4 * (m / 32 + 2), stream));
    const BucketIndex bi = make_bi(al, bucket_idx);
    dc_partition_kernel<KeyT><<<(unsigned)div_up((tiles + 1) * nch, (size_t)256), 256, 0, stream>>>(
        reals, m, al, nch, bi, tiles, part);
    const size_t smem = sizeof(KeyT) * (DC_HALO + DC_TILE + DC_POOL);
    auto kern = dc_sinks_kernel<KeyT>;
    static bool attr_set = false;
    if (!attr_set) {
        MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
        attr_set = true;
    }
    kern<<<(unsigned)tiles, DC_THREADS, smem, stream>>>(reals, m, al, nch, bi, part, incoming, heads, sink_keys,
                                                       sink_cap, d_num_sink);
    dc_sources_kernel<KeyT><<<(unsigned)div_up(div_up(m, (size_t)32), (size_t)256), 256, 0, stream>>>(
        reals, m, al, incoming, heads, src_fwd, src_cap, d_num_src);
    MTG_CUDA(cudaGetLastError());
    *launches += 3;
}

template <typename KeyT>
static KeyRange<KeyT> make_range(const KeyT *lo, const KeyT *hi) {
    KeyRange<KeyT> r;
    r.lo = lo ? *lo : WordTraits<KeyT>::zero();
    r.hi = hi ? *hi : WordTraits<KeyT>::zero();
    r.has_lo = lo ? 1 : 0;
    r.has_hi = hi ? 1 : 0;
    return r;
}

template <typename KeyT>
void source_prefix_counts(const KeyT *src_fwd, uint64_t d1, Alpha al, const KeyT *lo, const KeyT *hi,
                          uint32_t *per_entry, cudaStream_t stream, unsigned *launches) {
    if (d1 == 0) return;
    source_prefix_counts_kernel<KeyT><<<(unsigned)div_up(d1, (size_t)256), 256, 0, stream>>>(
        src_fwd, d1, al, make_range(lo, hi), per_entry);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

template <typename KeyT>
void generate_source_dummies(const KeyT *src_fwd, const uint64_t *offsets, uint64_t d1, Alpha al,
                             const KeyT *lo, const KeyT *hi, KeyT *out_keys, uint32_t *out_info,
                             unsigned long long *level_hist, cudaStream_t stream,
                             unsigned *launches) {
    if (d1 == 0) return;
    generate_source_dummies_kernel<KeyT><<<(unsigned)div_up(d1, (size_t)128), 128, 0, stream>>>(
        src_fwd, offsets, d1, al, make_range(lo, hi), out_keys, out_info, level_hist);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

template <typename KeyT>
void dummy_ranks(const KeyT *reals, uint64_t m, const KeyT *dummy_keys, uint64_t d, Alpha al,
                 const uint64_t *bucket_idx, uint64_t *rank, cudaStream_t stream, unsigned *launches) {
    if (d == 0) return;
    dummy_ranks_kernel<KeyT><<<(unsigned)div_up(d, (size_t)256), 256, 0, stream>>>(
        reals, m, dummy_keys, d, al, make_bi(al, bucket_idx), rank);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

template <typename KeyT>
void merge_rows(const KeyT *reals, const uint32_t *real_counts, uint64_t m, const KeyT *dummy_keys,
                const uint32_t *dummy_info, const uint64_t *rank, uint64_t d, Alpha al,
                KeyT *row_keys, uint8_t *row_info, uint32_t *row_counts, cudaStream_t stream,
                unsigned *launches, uint32_t row_off) {
    merge_dummies_kernel<KeyT><<<(unsigned)div_up(d ? d : 1, (size_t)256), 256, 0, stream>>>(
        dummy_keys, dummy_info, rank, d, al.K, row_off, row_keys, row_info, row_counts);
    *launches += 1;
    if (m) {
        merge_reals_kernel<KeyT><<<(unsigned)div_up(m, (size_t)MR_TILE), MR_THREADS, 0, stream>>>(
            reals, real_counts, m, rank, d, row_off, row_keys, row_info, row_counts);
        *launches += 1;
    }
    MTG_CUDA(cudaGetLastError());
}

template <typename KeyT>
void target_nodes(const KeyT *reals, uint64_t m, Alpha al, KeyT *out, cudaStream_t stream,
                  unsigned *launches) {
    if (m == 0) return;
    target_nodes_kernel<KeyT><<<(unsigned)div_up(m, (size_t)256), 256, 0, stream>>>(reals, m, al, out);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

size_t ranged_scratch_bytes(uint64_t m, uint64_t nt) {
    return 8 * (div_up(m, (size_t)RJ_TILE) + 2) + 4 * (nt / 32 + 2);
}

template <typename KeyT>
void ranged_dummy_candidates(const KeyT *reals, uint64_t m, const KeyT *targets, uint64_t nt, Alpha al,
                             void *scratch, KeyT *sink_keys, uint64_t sink_cap,
                             unsigned long long *d_num_sink, KeyT *src_fwd, uint64_t src_cap,
                             unsigned long long *d_num_src, cudaStream_t stream, unsigned *launches) {
    MTG_CUDA(cudaMemsetAsync(d_num_sink, 0, sizeof(unsigned long long), stream));
    MTG_CUDA(cudaMemsetAsync(d_num_src, 0, sizeof(unsigned long long), stream));
    const uint64_t tiles = div_up(m, (size_t)RJ_TILE);
    uint64_t *part = reinterpret_cast<uint64_t *>(scratch);
    uint32_t *hits = reinterpret_cast<uint32_t *>(part + tiles + 2);
    MTG_CUDA(cudaMemsetAsync(hits, 0, 4 * (nt / 32 + 2), stream));
    if (m) {
        rj_partition_kernel<KeyT><<<(unsigned)div_up(tiles + 1, (size_t)256), 256, 0, stream>>>(
            reals, m, targets, nt, al.L, tiles, part);
        const size_t smem = sizeof(KeyT) * (RJ_POOL + 2);
        auto kern = rj_join_kernel<KeyT>;
        static bool attr_set = false;
        if (!attr_set) {
            MTG_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
            attr_set = true;
        }
        kern<<<(unsigned)tiles, RJ_THREADS, smem, stream>>>(reals, m, targets, nt, al, part, hits, src_fwd, src_cap,
                                                           d_num_src);
        *launches += 2;
    }
    if (nt) {
        rj_sinks_kernel<KeyT><<<(unsigned)div_up(div_up(nt, (size_t)32), (size_t)256), 256, 0, stream>>>(
            targets, nt, al.L, hits, sink_keys, sink_cap, d_num_sink);
        *launches += 1;
    }
    MTG_CUDA(cudaGetLastError());
}

// bounds[b] = first position of the sorted array whose value >> shift is >= b, b = 0..2^B
template <typename KeyT>
void bucket_bounds(const KeyT *sorted, uint64_t n, uint32_t B, uint32_t shift, uint64_t *bounds,
                   cudaStream_t stream, unsigned *launches) {
    size_t nthreads = ((size_t)1 << B) + 1;
    bucket_index_kernel<KeyT><<<(unsigned)div_up(nthreads, (size_t)256), 256, 0, stream>>>(sorted, n, B, shift,
                                                                                        bounds);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

#define INSTANTIATE(KeyT)                                                                          \
    template void build_bucket_index<KeyT>(const KeyT *, uint64_t, Alpha, uint64_t *, cudaStream_t, \
                                           unsigned *);                                           \
    template void find_dummy_candidates<KeyT>(const KeyT *, uint64_t, Alpha, const uint64_t *, void *, KeyT *, uint64_t, \
                                              unsigned long long *, KeyT *, uint64_t,              \
                                              unsigned long long *, cudaStream_t, unsigned *);     \
    template void source_prefix_counts<KeyT>(const KeyT *, uint64_t, Alpha, const KeyT *, const KeyT *, \
                                             uint32_t *, cudaStream_t, unsigned *);                \
    template void generate_source_dummies<KeyT>(const KeyT *, const uint64_t *, uint64_t, Alpha,   \
                                                const KeyT *, const KeyT *, KeyT *, uint32_t *,    \
                                                unsigned long long *, cudaStream_t, unsigned *);   \
    template void dummy_ranks<KeyT>(const KeyT *, uint64_t, const KeyT *, uint64_t, Alpha,         \
                                    const uint64_t *, uint64_t *, cudaStream_t, unsigned *);                         \
    template void merge_rows<KeyT>(const KeyT *, const uint32_t *, uint64_t, const KeyT *,         \
                                   const uint32_t *, const uint64_t *, uint64_t, Alpha, KeyT *,    \
                                   uint8_t *, uint32_t *, cudaStream_t, unsigned *, uint32_t);     \
    template void target_nodes<KeyT>(const KeyT *, uint64_t, Alpha, KeyT *, cudaStream_t, unsigned *); \
    template void ranged_dummy_candidates<KeyT>(const KeyT *, uint64_t, const KeyT *, uint64_t, Alpha, \
                                                void *, KeyT *, uint64_t,               \
                                                unsigned long long *, KeyT *, uint64_t,            \
                                                unsigned long long *, cudaStream_t, unsigned *);   \
    template void bucket_bounds<KeyT>(const KeyT *, uint64_t, uint32_t, uint32_t, uint64_t *,      \
                                      cudaStream_t, unsigned *);
INSTANTIATE(uint64_t)
INSTANTIATE(u128)

}  // namespace mtg
