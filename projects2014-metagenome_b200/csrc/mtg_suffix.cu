// K1s: suffix-filtered chunks (`metagraph build --suffix S`).
// Replaces KmerExtractorBOSS::sequence_to_kmers (kmer/kmer_extractor.cpp:317-381) with
// KMerBOSS::match_suffix (kmer/kmer_boss.hpp:106-112) and the collector modes BASIC / BOTH
// (kmer/kmer_collector.cpp:26-67) behind BOSSChunkConstructor's suffix branch
// (boss_chunk_construct.cpp:961-981, :1002-1013), where the chunk is built straight from the
// sorted collector contents (no dummy recovery: every valid segment is framed as
// '$'^(K-1) seg '$' by the extractor itself).
//
// A framed k-mer is one of
//   A  $^nd s .. s | W      the window ending at a segment character (nd = missing left chars)
//   B  s .. s | $           the window ending at the '$' after the segment
// and, in mode BOTH, the same two shapes of the reverse-complemented segment. '$' only occurs as
// a prefix of the node or as W, so a k-mer is (tight word with '$' slots zero, nd, W-is-'$'),
// the row representation of mtg_emit.cu. The three fields are packed into one integer whose
// order is the order of the BOSS-alphabet words (node, more '$' first, '$' before letters):
//   composite = node bits << (L + 8) | ((127 - nd) << 1 | (W != '$')) << L | W
// so the existing radix sort / unique / count kernels apply unchanged.
#include "mtg_scan.cuh"
#include "mtg_stages.cuh"

namespace mtg {

namespace {

constexpr int SX_THREADS = 256;
constexpr int SX_PPT = 4;
constexpr int SX_TILE = SX_THREADS * SX_PPT;

struct SuffixFilter {
    uint32_t len;
    uint8_t code[MTG_SUFFIX_MAX];   // BOSS codes: 0 = '$', tight + 1 otherwise
};

template <typename KeyT>
__device__ __forceinline__ KeyT make_composite(KeyT node, uint32_t nd, bool w_dollar, uint32_t w, uint32_t L) {
    using WT = WordTraits<KeyT>;
    const uint32_t sortinfo = ((127u - nd) << 1) | (w_dollar ? 0u : 1u);
    return w_or(w_shl(node, L + 8), WT::from_u64(((uint64_t)sortinfo << L) | (w_dollar ? 0u : w)));
}

template <typename KeyT>
__device__ __forceinline__ void append2(bool emit, KeyT v, uint32_t cnt, KeyT *out, uint32_t *out_counts,
                                        uint64_t cap, unsigned long long *counter) {
    unsigned bal = __ballot_sync(0xffffffffu, emit);
    if (bal == 0) return;
    unsigned long long base = 0;
    if ((threadIdx.x & 31) == 0) base = atomicAdd(counter, (unsigned long long)__popc(bal));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (emit) {
        uint64_t dst = base + __popc(bal & lanemask_lt());
        if (dst < cap) {
            out[dst] = v;
            if (out_counts) out_counts[dst] = cnt;
        }
    }
}

// One thread per character position p of the packed reads. The four k-mer shapes anchored at p:
//   0: forward window ending at p                      (always, if p is in a segment of >= K chars)
//   1: forward window ending at the '$' after p        (p is the last char of its segment)
//   2: reverse-complement window "ending" at p         (BOTH; the rc segment read towards p)
//   3: reverse-complement window ending at its '$'     (BOTH; p is the first char of its segment)
template <typename KeyT>
__global__ void __launch_bounds__(SX_THREADS)
extract_suffix_kernel(const uint8_t *__restrict__ packed, uint64_t nbytes, ExtractLut lut, Alpha al,
                      SuffixFilter sf, int both, KeyT *__restrict__ out_keys,
                      uint32_t *__restrict__ out_counts, uint64_t cap, unsigned long long *out_counter,
                      const uint64_t *__restrict__ read_starts, const uint32_t *__restrict__ read_counts,
                      uint64_t n_reads) {
    using WT = WordTraits<KeyT>;
    extern __shared__ uint8_t sx_codes[];   // [SX_TILE + 2K], code of position tile_start - K + i
    const uint32_t K = al.K, L = al.L;
    const uint64_t tile_start = (uint64_t)blockIdx.x * SX_TILE;
    for (uint32_t i = threadIdx.x; i < SX_TILE + 2 * K; i += SX_THREADS) {
        const long long gp = (long long)tile_start - (long long)K + i;
        sx_codes[i] = gp >= 0 && (uint64_t)gp < nbytes ? lut.code[packed[gp]] : 0xff;
    }
    __syncthreads();

    for (int q = 0; q < SX_PPT; ++q) {
        const uint32_t r = q * SX_THREADS + threadIdx.x;
        const uint64_t gp = tile_start + r;
        const uint8_t *c0 = sx_codes + K + r;   // c0[0] = my character
        uint32_t jb = 0, jf = 0;
        bool in_seg = gp < nbytes && c0[0] != 0xff;
        if (in_seg) {
            while (jb < K - 1 && c0[-1 - (int)jb] != 0xff) ++jb;   // valid chars before p (clamped)
            while (jf < K - 1 && c0[1 + jf] != 0xff) ++jf;         // valid chars after p (clamped)
            in_seg = jb + 1 + jf >= K;                             // the segment holds >= K characters
        }
        uint32_t cnt = 1;
        if (in_seg && read_counts) {
            uint64_t lo = 0, hi = n_reads;   // read containing p: last start <= gp
            while (hi - lo > 1) {
                uint64_t mid = (lo + hi) >> 1;
                if (read_starts[mid] <= gp) lo = mid; else hi = mid;
            }
            cnt = read_counts[lo];
        }
        for (int shape = 0; shape < (both ? 4 : 2); ++shape) {
            bool emit = in_seg;
            if (shape == 1) emit = emit && jf == 0;
            if (shape == 3) emit = emit && jb == 0;
            KeyT comp = WT::zero();
            if (emit) {
                // BOSS code of s_i (i = 1..K), 0 = '$'
                const bool rc = shape >= 2, tail = shape & 1;
                const uint32_t avail = rc ? jf : jb;
                const uint32_t nd = tail ? 0 : K - 1 - avail;
                auto code_at = [&](uint32_t i) -> uint32_t {   // tight code of s_i, i > nd, i < K or W
                    if (!rc) return tail ? c0[(int)i - (int)K + 1] : c0[(int)i - (int)K];
                    return 3u - (tail ? c0[K - 1 - i] : c0[K - i]);
                };
                // suffix filter: s_{K-len} .. s_{K-1}
                for (uint32_t t = 0; t < sf.len && emit; ++t) {
                    const uint32_t i = K - sf.len + t;
                    const uint32_t b = i <= nd ? 0u : code_at(i) + 1;
                    emit = b == sf.code[t];
                }
                if (emit) {
                    KeyT node = WT::zero();
                    for (uint32_t i = nd + 1; i < K; ++i)
                        node = w_or(node, w_shl(WT::from_u64(code_at(i)), L * (i - 1)));
                    comp = make_composite(node, nd, tail, tail ? 0u : code_at(K), L);
                }
            }
            append2(emit, comp, cnt, out_keys, out_counts, cap, out_counter);
        }
    }
}

// keep[i] = 0 for a redundant dummy sink: W = '$', F > 0 and the next k-mer has the same node
// (initialize_chunk, boss_chunk.cpp:80-85)
template <typename KeyT>
__global__ void suffix_keep_kernel(const KeyT *__restrict__ comp, uint64_t m, Alpha al,
                                   uint32_t *__restrict__ keep) {
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= m) return;
    const uint32_t L = al.L, K = al.K;
    KeyT x = comp[i];
    const uint32_t si = (uint32_t)w_low64(w_shr(x, L)) & 0xffu;
    const uint32_t nd = 127u - (si >> 1);
    uint32_t k = 1;
    if (!(si & 1u) && nd < K - 1 && i + 1 < m) {
        KeyT y = comp[i + 1];
        // same node: node bits and nd equal (everything above the W$ bit)
        if (w_eq(w_shr(x, L + 1), w_shr(y, L + 1))) k = 0;
    }
    keep[i] = k;
}

template <typename KeyT>
__global__ void suffix_rows_kernel(const KeyT *__restrict__ comp, const uint32_t *__restrict__ counts,
                                   const uint32_t *__restrict__ keep, const uint64_t *__restrict__ offsets,
                                   uint64_t m, Alpha al, KeyT *__restrict__ row_keys,
                                   uint8_t *__restrict__ row_info, uint32_t *__restrict__ row_counts) {
    using WT = WordTraits<KeyT>;
    uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i == 0) {   // row 0: placeholder
        row_keys[0] = WT::zero();
        row_info[0] = (uint8_t)((al.K - 1) | INFO_W_DOLLAR);
        if (row_counts) row_counts[0] = 0;
    }
    if (i >= m || !keep[i]) return;
    const uint32_t L = al.L;
    KeyT x = comp[i];
    const uint32_t si = (uint32_t)w_low64(w_shr(x, L)) & 0xffu;
    const uint32_t nd = 127u - (si >> 1);
    const uint64_t pos = offsets[i] + 1;
    row_keys[pos] = w_or(w_shl(w_shr(x, L + 8), L), w_and(x, WT::low_mask(L)));
    row_info[pos] = (uint8_t)(nd | ((si & 1u) ? 0u : INFO_W_DOLLAR));
    if (row_counts) row_counts[pos] = counts[i];
}

}  // namespace

template <typename KeyT>
void extract_suffix_kmers(const ExtractArgs &a, Alpha al, const uint8_t *suffix_codes, uint32_t suffix_len,
                          KeyT *out_keys, uint64_t cap, cudaStream_t stream) {
    if (a.nbytes == 0) return;
    SuffixFilter sf;
    sf.len = suffix_len;
    for (uint32_t i = 0; i < MTG_SUFFIX_MAX; ++i) sf.code[i] = i < suffix_len ? suffix_codes[i] : 0;
    const size_t smem = SX_TILE + 2 * (size_t)al.K;
    extract_suffix_kernel<KeyT><<<(unsigned)div_up(a.nbytes, (size_t)SX_TILE), SX_THREADS, smem, stream>>>(
        a.packed, a.nbytes, *a.lut, al, sf, a.both ? 1 : 0, out_keys, a.out_counts, cap, a.out_counter,
        a.read_starts, a.read_counts, a.n_reads);
    MTG_CUDA(cudaGetLastError());
}

template <typename KeyT>
void suffix_keep_flags(const KeyT *comp, uint64_t m, Alpha al, uint32_t *keep, cudaStream_t stream,
                       unsigned *launches) {
    if (m == 0) return;
    suffix_keep_kernel<KeyT><<<(unsigned)div_up(m, (size_t)256), 256, 0, stream>>>(comp, m, al, keep);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

template <typename KeyT>
void suffix_rows(const KeyT *comp, const uint32_t *counts, const uint32_t *keep, const uint64_t *offsets,
                 uint64_t m, Alpha al, KeyT *row_keys, uint8_t *row_info, uint32_t *row_counts,
                 cudaStream_t stream, unsigned *launches) {
    suffix_rows_kernel<KeyT><<<(unsigned)div_up(m ? m : 1, (size_t)256), 256, 0, stream>>>(
        comp, counts, keep, offsets, m, al, row_keys, row_info, row_counts);
    MTG_CUDA(cudaGetLastError());
    *launches += 1;
}

#define INSTANTIATE(KeyT)                                                                              \
    template void extract_suffix_kmers<KeyT>(const ExtractArgs &, Alpha, const uint8_t *, uint32_t, KeyT *, \
                                             uint64_t, cudaStream_t);                                  \
    template void suffix_keep_flags<KeyT>(const KeyT *, uint64_t, Alpha, uint32_t *, cudaStream_t, unsigned *); \
    template void suffix_rows<KeyT>(const KeyT *, const uint32_t *, const uint32_t *, const uint64_t *,  \
                                    uint64_t, Alpha, KeyT *, uint8_t *, uint32_t *, cudaStream_t, unsigned *);
INSTANTIATE(uint64_t)
INSTANTIATE(u128)

}  // namespace mtg
