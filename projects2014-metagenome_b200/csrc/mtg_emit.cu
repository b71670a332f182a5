// K6: BOSS::Chunk assembly -- the fused adjacent-compare kernel.
// Replaces initialize_chunk (graph/representation/succinct/boss_chunk.cpp:31-133):
// one pass over the merged sorted rows emits `last`, `W` (with the "minus" flag),
// the F boundaries and the weights, bit-packed in sdsl::int_vector layout
// (LSB-first in little-endian 64-bit words; sdsl int_vector.hpp:594-603).
//
// Row representation (see mtg_stages.cuh): tight word with '$' slots zeroed + info
// byte (nd = number of leading '$' of the node, bit 7 = W is '$').
//   same node      : key>>L equal  and nd equal               (boss_chunk.cpp:78)
//   same x[2..k]   : key>>2L equal and max(nd,1) equal        (compare_suffix(.,.,1), :94)
//   minus flag     : an earlier row of the same x[2..k] group has the same W.
//     (last_kmer[curW] is only updated by unflagged rows, so "the previous unflagged
//      row with this W is in my group" == "some earlier row of my group has this W".)
// The redundant-dummy-sink skip (:80-85) cannot trigger on this path: sinks are only
// generated for nodes without outgoing real edges (boss_chunk_construct.cpp:54-98).
#include "mtg_stages.cuh"

namespace mtg {

namespace {

constexpr int EM_THREADS = 256;
constexpr int EM_IPT = 8;
constexpr int EM_TILE = EM_THREADS * EM_IPT;   // rows per block, multiple of 64

template <typename KeyT>
__global__ void __launch_bounds__(EM_THREADS)
emit_kernel(const KeyT *__restrict__ row_keys, const uint8_t *__restrict__ row_info,
            const uint32_t *__restrict__ row_counts, uint64_t n_rows, Alpha al,
            uint32_t weights_width, uint64_t *__restrict__ W_out, uint64_t *__restrict__ last_out,
            uint64_t *__restrict__ weights_out, unsigned long long *F) {
    using WT = WordTraits<KeyT>;
    __shared__ uint8_t sW[EM_TILE];
    __shared__ uint8_t sLast[EM_TILE];
    __shared__ uint32_t sCnt[EM_TILE];
    const uint32_t L = al.L, K = al.K, sigma = al.sigma;
    const KeyT maskL = WT::low_mask(L);
    const uint64_t tile_base = (uint64_t)blockIdx.x * EM_TILE;
    const uint32_t maxw = weights_width >= 32 ? 0xffffffffu : ((1u << weights_width) - 1);

    for (int it = 0; it < EM_IPT; ++it) {
        uint32_t r = it * EM_THREADS + threadIdx.x;
        uint64_t p = tile_base + r;
        uint32_t w = 0, last = 0, cnt = 0;
        if (p > 0 && p < n_rows) {
            KeyT x = row_keys[p];
            uint32_t info = row_info[p];
            uint32_t nd = info & 127;
            // last
            last = 1;
            if (p + 1 < n_rows) {
                uint32_t ninfo = row_info[p + 1];
                if ((ninfo & 127) == nd && w_eq(w_shr(row_keys[p + 1], L), w_shr(x, L))) last = 0;
            }
            // W with minus flag
            if (!(info & INFO_W_DOLLAR)) {
                uint32_t wc = (uint32_t)w_low64(w_and(x, maskL));
                w = wc + 1;
                KeyT grp = w_shr(x, 2 * L);
                uint32_t gnd = nd > 1 ? nd : 1;
                for (uint64_t q = p; q-- > 1;) {
                    uint32_t qi = row_info[q];
                    uint32_t qnd = qi & 127;
                    KeyT y = row_keys[q];
                    if ((qnd > 1 ? qnd : 1) != gnd || !w_eq(w_shr(y, 2 * L), grp)) break;
                    if (!(qi & INFO_W_DOLLAR) && (uint32_t)w_low64(w_and(y, maskL)) == wc) {
                        w += sigma;
                        break;
                    }
                }
                if (row_counts && nd == 0) {
                    uint32_t c = row_counts[p];
                    cnt = c < maxw ? c : maxw;
                }
            }
            // F boundaries: F[c] = (first row with F char >= c) - 1
            uint32_t f = nd == K - 1 ? 0 : (uint32_t)w_low64(w_and(w_shr(x, L * (K - 1)), maskL)) + 1;
            uint32_t fprev = 0;
            if (p > 1) {
                uint32_t pi = row_info[p - 1] & 127;
                fprev = pi == K - 1 ? 0
                        : (uint32_t)w_low64(w_and(w_shr(row_keys[p - 1], L * (K - 1)), maskL)) + 1;
            }
            for (uint32_t c = fprev + 1; c <= f; ++c) F[c] = p - 1;
        }
        sW[r] = (uint8_t)w;
        sLast[r] = (uint8_t)last;
        sCnt[r] = cnt;
    }
    __syncthreads();

    // ---- bit packing: EM_TILE rows -> EM_TILE*width/64 words per array ----
    const uint64_t rows_left = n_rows > tile_base ? n_rows - tile_base : 0;
    const uint32_t tile_rows = rows_left < (uint64_t)EM_TILE ? (uint32_t)rows_left : EM_TILE;
    {   // last: 1 bit
        const uint32_t nwords = (tile_rows + 63) / 64;
        for (uint32_t t = threadIdx.x; t < nwords; t += EM_THREADS) {
            uint64_t word = 0;
            for (int b = 0; b < 64; ++b) word |= (uint64_t)(sLast[t * 64 + b] & 1) << b;
            last_out[tile_base / 64 + t] = word;
        }
    }
    {   // W: al.W_width bits
        const uint32_t width = al.W_width;
        const uint32_t nwords = (uint32_t)(((uint64_t)tile_rows * width + 63) / 64);
        for (uint32_t t = threadIdx.x; t < nwords; t += EM_THREADS) {
            uint64_t word = 0;
            uint32_t bit0 = t * 64;
            uint32_t e = bit0 / width;
            int sh = (int)(e * width) - (int)bit0;   // <= 0 for the straddling first entry
            while (sh < 64 && e < (uint32_t)EM_TILE) {
                uint64_t v = sW[e];
                word |= sh >= 0 ? v << sh : v >> (-sh);
                ++e;
                sh += width;
            }
            W_out[tile_base / 64 * width + t] = word;
        }
    }
    if (weights_out) {
        const uint32_t width = weights_width;
        const uint32_t nwords = (uint32_t)(((uint64_t)tile_rows * width + 63) / 64);
        for (uint32_t t = threadIdx.x; t < nwords; t += EM_THREADS) {
            uint64_t word = 0;
            uint32_t bit0 = t * 64;
            uint32_t e = bit0 / width;
            int sh = (int)(e * width) - (int)bit0;
            while (sh < 64 && e < (uint32_t)EM_TILE) {
                uint64_t v = sCnt[e];
                word |= sh >= 0 ? v << sh : v >> (-sh);
                ++e;
                sh += width;
            }
            weights_out[tile_base / 64 * width + t] = word;
        }
    }
}

__global__ void init_F_kernel(unsigned long long *F, uint32_t sigma, uint64_t n_rows) {
    if (threadIdx.x < 32) F[threadIdx.x] = threadIdx.x == 0 || threadIdx.x >= sigma ? 0 : n_rows - 1;
}

}  // namespace

template <typename KeyT>
void emit_chunk(const KeyT *row_keys, const uint8_t *row_info, const uint32_t *row_counts,
                uint64_t n_rows, Alpha al, uint32_t weights_width, uint64_t *W, uint64_t *last,
                uint64_t *weights, unsigned long long *F, cudaStream_t stream, unsigned *launches) {
    init_F_kernel<<<1, 32, 0, stream>>>(F, al.sigma, n_rows);
    emit_kernel<KeyT><<<(unsigned)div_up(n_rows, (size_t)EM_TILE), EM_THREADS, 0, stream>>>(
        row_keys, row_info, row_counts, n_rows, al, weights_width, W, last, weights, F);
    MTG_CUDA(cudaGetLastError());
    *launches += 2;
}
template void emit_chunk<uint64_t>(const uint64_t *, const uint8_t *, const uint32_t *, uint64_t, Alpha,
                                   uint32_t, uint64_t *, uint64_t *, uint64_t *, unsigned long long *,
                                   cudaStream_t, unsigned *);
template void emit_chunk<u128>(const u128 *, const uint8_t *, const uint32_t *, uint64_t, Alpha, uint32_t,
                               uint64_t *, uint64_t *, uint64_t *, unsigned long long *, cudaStream_t,
                               unsigned *);

}  // namespace mtg
