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


// ---------------------------------------------------------------------------
// Fused merge + emit (DNA, L = 2): the rows are never materialised in HBM. A block owns FE_TILE
// consecutive OUTPUT rows; it stages them (plus FE_HALO rows before and one after) in shared
// memory straight from the sorted reals and the sorted dummies:
//   row of dummy j  = row_off + j + rank[j]        (rank[j] = reals sorting before it)
//   row of real  i  = row_off + i + #{j : rank[j] <= i}
// The dummies of the tile are found from a per-tile partition (fe_partition_kernel), a bitmap of
// the dummy slots gives every other slot its real by a popcount prefix. Then the same
// adjacent-compare as emit_kernel runs out of shared memory; `last` is packed by warp ballots and
// W (4 bits) by a byte -> nibble squeeze. The backward scan of the minus flag needs at most the
// rows of one x[2..k] group: <= 4*4 reals + 4 sinks + 4 sources = 24 < FE_HALO for DNA.
// ---------------------------------------------------------------------------
constexpr int FE_THREADS = 256;
constexpr int FE_TILE = 2048;
constexpr int FE_HALO = 32;
constexpr int FE_S = FE_TILE + FE_HALO + 1;
constexpr int FE_BMW = (FE_S + 31) / 32;   // bitmap words

__global__ void fe_partition_kernel(const uint64_t *__restrict__ rank, uint64_t d, uint32_t row_off,
                                    uint64_t ntiles, uint64_t *__restrict__ part) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= ntiles) return;
    uint64_t target = t * FE_TILE;
    target = target > FE_HALO ? target - FE_HALO : 0;
    // smallest j with row(j) >= target
    uint64_t lo = 0, hi = d;
    while (lo < hi) {
        uint64_t mid = (lo + hi) >> 1;
        if (row_off + mid + rank[mid] < target) lo = mid + 1; else hi = mid;
    }
    part[t] = lo;
}

template <typename KeyT>
__global__ void __launch_bounds__(FE_THREADS)
emit_merged_kernel(const KeyT *__restrict__ reals, const uint32_t *__restrict__ real_counts, uint64_t m,
                   const KeyT *__restrict__ dummy_keys, const uint32_t *__restrict__ dummy_info,
                   const uint64_t *__restrict__ rank, uint64_t d, const uint64_t *__restrict__ part,
                   uint32_t row_off, Alpha al, uint32_t weights_width, uint64_t *__restrict__ W_out,
                   uint64_t *__restrict__ last_out, uint64_t *__restrict__ weights_out,
                   unsigned long long *F) {
    using WT = WordTraits<KeyT>;
    __shared__ __align__(16) KeyT sKey[FE_S];
    __shared__ __align__(8) uint8_t sInfo[(FE_S + 7) & ~7];
    __shared__ uint32_t sbm[FE_BMW], spre[FE_BMW];
    __shared__ uint32_t sCnt[FE_TILE];
    __shared__ __align__(8) uint8_t sW[FE_TILE];
    constexpr uint32_t L = 2;
    const uint32_t K = al.K, sigma = al.sigma;
    const KeyT maskL = WT::low_mask(L);
    const uint64_t n_rows = m + d + row_off;
    const uint64_t p0 = (uint64_t)blockIdx.x * FE_TILE;
    const long long pstart = (long long)p0 - FE_HALO;
    const uint64_t pend = p0 + FE_TILE + 1 < n_rows ? p0 + FE_TILE + 1 : n_rows;   // rows staged: [pstart, pend)
    const uint64_t j0 = part[blockIdx.x];
    const unsigned tid = threadIdx.x, lane = tid & 31;

    for (uint32_t w = tid; w < FE_BMW; w += FE_THREADS) sbm[w] = 0;
    __syncthreads();
    // rows below row_off (the placeholder, the all-'$' k-mer) and the slots before row 0
    for (uint32_t s = tid; (long long)s + pstart < (long long)row_off && s < FE_S; s += FE_THREADS) {
        const long long p = pstart + s;
        sKey[s] = WT::zero();
        sInfo[s] = p >= 0 ? (uint8_t)((K - 1) | INFO_W_DOLLAR) : (uint8_t)0xff;
        atomicOr(&sbm[s >> 5], 1u << (s & 31));
    }
    for (uint64_t j = j0 + tid; j < d; j += FE_THREADS) {
        const uint64_t row = row_off + j + rank[j];
        if (row >= pend) break;
        const uint32_t s = (uint32_t)((long long)row - pstart);
        sKey[s] = dummy_keys[j];
        sInfo[s] = (uint8_t)dummy_info[j];
        atomicOr(&sbm[s >> 5], 1u << (s & 31));
    }
    __syncthreads();
    if (tid < 32) {   // exclusive popcount prefix over the bitmap words (3 consecutive words per lane)
        uint32_t c[3], sum = 0;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const uint32_t w = lane * 3 + i;
            c[i] = w < FE_BMW ? __popc(sbm[w]) : 0;
            sum += c[i];
        }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= (unsigned)o) inc += t;
        }
        uint32_t run = inc - sum;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const uint32_t w = lane * 3 + i;
            if (w < FE_BMW) spre[w] = run;
            run += c[i];
        }
    }
    __syncthreads();
    // every other slot holds a real: the i_first-th real is the first one of the tile
    const long long first_row = pstart > (long long)row_off ? pstart : (long long)row_off;
    const uint64_t i_first = (uint64_t)(first_row - row_off) - j0;
    for (uint32_t s = tid; s < FE_S; s += FE_THREADS) {
        const uint32_t word = sbm[s >> 5];
        const bool is_real = !((word >> (s & 31)) & 1);
        uint32_t c = 0;
        if (is_real) {
            const uint32_t bits_before = spre[s >> 5] + __popc(word & ((1u << (s & 31)) - 1));
            const uint64_t i = i_first + (s - bits_before);
            if (i < m) {
                sKey[s] = reals[i];
                sInfo[s] = 0;
                if (real_counts && s >= FE_HALO && s < FE_HALO + FE_TILE) c = real_counts[i];
            } else {
                sKey[s] = WT::zero();
                sInfo[s] = 0xff;   // past the last row
            }
        }
        if (s >= FE_HALO && s < FE_HALO + FE_TILE) sCnt[s - FE_HALO] = c;
    }
    __syncthreads();

    const uint32_t maxw = weights_width >= 32 ? 0xffffffffu : ((1u << weights_width) - 1);
    uint32_t *last32 = reinterpret_cast<uint32_t *>(last_out);
#pragma unroll 2
    for (int it = 0; it < FE_TILE / FE_THREADS; ++it) {
        const uint32_t r = it * FE_THREADS + tid;
        const uint32_t s = r + FE_HALO;
        const uint64_t p = p0 + r;
        uint32_t w = 0, last = 0, cnt = 0;
        if (p > 0 && p < n_rows) {
            const KeyT x = sKey[s];
            const uint32_t info = sInfo[s];
            const uint32_t nd = info & 127;
            last = 1;
            if (p + 1 < n_rows) {
                const uint32_t ninfo = sInfo[s + 1];
                if ((ninfo & 127) == nd && w_eq(w_shr(sKey[s + 1], L), w_shr(x, L))) last = 0;
            }
            if (!(info & INFO_W_DOLLAR)) {
                const uint32_t wc = (uint32_t)w_low64(w_and(x, maskL));
                w = wc + 1;
                const KeyT grp = w_shr(x, 2 * L);
                const uint32_t gnd = nd > 1 ? nd : 1;
                for (int sq = (int)s - 1; sq >= 0 && pstart + sq >= 1; --sq) {
                    const uint32_t qi = sInfo[sq];
                    const uint32_t qnd = qi & 127;
                    const KeyT y = sKey[sq];
                    if ((qnd > 1 ? qnd : 1) != gnd || !w_eq(w_shr(y, 2 * L), grp)) break;
                    if (!(qi & INFO_W_DOLLAR) && (uint32_t)w_low64(w_and(y, maskL)) == wc) {
                        w += sigma;
                        break;
                    }
                }
                if (real_counts && nd == 0) {
                    const uint32_t c = sCnt[r];
                    cnt = c < maxw ? c : maxw;
                }
            }
            const uint32_t f = nd == K - 1 ? 0 : (uint32_t)w_low64(w_and(w_shr(x, L * (K - 1)), maskL)) + 1;
            uint32_t fprev = 0;
            if (p > 1) {
                const uint32_t pi = sInfo[s - 1] & 127;
                fprev = pi == K - 1 ? 0 : (uint32_t)w_low64(w_and(w_shr(sKey[s - 1], L * (K - 1)), maskL)) + 1;
            }
            for (uint32_t c = fprev + 1; c <= f; ++c) F[c] = p - 1;
        }
        const unsigned bal = __ballot_sync(0xffffffffu, last);
        if (lane == 0) last32[(p0 + r) >> 5] = bal;
        sW[r] = (uint8_t)w;
        if (weights_out) sCnt[r] = cnt;
    }
    __syncthreads();
    {   // W: 8 rows (one byte each, < 16) -> one u32 of nibbles
        uint64_t b = reinterpret_cast<const uint64_t *>(sW)[tid];
        b = (b | (b >> 4)) & 0x00ff00ff00ff00ffull;
        b = (b | (b >> 8)) & 0x0000ffff0000ffffull;
        b = (b | (b >> 16)) & 0x00000000ffffffffull;
        reinterpret_cast<uint32_t *>(W_out)[(p0 >> 3) + tid] = (uint32_t)b;
    }
    if (weights_out) {
        const uint32_t width = weights_width;
        const uint64_t rows_left = n_rows > p0 ? n_rows - p0 : 0;
        const uint32_t tile_rows = rows_left < (uint64_t)FE_TILE ? (uint32_t)rows_left : FE_TILE;
        const uint32_t nwords = (uint32_t)(((uint64_t)tile_rows * width + 63) / 64);
        for (uint32_t t = tid; t < nwords; t += FE_THREADS) {
            uint64_t word = 0;
            uint32_t bit0 = t * 64;
            uint32_t e = bit0 / width;
            int sh = (int)(e * width) - (int)bit0;
            while (sh < 64 && e < (uint32_t)FE_TILE) {
                uint64_t v = sCnt[e];
                word |= sh >= 0 ? v << sh : v >> (-sh);
                ++e;
                sh += width;
            }
            weights_out[p0 / 64 * width + t] = word;
        }
    }
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


size_t emit_merged_scratch_words(uint64_t n_rows) { return div_up(n_rows, (size_t)FE_TILE) + 1; }

template <typename KeyT>
void emit_merged(const KeyT *reals, const uint32_t *real_counts, uint64_t m, const KeyT *dummy_keys,
                 const uint32_t *dummy_info, const uint64_t *rank, uint64_t d, uint32_t row_off, Alpha al,
                 uint32_t weights_width, uint64_t *part, uint64_t *W, uint64_t *last, uint64_t *weights,
                 unsigned long long *F, cudaStream_t stream, unsigned *launches) {
    const uint64_t n_rows = m + d + row_off;
    const uint64_t ntiles = div_up(n_rows, (size_t)FE_TILE);
    init_F_kernel<<<1, 32, 0, stream>>>(F, al.sigma, n_rows);
    fe_partition_kernel<<<(unsigned)div_up(ntiles, (size_t)256), 256, 0, stream>>>(rank, d, row_off, ntiles, part);
    emit_merged_kernel<KeyT><<<(unsigned)ntiles, FE_THREADS, 0, stream>>>(
        reals, real_counts, m, dummy_keys, dummy_info, rank, d, part, row_off, al, weights_width, W, last,
        weights, F);
    MTG_CUDA(cudaGetLastError());
    *launches += 3;
}
template void emit_merged<uint64_t>(const uint64_t *, const uint32_t *, uint64_t, const uint64_t *,
                                    const uint32_t *, const uint64_t *, uint64_t, uint32_t, Alpha, uint32_t,
                                    uint64_t *, uint64_t *, uint64_t *, uint64_t *, unsigned long long *,
                                    cudaStream_t, unsigned *);
template void emit_merged<u128>(const u128 *, const uint32_t *, uint64_t, const u128 *, const uint32_t *,
                                const uint64_t *, uint64_t, uint32_t, Alpha, uint32_t, uint64_t *, uint64_t *,
                                uint64_t *, uint64_t *, unsigned long long *, cudaStream_t, unsigned *);

}  // namespace mtg
