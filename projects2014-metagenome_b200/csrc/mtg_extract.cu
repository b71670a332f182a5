// K1: reads -> packed (k+1)-mers in KMerBOSS layout, tight alphabet.
// Replaces KmerExtractorT<L>::sequence_to_kmers (kmer/kmer_extractor.cpp:472-507),
// ::sequence_to_kmers slide variants (:87-108, :166-196), KMerBOSS ctor/to_next/
// to_prev (kmer/kmer_boss.hpp:123-186) and utils::drag_and_mark_segments
// (common/algorithms.hpp:50-67) for the default (no --suffix) build.
//
// Layout: reads are packed back to back with one 0x00 separator after each read.
// A block stages TILE_CHARS (+halo) characters: every thread encodes 64 chars
// (4 x 128-bit loads) through the alphabet LUT into an L-bit little-endian bit
// stream in shared memory plus an "invalid character" bitmap. A window starting
// at char p is then just bits [L*p, L*(p+K)) of the stream: v = sum s_i << L(i-1).
// The KMerBOSS word is the rotation  ((v << L) & mask) | (v >> L(K-1))
// (kmer_boss.hpp:17-28) -- no rolling state, no redundant work per window.
// Canonical mode (DNA, complement = code ^ 3): rc(v) = reverse 2-bit groups of ~v.
#include "mtg_scan.cuh"
#include "mtg_stages.cuh"

namespace mtg {

namespace {

constexpr int EX_THREADS = 256;
constexpr int EX_GROUP = 64;                       // chars encoded per thread
constexpr int EX_TILE = EX_THREADS * EX_GROUP;     // windows per block
constexpr int EX_WARPS = EX_THREADS / 32;
constexpr int EX_WPW = EX_TILE / EX_WARPS;         // windows per warp

// extract nb (<=64) bits at bit offset `off` of a little-endian u32 stream
__device__ __forceinline__ uint64_t stream_bits64(const uint32_t *w, uint32_t off) {
    uint32_t q = off >> 5, s = off & 31;
    uint32_t w0 = w[q], w1 = w[q + 1], w2 = w[q + 2];
    uint32_t lo = __funnelshift_r(w0, w1, s);
    uint32_t hi = __funnelshift_r(w1, w2, s);
    return ((uint64_t)hi << 32) | lo;
}

template <typename KeyT> struct Win;
template <> struct Win<uint64_t> {
    __device__ static __forceinline__ uint64_t load(const uint32_t *w, uint32_t off, unsigned nbits) {
        return stream_bits64(w, off) & WordTraits<uint64_t>::low_mask(nbits);
    }
};
template <> struct Win<u128> {
    __device__ static __forceinline__ u128 load(const uint32_t *w, uint32_t off, unsigned nbits) {
        u128 r = make_u128(stream_bits64(w, off), stream_bits64(w, off + 64));
        return w_and(r, WordTraits<u128>::low_mask(nbits));
    }
};

// reverse the order of the 2-bit groups of a word
__device__ __forceinline__ uint64_t rev2(uint64_t x) {
    x = __brevll(x);
    return ((x >> 1) & 0x5555555555555555ull) | ((x & 0x5555555555555555ull) << 1);
}
__device__ __forceinline__ u128 rev2(u128 x) { return make_u128(rev2(x.hi), rev2(x.lo)); }

template <typename KeyT>
__device__ __forceinline__ KeyT boss_rotate(KeyT v, unsigned L, unsigned K) {
    // v = sum_{i=1..K} s_i << L(i-1)  ->  sum_{i<K} s_i << L*i | s_K
    KeyT hi = w_and(w_shl(v, L), WordTraits<KeyT>::low_mask(L * K));
    return w_or(hi, w_shr(v, L * (K - 1)));
}

template <typename KeyT, int L, int CANON, bool HAS_COUNT>   // CANON: 0 forward, 1 min(fwd, rc), 2 both strands
__global__ void __launch_bounds__(EX_THREADS)
extract_kernel(const uint8_t *__restrict__ packed, uint64_t nbytes, ExtractLut lut, uint32_t K,
               KeyT *__restrict__ out_keys, uint32_t *__restrict__ out_counts,
               unsigned long long *__restrict__ out_counter,
               const uint64_t *__restrict__ read_starts, const uint32_t *__restrict__ read_counts,
               uint64_t n_reads) {
    constexpr int GW = 2 * L;                      // u32 words per 64-char group
    extern __shared__ __align__(16) uint32_t ex_smem[];
    const uint32_t halo_groups = (K - 1 + EX_GROUP - 1) / EX_GROUP;
    const uint32_t ngroups = EX_THREADS + halo_groups;
    uint32_t *codes = ex_smem;                         // [(ngroups) * GW + 4]
    uint32_t *inval = codes + ngroups * GW + 4;        // [(ngroups) * 2 + 4]
    __shared__ uint8_t sh_lut[256];
    sh_lut[threadIdx.x] = lut.code[threadIdx.x];
    __syncthreads();

    const uint64_t tile_start = (uint64_t)blockIdx.x * EX_TILE;
    // ---- stage + encode ----
    for (uint32_t g = threadIdx.x; g < ngroups + 1; g += EX_THREADS) {
        if (g == ngroups) {   // zero the slack words read by the funnel shifts
            for (int i = 0; i < 4; ++i) { codes[ngroups * GW + i] = 0; inval[ngroups * 2 + i] = ~0u; }
            break;
        }
        uint64_t gpos = tile_start + (uint64_t)g * EX_GROUP;
        uint64_t acc = 0;
        int nb = 0, wi = 0;
        uint32_t bad[2] = {0, 0};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            uint4 raw = make_uint4(0, 0, 0, 0);
            uint64_t pos = gpos + q * 16;
            if (pos + 16 <= nbytes) {
                raw = *reinterpret_cast<const uint4 *>(packed + pos);
            } else if (pos < nbytes) {
                uint8_t tmp[16];
                for (int i = 0; i < 16; ++i) tmp[i] = pos + i < nbytes ? packed[pos + i] : 0;
                raw = *reinterpret_cast<uint4 *>(tmp);
            }
            uint32_t rw[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
            for (int i = 0; i < 16; ++i) {
                uint32_t c = (rw[i >> 2] >> ((i & 3) * 8)) & 0xff;
                uint32_t code = sh_lut[c];
                if (code == 0xff) { bad[q >> 1] |= 1u << ((q & 1) * 16 + i); code = 0; }
                acc |= (uint64_t)code << nb;
                nb += L;
                if (nb >= 32) {
                    codes[g * GW + wi++] = (uint32_t)acc;
                    acc >>= 32;
                    nb -= 32;
                }
            }
        }
        inval[g * 2] = bad[0];
        inval[g * 2 + 1] = bad[1];
    }
    __syncthreads();

    // ---- window validity bitmap: window p is valid iff no invalid char in [p, p+K) ----
    // wvalid[q] bit s <-> window 32q+s; computed word-wise by smearing the invalid-char bitmap
    // over K positions (K <= 97 always: 32 + K - 1 bits fit 128). No warp votes are needed
    // afterwards: the output slot of a window is prefix[q] + popc(wvalid[q] below my lane).
    uint32_t *wvalid = inval + ngroups * 2 + 4;          // [EX_TILE / 32]
    uint32_t *wprefix = wvalid + EX_TILE / 32;           // [EX_TILE / 32]
    __shared__ uint32_t scan_tmp[EX_THREADS / 32 + 1];
    __shared__ unsigned long long sh_base;
    uint32_t my_cnt = 0, w0 = 0, w1 = 0;
    {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t q = threadIdx.x * 2 + h;
            unsigned __int128 x = (unsigned __int128)inval[q] | ((unsigned __int128)inval[q + 1] << 32)
                                  | ((unsigned __int128)inval[q + 2] << 64) | ((unsigned __int128)inval[q + 3] << 96);
            unsigned __int128 r = x;
            uint32_t covered = 1;
            while (covered * 2 <= K) { r |= r >> covered; covered *= 2; }
            if (K > covered) r |= r >> (K - covered);
            uint32_t ok = ~(uint32_t)r;
            if (h == 0) w0 = ok; else w1 = ok;
            wvalid[q] = ok;
        }
        my_cnt = __popc(w0) + __popc(w1);
    }
    uint32_t excl = block_exclusive_scan<EX_THREADS>(my_cnt, scan_tmp);
    wprefix[threadIdx.x * 2] = excl;
    wprefix[threadIdx.x * 2 + 1] = excl + __popc(w0);
    const uint32_t block_total = scan_tmp[EX_THREADS / 32];
    if (threadIdx.x == 0)
        sh_base = block_total ? atomicAdd(out_counter, (unsigned long long)block_total * (CANON == 2 ? 2 : 1))
                              : 0ull;
    __syncthreads();
    if (block_total == 0) return;
    const unsigned long long base = sh_base;
    const unsigned lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t wbase = warp * EX_WPW;
    const unsigned lt = lanemask_lt();

    const unsigned KL = K * L;
#pragma unroll 4
    for (int it = 0; it < EX_WPW / 32; ++it) {
        const uint32_t q = (wbase >> 5) + it;
        const uint32_t word = wvalid[q];
        if ((word >> lane) & 1) {
            uint32_t p = wbase + it * 32 + lane;
            KeyT v = Win<KeyT>::load(codes, p * L, KL);
            KeyT key = boss_rotate(v, L, K);
            KeyT rkey = key;
            if (CANON) {
                // complement = ^3 on every 2-bit group, then reverse the groups
                KeyT r = w_shr(rev2(w_not(v)), WordTraits<KeyT>::BITS - KL);
                rkey = boss_rotate(r, L, K);
                if (CANON == 1 && w_lt(rkey, key)) key = rkey;   // forward <= reverse keeps forward
            }
            uint64_t dst = base + (uint64_t)(wprefix[q] + __popc(word & lt)) * (CANON == 2 ? 2 : 1);
            out_keys[dst] = key;
            if (CANON == 2) out_keys[dst + 1] = rkey;   // a palindrome is emitted twice (its count doubles)
            if (HAS_COUNT) {
                // read containing this window: last start <= global position
                uint64_t gp = tile_start + p;
                uint64_t lo = 0, hi = n_reads;   // invariant: starts[lo] <= gp < starts[hi]
                while (hi - lo > 1) {
                    uint64_t mid = (lo + hi) >> 1;
                    if (read_starts[mid] <= gp) lo = mid; else hi = mid;
                }
                out_counts[dst] = read_counts[lo];
                if (CANON == 2) out_counts[dst + 1] = read_counts[lo];
            }
        }
    }
}

template <typename KeyT, int L, int CANON>
void launch_extract(const ExtractArgs &a, KeyT *out_keys, cudaStream_t stream) {
    const uint32_t halo_groups = (a.K - 1 + EX_GROUP - 1) / EX_GROUP;
    const uint32_t ngroups = EX_THREADS + halo_groups;
    size_t smem = sizeof(uint32_t) * ((size_t)ngroups * 2 * L + 4 + (size_t)ngroups * 2 + 4 + 2 * (EX_TILE / 32));
    unsigned grid = (unsigned)div_up(a.nbytes, (size_t)EX_TILE);
    if (a.read_counts) {
        auto k = extract_kernel<KeyT, L, CANON, true>;
        k<<<grid, EX_THREADS, smem, stream>>>(a.packed, a.nbytes, *a.lut, a.K, out_keys, a.out_counts,
                                             a.out_counter, a.read_starts, a.read_counts, a.n_reads);
    } else {
        auto k = extract_kernel<KeyT, L, CANON, false>;
        k<<<grid, EX_THREADS, smem, stream>>>(a.packed, a.nbytes, *a.lut, a.K, out_keys, nullptr,
                                             a.out_counter, nullptr, nullptr, 0);
    }
    MTG_CUDA(cudaGetLastError());
}

}  // namespace

namespace {
// Raw concatenated reads + offsets -> packed layout (read r at raw offset o_r - o_0 moves to
// o_r - o_0 + r, followed by its 0x00 separator). One warp per read, 32 consecutive bytes per step.
__global__ void __launch_bounds__(256)
pack_reads_kernel(const uint8_t *__restrict__ raw, const uint64_t *__restrict__ offsets, uint64_t n_reads,
                  uint8_t *__restrict__ packed, uint64_t packed_bytes, uint64_t *__restrict__ starts) {
    const uint64_t warp = ((uint64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const unsigned lane = threadIdx.x & 31;
    const uint64_t nwarps = ((uint64_t)gridDim.x * blockDim.x) >> 5;
    const uint64_t o0 = offsets[0];
    for (uint64_t r = warp; r < n_reads; r += nwarps) {
        const uint64_t a = offsets[r] - o0, b = offsets[r + 1] - o0;
        uint8_t *dst = packed + a + r;
        for (uint64_t i = a + lane; i < b; i += 32) dst[i - a] = raw[i];
        if (lane == 0) {
            dst[b - a] = 0;
            if (starts) starts[r] = a + r;
        }
    }
    // zero the alignment tail
    if (blockIdx.x == 0 && threadIdx.x < 16) {
        const uint64_t end = offsets[n_reads] - o0 + n_reads;
        if (end + threadIdx.x < packed_bytes) packed[end + threadIdx.x] = 0;
    }
}
}  // namespace

void pack_reads_device(const uint8_t *raw, const uint64_t *offsets, uint64_t n_reads, uint8_t *packed,
                       uint64_t packed_bytes, uint64_t *starts, cudaStream_t stream) {
    if (n_reads == 0) return;
    const uint64_t want = div_up(n_reads * 32, (size_t)256);
    const unsigned grid = (unsigned)(want < 148 * 32 ? want : 148 * 32);
    pack_reads_kernel<<<grid, 256, 0, stream>>>(raw, offsets, n_reads, packed, packed_bytes, starts);
    MTG_CUDA(cudaGetLastError());
}

template <typename KeyT>
void extract_kmers(const ExtractArgs &a, KeyT *out_keys, cudaStream_t stream) {
    if (a.nbytes == 0) return;
    if (a.L == 2) {
        if (a.canonical && a.both) launch_extract<KeyT, 2, 2>(a, out_keys, stream);
        else if (a.canonical) launch_extract<KeyT, 2, 1>(a, out_keys, stream);
        else launch_extract<KeyT, 2, 0>(a, out_keys, stream);
    } else if (a.L == 5) {
        launch_extract<KeyT, 5, 0>(a, out_keys, stream);
    } else {
        launch_extract<KeyT, 3, 0>(a, out_keys, stream);
    }
}

template void extract_kmers<uint64_t>(const ExtractArgs &, uint64_t *, cudaStream_t);
template void extract_kmers<u128>(const ExtractArgs &, u128 *, cudaStream_t);

}  // namespace mtg
