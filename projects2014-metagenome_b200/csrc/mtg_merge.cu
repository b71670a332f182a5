// Merge-path merge of two sorted k-mer arrays (with optional 32-bit count payload).
// Used by the canonical-mode completion: the reference appends the reverse complements and
// re-sorts the doubled array (add_reverse_complements, boss_chunk_construct.cpp:179-222,
// ips4o sort at :220); here only the appended half is sorted and then merged with the
// already sorted canonical half -- same resulting array, half the sort work.
#include "mtg_stages.cuh"

namespace mtg {

namespace {

constexpr int MG_THREADS = 256;
template <typename KeyT> struct MergeTile { static constexpr int IPT = sizeof(KeyT) == 8 ? 8 : 4; };

// merge-path split of diagonal `diag`: number of elements taken from A (A first on ties)
template <typename KeyT, typename IdxT>
__device__ __forceinline__ IdxT merge_path(const KeyT *A, IdxT na, const KeyT *B, IdxT nb, IdxT diag) {
    IdxT lo = diag > nb ? diag - nb : 0, hi = diag < na ? diag : na;
    while (lo < hi) {
        IdxT mid = (lo + hi) >> 1;
        if (!w_lt(B[diag - 1 - mid], A[mid])) lo = mid + 1; else hi = mid;   // A[mid] <= B[..]
    }
    return lo;
}

template <typename KeyT>
__global__ void merge_partition_kernel(const KeyT *__restrict__ A, uint64_t na, const KeyT *__restrict__ B,
                                       uint64_t nb, uint64_t tile, uint64_t ntiles,
                                       uint64_t *__restrict__ part) {
    uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t > ntiles) return;
    uint64_t diag = t * tile;
    if (diag > na + nb) diag = na + nb;
    part[t] = merge_path<KeyT, uint64_t>(A, na, B, nb, diag);
}

template <typename KeyT, bool HAS_VAL>
__global__ void __launch_bounds__(MG_THREADS)
merge_tiles_kernel(const KeyT *__restrict__ A, uint64_t na, const KeyT *__restrict__ B, uint64_t nb,
                   const uint32_t *__restrict__ Ac, const uint32_t *__restrict__ Bc,
                   const uint64_t *__restrict__ part, KeyT *__restrict__ out,
                   uint32_t *__restrict__ outc) {
    constexpr int MG_IPT = MergeTile<KeyT>::IPT;
    constexpr int TILE = MG_THREADS * MG_IPT;
    // Shared-memory indices are padded (PI: one slot every 4, PO: one every MG_IPT): thread t walks
    // the inputs around index IPT/2 * t and writes outputs IPT * t .. IPT * t + IPT - 1, and without
    // the padding those strides put 8 to 16 lanes of a warp on the same banks (the ncu capture of
    // the unpadded kernel: L1TEX 98 %, 64 % of the shared wavefronts were conflicts).
    auto PI = [](uint32_t i) { return i + (i >> 2); };
    auto PO = [](uint32_t i) { return i + i / MG_IPT; };
    __shared__ KeyT s_in[TILE + 2 + (TILE + 2) / 4 + 1];
    __shared__ KeyT s_out[TILE + TILE / MG_IPT];
    __shared__ uint32_t s_src[HAS_VAL ? TILE : 1];
    const uint64_t a0 = part[blockIdx.x], a1 = part[blockIdx.x + 1];
    const uint64_t d0 = (uint64_t)blockIdx.x * TILE;
    uint64_t d1 = d0 + TILE;
    if (d1 > na + nb) d1 = na + nb;
    const uint64_t b0 = d0 - a0, b1 = d1 - a1;
    const uint32_t la = (uint32_t)(a1 - a0), lb = (uint32_t)(b1 - b0), total = la + lb;
    for (uint32_t i = threadIdx.x; i < total; i += MG_THREADS)
        s_in[PI(i)] = i < la ? A[a0 + i] : B[b0 + (i - la)];
    __syncthreads();
    uint32_t diag = threadIdx.x * MG_IPT;
    if (diag > total) diag = total;
    uint32_t ia;
    {   // merge-path split of the thread's diagonal (A first on ties), on the padded layout
        uint32_t lo = diag > lb ? diag - lb : 0, hi = diag < la ? diag : la;
        while (lo < hi) {
            const uint32_t mid = (lo + hi) >> 1;
            if (!w_lt(s_in[PI(la + diag - 1 - mid)], s_in[PI(mid)])) lo = mid + 1; else hi = mid;
        }
        ia = lo;
    }
    uint32_t ib = diag - ia;
    // the two current heads live in registers: one shared-memory read per output
    KeyT ka = ia < la ? s_in[PI(ia)] : WordTraits<KeyT>::zero();
    KeyT kb = ib < lb ? s_in[PI(la + ib)] : WordTraits<KeyT>::zero();
#pragma unroll
    for (int i = 0; i < MG_IPT; ++i) {
        const uint32_t o = diag + i;
        if (o < total) {
            const bool takeA = ib >= lb || (ia < la && !w_lt(kb, ka));
            s_out[PO(o)] = takeA ? ka : kb;
            if (HAS_VAL) s_src[o] = takeA ? ia : (0x80000000u | ib);
            if (takeA) {
                ++ia;
                if (ia < la) ka = s_in[PI(ia)];
            } else {
                ++ib;
                if (ib < lb) kb = s_in[PI(la + ib)];
            }
        }
    }
    __syncthreads();
    for (uint32_t i = threadIdx.x; i < total; i += MG_THREADS) {
        out[d0 + i] = s_out[PO(i)];
        if (HAS_VAL) {
            uint32_t src = s_src[i];
            outc[d0 + i] = (src & 0x80000000u) ? Bc[b0 + (src & 0x7fffffffu)] : Ac[a0 + src];
        }
    }
}

}  // namespace

size_t merge_partition_words(uint64_t n_total) { return div_up(n_total, (size_t)MG_THREADS * 4) + 2; }

template <typename KeyT>
void merge_sorted(const KeyT *A, uint64_t na, const KeyT *B, uint64_t nb, const uint32_t *Ac,
                  const uint32_t *Bc, KeyT *out, uint32_t *outc, uint64_t *part, cudaStream_t stream,
                  unsigned *launches) {
    const uint64_t total = na + nb;
    if (total == 0) return;
    constexpr uint64_t TILE = MG_THREADS * MergeTile<KeyT>::IPT;
    const uint64_t ntiles = div_up(total, (size_t)TILE);
    merge_partition_kernel<KeyT><<<(unsigned)div_up(ntiles + 1, (size_t)128), 128, 0, stream>>>(
        A, na, B, nb, TILE, ntiles, part);
    if (outc)   // (Ac / Bc may be NULL when their array is empty)
        merge_tiles_kernel<KeyT, true><<<(unsigned)ntiles, MG_THREADS, 0, stream>>>(A, na, B, nb, Ac, Bc, part,
                                                                                   out, outc);
    else
        merge_tiles_kernel<KeyT, false><<<(unsigned)ntiles, MG_THREADS, 0, stream>>>(A, na, B, nb, nullptr,
                                                                                    nullptr, part, out, nullptr);
    MTG_CUDA(cudaGetLastError());
    *launches += 2;
}
template void merge_sorted<uint64_t>(const uint64_t *, uint64_t, const uint64_t *, uint64_t, const uint32_t *,
                                     const uint32_t *, uint64_t *, uint32_t *, uint64_t *, cudaStream_t,
                                     unsigned *);
template void merge_sorted<u128>(const u128 *, uint64_t, const u128 *, uint64_t, const uint32_t *,
                                 const uint32_t *, u128 *, uint32_t *, uint64_t *, cudaStream_t, unsigned *);

}  // namespace mtg
