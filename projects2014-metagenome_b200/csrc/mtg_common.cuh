// Common device/host helpers for the B200 (sm_100a) BOSS-construction kernels.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>

namespace mtg {

// ---------------------------------------------------------------------------
// 128-bit k-mer word (stands in for sdsl::uint128_t; K*L in (64,128]).
// ---------------------------------------------------------------------------
struct __align__(16) u128 {
    uint64_t lo, hi;
};

__host__ __device__ __forceinline__ u128 make_u128(uint64_t lo, uint64_t hi = 0) {
    u128 r; r.lo = lo; r.hi = hi; return r;
}

// Word traits / generic operations usable with uint64_t and u128.
__host__ __device__ __forceinline__ bool w_eq(uint64_t a, uint64_t b) { return a == b; }
__host__ __device__ __forceinline__ bool w_eq(u128 a, u128 b) { return a.lo == b.lo && a.hi == b.hi; }
__host__ __device__ __forceinline__ bool w_lt(uint64_t a, uint64_t b) { return a < b; }
__host__ __device__ __forceinline__ bool w_lt(u128 a, u128 b) {
    return a.hi < b.hi || (a.hi == b.hi && a.lo < b.lo);
}
__host__ __device__ __forceinline__ uint64_t w_shr(uint64_t a, unsigned n) { return n >= 64 ? 0 : a >> n; }
__host__ __device__ __forceinline__ u128 w_shr(u128 a, unsigned n) {
    if (n == 0) return a;
    if (n >= 128) return make_u128(0, 0);
    if (n >= 64) return make_u128(a.hi >> (n - 64), 0);
    return make_u128((a.lo >> n) | (a.hi << (64 - n)), a.hi >> n);
}
__host__ __device__ __forceinline__ uint64_t w_shl(uint64_t a, unsigned n) { return n >= 64 ? 0 : a << n; }
__host__ __device__ __forceinline__ u128 w_shl(u128 a, unsigned n) {
    if (n == 0) return a;
    if (n >= 128) return make_u128(0, 0);
    if (n >= 64) return make_u128(0, a.lo << (n - 64));
    return make_u128(a.lo << n, (a.hi << n) | (a.lo >> (64 - n)));
}
__host__ __device__ __forceinline__ uint64_t w_or(uint64_t a, uint64_t b) { return a | b; }
__host__ __device__ __forceinline__ u128 w_or(u128 a, u128 b) { return make_u128(a.lo | b.lo, a.hi | b.hi); }
__host__ __device__ __forceinline__ uint64_t w_and(uint64_t a, uint64_t b) { return a & b; }
__host__ __device__ __forceinline__ u128 w_and(u128 a, u128 b) { return make_u128(a.lo & b.lo, a.hi & b.hi); }
__host__ __device__ __forceinline__ uint64_t w_xor(uint64_t a, uint64_t b) { return a ^ b; }
__host__ __device__ __forceinline__ u128 w_xor(u128 a, u128 b) { return make_u128(a.lo ^ b.lo, a.hi ^ b.hi); }
__host__ __device__ __forceinline__ uint64_t w_not(uint64_t a) { return ~a; }
__host__ __device__ __forceinline__ u128 w_not(u128 a) { return make_u128(~a.lo, ~a.hi); }
__host__ __device__ __forceinline__ uint64_t w_low64(uint64_t a) { return a; }
__host__ __device__ __forceinline__ uint64_t w_low64(u128 a) { return a.lo; }

template <typename T> struct WordTraits;
template <> struct WordTraits<uint64_t> {
    static constexpr int BITS = 64;
    __host__ __device__ static __forceinline__ uint64_t from_u64(uint64_t x) { return x; }
    __host__ __device__ static __forceinline__ uint64_t zero() { return 0; }
    // n ones in the low bits, n in [0,64]
    __host__ __device__ static __forceinline__ uint64_t low_mask(unsigned n) {
        return n >= 64 ? ~0ull : ((1ull << n) - 1);
    }
};
template <> struct WordTraits<u128> {
    static constexpr int BITS = 128;
    __host__ __device__ static __forceinline__ u128 from_u64(uint64_t x) { return make_u128(x, 0); }
    __host__ __device__ static __forceinline__ u128 zero() { return make_u128(0, 0); }
    __host__ __device__ static __forceinline__ u128 low_mask(unsigned n) {
        if (n >= 128) return make_u128(~0ull, ~0ull);
        if (n >= 64) return make_u128(~0ull, n == 64 ? 0 : ((1ull << (n - 64)) - 1));
        return make_u128((1ull << n) - 1, 0);
    }
};

// bits [b, b+nb) of a word, nb <= 32
__host__ __device__ __forceinline__ uint32_t w_bits(uint64_t a, unsigned b, unsigned nb) {
    return (uint32_t)(w_shr(a, b)) & ((1u << nb) - 1);
}
__host__ __device__ __forceinline__ uint32_t w_bits(u128 a, unsigned b, unsigned nb) {
    return (uint32_t)(w_shr(a, b).lo) & ((1u << nb) - 1);
}

// ---------------------------------------------------------------------------
// Relaxed gpu-scope loads/stores for decoupled look-back status words.
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long ld_relaxed_u64(const unsigned long long *p) {
    unsigned long long v;
    asm volatile("ld.relaxed.gpu.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_u64(unsigned long long *p, unsigned long long v) {
    asm volatile("st.relaxed.gpu.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}

__device__ __forceinline__ unsigned lane_id() { return threadIdx.x & 31; }
__device__ __forceinline__ unsigned lanemask_lt() {
    unsigned m;
    asm("mov.u32 %0, %%lanemask_lt;" : "=r"(m));
    return m;
}

// ---------------------------------------------------------------------------
// TMA 1-D bulk copy global -> shared with mbarrier completion (cp.async.bulk, SASS UBLKCP):
// one thread issues the copy of a whole tile, no registers hold data in flight, so a few
// resident blocks per SM keep >100 KB of loads outstanding. src, dst and bytes must be
// multiples of 16.
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(unsigned long long *bar, uint32_t arrivals) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(arrivals) : "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void bulk_load_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                              unsigned long long *bar) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
// several copies on one barrier: mbar_expect(total bytes) once, then bulk_copy_g2s per piece
__device__ __forceinline__ void mbar_expect(unsigned long long *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bulk_copy_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                              unsigned long long *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(unsigned long long *bar, uint32_t phase) {
    uint32_t ok = 0;
    while (!ok) {
        asm volatile(
            "{\n.reg .pred p;\nmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\nselp.u32 %0, 1, 0, p;\n}"
            : "=r"(ok)
            : "r"(smem_u32(bar)), "r"(phase)
            : "memory");
    }
}

// ---------------------------------------------------------------------------
// Error handling (host)
// ---------------------------------------------------------------------------
struct CudaError {
    cudaError_t err;
    const char *file;
    int line;
};
#define MTG_CUDA(call)                                                   \
    do {                                                                 \
        cudaError_t e__ = (call);                                        \
        if (e__ != cudaSuccess) throw ::mtg::CudaError{e__, __FILE__, __LINE__}; \
    } while (0)

static inline size_t div_up(size_t a, size_t b) { return (a + b - 1) / b; }

}  // namespace mtg
