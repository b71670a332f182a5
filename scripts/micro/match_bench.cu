// Micro-benchmark: throughput of __match_any_sync vs the 8-ballot equivalent vs smem atomics.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ unsigned ballot_match8(unsigned d) {
    unsigned peers = 0xffffffffu;
#pragma unroll
    for (int b = 0; b < 8; ++b) {
        unsigned m = __ballot_sync(0xffffffffu, (d >> b) & 1);
        peers &= ((d >> b) & 1) ? m : ~m;
    }
    return peers;
}
template <int MODE>
__global__ void k(unsigned *out, int iters, unsigned seed) {
    __shared__ unsigned hist[8][256];
    unsigned x = threadIdx.x * 2654435761u + blockIdx.x * 40503u + seed;
    unsigned acc = 0;
    for (int i = threadIdx.x; i < 8 * 256; i += blockDim.x) (&hist[0][0])[i] = 0;
    __syncthreads();
    for (int i = 0; i < iters; ++i) {
        x = x * 1664525u + 1013904223u;
        unsigned d = x >> 24;
        if (MODE == 0) acc += __popc(__match_any_sync(0xffffffffu, d));
        else if (MODE == 1) acc += __popc(ballot_match8(d));
        else if (MODE == 2) atomicAdd(&hist[threadIdx.x >> 5][d], 1u);
        else if (MODE == 3) acc += atomicAdd(&hist[threadIdx.x >> 5][d], 1u);
        else acc += d;
    }
    if (MODE == 2) acc = hist[threadIdx.x >> 5][threadIdx.x & 255];
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int MODE> void run(const char *name) {
    unsigned *out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    int iters = 4096;
    k<MODE><<<148 * 8, 256>>>(out, iters, 1);
    cudaEventRecord(a);
    k<MODE><<<148 * 8, 256>>>(out, iters, 2);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double warp_instr_per_sm = 8.0 * 8 * iters;   // 8 blocks x 8 warps per SM
    printf("%-28s %.3f ms  -> %.1f cycles per warp-op per SM (at 1.965 GHz)\n", name, ms,
           ms * 1e-3 * 1.965e9 / warp_instr_per_sm);
    cudaFree(out);
}
int main() {
    run<4>("baseline (lcg only)");
    run<0>("match_any");
    run<1>("8 x ballot");
    run<2>("smem atomicAdd (no return)");
    run<3>("smem atomicAdd (return)");
    return 0;
}
