// All-to-all over NVLink as ONE kernel of peer stores: every rank maps the receive buffers of
// its peers (symmetric memory, host side in multi_gpu.py) and this kernel pushes the rank's
// bucket-grouped slices straight into them -- coalesced 16-byte stores to peer addresses, all
// destinations served concurrently (each block owns one destination), no staging copy and no
// per-peer copy-engine launches. Replaces the bulk exchange step of the reference's disk-swap
// partitioning (boss_chunk_construct.cpp:462-525 routing at :514-517) between GPUs.
#include "../../include/mtg_gpu.h"
#include "mtg_common.cuh"

namespace mtg {
namespace {

constexpr int PP_MAX_DEST = 16;
constexpr int PP_THREADS = 512;

struct PushPlan {
    uint32_t n;
    unsigned char *dst[PP_MAX_DEST];       // peer-mapped destination (already offset)
    const unsigned char *src[PP_MAX_DEST];
    uint64_t bytes[PP_MAX_DEST];
};

template <typename VecT>
__global__ void __launch_bounds__(PP_THREADS)
peer_push_kernel(PushPlan plan, uint32_t blocks_per_dest) {
    const uint32_t d = blockIdx.x / blocks_per_dest, b = blockIdx.x % blocks_per_dest;
    if (d >= plan.n) return;
    const VecT *__restrict__ src = reinterpret_cast<const VecT *>(plan.src[d]);
    VecT *__restrict__ dst = reinterpret_cast<VecT *>(plan.dst[d]);
    const uint64_t n = plan.bytes[d] / sizeof(VecT);
    const uint64_t stride = (uint64_t)blocks_per_dest * PP_THREADS;
    uint64_t i = (uint64_t)b * PP_THREADS + threadIdx.x;
    // 4 independent loads in flight per thread
    for (; i + 3 * stride < n; i += 4 * stride) {
        VecT v0 = src[i], v1 = src[i + stride], v2 = src[i + 2 * stride], v3 = src[i + 3 * stride];
        dst[i] = v0; dst[i + stride] = v1; dst[i + 2 * stride] = v2; dst[i + 3 * stride] = v3;
    }
    for (; i < n; i += stride) dst[i] = src[i];
}

}  // namespace
}  // namespace mtg

extern "C" int mtg_peer_push(void *cuda_stream, uint32_t n_dest, void *const *dst_ptrs,
                             const void *const *src_ptrs, const uint64_t *nbytes) {
    using namespace mtg;
    int dev = 0, sms = 148;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaStream_t s = (cudaStream_t)cuda_stream;
    // any number of (destination, item) ranges: one launch per PP_MAX_DEST of them (8 GPUs x 4 items
    // of the counting canonical build = 32 ranges)
    uint32_t i = 0;
    while (i < n_dest) {
        PushPlan plan;
        plan.n = 0;
        uint64_t align = 0;
        for (; i < n_dest && plan.n < (uint32_t)PP_MAX_DEST; ++i) {
            if (!nbytes[i]) continue;
            plan.dst[plan.n] = (unsigned char *)dst_ptrs[i];
            plan.src[plan.n] = (const unsigned char *)src_ptrs[i];
            plan.bytes[plan.n] = nbytes[i];
            align |= (uint64_t)(uintptr_t)dst_ptrs[i] | (uint64_t)(uintptr_t)src_ptrs[i] | nbytes[i];
            ++plan.n;
        }
        if (plan.n == 0) continue;
        const uint32_t bpd = (uint32_t)((sms * 4 + plan.n - 1) / plan.n);
        const unsigned grid = bpd * plan.n;
        if ((align & 15) == 0) peer_push_kernel<uint4><<<grid, PP_THREADS, 0, s>>>(plan, bpd);
        else if ((align & 7) == 0) peer_push_kernel<uint2><<<grid, PP_THREADS, 0, s>>>(plan, bpd);
        else if ((align & 3) == 0) peer_push_kernel<uint32_t><<<grid, PP_THREADS, 0, s>>>(plan, bpd);
        else peer_push_kernel<unsigned char><<<grid, PP_THREADS, 0, s>>>(plan, bpd);
        if (cudaGetLastError() != cudaSuccess) return MTG_ERR_CUDA;
    }
    return MTG_OK;
}
