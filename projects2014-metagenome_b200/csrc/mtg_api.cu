// C ABI (include/mtg_gpu.h) and the pipeline driver of the B200 BOSS chunk build.
// Host-side counterpart of BOSSChunkConstructor / KmerCollector
// (boss_chunk_construct.cpp:948-1048, kmer/kmer_collector.cpp:129-234) and of
// construct_boss_chunk (boss_chunk_construct.cpp:248-356).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <deque>
#include <map>
#include <memory>
#include <functional>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "../../include/mtg_gpu.h"
#include "mtg_stages.cuh"

using namespace mtg;

namespace {

thread_local std::string g_create_error;
thread_local std::string g_oom_detail;   // what the arena could not get (appended to the OOM message)

// Device memory arena: a few large cudaMalloc slabs kept for the lifetime of the context,
// carved first-fit with coalescing. All work of a context is ordered on one stream, so a
// block can be reused as soon as it is released. (The CUDA stream-ordered pool showed
// 100+ ms stalls whenever it had to grow or re-map multi-GB blocks.)
struct Arena {
    struct Slab { char *base; size_t size; };
    std::vector<Slab> slabs;
    std::map<char *, size_t> free_blocks;   // start -> size
    std::map<char *, size_t> used;          // start -> size
    std::mutex mu;
    size_t total = 0;
    static constexpr size_t ALIGN = 512;

    void *alloc(size_t n) {
        n = (std::max<size_t>(n, 1) + ALIGN - 1) & ~(ALIGN - 1);
        std::lock_guard<std::mutex> lk(mu);
        for (int attempt = 0; attempt < 2; ++attempt) {
            // best fit
            auto best = free_blocks.end();
            for (auto it = free_blocks.begin(); it != free_blocks.end(); ++it)
                if (it->second >= n && (best == free_blocks.end() || it->second < best->second)) best = it;
            if (best != free_blocks.end()) {
                char *p = best->first;
                size_t sz = best->second;
                free_blocks.erase(best);
                if (sz > n) free_blocks[p + n] = sz - n;
                used[p] = n;
                return p;
            }
            // grow: one new slab (at least 256 MB, or the request)
            size_t want = std::max<size_t>(n, (size_t)256 << 20);
            void *p = nullptr;
            cudaError_t e = cudaMalloc(&p, want);
            if (e != cudaSuccess && want > n) { cudaGetLastError(); want = n; e = cudaMalloc(&p, want); }
            if (e != cudaSuccess) {
                // slabs are never split across requests of other sizes: give the completely free ones
                // back to the driver (cudaFree synchronises the device) and ask again
                cudaGetLastError();
                if (trim()) e = cudaMalloc(&p, want);
            }
            if (e != cudaSuccess) {
                cudaGetLastError();
                size_t free_b = 0, total_b = 0, largest = 0, free_sum = 0;
                cudaMemGetInfo(&free_b, &total_b);
                for (auto &fb : free_blocks) { largest = std::max(largest, fb.second); free_sum += fb.second; }
                char msg[256];
                snprintf(msg, sizeof(msg), " [request %.2f GB; arena %.2f GB in %zu slabs, %.2f GB of it free (largest block %.2f GB); "
                         "device free %.2f of %.2f GB]", n / 1e9, total / 1e9, slabs.size(), free_sum / 1e9, largest / 1e9,
                         free_b / 1e9, total_b / 1e9);
                g_oom_detail = msg;
                throw CudaError{cudaErrorMemoryAllocation, __FILE__, __LINE__};
            }
            slabs.push_back(Slab{(char *)p, want});
            total += want;
            free_blocks[(char *)p] = want;
        }
        throw CudaError{cudaErrorMemoryAllocation, __FILE__, __LINE__};
    }
    // caller holds mu: cudaFree every slab that is one free block; true if something was released
    bool trim() {
        bool any = false;
        for (size_t i = 0; i < slabs.size();) {
            auto it = free_blocks.find(slabs[i].base);
            if (it != free_blocks.end() && it->second == slabs[i].size) {
                free_blocks.erase(it);
                cudaFree(slabs[i].base);
                total -= slabs[i].size;
                slabs.erase(slabs.begin() + i);
                any = true;
            } else {
                ++i;
            }
        }
        return any;
    }
    void free(void *ptr) {
        if (!ptr) return;
        std::lock_guard<std::mutex> lk(mu);
        char *p = (char *)ptr;
        auto u = used.find(p);
        if (u == used.end()) return;
        size_t n = u->second;
        used.erase(u);
        auto it = free_blocks.emplace(p, n).first;
        // coalesce with the next / previous block when they are adjacent in the same slab
        auto same_slab = [&](char *a, char *b) {
            for (auto &sl : slabs)
                if (a >= sl.base && a < sl.base + sl.size) return b >= sl.base && b < sl.base + sl.size;
            return false;
        };
        auto nx = std::next(it);
        if (nx != free_blocks.end() && it->first + it->second == nx->first && same_slab(it->first, nx->first)) {
            it->second += nx->second;
            free_blocks.erase(nx);
        }
        if (it != free_blocks.begin()) {
            auto pv = std::prev(it);
            if (pv->first + pv->second == it->first && same_slab(pv->first, it->first)) {
                pv->second += it->second;
                free_blocks.erase(it);
            }
        }
    }
    ~Arena() {
        for (auto &sl : slabs) cudaFree(sl.base);
    }
};

thread_local std::shared_ptr<Arena> t_arena;   // arena of the context the thread works for

struct DevBuf {   // arena-backed device allocation (stream-ordered by construction)
    void *p = nullptr;
    size_t bytes = 0;
    std::shared_ptr<Arena> arena;
    DevBuf() {}
    DevBuf(size_t n, cudaStream_t st) { alloc(n, st); }
    DevBuf(const DevBuf &) = delete;
    DevBuf &operator=(const DevBuf &) = delete;
    DevBuf(DevBuf &&o) noexcept { p = o.p; bytes = o.bytes; arena = std::move(o.arena); o.p = nullptr; }
    DevBuf &operator=(DevBuf &&o) noexcept {
        if (this != &o) { release(); p = o.p; bytes = o.bytes; arena = std::move(o.arena); o.p = nullptr; }
        return *this;
    }
    void alloc(size_t n, cudaStream_t) {
        release();
        arena = t_arena;
        bytes = n ? n : 16;
        p = arena->alloc(bytes);
    }
    void release() {
        if (p && arena) arena->free(p);
        p = nullptr;
    }
    ~DevBuf() { release(); }
    template <typename T> T *as() const { return reinterpret_cast<T *>(p); }
};

// Pinned staging buffer of mtg_push_reads. FILLING: a pushing thread owns it (it fills it outside the
// context mutex); IN_FLIGHT: its H2D copies are enqueued and `ev` was recorded behind them FOR THIS USE --
// only then does the event say anything about the buffer (a never-recorded or previously completed
// event reads "done"). The pool is a deque: pointers into it stay valid while other threads add buffers.
struct PinnedBuf {
    enum State { FREE, FILLING, IN_FLIGHT };
    void *p = nullptr;
    size_t bytes = 0;
    cudaEvent_t ev = nullptr;
    State state = FREE;
};

struct ReadSegment {
    const uint8_t *d_packed = nullptr;
    size_t nbytes = 0;
    uint64_t n_windows_max = 0;
    DevBuf own_packed, starts, counts;
    uint64_t n_reads = 0;
    bool has_starts = false, has_counts = false;
};

struct ChunkImpl {
    void *pinned = nullptr;
    size_t pinned_bytes = 0;
    DevBuf dev;
};

// Small process-wide cache of pinned result buffers (pinning hundreds of MB per build
// costs more than the build itself).
struct PinnedCache {
    std::mutex mu;
    std::vector<std::pair<void *, size_t>> free_list;
    void *acquire(size_t bytes, size_t *got) {
        {
            std::lock_guard<std::mutex> lk(mu);
            for (size_t i = 0; i < free_list.size(); ++i)
                if (free_list[i].second >= bytes) {
                    void *p = free_list[i].first;
                    *got = free_list[i].second;
                    free_list.erase(free_list.begin() + i);
                    return p;
                }
        }
        void *p = nullptr;
        if (cudaMallocHost(&p, bytes) != cudaSuccess) { cudaGetLastError(); return nullptr; }
        *got = bytes;
        return p;
    }
    void release(void *p, size_t bytes) {
        std::lock_guard<std::mutex> lk(mu);
        if (free_list.size() >= 2) {   // keep the two largest
            size_t mi = 0;
            for (size_t i = 1; i < free_list.size(); ++i)
                if (free_list[i].second < free_list[mi].second) mi = i;
            if (free_list[mi].second >= bytes) { cudaFreeHost(p); return; }
            cudaFreeHost(free_list[mi].first);
            free_list.erase(free_list.begin() + mi);
        }
        free_list.emplace_back(p, bytes);
    }
};
PinnedCache g_pinned_cache;

struct RealSet {   // sorted distinct real k-mers (both strands in canonical mode)
    DevBuf keys, counts;
    uint64_t m = 0, n_input = 0;
};

}  // namespace

struct mtg_ctx {
    mtg_params prm;
    Alpha al;
    uint32_t L2;
    ExtractLut lut;
    int device = 0;
    cudaStream_t stream = nullptr, own_stream = nullptr;
    std::mutex mu;
    std::vector<ReadSegment> segs;
    std::deque<PinnedBuf> pinned;
    std::shared_ptr<Arena> arena = std::make_shared<Arena>();
    bool any_explicit_counts = false;
    uint32_t cmax = 0;   // container count type max (boss_chunk_construct.cpp:1147-1160)
    std::vector<uint8_t> suffix;   // --suffix filter in BOSS codes (0 = '$'); empty = default build
    std::string error;
    unsigned launches = 0;
    // Super-batching (SortedSet::shrink_data, common/sorted_sets/sorted_set.hpp:69-82: "buffer full ->
    // sort + dedupe what is there, go on appending"): once the pushed reads hold batch_limit k-mer
    // windows they are extracted, sorted and folded into a sorted run of distinct k-mers (+ counts) and
    // the reads are dropped; mtg_build_chunk merges the runs (saturating count sums).
    std::vector<RealSet> runs;
    uint64_t pending_windows = 0, batch_limit = 0, n_batches = 0;
    // Pipelined host pushes: the reads of a segment go up on copy_stream into one of two staging
    // buffers while the device folds the previous batch (extract + sort + fold of >= push_batch k-mer
    // windows) on `stream`; the batch of a segment is cut when the NEXT segment's copy is in flight
    // (or by mtg_build_chunk).
    cudaStream_t copy_stream = nullptr;
    cudaEvent_t copy_done = nullptr;
    struct Stage {
        void *d_raw = nullptr, *d_off = nullptr;
        size_t raw_cap = 0, off_cap = 0;
        cudaEvent_t consumed = nullptr;   // the pack kernel that read this staging buffer is done
    } stage[2];
    int stage_idx = 0;
    bool flush_pending = false;
    uint64_t push_batch = 0;
    // range-partitioned (multi-GPU) build state, see mtg_mg_* below
    struct Mg {
        DevBuf keys, targets, tset, sinks, src, rkeys, rcounts;
        RealSet reals;
        uint64_t nt = 0, n_sink = 0, n_src = 0, n_input = 0, n_routed = 0, n_routed_targets = 0;
        RealSet acc;        // mtg_mg_accumulate: the folded k-mers of the earlier rounds of this range
        bool has_acc = false;
    } mg;
};

namespace {

int fail(mtg_ctx *ctx, int code, const std::string &msg) {
    if (ctx) ctx->error = msg; else g_create_error = msg;
    return code;
}
int fail_cuda(mtg_ctx *ctx, const CudaError &e) {
    std::string msg = std::string("CUDA error: ") + cudaGetErrorString(e.err) + " at " + e.file + ":"
                      + std::to_string(e.line);
    if (e.err == cudaErrorMemoryAllocation) { msg += g_oom_detail; g_oom_detail.clear(); }
    cudaGetLastError();
    return fail(ctx, e.err == cudaErrorMemoryAllocation ? MTG_ERR_OOM : MTG_ERR_CUDA, msg);
}

// alphabets (kmer/alphabets.hpp:27-149), rebuilt from their definitions
void build_lut(uint32_t alphabet, ExtractLut *lut) {
    if (alphabet == MTG_ALPHABET_DNA) {
        for (int i = 0; i < 256; ++i) lut->code[i] = 0xff;
        const char *up = "ACGT", *lo = "acgt";
        for (int c = 0; c < 4; ++c) { lut->code[(int)up[c]] = c; lut->code[(int)lo[c]] = c; }
        lut->code[(int)'U'] = lut->code[(int)'u'] = 3;
    } else {
        // every character is a letter of the protein alphabet ('X' = 25 for unknowns);
        // chars >= 0x80 map like table[0] (kmer_extractor.cpp:30-34)
        for (int i = 0; i < 256; ++i) lut->code[i] = 25;
        const char *alph = "ABCDEFGHIJKLMNOPQRSTUVWYZX";
        for (int c = 0; c < 26; ++c) {
            lut->code[(int)alph[c]] = c;
            lut->code[(int)alph[c] - 'A' + 'a'] = c;
        }
    }
    lut->code[0] = 0xff;   // read separator of the packed layout
}

// caller holds ctx->mu
PinnedBuf *get_pinned(mtg_ctx *ctx, size_t bytes) {
    for (auto &b : ctx->pinned)
        if (b.state == PinnedBuf::IN_FLIGHT && cudaEventQuery(b.ev) == cudaSuccess) b.state = PinnedBuf::FREE;
    for (auto &b : ctx->pinned)
        if (b.state == PinnedBuf::FREE && b.bytes >= bytes) { b.state = PinnedBuf::FILLING; return &b; }
    for (auto &b : ctx->pinned)
        if (b.state == PinnedBuf::FREE) {   // grow an idle one
            cudaFreeHost(b.p);
            b.p = nullptr;
            b.bytes = 0;
            MTG_CUDA(cudaMallocHost(&b.p, bytes));
            b.bytes = bytes;
            b.state = PinnedBuf::FILLING;
            return &b;
        }
    PinnedBuf nb;
    MTG_CUDA(cudaMallocHost(&nb.p, bytes));
    nb.bytes = bytes;
    MTG_CUDA(cudaEventCreateWithFlags(&nb.ev, cudaEventDisableTiming));
    nb.state = PinnedBuf::FILLING;
    ctx->pinned.push_back(nb);
    return &ctx->pinned.back();
}

void pack_range(const char *bases, const uint64_t *offsets, size_t r0, size_t r1, char *dst) {
    for (size_t r = r0; r < r1; ++r) {
        size_t len = offsets[r + 1] - offsets[r];
        memcpy(dst, bases + offsets[r], len);
        dst[len] = 0;
        dst += len + 1;
    }
}

void pack_parallel(const char *bases, const uint64_t *offsets, size_t r0, size_t r1, char *dst) {
    size_t bytes = offsets[r1] - offsets[r0] + (r1 - r0);
    unsigned nt = (unsigned)std::min<size_t>(std::max<size_t>(bytes >> 22, 1),
                                             std::min(16u, std::max(1u, std::thread::hardware_concurrency())));
    if (nt <= 1) { pack_range(bases, offsets, r0, r1, dst); return; }
    std::vector<std::thread> th;
    size_t per = (r1 - r0 + nt - 1) / nt;
    for (unsigned t = 0; t < nt; ++t) {
        size_t a = r0 + t * per, b = std::min(r1, a + per);
        if (a >= b) break;
        char *d = dst + (offsets[a] - offsets[r0]) + (a - r0);
        th.emplace_back(pack_range, bases, offsets, a, b, d);
    }
    for (auto &t : th) t.join();
}

// number of K-windows of reads [r0, r1): sum of max(0, len - K + 1), multi-threaded for big pushes
uint64_t count_windows(const uint64_t *offsets, size_t r0, size_t r1, uint32_t K) {
    auto part = [&](size_t a, size_t b) {
        uint64_t n = 0;
        for (size_t r = a; r < b; ++r) {
            const uint64_t len = offsets[r + 1] - offsets[r];
            if (len >= K) n += len - K + 1;
        }
        return n;
    };
    const size_t nr = r1 - r0;
    unsigned nt = (unsigned)std::min<size_t>(nr >> 18, std::min(16u, std::max(1u, std::thread::hardware_concurrency())));
    if (nt <= 1) return part(r0, r1);
    std::vector<uint64_t> sums(nt, 0);
    std::vector<std::thread> th;
    const size_t per = (nr + nt - 1) / nt;
    for (unsigned t = 0; t < nt; ++t) {
        const size_t a = r0 + t * per, b = std::min(r1, a + per);
        if (a >= b) break;
        th.emplace_back([&, t, a, b]() { sums[t] = part(a, b); });
    }
    for (auto &x : th) x.join();
    uint64_t n = 0;
    for (uint64_t v : sums) n += v;
    return n;
}

// multi-threaded memcpy (staging of pageable caller memory)
void copy_parallel(char *dst, const char *src, size_t bytes) {
    unsigned nt = (unsigned)std::min<size_t>(std::max<size_t>(bytes >> 22, 1),
                                             std::min(16u, std::max(1u, std::thread::hardware_concurrency())));
    if (nt <= 1) { memcpy(dst, src, bytes); return; }
    std::vector<std::thread> th;
    size_t per = ((bytes + nt - 1) / nt + 63) & ~(size_t)63;
    for (unsigned t = 0; t < nt; ++t) {
        size_t a = (size_t)t * per, b = std::min(bytes, a + per);
        if (a >= b) break;
        th.emplace_back([=]() { memcpy(dst + a, src + a, b - a); });
    }
    for (auto &t : th) t.join();
}

}  // namespace

extern "C" {

const char *mtg_last_error(const mtg_ctx *ctx) {
    return ctx ? ctx->error.c_str() : g_create_error.c_str();
}

int mtg_create(const mtg_params *p, mtg_ctx **out) {
    if (!p || !out) return fail(nullptr, MTG_ERR_BAD_ARG, "null argument");
    *out = nullptr;
    if (p->alphabet != MTG_ALPHABET_DNA && p->alphabet != MTG_ALPHABET_PROTEIN)
        return fail(nullptr, MTG_ERR_BAD_ARG, "unknown alphabet");
    const uint32_t L1 = p->alphabet == MTG_ALPHABET_DNA ? 2 : 5;
    const uint32_t L2 = p->alphabet == MTG_ALPHABET_DNA ? 3 : 5;
    const uint32_t sigma = p->alphabet == MTG_ALPHABET_DNA ? 5 : 27;
    // boss_chunk_construct.cpp:1057-1062
    if (p->k < 1 || p->k > 256 / L2 - 1)
        return fail(nullptr, MTG_ERR_BAD_ARG, "For succinct graph, k must be between 2 and "
                                               + std::to_string(256 / L2));
    if (p->bits_per_count > 32)   // :1157-1159
        return fail(nullptr, MTG_ERR_BAD_ARG, "Error: trying to allocate too many bits per k-mer count");
    const size_t suffix_len = p->filter_suffix ? strlen(p->filter_suffix) : 0;
    if (p->both_strands && p->alphabet != MTG_ALPHABET_DNA)
        return fail(nullptr, MTG_ERR_BAD_ARG, "canonical mode requires a DNA alphabet");
    const uint32_t K = p->k + 1;
    if (K * L1 + (suffix_len ? 8 : 0) > 128)
        return fail(nullptr, MTG_ERR_UNSUPPORTED, "k-mers wider than 128 bits are not supported yet");
    if (suffix_len >= K || suffix_len > MTG_SUFFIX_MAX)   // kmer_extractor.cpp:326
        return fail(nullptr, MTG_ERR_BAD_ARG, "the suffix must be shorter than k + 1 (it does not include the last character)");
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) {
        cudaGetLastError();
        return fail(nullptr, MTG_ERR_CUDA, "no CUDA device available (there is no CPU fallback)");
    }
    if (p->device < 0 || p->device >= ndev) return fail(nullptr, MTG_ERR_BAD_ARG, "bad device ordinal");
    mtg_ctx *ctx = new mtg_ctx();
    try {
        ctx->prm = *p;
        ctx->prm.filter_suffix = nullptr;
        ctx->device = p->device;
        ctx->al = Alpha{K, L1, sigma, 0};
        uint32_t hi = 0;
        for (uint32_t v = 2 * sigma - 1; v > 1; v >>= 1) ++hi;   // sdsl::bits::hi
        ctx->al.W_width = hi + 1;                                // boss_chunk.cpp:388-390
        ctx->L2 = L2;
        ctx->cmax = p->bits_per_count == 0 ? 0xffffffffu
                    : p->bits_per_count <= 8 ? 0xffu : p->bits_per_count <= 16 ? 0xffffu : 0xffffffffu;
        build_lut(p->alphabet, &ctx->lut);
        for (size_t i = 0; i < suffix_len; ++i) {   // encode_filter_suffix_boss (boss_chunk_construct.cpp:934-945)
            const unsigned char ch = (unsigned char)p->filter_suffix[i];
            uint8_t code = ch == '$' ? 0xfe : ctx->lut.code[ch];
            if (code == 0xff) {
                delete ctx;
                return fail(nullptr, MTG_ERR_BAD_ARG, "the suffix holds a character outside the alphabet");
            }
            ctx->suffix.push_back(code == 0xfe ? 0 : code + 1);
        }
        MTG_CUDA(cudaSetDevice(ctx->device));
        cudaDeviceProp prop;
        MTG_CUDA(cudaGetDeviceProperties(&prop, ctx->device));
        if (prop.major < 10) {
            delete ctx;
            return fail(nullptr, MTG_ERR_CUDA, "device is not sm_100 (kernels are built for sm_100a only)");
        }
        MTG_CUDA(cudaStreamCreateWithFlags(&ctx->own_stream, cudaStreamNonBlocking));
        ctx->stream = ctx->own_stream;
        // super-batch size: the sort ping-pongs two key arrays, which may take ~35 % of the device
        // memory; the rest is for the reads, the folded runs and the later stages. MTG_BATCH_KMERS
        // overrides (the tests force the batched path on small inputs with it)
        // per k-mer of a batch: the two key arrays of the sort, the fold's output and -- in counting mode
        // -- the run starts (8 B) and counts; a quarter of the device memory for all of that
        const uint64_t kb = ctx->al.K * ctx->al.L <= 64 ? 8 : 16;
        const double per_kmer = 3.0 * (double)kb + (p->bits_per_count ? 16.0 : 0.0);
        ctx->batch_limit = (uint64_t)(0.25 * (double)prop.totalGlobalMem / per_kmer);
        if (const char *e = getenv("MTG_BATCH_KMERS")) ctx->batch_limit = strtoull(e, nullptr, 10);
        // host pushes are folded in batches of ~280 M k-mer windows (a quarter of BASELINE cfg #2) so that
        // the sort of one batch runs under the H2D copy of the next
        ctx->push_batch = 280000000ull;
        if (const char *e = getenv("MTG_PUSH_BATCH_KMERS")) ctx->push_batch = strtoull(e, nullptr, 10);
        if (ctx->batch_limit && ctx->push_batch > ctx->batch_limit) ctx->push_batch = ctx->batch_limit;
        if (p->reserved & MTG_FLAG_NO_BATCHING) ctx->batch_limit = ctx->push_batch = 0;
        MTG_CUDA(cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        MTG_CUDA(cudaEventCreateWithFlags(&ctx->copy_done, cudaEventDisableTiming));
        for (auto &st : ctx->stage) MTG_CUDA(cudaEventCreateWithFlags(&st.consumed, cudaEventDisableTiming));
    } catch (const CudaError &e) {
        int rc = fail_cuda(nullptr, e);
        delete ctx;
        return rc;
    }
    *out = ctx;
    return MTG_OK;
}

void mtg_destroy(mtg_ctx *ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    ctx->segs.clear();
    for (auto &b : ctx->pinned) {
        if (b.p) cudaFreeHost(b.p);
        if (b.ev) cudaEventDestroy(b.ev);
    }
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    if (ctx->copy_done) cudaEventDestroy(ctx->copy_done);
    for (auto &st : ctx->stage) {
        if (st.d_raw) cudaFree(st.d_raw);
        if (st.d_off) cudaFree(st.d_off);
        if (st.consumed) cudaEventDestroy(st.consumed);
    }
    ctx->runs.clear();
    if (ctx->own_stream) cudaStreamDestroy(ctx->own_stream);
    delete ctx;
}

size_t mtg_packed_size(const uint64_t *offsets, size_t n_reads) {
    size_t b = (size_t)(offsets[n_reads] - offsets[0]) + n_reads;
    return (b + 15) & ~(size_t)15;
}

void mtg_pack_reads(const char *bases, const uint64_t *offsets, size_t n_reads, char *dst) {
    size_t raw = (size_t)(offsets[n_reads] - offsets[0]) + n_reads;
    pack_parallel(bases, offsets, 0, n_reads, dst);
    memset(dst + raw, 0, mtg_packed_size(offsets, n_reads) - raw);
}

}  // extern "C"
namespace { void maybe_flush_batch(mtg_ctx *ctx, uint64_t new_windows, bool force = false); }
extern "C" {

int mtg_push_reads(mtg_ctx *ctx, const char *bases, const uint64_t *offsets, const uint64_t *counts,
                   size_t n_reads) {
    if (!ctx || (!bases && n_reads) || !offsets) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
    if (n_reads == 0) return MTG_OK;
    const uint32_t K = ctx->al.K;
    try {
        MTG_CUDA(cudaSetDevice(ctx->device));
        t_arena = ctx->arena;
        // The reads go up raw (as the caller holds them: concatenated bases + offsets) and are
        // brought into the packed layout by a kernel. Pinned caller memory is copied from directly
        // (the call returns after the copies: the caller's buffers are free again); pageable memory
        // is staged through pinned buffers in <= 64 MB pieces so that staging and H2D overlap.
        cudaPointerAttributes attr;
        const bool pinned_src = cudaPointerGetAttributes(&attr, bases) == cudaSuccess && attr.type == cudaMemoryTypeHost;
        cudaGetLastError();
        // segment size: pinned caller memory goes up in pieces of about one fold batch (~400 MB), pageable
        // memory through 64 MB pinned staging pieces
        const size_t CHUNK = pinned_src ? (size_t)400 << 20 : (size_t)64 << 20;
        size_t r0 = 0;
        while (r0 < n_reads) {
            // segment end: first r1 with packed size (raw bytes + one separator per read) >= CHUNK
            size_t lo = r0 + 1, hi = n_reads;
            while (lo < hi) {
                size_t mid = (lo + hi) >> 1;
                if ((offsets[mid] - offsets[r0]) + (mid - r0) >= CHUNK) hi = mid; else lo = mid + 1;
            }
            const size_t r1 = lo;
            const size_t bytes = (offsets[r1] - offsets[r0]) + (r1 - r0);
            const uint64_t nwin = count_windows(offsets, r0, r1, K);
            const size_t nr = r1 - r0;
            const size_t raw_bytes = offsets[r1] - offsets[r0];
            const size_t padded = (bytes + 15) & ~(size_t)15;
            const bool want_starts = ctx->prm.bits_per_count > 0;
            const size_t off_bytes = (nr + 1) * 8, cnt_bytes = counts ? nr * 4 : 0;
            const size_t stage_bytes = (pinned_src ? 0 : ((raw_bytes + 15) & ~(size_t)15)) + off_bytes + cnt_bytes + 16;
            PinnedBuf *pb;
            ReadSegment seg;
            {
                std::lock_guard<std::mutex> lk(ctx->mu);
                pb = get_pinned(ctx, stage_bytes);
            }
            char *h = (char *)pb->p;
            char *h_raw = h;
            if (!pinned_src) {
                copy_parallel(h, bases + offsets[r0], raw_bytes);
                h += (raw_bytes + 15) & ~(size_t)15;
            }
            uint64_t *h_off = (uint64_t *)h;
            copy_parallel((char *)h_off, (const char *)(offsets + r0), off_bytes);
            uint32_t *h_counts = (uint32_t *)(h + off_bytes);
            if (counts)
                for (size_t r = r0; r < r1; ++r)   // kmer_collector.cpp:92
                    h_counts[r - r0] = (uint32_t)std::min<uint64_t>(counts[r], ctx->cmax);
            {
                std::lock_guard<std::mutex> lk(ctx->mu);
                cudaStream_t s = ctx->stream, cs = ctx->copy_stream;
                mtg_ctx::Stage &st = ctx->stage[ctx->stage_idx];
                ctx->stage_idx ^= 1;
                if (st.raw_cap < raw_bytes || st.off_cap < off_bytes) {   // (re)size this staging buffer
                    MTG_CUDA(cudaEventSynchronize(st.consumed));
                    if (st.raw_cap < raw_bytes) {
                        if (st.d_raw) cudaFree(st.d_raw);
                        st.d_raw = nullptr;
                        st.raw_cap = 0;
                        const size_t cap = std::max<size_t>(raw_bytes + (raw_bytes >> 3), 1 << 20);
                        MTG_CUDA(cudaMalloc(&st.d_raw, cap));
                        st.raw_cap = cap;
                    }
                    if (st.off_cap < off_bytes) {
                        if (st.d_off) cudaFree(st.d_off);
                        st.d_off = nullptr;
                        st.off_cap = 0;
                        const size_t cap = std::max<size_t>(off_bytes + (off_bytes >> 3), 1 << 16);
                        MTG_CUDA(cudaMalloc(&st.d_off, cap));
                        st.off_cap = cap;
                    }
                }
                // the copies go up on their own stream, behind the pack kernel that last read this buffer
                MTG_CUDA(cudaStreamWaitEvent(cs, st.consumed, 0));
                MTG_CUDA(cudaMemcpyAsync(st.d_raw, pinned_src ? bases + offsets[r0] : h_raw, raw_bytes,
                                         cudaMemcpyHostToDevice, cs));
                MTG_CUDA(cudaMemcpyAsync(st.d_off, h_off, off_bytes, cudaMemcpyHostToDevice, cs));
                MTG_CUDA(cudaEventRecord(ctx->copy_done, cs));
                MTG_CUDA(cudaEventRecord(pb->ev, cs));
                pb->state = PinnedBuf::IN_FLIGHT;
                // ... and while they are in flight the device folds the batch that is due
                if (ctx->flush_pending) {
                    ctx->flush_pending = false;
                    maybe_flush_batch(ctx, 0, true);
                }
                MTG_CUDA(cudaStreamWaitEvent(s, ctx->copy_done, 0));
                seg.own_packed.alloc(padded, s);
                if (want_starts) {
                    seg.starts.alloc(nr * 8, s);
                    seg.has_starts = true;
                }
                if (counts) {
                    seg.counts.alloc(nr * 4, s);
                    MTG_CUDA(cudaMemcpyAsync(seg.counts.p, h_counts, nr * 4, cudaMemcpyHostToDevice, s));
                    seg.has_counts = true;
                    ctx->any_explicit_counts = true;
                    MTG_CUDA(cudaEventRecord(pb->ev, s));   // the staging buffer is also read on this stream
                }
                pack_reads_device((const uint8_t *)st.d_raw, (const uint64_t *)st.d_off, nr, seg.own_packed.as<uint8_t>(),
                                  padded, want_starts ? seg.starts.as<uint64_t>() : nullptr, s);
                ctx->launches += 1;
                MTG_CUDA(cudaEventRecord(st.consumed, s));
                seg.d_packed = seg.own_packed.as<uint8_t>();
                seg.nbytes = padded;
                seg.n_windows_max = nwin;
                seg.n_reads = nr;
                ctx->segs.push_back(std::move(seg));
                ctx->pending_windows += nwin;
                if (ctx->push_batch && ctx->pending_windows >= ctx->push_batch && ctx->suffix.empty())
                    ctx->flush_pending = true;   // cut when the next segment's copy is in flight (or never)
                if (pinned_src) MTG_CUDA(cudaEventSynchronize(ctx->copy_done));   // the caller's memory is free again
            }
            r0 = r1;
        }
    } catch (const CudaError &e) {
        return fail_cuda(ctx, e);
    }
    return MTG_OK;
}

void *mtg_alloc_pinned(size_t bytes) {
    void *p = nullptr;
    if (cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) {
        cudaGetLastError();
        return nullptr;
    }
    return p;
}

void mtg_free_pinned(void *p) {
    if (p) cudaFreeHost(p);
}

int mtg_push_packed_device(mtg_ctx *ctx, const void *d_packed, size_t n_bytes, uint64_t n_windows_max) {
    if (!ctx || !d_packed) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
    if (((uintptr_t)d_packed & 15) != 0) return fail(ctx, MTG_ERR_BAD_ARG, "packed reads must be 16-byte aligned");
    std::lock_guard<std::mutex> lk(ctx->mu);
    try {
        MTG_CUDA(cudaSetDevice(ctx->device));
        t_arena = ctx->arena;
        // a batch never exceeds the limit by a whole segment: what is pending is folded first
        if (ctx->batch_limit && ctx->pending_windows && ctx->pending_windows + n_windows_max > ctx->batch_limit)
            maybe_flush_batch(ctx, 0, true);
        ReadSegment seg;
        seg.d_packed = (const uint8_t *)d_packed;
        seg.nbytes = n_bytes;
        seg.n_windows_max = n_windows_max;
        ctx->segs.push_back(std::move(seg));
        maybe_flush_batch(ctx, n_windows_max);   // a full batch is folded into a run here (the buffer is read now)
    } catch (const CudaError &e) {
        return fail_cuda(ctx, e);
    }
    return MTG_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Pipeline driver
// ---------------------------------------------------------------------------
namespace {

struct Timer {
    cudaEvent_t ev[8];
    int n = 0;
    cudaStream_t s;
    explicit Timer(cudaStream_t st) : s(st) {
        for (auto &e : ev) cudaEventCreate(&e);
    }
    ~Timer() { for (auto &e : ev) cudaEventDestroy(e); }
    void mark() { cudaEventRecord(ev[n++], s); }
    float ms(int a, int b) { float t = 0; cudaEventElapsedTime(&t, ev[a], ev[b]); return t; }
};

unsigned long long read_counter(const unsigned long long *d, cudaStream_t s) {
    unsigned long long h = 0;
    MTG_CUDA(cudaMemcpyAsync(&h, d, sizeof(h), cudaMemcpyDeviceToHost, s));
    MTG_CUDA(cudaStreamSynchronize(s));
    return h;
}

struct SortProfile {   // CUDA events around every kernel of the main sort
    cudaEvent_t ev[2 * MAX_SORT_PASSES + 2];
    int npasses = 0;
    uint64_t n = 0;
    uint32_t bytes_per_key = 0;
    SortProfile() { for (auto &e : ev) cudaEventCreate(&e); }
    ~SortProfile() { for (auto &e : ev) cudaEventDestroy(e); }
};


template <typename KeyT>
void sort_keys(DevBuf &a, DevBuf &b, DevBuf *va, DevBuf *vb, uint64_t n, const SortPlan &plan,
               cudaStream_t s, unsigned *launches, cudaEvent_t *events = nullptr,
               const void *first_src = nullptr) {
    // sorts; on return `a` (and *va) hold the sorted data. first_src: the unsorted keys are read
    // from there (first pass and histogram) instead of `a`
    SortWorkspace ws;
    DevBuf status(sizeof(unsigned long long) * sort_status_words<KeyT>(n), s);
    DevBuf hist(sizeof(unsigned long long) * MAX_SORT_PASSES * RADIX, s);
    DevBuf ctr(sizeof(unsigned int) * MAX_SORT_PASSES, s);
    ws.status = status.as<unsigned long long>();
    ws.hist = hist.as<unsigned long long>();
    ws.tile_counters = ctr.as<unsigned int>();
    ws.events = events;
    ws.first_src = plan.n ? first_src : nullptr;
    if (first_src && !plan.n && n)
        MTG_CUDA(cudaMemcpyAsync(a.p, first_src, sizeof(KeyT) * n, cudaMemcpyDeviceToDevice, s));
    KeyT *keys[2] = {a.as<KeyT>(), b.as<KeyT>()};
    uint32_t *vals[2] = {va ? va->as<uint32_t>() : nullptr, vb ? vb->as<uint32_t>() : nullptr};
    int cur = radix_sort<KeyT>(keys, va ? vals : nullptr, n, plan, ws, s, launches);
    if (cur == 1) {
        std::swap(a, b);
        if (va) std::swap(*va, *vb);
    }
}

// K1 over every pushed read segment; returns through *d_ctr (device) the number of k-mers
template <typename KeyT>
void run_extraction(mtg_ctx *ctx, bool both, bool payload, KeyT *out_keys, uint32_t *out_counts,
                    unsigned long long *d_ctr) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    std::vector<DevBuf> ones;
    for (auto &seg : ctx->segs) {
        ExtractArgs a;
        a.packed = seg.d_packed;
        a.nbytes = seg.nbytes;
        a.lut = &ctx->lut;
        a.K = al.K;
        a.L = al.L;
        a.canonical = ctx->prm.both_strands != 0;
        a.both = both;
        a.out_counts = payload ? out_counts : nullptr;
        a.out_counter = d_ctr;
        a.read_starts = nullptr;
        a.read_counts = nullptr;
        a.n_reads = 0;
        if (payload) {
            if (!seg.has_starts) throw CudaError{cudaErrorInvalidValue, "device-packed reads cannot be mixed with explicit counts", 0};
            a.read_starts = seg.starts.as<uint64_t>();
            a.n_reads = seg.n_reads;
            if (seg.has_counts) {
                a.read_counts = seg.counts.as<uint32_t>();
            } else {
                ones.emplace_back(4 * seg.n_reads, s);
                fill_u32(ones.back().as<uint32_t>(), 1, seg.n_reads, s, &ctx->launches);
                a.read_counts = ones.back().as<uint32_t>();
            }
        }
        extract_kmers<KeyT>(a, out_keys, s);
        ctx->launches += 1;
    }
    MTG_CUDA(cudaStreamSynchronize(s));   // `ones` must outlive the kernels
}

// extraction + sort + unique/count (+ rc completion): KmerCollector::data()
// (kmer_collector.hpp:94) followed by add_reverse_complements.
// With ext_keys (multi-GPU: the k-mers routed to this key range, both strands already present)
// extraction and rc completion are skipped.
template <typename KeyT>
void complete_reverse_complements(mtg_ctx *ctx, DevBuf &keysB, DevBuf &counts, uint64_t m, RealSet *rs);

template <typename KeyT>
void collect_reals(mtg_ctx *ctx, RealSet *rs, Timer *tm, SortProfile *prof = nullptr,
                   const void *ext_keys = nullptr, uint64_t ext_n = 0, bool skip_rc = false) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const bool counting = ctx->prm.bits_per_count > 0;
    const bool payload = counting && ctx->any_explicit_counts && !ext_keys;
    uint64_t n_ub = ext_n;
    if (!ext_keys)
        for (auto &seg : ctx->segs) n_ub += seg.n_windows_max;

    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();   // [0] extract, [1] unique, [2] rc
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64, s));
    DevBuf keysA(sizeof(KeyT) * n_ub, s), keysB(sizeof(KeyT) * n_ub, s);
    DevBuf valsA, valsB;
    if (payload) { valsA.alloc(4 * n_ub, s); valsB.alloc(4 * n_ub, s); }

    if (tm) tm->mark();
    if (ext_keys) {   // sorted straight out of the caller's (receive) buffer, no staging copy
        unsigned long long hn = ext_n;
        MTG_CUDA(cudaMemcpyAsync(d_ctr, &hn, 8, cudaMemcpyHostToDevice, s));
    } else {
        run_extraction<KeyT>(ctx, false, payload, keysA.as<KeyT>(), valsA.as<uint32_t>(), d_ctr);
    }
    const uint64_t n = read_counter(d_ctr, s);
    rs->n_input = n;
    if (tm) tm->mark();

    SortPlan plan;
    plan.n = 0;
    plan_add_key_bits(plan, 0, al.K * al.L);
    // Sequencing inputs repeat every k-mer ~coverage times. An LSD sort over the low `split`
    // bits already brings the copies of a k-mer next to each other once 2^split >> n (equal keys
    // share every digit and the passes are stable), so adjacent duplicates are folded (counts
    // summed, saturating) after the low passes and the high passes run on the folded array.
    // Folding adjacent equal keys is always safe; the final unique/count-reduce completes it.
    int split_pass = 0;
    {
        int need_bits = 2;
        while (need_bits < 64 && (n >> need_bits) != 0) ++need_bits;   // ~log2(n) + 1
        int p = (need_bits + 1 + RADIX_BITS - 1) / RADIX_BITS;
        const char *mn = getenv("MTG_MIDSORT_MIN");   // tests force the folded path on small inputs
        const uint64_t min_n = mn ? strtoull(mn, nullptr, 10) : (1ull << 22);
        if (n >= min_n && n > 0 && p < plan.n) split_pass = p;
    }
    SortPlan plan_lo = plan, plan_hi;
    plan_lo.unstable_first = true;   // equal k-mers are folded (counts summed, saturating: order-free)
    plan_hi.n = 0;
    if (split_pass) {
        plan_lo.n = split_pass;
        for (int p = split_pass; p < plan.n; ++p) plan_hi.p[plan_hi.n++] = plan.p[p];
    }
    if (prof) {
        prof->npasses = plan_lo.n;
        prof->n = n;
        prof->bytes_per_key = 2 * (sizeof(KeyT) + (payload ? 4 : 0));
    }
    sort_keys<KeyT>(keysA, keysB, payload ? &valsA : nullptr, payload ? &valsB : nullptr, n, plan_lo, s,
                    &ctx->launches, prof && n ? prof->ev : nullptr, ext_keys);

    auto fold = [&](DevBuf &src, uint64_t cnt, DevBuf &dst, const uint32_t *src_vals, DevBuf *dst_counts,
                    uint64_t counts_cap) -> uint64_t {
        // dst = distinct adjacent keys of src; *dst_counts = saturating run sums (counting mode)
        DevBuf starts;
        if (counting) starts.alloc(8 * (cnt + 1), s);
        DevBuf status(sizeof(unsigned long long) * unique_status_words(cnt), s);
        DevBuf tctr(sizeof(unsigned int), s);
        unique_keys<KeyT>(src.as<KeyT>(), cnt, dst.as<KeyT>(), counting ? starts.as<uint64_t>() : nullptr,
                          d_ctr + 1, status.as<unsigned long long>(), tctr.as<unsigned int>(), s,
                          &ctx->launches);
        const uint64_t out = read_counter(d_ctr + 1, s);
        if (counting) {
            dst_counts->alloc(4 * std::max<uint64_t>(counts_cap ? counts_cap : out, 1), s);
            run_counts(starts.as<uint64_t>(), src_vals, out, ctx->cmax, dst_counts->as<uint32_t>(), s,
                       &ctx->launches);
        }
        return out;
    };

    const bool canon = ctx->prm.both_strands != 0;
    uint64_t m = 0;
    DevBuf counts;
    if (split_pass) {
        DevBuf c1, c1b;
        const uint64_t m1 = fold(keysA, n, keysB, payload ? valsA.as<uint32_t>() : nullptr, &c1, 0);
        valsA.release();
        valsB.release();
        if (counting) c1b.alloc(4 * std::max<uint64_t>(m1, 1), s);
        if (tm) tm->mark();   // "sort" = low passes + fold; the high passes are booked under "unique"
        sort_keys<KeyT>(keysB, keysA, counting ? &c1 : nullptr, counting ? &c1b : nullptr, m1, plan_hi, s,
                        &ctx->launches);
        c1b.release();
        m = fold(keysB, m1, keysA, counting ? c1.as<uint32_t>() : nullptr, &counts, 0);
        keysB.release();
        std::swap(keysA, keysB);   // result in keysB, like the single-sort path
    } else {
        if (tm) tm->mark();
        m = fold(keysA, n, keysB, payload ? valsA.as<uint32_t>() : nullptr, &counts, 0);
        keysA.release();
        valsA.release();
        valsB.release();
    }
    const uint64_t cap = canon ? 2 * m : m;
    if (tm) tm->mark();

    if (!canon || ext_keys || skip_rc) {
        rs->keys = std::move(keysB);
        rs->counts = std::move(counts);
        rs->m = m;
        if (tm) tm->mark();
        return;
    }
    (void)cap;
    complete_reverse_complements<KeyT>(ctx, keysB, counts, m, rs);
    if (tm) tm->mark();
}

// canonical: reverse complements of the non-palindromes, sorted on their own and merged with
// the (already sorted) canonical half == add_reverse_complements' re-sort of the doubled array
// (boss_chunk_construct.cpp:179-222). keysB / counts: the m sorted distinct canonical k-mers.
template <typename KeyT>
void complete_reverse_complements(mtg_ctx *ctx, DevBuf &keysB, DevBuf &counts, uint64_t m, RealSet *rs) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const bool counting = ctx->prm.bits_per_count > 0;
    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64, s));
    SortPlan plan;
    plan.n = 0;
    plan_add_key_bits(plan, 0, al.K * al.L);
    DevBuf rb0(sizeof(KeyT) * std::max<uint64_t>(m, 1), s), rb1(sizeof(KeyT) * std::max<uint64_t>(m, 1), s);
    DevBuf cb0, cb1;
    if (counting) { cb0.alloc(4 * std::max<uint64_t>(m, 1), s); cb1.alloc(4 * std::max<uint64_t>(m, 1), s); }
    append_reverse_complements<KeyT>(keysB.as<KeyT>(), rb0.as<KeyT>(), counting ? counts.as<uint32_t>() : nullptr,
                                     counting ? cb0.as<uint32_t>() : nullptr, m, al, ctx->cmax, d_ctr + 2, s,
                                     &ctx->launches);
    const uint64_t mb = read_counter(d_ctr + 2, s);
    // (no unstable first pass here: the low digits of rc(x) come from x's leading characters, i.e. they
    // arrive in long runs of equal digits, where the returning atomics of the unstable ranking serialise)
    sort_keys<KeyT>(rb0, rb1, counting ? &cb0 : nullptr, counting ? &cb1 : nullptr, mb, plan, s,
                    &ctx->launches);
    rb1.release();
    cb1.release();
    const uint64_t m2 = m + mb;
    DevBuf merged(sizeof(KeyT) * std::max<uint64_t>(m2, 1), s), merged_counts;
    if (counting) merged_counts.alloc(4 * std::max<uint64_t>(m2, 1), s);
    {
        DevBuf part(8 * merge_partition_words(m2), s);
        merge_sorted<KeyT>(keysB.as<KeyT>(), m, rb0.as<KeyT>(), mb, counting ? counts.as<uint32_t>() : nullptr,
                           counting ? cb0.as<uint32_t>() : nullptr, merged.as<KeyT>(),
                           counting ? merged_counts.as<uint32_t>() : nullptr, part.as<uint64_t>(), s,
                           &ctx->launches);
    }
    rs->keys = std::move(merged);
    rs->counts = std::move(merged_counts);
    rs->m = m2;
}

// Two sorted runs of distinct k-mers (+ counts) -> one: merge-path merge, then equal neighbours are
// folded with saturating count sums (sorted_multiset.cpp:70-82). This is the merge step of the
// super-batched build (SortedSet::shrink_data semantics) and of mtg_mg_accumulate.
template <typename KeyT>
void merge_fold_runs(mtg_ctx *ctx, RealSet &a, RealSet &b, RealSet *out) {
    cudaStream_t s = ctx->stream;
    const bool counting = ctx->prm.bits_per_count > 0;
    const uint64_t n = a.m + b.m;
    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64, s));
    DevBuf mk(sizeof(KeyT) * std::max<uint64_t>(n, 1) + 16, s), mc;
    if (counting) mc.alloc(4 * std::max<uint64_t>(n, 1), s);
    {
        DevBuf part(8 * merge_partition_words(n), s);
        merge_sorted<KeyT>(a.keys.as<KeyT>(), a.m, b.keys.as<KeyT>(), b.m, counting ? a.counts.as<uint32_t>() : nullptr,
                           counting ? b.counts.as<uint32_t>() : nullptr, mk.as<KeyT>(),
                           counting ? mc.as<uint32_t>() : nullptr, part.as<uint64_t>(), s, &ctx->launches);
    }
    const uint64_t n_input = a.n_input + b.n_input;
    a.keys.release(); a.counts.release(); b.keys.release(); b.counts.release();
    // fold in chunks, so that the run-start scratch (8 B per key of a chunk) stays small next to arrays
    // of billions of keys; a chunk never ends between two equal keys
    DevBuf dst(sizeof(KeyT) * std::max<uint64_t>(n, 1) + 16, s), dc;
    if (counting) dc.alloc(4 * std::max<uint64_t>(n, 1), s);
    uint64_t chunk = 1ull << 29;
    if (const char *e = getenv("MTG_FOLD_CHUNK")) chunk = std::max<uint64_t>(2, strtoull(e, nullptr, 10)) & ~1ull;
    DevBuf starts, status(sizeof(unsigned long long) * unique_status_words(std::min(n, chunk + 1)), s), tctr(sizeof(unsigned int), s);
    if (counting) starts.alloc(8 * (std::min(n, chunk + 1) + 1), s);
    uint64_t m = 0;
    for (uint64_t lo = 0; lo < n;) {
        uint64_t hi = std::min(n, lo + chunk);
        if (hi < n) {   // do not split a pair
            KeyT edge[2];
            MTG_CUDA(cudaMemcpyAsync(edge, mk.as<KeyT>() + hi - 1, 2 * sizeof(KeyT), cudaMemcpyDeviceToHost, s));
            MTG_CUDA(cudaStreamSynchronize(s));
            if (w_eq(edge[0], edge[1])) ++hi;
        }
        unique_keys<KeyT>(mk.as<KeyT>() + lo, hi - lo, dst.as<KeyT>() + m, counting ? starts.as<uint64_t>() : nullptr,
                          d_ctr, status.as<unsigned long long>(), tctr.as<unsigned int>(), s, &ctx->launches);
        const uint64_t mc_chunk = read_counter(d_ctr, s);
        if (counting)
            run_counts(starts.as<uint64_t>(), mc.as<uint32_t>() + lo, mc_chunk, ctx->cmax, dc.as<uint32_t>() + m, s,
                       &ctx->launches);
        m += mc_chunk;
        lo = hi;
    }
    out->keys = std::move(dst);
    out->counts = std::move(dc);
    out->m = m;
    out->n_input = n_input;
}

// the reads pushed so far -> one sorted run of distinct (canonical) k-mers; the reads are dropped
template <typename KeyT>
void flush_batch(mtg_ctx *ctx) {
    if (ctx->segs.empty()) return;
    RealSet rs;
    collect_reals<KeyT>(ctx, &rs, nullptr, nullptr, nullptr, 0, /*skip_rc=*/true);
    ctx->runs.push_back(std::move(rs));
    ctx->segs.clear();
    ctx->any_explicit_counts = false;
    ctx->pending_windows = 0;
    ctx->n_batches += 1;
    // keep the runs geometric (like the levels of a log-structured merge): the newest two are merged
    // while the older is not more than twice the newer, so what is held is < 2x the distinct k-mers
    while (ctx->runs.size() >= 2) {
        RealSet &older = ctx->runs[ctx->runs.size() - 2], &newer = ctx->runs.back();
        if (older.m > 2 * newer.m) break;
        RealSet merged;
        merge_fold_runs<KeyT>(ctx, older, newer, &merged);
        ctx->runs.pop_back();
        ctx->runs.back() = std::move(merged);
    }
}

// pairwise merge rounds over the runs -> one run
template <typename KeyT>
void merge_all_runs(mtg_ctx *ctx, std::vector<RealSet> &runs, RealSet *out) {
    while (runs.size() > 1) {   // the runs are geometric, oldest = largest first: fold from the small end
        RealSet m;
        merge_fold_runs<KeyT>(ctx, runs[runs.size() - 2], runs.back(), &m);
        runs.pop_back();
        runs.back() = std::move(m);
    }
    *out = std::move(runs[0]);
    runs.clear();
}

// sources-1 nodes (forward-lex, unsorted) -> all dummy source k-mers; the result arrays hold
// [n_sink dummy sinks | d_src dummy sources] with their info bytes. d_ctr + 8 = level histogram.
template <typename KeyT>
void expand_sources(mtg_ctx *ctx, DevBuf &src, uint64_t d1, DevBuf &sink, uint64_t n_sink,
                    unsigned long long *d_ctr, DevBuf &dk0, DevBuf &di0, uint64_t *d_out,
                    const KeyT *lo = nullptr, const KeyT *hi = nullptr) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint32_t K = al.K, L = al.L;
    DevBuf srcB;
    // sources-1 in forward-lex order, then all their distinct prefixes
    srcB.alloc(sizeof(KeyT) * std::max<uint64_t>(d1, 1), s);
    {
        SortPlan plan;
        plan.n = 0;
        plan_add_key_bits(plan, 0, (K - 1) * L);
        sort_keys<KeyT>(src, srcB, nullptr, nullptr, d1, plan, s, &ctx->launches);
    }
    srcB.release();
    DevBuf per_entry(4 * std::max<uint64_t>(d1, 1), s), offsets(8 * (d1 + 1), s);
    source_prefix_counts<KeyT>(src.as<KeyT>(), d1, al, lo, hi, per_entry.as<uint32_t>(), s, &ctx->launches);
    {
        DevBuf status(8 * (d1 / 2048 + 2), s), tctr(4, s);
        exclusive_scan_u32(per_entry.as<uint32_t>(), d1, offsets.as<uint64_t>(),
                           status.as<unsigned long long>(), tctr.as<unsigned int>(), s, &ctx->launches);
    }
    const uint64_t d_src = read_counter(offsets.as<unsigned long long>() + d1, s);
    const uint64_t d = n_sink + d_src;
    dk0.alloc(sizeof(KeyT) * std::max<uint64_t>(d, 1), s);
    di0.alloc(4 * std::max<uint64_t>(d, 1), s);
    MTG_CUDA(cudaMemcpyAsync(dk0.p, sink.p, sizeof(KeyT) * n_sink, cudaMemcpyDeviceToDevice, s));
    fill_u32(di0.as<uint32_t>(), INFO_W_DOLLAR, n_sink, s, &ctx->launches);
    generate_source_dummies<KeyT>(src.as<KeyT>(), offsets.as<uint64_t>(), d1, al, lo, hi,
                                  dk0.as<KeyT>() + n_sink, di0.as<uint32_t>() + n_sink, d_ctr + 8, s,
                                  &ctx->launches);
    sink.release(); src.release(); per_entry.release(); offsets.release();
    *d_out = d;
}

// rows -> W / last / F / weights (K6) -> mtg_chunk (pinned host copy or device-resident)
// launches the kernels that write W / last / weights / F (device pointers) of an E-row chunk
using EmitFn = std::function<void(uint64_t *, uint64_t *, uint64_t *, unsigned long long *)>;

template <typename KeyT>
void emit_deliver(mtg_ctx *ctx, const EmitFn &emit, uint64_t E,
                  const unsigned long long *d_levels, Timer &tm, SortProfile &prof, mtg_chunk *out,
                  bool device_output, uint64_t n_input, uint64_t n_real, uint64_t n_sink) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint32_t K = al.K;
    const bool counting = ctx->prm.bits_per_count > 0;
    const uint32_t ww = ctx->prm.bits_per_count;
    const size_t tiles = div_up(E, (size_t)2048);
    const size_t W_words = tiles * 32 * al.W_width, last_words = tiles * 32, wt_words = tiles * 32 * ww;
    DevBuf outbuf(8 * (W_words + last_words + wt_words + MTG_MAX_SIGMA), s);
    uint64_t *dW = outbuf.as<uint64_t>(), *dLast = dW + W_words, *dWt = dLast + last_words;
    unsigned long long *dF = (unsigned long long *)(dWt + wt_words);
    emit(dW, dLast, counting ? dWt : nullptr, dF);
    tm.mark();   // 6: emit done

    // ---- results to pinned host memory ----
    ChunkImpl *impl = new ChunkImpl();
    const size_t arr_words = device_output ? 0 : W_words + last_words + wt_words;
    size_t hbytes = 8 * (arr_words + MTG_MAX_SIGMA + MTG_MAX_K + 16);
    impl->pinned = g_pinned_cache.acquire(hbytes, &impl->pinned_bytes);
    if (!impl->pinned) {
        delete impl;
        throw CudaError{cudaErrorMemoryAllocation, __FILE__, __LINE__};
    }
    uint64_t *h = (uint64_t *)impl->pinned;
    MTG_CUDA(cudaMemcpyAsync(h, dW + (device_output ? W_words + last_words + wt_words : 0),
                             8 * (arr_words + MTG_MAX_SIGMA), cudaMemcpyDeviceToHost, s));
    uint64_t *h_levels = h + arr_words + MTG_MAX_SIGMA;
    MTG_CUDA(cudaMemcpyAsync(h_levels, d_levels, 8 * MTG_MAX_K, cudaMemcpyDeviceToHost, s));
    MTG_CUDA(cudaStreamSynchronize(s));

    memset(out, 0, sizeof(*out));
    out->n_rows = E;
    out->sigma = al.sigma;
    out->k = K - 1;
    out->W_width = al.W_width;
    out->weights_width = ww;
    if (device_output) {
        out->W = dW;
        out->last = dLast;
        out->weights = counting ? dWt : nullptr;
        out->device_resident = 1;
        impl->dev = std::move(outbuf);
    } else {
        out->W = h;
        out->last = h + W_words;
        out->weights = counting ? h + W_words + last_words : nullptr;
    }
    const uint64_t *hF = h + arr_words;
    if (prof.n) {
        float t = 0;
        for (int p = 0; p < prof.npasses; ++p) {
            cudaEventElapsedTime(&t, prof.ev[2 * p], prof.ev[2 * p + 1]);
            out->ms_onesweep_sum += t;
        }
        cudaEventElapsedTime(&t, prof.ev[2 * MAX_SORT_PASSES], prof.ev[2 * MAX_SORT_PASSES + 1]);
        out->ms_sort_histogram = t;
        out->n_onesweep = prof.npasses;
        out->onesweep_bytes_per_key = prof.bytes_per_key;
        out->onesweep_keys = prof.n;
    }
    for (uint32_t c = 0; c < al.sigma; ++c) out->F[c] = hF[c];
    out->n_input_kmers = n_input;
    out->n_real = n_real;
    out->n_dummy_sink = n_sink;
    for (uint32_t c = 0; c + 1 < K && c < MTG_MAX_K; ++c) out->n_dummy_source[c] = h_levels[c];
    out->ms_extract = tm.ms(0, 1);
    out->ms_sort = tm.ms(1, 2);
    out->ms_unique = tm.ms(2, 3);
    out->ms_rc = tm.ms(3, 4);
    out->ms_dummy = tm.ms(4, 5);
    out->ms_emit = tm.ms(5, 6);
    out->ms_total = tm.ms(0, 6);
    out->n_kernel_launches = ctx->launches;
    out->impl = impl;
}

// dummy k-mers (sinks + sources, any order) + sorted reals -> rows -> W / last / F / weights.
// row_off = 2: rows 0 (placeholder) and 1 (the all-'$' k-mer) precede the merge; 1: placeholder
// only (a key range other than the first one, multi-GPU).
template <typename KeyT>
void finish_chunk(mtg_ctx *ctx, RealSet &rs, DevBuf &dk0, DevBuf &di0, uint64_t d, uint64_t n_sink,
                  unsigned long long *d_ctr, const uint64_t *d_bidx, Timer &tm, SortProfile &prof,
                  mtg_chunk *out, bool device_output, uint32_t row_off) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint32_t K = al.K, L = al.L;
    const bool counting = ctx->prm.bits_per_count > 0;
    const uint64_t m = rs.m;
    const KeyT *reals = rs.keys.as<KeyT>();
    DevBuf dk1(sizeof(KeyT) * std::max<uint64_t>(d, 1), s), di1(4 * std::max<uint64_t>(d, 1), s);
    // final order: node bits, then more '$' first, then W  (LSD: W, nd, node)
    flip_info(di0.as<uint32_t>(), d, s, &ctx->launches);
    {
        SortPlan plan;
        plan.n = 0;
        plan.p[plan.n++] = DigitPass{0, 0, (int)L};
        plan.p[plan.n++] = DigitPass{1, 0, 7};
        plan_add_key_bits(plan, L, K * L);
        sort_keys<KeyT>(dk0, dk1, &di0, &di1, d, plan, s, &ctx->launches);
    }
    flip_info(di0.as<uint32_t>(), d, s, &ctx->launches);
    dk1.release(); di1.release();
    DevBuf rank(8 * std::max<uint64_t>(d, 1), s);
    dummy_ranks<KeyT>(reals, m, dk0.as<KeyT>(), d, al, d_bidx, rank.as<uint64_t>(), s, &ctx->launches);
    tm.mark();   // 5: dummies done

    // ---- merge + emit ----
    const uint64_t E = m + d + row_off;
    const uint32_t ww = ctx->prm.bits_per_count;
    if (L == 2 && al.W_width == 4 && !getenv("MTG_EMIT_UNFUSED")) {
        // DNA: one kernel merges and emits out of shared memory, the rows are never written to HBM
        DevBuf part(8 * emit_merged_scratch_words(E), s);
        const uint32_t *rc = counting ? rs.counts.as<uint32_t>() : nullptr;
        emit_deliver<KeyT>(ctx, [&](uint64_t *dW, uint64_t *dLast, uint64_t *dWt, unsigned long long *dF) {
            emit_merged<KeyT>(reals, rc, m, dk0.as<KeyT>(), di0.as<uint32_t>(), rank.as<uint64_t>(), d, row_off,
                              al, ww, part.as<uint64_t>(), dW, dLast, dWt, dF, s, &ctx->launches);
        }, E, d_ctr + 8, tm, prof, out, device_output, rs.n_input, m, n_sink);
        return;
    }
    DevBuf row_keys(sizeof(KeyT) * E, s), row_info(E, s), row_counts;
    if (counting) row_counts.alloc(4 * E, s);
    merge_rows<KeyT>(reals, counting ? rs.counts.as<uint32_t>() : nullptr, m, dk0.as<KeyT>(),
                     di0.as<uint32_t>(), rank.as<uint64_t>(), d, al, row_keys.as<KeyT>(),
                     row_info.as<uint8_t>(), counting ? row_counts.as<uint32_t>() : nullptr, s,
                     &ctx->launches, row_off);
    rs.keys.release(); rs.counts.release(); dk0.release(); di0.release(); rank.release();
    emit_deliver<KeyT>(ctx, [&](uint64_t *dW, uint64_t *dLast, uint64_t *dWt, unsigned long long *dF) {
        emit_chunk<KeyT>(row_keys.as<KeyT>(), row_info.as<uint8_t>(), counting ? row_counts.as<uint32_t>() : nullptr,
                         E, al, ww, dW, dLast, dWt, dF, s, &ctx->launches);
    }, E, d_ctr + 8, tm, prof, out, device_output, rs.n_input, m, n_sink);
}

template <typename KeyT>
void build_impl(mtg_ctx *ctx, mtg_chunk *out, bool device_output) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    ctx->launches = 0;
    Timer tm(s);
    RealSet rs;
    SortProfile prof;
    if (ctx->runs.empty()) {
        collect_reals<KeyT>(ctx, &rs, &tm, &prof);   // marks 0..4: start, extract, sort, unique, rc
    } else {
        // super-batched input: last batch, merge of the runs, then the second strand
        tm.mark();
        flush_batch<KeyT>(ctx);
        tm.mark();
        tm.mark();
        RealSet all;
        merge_all_runs<KeyT>(ctx, ctx->runs, &all);
        tm.mark();
        if (ctx->prm.both_strands) {
            complete_reverse_complements<KeyT>(ctx, all.keys, all.counts, all.m, &rs);
            rs.n_input = all.n_input;
        } else {
            rs = std::move(all);
        }
        tm.mark();
    }
    const uint64_t m = rs.m;
    const KeyT *reals = rs.keys.as<KeyT>();

    // ---- dummy k-mers ----
    DevBuf ctr(64 + 8 * MTG_MAX_K, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();   // [0] sinks [1] sources-1 [8..] levels
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64 + 8 * MTG_MAX_K, s));
    uint64_t cap = std::max<uint64_t>(1024, m / 8);
    DevBuf bidx;
    const bool use_index = m >= (1u << 16);
    if (use_index) {
        bidx.alloc(8 * bucket_index_words(al), s);
        build_bucket_index<KeyT>(reals, m, al, bidx.as<uint64_t>(), s, &ctx->launches);
    }
    const uint64_t *d_bidx = use_index ? bidx.as<uint64_t>() : nullptr;
    DevBuf sink, src, srcB;
    DevBuf dscratch(dummy_scratch_bytes(m, al), s);
    uint64_t n_sink = 0, d1 = 0;
    while (true) {
        sink.alloc(sizeof(KeyT) * cap, s);
        src.alloc(sizeof(KeyT) * cap, s);
        find_dummy_candidates<KeyT>(reals, m, al, d_bidx, dscratch.p, sink.as<KeyT>(), cap, d_ctr, src.as<KeyT>(), cap,
                                    d_ctr + 1, s, &ctx->launches);
        n_sink = read_counter(d_ctr, s);
        d1 = read_counter(d_ctr + 1, s);
        if (n_sink <= cap && d1 <= cap) break;
        cap = std::max(n_sink, d1);
    }
    dscratch.release();
    uint64_t d = 0;
    DevBuf dk0, di0;
    expand_sources<KeyT>(ctx, src, d1, sink, n_sink, d_ctr, dk0, di0, &d);
    finish_chunk<KeyT>(ctx, rs, dk0, di0, d, n_sink, d_ctr, d_bidx, tm, prof, out, device_output, 2);
}

// `build --suffix S`: framed k-mers matching the filter -> sort -> unique / counts -> rows ->
// chunk (BOSSChunkConstructor's KmerExtractorBOSS branch, boss_chunk_construct.cpp:1002-1013).
template <typename KeyT>
void build_suffix_impl(mtg_ctx *ctx, mtg_chunk *out, bool device_output) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const bool counting = ctx->prm.bits_per_count > 0;
    const bool payload = counting && ctx->any_explicit_counts;
    const bool both = ctx->prm.both_strands != 0;   // Mode::BOTH (:961-971)
    ctx->launches = 0;
    Timer tm(s);
    SortProfile prof;
    bool all_dollar = true;
    for (uint8_t c : ctx->suffix) all_dollar &= c == 0;

    DevBuf ctr(64 + 8 * MTG_MAX_K, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64 + 8 * MTG_MAX_K, s));
    uint64_t nbytes = 0;
    for (auto &seg : ctx->segs) nbytes += seg.nbytes;
    uint64_t cap = std::max<uint64_t>(1024, nbytes / 2);
    DevBuf keysA, keysB, valsA, valsB;
    std::vector<DevBuf> ones;
    uint64_t n = 0;
    tm.mark();
    while (true) {
        keysA.alloc(sizeof(KeyT) * (cap + 1), s);
        if (payload) valsA.alloc(4 * (cap + 1), s);
        MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 8, s));
        for (auto &seg : ctx->segs) {
            ExtractArgs a;
            a.packed = seg.d_packed;
            a.nbytes = seg.nbytes;
            a.lut = &ctx->lut;
            a.K = al.K;
            a.L = al.L;
            a.canonical = false;
            a.both = both;
            a.out_counts = payload ? valsA.as<uint32_t>() : nullptr;
            a.out_counter = d_ctr;
            a.read_starts = nullptr;
            a.read_counts = nullptr;
            a.n_reads = 0;
            if (payload) {
                if (!seg.has_starts)
                    throw CudaError{cudaErrorInvalidValue, "device-packed reads cannot be mixed with explicit counts", 0};
                a.read_starts = seg.starts.as<uint64_t>();
                a.n_reads = seg.n_reads;
                if (seg.has_counts) {
                    a.read_counts = seg.counts.as<uint32_t>();
                } else {
                    ones.emplace_back(4 * seg.n_reads, s);
                    fill_u32(ones.back().as<uint32_t>(), 1, seg.n_reads, s, &ctx->launches);
                    a.read_counts = ones.back().as<uint32_t>();
                }
            }
            extract_suffix_kmers<KeyT>(a, al, ctx->suffix.data(), (uint32_t)ctx->suffix.size(), keysA.as<KeyT>(),
                                       cap, s);
            ctx->launches += 1;
        }
        n = read_counter(d_ctr, s);
        ones.clear();
        if (n <= cap) break;
        cap = n;
    }
    const uint64_t n_input = n;
    if (all_dollar) {   // kmer_collector_.add_kmer(all '$'), count 1 (:978-981)
        const uint32_t sortinfo = (127u - (al.K - 1)) << 1;
        KeyT dollar = w_shl(WordTraits<KeyT>::from_u64(sortinfo), al.L);
        MTG_CUDA(cudaMemcpyAsync(keysA.as<KeyT>() + n, &dollar, sizeof(KeyT), cudaMemcpyHostToDevice, s));
        if (payload) {
            const uint32_t one = 1;
            MTG_CUDA(cudaMemcpyAsync(valsA.as<uint32_t>() + n, &one, 4, cudaMemcpyHostToDevice, s));
        }
        MTG_CUDA(cudaStreamSynchronize(s));
        ++n;
    }
    tm.mark();   // 1: extraction done

    keysB.alloc(sizeof(KeyT) * std::max<uint64_t>(n, 1), s);
    if (payload) valsB.alloc(4 * std::max<uint64_t>(n, 1), s);
    SortPlan plan;
    plan.n = 0;
    plan_add_key_bits(plan, 0, al.K * al.L + 8);
    sort_keys<KeyT>(keysA, keysB, payload ? &valsA : nullptr, payload ? &valsB : nullptr, n, plan, s,
                    &ctx->launches);
    tm.mark();   // 2: sorted

    DevBuf counts, starts;
    if (counting) starts.alloc(8 * (n + 1), s);
    {
        DevBuf status(sizeof(unsigned long long) * unique_status_words(n), s), tctr(sizeof(unsigned int), s);
        unique_keys<KeyT>(keysA.as<KeyT>(), n, keysB.as<KeyT>(), counting ? starts.as<uint64_t>() : nullptr,
                          d_ctr + 1, status.as<unsigned long long>(), tctr.as<unsigned int>(), s, &ctx->launches);
    }
    const uint64_t m = read_counter(d_ctr + 1, s);
    if (counting) {
        counts.alloc(4 * std::max<uint64_t>(m, 1), s);
        run_counts(starts.as<uint64_t>(), payload ? valsA.as<uint32_t>() : nullptr, m, ctx->cmax,
                   counts.as<uint32_t>(), s, &ctx->launches);
    }
    keysA.release(); valsA.release(); valsB.release(); starts.release();
    tm.mark();   // 3: distinct k-mers (+ counts)
    tm.mark();   // 4: (no reverse-complement completion in suffix mode)

    DevBuf keep(4 * std::max<uint64_t>(m, 1), s), offs(8 * (m + 1), s);
    suffix_keep_flags<KeyT>(keysB.as<KeyT>(), m, al, keep.as<uint32_t>(), s, &ctx->launches);
    {
        DevBuf status(8 * (m / 2048 + 2), s), tctr(4, s);
        exclusive_scan_u32(keep.as<uint32_t>(), m, offs.as<uint64_t>(), status.as<unsigned long long>(),
                           tctr.as<unsigned int>(), s, &ctx->launches);
    }
    const uint64_t kept = read_counter(offs.as<unsigned long long>() + m, s);
    const uint64_t E = kept + 1;
    DevBuf row_keys(sizeof(KeyT) * E, s), row_info(E, s), row_counts;
    if (counting) row_counts.alloc(4 * E, s);
    suffix_rows<KeyT>(keysB.as<KeyT>(), counting ? counts.as<uint32_t>() : nullptr, keep.as<uint32_t>(),
                      offs.as<uint64_t>(), m, al, row_keys.as<KeyT>(), row_info.as<uint8_t>(),
                      counting ? row_counts.as<uint32_t>() : nullptr, s, &ctx->launches);
    keysB.release(); counts.release(); keep.release(); offs.release();
    tm.mark();   // 5: rows
    const uint32_t ww = ctx->prm.bits_per_count;
    emit_deliver<KeyT>(ctx, [&](uint64_t *dW, uint64_t *dLast, uint64_t *dWt, unsigned long long *dF) {
        emit_chunk<KeyT>(row_keys.as<KeyT>(), row_info.as<uint8_t>(), counting ? row_counts.as<uint32_t>() : nullptr,
                         E, al, ww, dW, dLast, dWt, dF, s, &ctx->launches);
    }, E, d_ctr + 8, tm, prof, out, device_output, n_input, m, 0);
}

void clear_ctx(mtg_ctx *ctx) {
    ctx->segs.clear();
    ctx->runs.clear();
    ctx->pending_windows = 0;
    ctx->flush_pending = false;
    ctx->any_explicit_counts = false;
}

// caller holds ctx->mu: cut a batch when the pending reads hold batch_limit k-mer windows
void maybe_flush_batch(mtg_ctx *ctx, uint64_t new_windows, bool force) {
    ctx->pending_windows += new_windows;
    if (!ctx->suffix.empty()) return;
    if (!force && (!ctx->batch_limit || ctx->pending_windows < ctx->batch_limit)) return;
    if (ctx->al.K * ctx->al.L <= 64) flush_batch<uint64_t>(ctx);
    else flush_batch<u128>(ctx);
}

}  // namespace

extern "C" {

static int build_chunk_common(mtg_ctx *ctx, mtg_chunk *out, bool device_output) {
    if (!ctx || !out) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = MTG_OK;
    try {
        MTG_CUDA(cudaSetDevice(ctx->device));
        t_arena = ctx->arena;
        if (!ctx->suffix.empty()) {
            if (ctx->al.K * ctx->al.L + 8 <= 64) build_suffix_impl<uint64_t>(ctx, out, device_output);
            else build_suffix_impl<u128>(ctx, out, device_output);
        } else if (ctx->al.K * ctx->al.L <= 64) build_impl<uint64_t>(ctx, out, device_output);
        else build_impl<u128>(ctx, out, device_output);
    } catch (const CudaError &e) {
        rc = fail_cuda(ctx, e);
        cudaStreamSynchronize(ctx->stream);
    }
    clear_ctx(ctx);
    return rc;
}

int mtg_build_chunk(mtg_ctx *ctx, mtg_chunk *out) { return build_chunk_common(ctx, out, false); }
int mtg_build_chunk_device(mtg_ctx *ctx, mtg_chunk *out) { return build_chunk_common(ctx, out, true); }

int mtg_set_stream(mtg_ctx *ctx, void *cuda_stream) {
    if (!ctx) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (!ctx->segs.empty()) return fail(ctx, MTG_ERR_STATE, "mtg_set_stream must precede the first push");
    cudaStreamSynchronize(ctx->stream);
    ctx->stream = cuda_stream ? (cudaStream_t)cuda_stream : ctx->own_stream;
    return MTG_OK;
}

void mtg_free_chunk(mtg_chunk *chunk) {
    if (!chunk || !chunk->impl) return;
    ChunkImpl *impl = (ChunkImpl *)chunk->impl;
    if (impl->pinned) g_pinned_cache.release(impl->pinned, impl->pinned_bytes);
    impl->dev.release();
    delete impl;
    chunk->impl = nullptr;
    chunk->W = chunk->last = chunk->weights = nullptr;
}

int mtg_debug_sorted_kmers(mtg_ctx *ctx, uint64_t **words, uint32_t **counts, uint64_t *n) {
    if (!ctx || !words || !n) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
    std::lock_guard<std::mutex> lk(ctx->mu);
    int rc = MTG_OK;
    try {
        MTG_CUDA(cudaSetDevice(ctx->device));
        t_arena = ctx->arena;
        RealSet rs;
        const bool wide = ctx->al.K * ctx->al.L > 64;
        if (wide) collect_reals<u128>(ctx, &rs, nullptr);
        else collect_reals<uint64_t>(ctx, &rs, nullptr);
        *n = rs.m;
        *words = (uint64_t *)malloc(16 * std::max<uint64_t>(rs.m, 1));
        if (wide) {
            MTG_CUDA(cudaMemcpyAsync(*words, rs.keys.p, 16 * rs.m, cudaMemcpyDeviceToHost, ctx->stream));
            MTG_CUDA(cudaStreamSynchronize(ctx->stream));
        } else {
            std::vector<uint64_t> tmp(rs.m);
            MTG_CUDA(cudaMemcpyAsync(tmp.data(), rs.keys.p, 8 * rs.m, cudaMemcpyDeviceToHost, ctx->stream));
            MTG_CUDA(cudaStreamSynchronize(ctx->stream));
            for (uint64_t i = 0; i < rs.m; ++i) { (*words)[2 * i] = tmp[i]; (*words)[2 * i + 1] = 0; }
        }
        if (counts) {
            *counts = nullptr;
            if (ctx->prm.bits_per_count) {
                *counts = (uint32_t *)malloc(4 * std::max<uint64_t>(rs.m, 1));
                MTG_CUDA(cudaMemcpyAsync(*counts, rs.counts.p, 4 * rs.m, cudaMemcpyDeviceToHost, ctx->stream));
                MTG_CUDA(cudaStreamSynchronize(ctx->stream));
            }
        }
    } catch (const CudaError &e) {
        rc = fail_cuda(ctx, e);
    }
    clear_ctx(ctx);
    return rc;
}

void mtg_debug_free(void *p) { free(p); }

static int debug_sort(int device, void *keys, uint32_t *vals, uint64_t n, uint32_t key_bits, bool wide) {
    try {
        MTG_CUDA(cudaSetDevice(device));
        t_arena = std::make_shared<Arena>();
        cudaStream_t s;
        MTG_CUDA(cudaStreamCreate(&s));
        {
            size_t kb = wide ? 16 : 8;
            DevBuf a(kb * n, s), b(kb * n, s), va, vb;
            MTG_CUDA(cudaMemcpyAsync(a.p, keys, kb * n, cudaMemcpyHostToDevice, s));
            if (vals) {
                va.alloc(4 * n, s); vb.alloc(4 * n, s);
                MTG_CUDA(cudaMemcpyAsync(va.p, vals, 4 * n, cudaMemcpyHostToDevice, s));
            }
            SortPlan plan;
            plan.n = 0;
            plan_add_key_bits(plan, 0, key_bits);
            unsigned launches = 0;
            if (wide) sort_keys<u128>(a, b, vals ? &va : nullptr, vals ? &vb : nullptr, n, plan, s, &launches);
            else sort_keys<uint64_t>(a, b, vals ? &va : nullptr, vals ? &vb : nullptr, n, plan, s, &launches);
            MTG_CUDA(cudaMemcpyAsync(keys, a.p, kb * n, cudaMemcpyDeviceToHost, s));
            if (vals) MTG_CUDA(cudaMemcpyAsync(vals, va.p, 4 * n, cudaMemcpyDeviceToHost, s));
            MTG_CUDA(cudaStreamSynchronize(s));
        }
        cudaStreamSynchronize(s);
        cudaStreamDestroy(s);
        t_arena.reset();
    } catch (const CudaError &e) {
        return fail_cuda(nullptr, e);
    }
    return MTG_OK;
}

// Device-resident timing of the radix sort alone on pseudo-random keys (sort tuning):
// ms_out[0] = histogram kernel, ms_out[1] = mean onesweep pass, ms_out[2] = whole sort; best of iters.
int mtg_debug_sort_bench(int device, uint64_t n, uint32_t key_bits, int with_vals, int iters,
                         float *ms_out) {
    const bool wide = key_bits > 64;
    try {
        MTG_CUDA(cudaSetDevice(device));
        t_arena = std::make_shared<Arena>();
        cudaStream_t s;
        MTG_CUDA(cudaStreamCreate(&s));
        {
            size_t kb = wide ? 16 : 8;
            DevBuf a(kb * n, s), b(kb * n, s), va, vb;
            if (with_vals) { va.alloc(4 * n, s); vb.alloc(4 * n, s); }
            SortPlan plan;
            plan.n = 0;
            plan_add_key_bits(plan, 0, key_bits);
            plan.unstable_first = getenv("MTG_BENCH_UNSTABLE_FIRST") != nullptr;
            uint32_t top = wide ? key_bits - 64 : key_bits;
            uint64_t hi_mask = top >= 64 ? ~0ull : ((1ull << top) - 1);
            cudaEvent_t ev[2 * MAX_SORT_PASSES + 2], e0, e1;
            for (auto &e : ev) cudaEventCreate(&e);
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            ms_out[0] = ms_out[1] = ms_out[2] = ms_out[3] = ms_out[4] = 1e30f;
            for (int it = 0; it < iters; ++it) {
                DevBuf &cur = a;
                fill_random_words(cur.as<uint64_t>(), n * (wide ? 2 : 1), it + 1, hi_mask, wide, s);
                if (with_vals) fill_random_words(va.as<uint64_t>(), n / 2, it + 7, ~0ull, false, s);
                unsigned launches = 0;
                cudaEventRecord(e0, s);
                if (wide) sort_keys<u128>(a, b, with_vals ? &va : nullptr, with_vals ? &vb : nullptr, n, plan, s, &launches, ev);
                else sort_keys<uint64_t>(a, b, with_vals ? &va : nullptr, with_vals ? &vb : nullptr, n, plan, s, &launches, ev);
                cudaEventRecord(e1, s);
                MTG_CUDA(cudaStreamSynchronize(s));
                float t = 0, sum = 0;
                for (int p = 0; p < plan.n; ++p) { cudaEventElapsedTime(&t, ev[2 * p], ev[2 * p + 1]); sum += t; }
                ms_out[1] = std::min(ms_out[1], sum / plan.n);
                cudaEventElapsedTime(&t, ev[0], ev[1]);
                ms_out[3] = std::min(ms_out[3], t);
                if (plan.n > 1) ms_out[4] = std::min(ms_out[4], (sum - t) / (plan.n - 1));
                cudaEventElapsedTime(&t, ev[2 * MAX_SORT_PASSES], ev[2 * MAX_SORT_PASSES + 1]);
                ms_out[0] = std::min(ms_out[0], t);
                cudaEventElapsedTime(&t, e0, e1);
                ms_out[2] = std::min(ms_out[2], t);
            }
            for (auto &e : ev) cudaEventDestroy(e);
            cudaEventDestroy(e0); cudaEventDestroy(e1);
        }
        cudaStreamSynchronize(s);
        cudaStreamDestroy(s);
        t_arena.reset();
    } catch (const CudaError &e) {
        return fail_cuda(nullptr, e);
    }
    return MTG_OK;
}

// tuning hook: stable-pass kernel (3 = peer masks, 5 = atomic slot + repair) and whether the first pass
// of an order-free sort may rank unstably; -1 keeps a setting
void mtg_debug_sort_config(int variant, int unstable_first, int shape, int prefetch) {
    set_sort_config(variant, unstable_first, shape, prefetch);
}

// Device-resident timing of the unique / fold kernel alone: keys[i] = i / dup (u64), best of iters.
int mtg_debug_unique_bench(int device, uint64_t n, uint32_t dup, int with_starts, int iters, float *ms_out) {
    try {
        MTG_CUDA(cudaSetDevice(device));
        t_arena = std::make_shared<Arena>();
        cudaStream_t s;
        MTG_CUDA(cudaStreamCreate(&s));
        {
            DevBuf a(8 * n, s), b(8 * n, s), starts, ctr(64, s), tctr(4, s);
            DevBuf status(sizeof(unsigned long long) * unique_status_words(n), s);
            if (with_starts) starts.alloc(8 * (n + 1), s);
            std::vector<uint64_t> h(1 << 20);
            fill_random_words(a.as<uint64_t>(), n, 1, ~0ull, false, s);   // touch
            fill_iota_div(a.as<uint64_t>(), n, dup ? dup : 1, s);
            cudaEvent_t e0, e1;
            cudaEventCreate(&e0); cudaEventCreate(&e1);
            ms_out[0] = 1e30f;
            unsigned launches = 0;
            for (int it = 0; it < iters; ++it) {
                cudaEventRecord(e0, s);
                unique_keys<uint64_t>(a.as<uint64_t>(), n, b.as<uint64_t>(), with_starts ? starts.as<uint64_t>() : nullptr,
                                      ctr.as<unsigned long long>(), status.as<unsigned long long>(),
                                      tctr.as<unsigned int>(), s, &launches);
                cudaEventRecord(e1, s);
                MTG_CUDA(cudaStreamSynchronize(s));
                float t = 0;
                cudaEventElapsedTime(&t, e0, e1);
                ms_out[0] = std::min(ms_out[0], t);
            }
            unsigned long long m = 0;
            MTG_CUDA(cudaMemcpy(&m, ctr.p, 8, cudaMemcpyDeviceToHost));
            ms_out[1] = (float)m;
            cudaEventDestroy(e0); cudaEventDestroy(e1);
        }
        cudaStreamSynchronize(s);
        cudaStreamDestroy(s);
        t_arena.reset();
    } catch (const CudaError &e) {
        return fail_cuda(nullptr, e);
    }
    return MTG_OK;
}

int mtg_debug_sort_u64(int device, uint64_t *keys, uint32_t *vals, uint64_t n, uint32_t key_bits) {
    return debug_sort(device, keys, vals, n, key_bits, false);
}
int mtg_debug_sort_u128(int device, uint64_t *keys, uint32_t *vals, uint64_t n, uint32_t key_bits) {
    return debug_sort(device, keys, vals, n, key_bits, true);
}

}  // extern "C"

// ---------------------------------------------------------------------------
// Range-partitioned (multi-GPU) build, SURVEY.md 8e. Each context owns the k-mers whose top
// `part_bits` bits fall in its bucket range; the caller (multi_gpu.py) moves data between
// contexts with NCCL all-to-all / all-gather. part_bits <= min(8, (K-2)*L) so that no group
// of rows sharing kmer[2..K] straddles two ranges.
// ---------------------------------------------------------------------------
namespace {

template <typename KeyT>
void mg_extract_impl(mtg_ctx *ctx, uint32_t pb, uint64_t *bucket_counts, void **d_keys, uint64_t *n_out) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    // canonical mode: only min(x, rc(x)) is routed and reduced here; the second strand is
    // produced from the DISTINCT canonical k-mers afterwards (mtg_mg_strands), like
    // add_reverse_complements does after the collector (boss_chunk_construct.cpp:179-222)
    const bool both = false;
    uint64_t n_ub = 0;
    for (auto &seg : ctx->segs) n_ub += seg.n_windows_max;
    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64, s));
    DevBuf a(sizeof(KeyT) * std::max<uint64_t>(n_ub, 1), s), b(sizeof(KeyT) * std::max<uint64_t>(n_ub, 1), s);
    ctx->launches = 0;
    run_extraction<KeyT>(ctx, both, false, a.as<KeyT>(), nullptr, d_ctr);
    const uint64_t n = read_counter(d_ctr, s);
    for (uint32_t i = 0; i < (1u << pb); ++i) bucket_counts[i] = 0;
    if (pb == 0 || n == 0) {
        bucket_counts[0] = n;
    } else {
        // one MSD pass on the top pb bits groups the k-mers by bucket; the digit offsets of
        // the pass are the bucket boundaries
        SortPlan plan;
        plan.n = 0;
        plan.p[plan.n++] = DigitPass{0, (int)(al.K * al.L - pb), (int)pb};
        plan.unstable_first = true;   // grouping only: mtg_mg_collect sorts
        SortWorkspace ws;
        DevBuf status(sizeof(unsigned long long) * sort_status_words<KeyT>(n), s);
        DevBuf hist(sizeof(unsigned long long) * MAX_SORT_PASSES * RADIX, s);
        DevBuf tc(sizeof(unsigned int) * MAX_SORT_PASSES, s);
        ws.status = status.as<unsigned long long>();
        ws.hist = hist.as<unsigned long long>();
        ws.tile_counters = tc.as<unsigned int>();
        KeyT *keys[2] = {a.as<KeyT>(), b.as<KeyT>()};
        int cur = radix_sort<KeyT>(keys, nullptr, n, plan, ws, s, &ctx->launches);
        std::vector<unsigned long long> off(RADIX);
        MTG_CUDA(cudaMemcpyAsync(off.data(), ws.hist, 8 * RADIX, cudaMemcpyDeviceToHost, s));
        MTG_CUDA(cudaStreamSynchronize(s));
        const uint32_t nb = 1u << pb;
        for (uint32_t i = 0; i < nb; ++i) bucket_counts[i] = (i + 1 < nb ? off[i + 1] : n) - off[i];
        if (cur == 1) std::swap(a, b);
    }
    ctx->mg.keys = std::move(a);
    ctx->mg.n_input += both ? n / 2 : n;
    *d_keys = ctx->mg.keys.p;
    *n_out = n;
    ctx->segs.clear();
}

// Phase 1 split for the fused exchange: mg_count extracts and counts the k-mers per routing bucket
// (they stay unordered in the context); mg_route then scatters them ONCE, each bucket straight to
// the address the caller gives for it -- a slice of the owner GPU's peer-mapped receive buffer.
template <typename KeyT>
void mg_count_impl(mtg_ctx *ctx, uint32_t pb, uint64_t *bucket_counts, uint64_t *n_out) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    uint64_t n_ub = 0;
    for (auto &seg : ctx->segs) n_ub += seg.n_windows_max;
    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64, s));
    DevBuf a(sizeof(KeyT) * std::max<uint64_t>(n_ub, 1), s);
    ctx->launches = 0;
    run_extraction<KeyT>(ctx, false, false, a.as<KeyT>(), nullptr, d_ctr);
    const uint64_t n = read_counter(d_ctr, s);
    const uint32_t nb = 1u << pb;
    for (uint32_t i = 0; i < nb; ++i) bucket_counts[i] = 0;
    if (pb == 0 || n == 0) {
        bucket_counts[0] = n;
    } else {
        DevBuf hist(sizeof(unsigned long long) * RADIX, s);
        digit_histogram<KeyT>(a.as<KeyT>(), n, DigitPass{0, (int)(al.K * al.L - pb), (int)pb},
                              hist.as<unsigned long long>(), s, &ctx->launches);
        std::vector<unsigned long long> h(RADIX);
        MTG_CUDA(cudaMemcpyAsync(h.data(), hist.p, 8 * RADIX, cudaMemcpyDeviceToHost, s));
        MTG_CUDA(cudaStreamSynchronize(s));
        for (uint32_t i = 0; i < nb; ++i) bucket_counts[i] = h[i];
    }
    ctx->mg.keys = std::move(a);
    ctx->mg.n_input += n;      // summed over the rounds of a super-batched build (mtg_mg_accumulate)
    ctx->mg.n_routed = n;
    *n_out = n;
    ctx->segs.clear();
}

// one scatter kernel: key with digit dp == b -> array at owner_dst[bucket_owner[b]]
template <typename KeyT>
void route_buffer(mtg_ctx *ctx, const KeyT *keys, uint64_t n, DigitPass dp, uint32_t nb, const uint32_t *bucket_owner,
                  uint32_t n_owners, const uint64_t *owner_dst, bool stable) {
    cudaStream_t s = ctx->stream;
    if (!n) return;
    // one staging block: RADIX addresses followed by RADIX owner bytes
    std::vector<unsigned long long> h(RADIX + RADIX / 8, 0ull);
    unsigned char *ho = reinterpret_cast<unsigned char *>(h.data() + RADIX);
    for (uint32_t i = 0; i < n_owners; ++i) h[i] = owner_dst[i];
    for (uint32_t i = 0; i < nb; ++i) ho[i] = (unsigned char)bucket_owner[i];
    DevBuf route(8 * (RADIX + RADIX / 8), s), tc(sizeof(unsigned int), s);
    DevBuf status(sizeof(unsigned long long) * sort_status_words<KeyT>(n), s);
    MTG_CUDA(cudaMemcpyAsync(route.p, h.data(), 8 * (RADIX + RADIX / 8), cudaMemcpyHostToDevice, s));
    route_scatter<KeyT>(keys, n, dp, reinterpret_cast<const unsigned char *>(route.as<unsigned long long>() + RADIX),
                        route.as<unsigned long long>(), status.as<unsigned long long>(), tc.as<unsigned int>(), s,
                        &ctx->launches, stable);
    MTG_CUDA(cudaStreamSynchronize(s));   // `h` and the scratch outlive the kernel
}

template <typename KeyT>
void mg_route_impl(mtg_ctx *ctx, uint32_t pb, const uint32_t *bucket_owner, uint32_t n_owners,
                   const uint64_t *owner_dst) {
    const Alpha al = ctx->al;
    // pb = 0: one bucket, a 1-bit digit above the key bits (always 0) keeps the pass generic
    const DigitPass dp = pb ? DigitPass{0, (int)(al.K * al.L - pb), (int)pb}
                            : DigitPass{0, (int)(8 * sizeof(KeyT) - 1), 1};
    // the k-mers arrive unordered and are sorted by mtg_mg_collect: the scatter need not be stable
    route_buffer<KeyT>(ctx, ctx->mg.keys.as<KeyT>(), ctx->mg.n_routed, dp, 1u << pb, bucket_owner, n_owners, owner_dst,
                       false);
    ctx->mg.keys.release();
    ctx->mg.n_routed = 0;
}

// Exchange #3 the same way: target nodes of this range's reals, counted per routing bucket, then
// scattered by ONE kernel into the owners' buffers. The caller's "owners" are SEGMENTS (owner rank x
// leading character): within a segment the stable scatter keeps the order of the reals, i.e. the
// targets of one edge label, which are sorted -- every (source, segment) slice arrives as a sorted run.
template <typename KeyT>
void mg_targets_count_impl(mtg_ctx *ctx, uint32_t pb, uint64_t *bucket_counts, uint64_t *n_out) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint64_t m = ctx->mg.reals.m;
    DevBuf t0(sizeof(KeyT) * std::max<uint64_t>(m, 1), s);
    target_nodes<KeyT>(ctx->mg.reals.keys.as<KeyT>(), m, al, t0.as<KeyT>(), s, &ctx->launches);
    const uint32_t nb = 1u << pb;
    for (uint32_t i = 0; i < nb; ++i) bucket_counts[i] = 0;
    if (pb == 0 || m == 0) {
        bucket_counts[0] = m;
    } else {
        DevBuf hist(sizeof(unsigned long long) * RADIX, s);
        digit_histogram<KeyT>(t0.as<KeyT>(), m, DigitPass{0, (int)((al.K - 1) * al.L - pb), (int)pb},
                              hist.as<unsigned long long>(), s, &ctx->launches);
        std::vector<unsigned long long> h(RADIX);
        MTG_CUDA(cudaMemcpyAsync(h.data(), hist.p, 8 * RADIX, cudaMemcpyDeviceToHost, s));
        MTG_CUDA(cudaStreamSynchronize(s));
        for (uint32_t i = 0; i < nb; ++i) bucket_counts[i] = h[i];
    }
    ctx->mg.targets = std::move(t0);
    ctx->mg.n_routed_targets = m;
    *n_out = m;
}

template <typename KeyT>
void mg_targets_route_impl(mtg_ctx *ctx, uint32_t pb, const uint32_t *bucket_owner, uint32_t n_owners,
                           const uint64_t *owner_dst) {
    const Alpha al = ctx->al;
    const DigitPass dp = pb ? DigitPass{0, (int)((al.K - 1) * al.L - pb), (int)pb}
                            : DigitPass{0, (int)(8 * sizeof(KeyT) - 1), 1};
    // stable: every (source, segment) slice must arrive as the sorted run it is
    route_buffer<KeyT>(ctx, ctx->mg.targets.as<KeyT>(), ctx->mg.n_routed_targets, dp, 1u << pb, bucket_owner, n_owners,
                       owner_dst, true);
    ctx->mg.targets.release();
    ctx->mg.n_routed_targets = 0;
}

template <typename KeyT>
void mg_targets_impl(mtg_ctx *ctx, uint32_t pb, uint64_t *bucket_counts, void **d_targets, uint64_t *n_out) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint32_t K = al.K, L = al.L;
    const uint64_t m = ctx->mg.reals.m;
    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    DevBuf t0(sizeof(KeyT) * std::max<uint64_t>(m, 1), s), t1(sizeof(KeyT) * std::max<uint64_t>(m, 1), s);
    target_nodes<KeyT>(ctx->mg.reals.keys.as<KeyT>(), m, al, t0.as<KeyT>(), s, &ctx->launches);
    {   // per edge label the targets are already sorted: one stable pass on the leading char
        SortPlan plan;
        plan.n = 0;
        plan.p[plan.n++] = DigitPass{0, (int)((K - 2) * L), (int)L};
        sort_keys<KeyT>(t0, t1, nullptr, nullptr, m, plan, s, &ctx->launches);
    }
    {
        DevBuf status(sizeof(unsigned long long) * unique_status_words(m), s), tctr(sizeof(unsigned int), s);
        unique_keys<KeyT>(t0.as<KeyT>(), m, t1.as<KeyT>(), nullptr, d_ctr, status.as<unsigned long long>(),
                          tctr.as<unsigned int>(), s, &ctx->launches);
    }
    const uint64_t nt = read_counter(d_ctr, s);
    const uint32_t nb = 1u << pb;
    std::vector<uint64_t> bounds(nb + 1, 0);
    if (pb == 0) {
        bounds[1] = nt;
    } else {
        DevBuf db(8 * (nb + 1), s);
        bucket_bounds<KeyT>(t1.as<KeyT>(), nt, pb, (K - 1) * L - pb, db.as<uint64_t>(), s, &ctx->launches);
        MTG_CUDA(cudaMemcpyAsync(bounds.data(), db.p, 8 * (nb + 1), cudaMemcpyDeviceToHost, s));
        MTG_CUDA(cudaStreamSynchronize(s));
    }
    for (uint32_t i = 0; i < nb; ++i) bucket_counts[i] = bounds[i + 1] - bounds[i];
    ctx->mg.targets = std::move(t1);
    *d_targets = ctx->mg.targets.p;
    *n_out = nt;
}

// `a` holds consecutive sorted runs: pairwise merge-path rounds (ceil(log2(#runs)) passes over the
// data instead of a full radix sort). On return `a` holds the sorted array.
template <typename KeyT>
// The runs start in `a` (or, with ext_a, in the caller's receive buffer: the first level of the
// merge tree then reads them from there and no staging copy is made); the result ends in `a`.
// group > 0: every `group` consecutive runs form a set whose keys all sort before those of the next
// set (targets of one leading character): only the runs inside a set are merged.
void merge_runs(mtg_ctx *ctx, DevBuf &a, DevBuf &b, std::vector<uint64_t> runs, DevBuf *ca = nullptr,
                DevBuf *cb = nullptr, const void *ext_a = nullptr, const uint32_t *ext_ca = nullptr,
                uint32_t group = 0) {
    cudaStream_t s = ctx->stream;
    uint64_t total = 0;
    for (uint64_t r : runs) total += r;
    if (group == 0 || group > runs.size()) group = (uint32_t)std::max<size_t>(runs.size(), 1);
    std::vector<std::vector<uint64_t>> sets;
    for (size_t i = 0; i < runs.size(); i += group)
        sets.emplace_back(runs.begin() + i, runs.begin() + std::min(runs.size(), i + group));
    auto unmerged = [&]() {
        for (auto &g : sets)
            if (g.size() > 1) return true;
        return false;
    };
    if (!unmerged()) {
        if (ext_a) {
            MTG_CUDA(cudaMemcpyAsync(a.p, ext_a, sizeof(KeyT) * total, cudaMemcpyDeviceToDevice, s));
            if (ca) MTG_CUDA(cudaMemcpyAsync(ca->p, ext_ca, 4 * total, cudaMemcpyDeviceToDevice, s));
        }
        return;
    }
    DevBuf part(8 * merge_partition_words(total), s);
    bool from_ext = ext_a != nullptr;
    while (unmerged()) {
        uint64_t off = 0;
        const KeyT *ik = from_ext ? static_cast<const KeyT *>(ext_a) : a.as<KeyT>();
        const uint32_t *ic = !ca ? nullptr : from_ext ? ext_ca : ca->as<uint32_t>();
        KeyT *ok = from_ext ? a.as<KeyT>() : b.as<KeyT>();
        uint32_t *oc = !ca ? nullptr : from_ext ? ca->as<uint32_t>() : cb->as<uint32_t>();
        for (auto &g : sets) {
            std::vector<uint64_t> next;
            for (size_t i = 0; i < g.size(); i += 2) {
                const uint64_t la = g[i], lb = i + 1 < g.size() ? g[i + 1] : 0;
                if (lb == 0 || la == 0) {
                    if (la + lb) {
                        MTG_CUDA(cudaMemcpyAsync(ok + off, ik + off, sizeof(KeyT) * (la + lb), cudaMemcpyDeviceToDevice, s));
                        if (ca) MTG_CUDA(cudaMemcpyAsync(oc + off, ic + off, 4 * (la + lb), cudaMemcpyDeviceToDevice, s));
                    }
                } else {
                    merge_sorted<KeyT>(ik + off, la, ik + off + la, lb, ca ? ic + off : nullptr,
                                       ca ? ic + off + la : nullptr, ok + off, ca ? oc + off : nullptr,
                                       part.as<uint64_t>(), s, &ctx->launches);
                }
                next.push_back(la + lb);
                off += la + lb;
            }
            g.swap(next);
        }
        if (!from_ext) {
            std::swap(a, b);
            if (ca) std::swap(*ca, *cb);
        }
        from_ext = false;
    }
}

template <typename KeyT>
void mg_dummies_impl(mtg_ctx *ctx, const void *d_in, uint64_t n_in, const uint64_t *run_lengths, uint32_t n_runs,
                     uint32_t runs_per_set, void **d_src, uint64_t *n_src, uint64_t *n_sink_out) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint32_t K = al.K, L = al.L;
    const uint64_t m = ctx->mg.reals.m;
    const KeyT *reals = ctx->mg.reals.keys.as<KeyT>();
    ctx->mg.targets.release();
    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    // targets received from every range: sort, fold -> T_i
    DevBuf t0(sizeof(KeyT) * std::max<uint64_t>(n_in, 1), s), t1(sizeof(KeyT) * std::max<uint64_t>(n_in, 1), s);
    {
        uint64_t run_total = 0;
        for (uint32_t i = 0; i < n_runs; ++i) run_total += run_lengths[i];
        if (run_lengths && n_runs && run_total == n_in) {
            // every source sends its targets sorted: merge the runs (read from the receive buffer)
            merge_runs<KeyT>(ctx, t0, t1, std::vector<uint64_t>(run_lengths, run_lengths + n_runs), nullptr, nullptr,
                             d_in, nullptr, runs_per_set);
        } else {
            SortPlan plan;
            plan.n = 0;
            plan_add_key_bits(plan, 0, (K - 1) * L);
            sort_keys<KeyT>(t0, t1, nullptr, nullptr, n_in, plan, s, &ctx->launches, nullptr, d_in);
        }
        DevBuf status(sizeof(unsigned long long) * unique_status_words(n_in), s), tctr(sizeof(unsigned int), s);
        unique_keys<KeyT>(t0.as<KeyT>(), n_in, t1.as<KeyT>(), nullptr, d_ctr, status.as<unsigned long long>(),
                          tctr.as<unsigned int>(), s, &ctx->launches);
    }
    const uint64_t nt = read_counter(d_ctr, s);
    t0.release();
    DevBuf scratch(ranged_scratch_bytes(m, nt), s);
    uint64_t cap = std::max<uint64_t>(1024, std::max(m, nt) / 8);
    DevBuf sink, src;
    uint64_t n_sink = 0, d1 = 0;
    while (true) {
        sink.alloc(sizeof(KeyT) * cap, s);
        src.alloc(sizeof(KeyT) * cap, s);
        ranged_dummy_candidates<KeyT>(reals, m, t1.as<KeyT>(), nt, al, scratch.p, sink.as<KeyT>(), cap, d_ctr,
                                      src.as<KeyT>(), cap, d_ctr + 1, s, &ctx->launches);
        n_sink = read_counter(d_ctr, s);
        d1 = read_counter(d_ctr + 1, s);
        if (n_sink <= cap && d1 <= cap) break;
        cap = std::max(n_sink, d1);
    }
    ctx->mg.sinks = std::move(sink);
    ctx->mg.src = std::move(src);
    ctx->mg.n_sink = n_sink;
    ctx->mg.n_src = d1;
    *d_src = ctx->mg.src.p;
    *n_src = d1;
    *n_sink_out = n_sink;
}

}  // namespace

namespace {

// Canonical mode, after mg_collect: X = distinct canonical k-mers of this rank's stage-1 range
// (sorted; palindrome counts doubled), R = reverse complements of the non-palindromes grouped by
// routing bucket. Both are routed to the owners of the FINAL ranges (planned on x + r counts).
template <typename KeyT>
void mg_strands_impl(mtg_ctx *ctx, uint32_t pb, uint64_t *x_counts, uint64_t *r_counts, void **d_x, void **d_xc,
                     uint64_t *n_x, void **d_r, void **d_rc, uint64_t *n_r) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint32_t KL = al.K * al.L, nb = 1u << pb;
    const bool counting = ctx->prm.bits_per_count > 0;
    RealSet &rs = ctx->mg.reals;
    const uint64_t m = rs.m;
    DevBuf ctr(64, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    DevBuf rb0(sizeof(KeyT) * std::max<uint64_t>(m, 1), s), rb1(sizeof(KeyT) * std::max<uint64_t>(m, 1), s);
    DevBuf cb0, cb1;
    if (counting) { cb0.alloc(4 * std::max<uint64_t>(m, 1), s); cb1.alloc(4 * std::max<uint64_t>(m, 1), s); }
    append_reverse_complements<KeyT>(rs.keys.as<KeyT>(), rb0.as<KeyT>(), counting ? rs.counts.as<uint32_t>() : nullptr,
                                     counting ? cb0.as<uint32_t>() : nullptr, m, al, ctx->cmax, d_ctr, s,
                                     &ctx->launches);
    const uint64_t mb = read_counter(d_ctr, s);
    for (uint32_t i = 0; i < nb; ++i) x_counts[i] = r_counts[i] = 0;
    if (pb == 0) {
        x_counts[0] = m;
        r_counts[0] = mb;
        SortPlan plan;
        plan.n = 0;
        plan_add_key_bits(plan, 0, KL);
        sort_keys<KeyT>(rb0, rb1, counting ? &cb0 : nullptr, counting ? &cb1 : nullptr, mb, plan, s, &ctx->launches);
    } else {
        if (mb) {
            // R is sorted HERE, where the work is balanced (every stage-1 owner holds about the
            // same number of canonical k-mers); the reverse complements concentrate in the high
            // final ranges, whose owners then only merge sorted runs
            SortPlan plan;
            plan.n = 0;
            plan_add_key_bits(plan, 0, KL);
            sort_keys<KeyT>(rb0, rb1, counting ? &cb0 : nullptr, counting ? &cb1 : nullptr, mb, plan, s,
                            &ctx->launches);
            std::vector<uint64_t> rbounds(nb + 1, 0);
            DevBuf rdb(8 * (nb + 1), s);
            bucket_bounds<KeyT>(rb0.as<KeyT>(), mb, pb, KL - pb, rdb.as<uint64_t>(), s, &ctx->launches);
            MTG_CUDA(cudaMemcpyAsync(rbounds.data(), rdb.p, 8 * (nb + 1), cudaMemcpyDeviceToHost, s));
            MTG_CUDA(cudaStreamSynchronize(s));
            for (uint32_t i = 0; i < nb; ++i) r_counts[i] = rbounds[i + 1] - rbounds[i];
        }
        std::vector<uint64_t> bounds(nb + 1, 0);
        DevBuf db(8 * (nb + 1), s);
        bucket_bounds<KeyT>(rs.keys.as<KeyT>(), m, pb, KL - pb, db.as<uint64_t>(), s, &ctx->launches);
        MTG_CUDA(cudaMemcpyAsync(bounds.data(), db.p, 8 * (nb + 1), cudaMemcpyDeviceToHost, s));
        MTG_CUDA(cudaStreamSynchronize(s));
        for (uint32_t i = 0; i < nb; ++i) x_counts[i] = bounds[i + 1] - bounds[i];
    }
    ctx->mg.rkeys = std::move(rb0);
    ctx->mg.rcounts = std::move(cb0);
    *d_x = rs.keys.p;
    *d_xc = counting ? rs.counts.p : nullptr;
    *n_x = m;
    *d_r = ctx->mg.rkeys.p;
    *d_rc = counting ? ctx->mg.rcounts.p : nullptr;
    *n_r = mb;
}

// Final-range owner: X received from all ranks in rank order (= sorted as a whole, the stage-1
// ranges are increasing key ranges), R in any order. Sorts R and merges -> the sorted distinct
// real k-mers of this range, both strands (what add_reverse_complements leaves behind).
template <typename KeyT>
void mg_merge_impl(mtg_ctx *ctx, const void *d_x, const uint32_t *d_xc, uint64_t nx, const void *d_r,
                   const uint32_t *d_rc, uint64_t nr, const uint64_t *r_runs, uint32_t n_runs, uint64_t *n_real) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const bool counting = ctx->prm.bits_per_count > 0;
    DevBuf rb0(sizeof(KeyT) * std::max<uint64_t>(nr, 1), s), rb1(sizeof(KeyT) * std::max<uint64_t>(nr, 1), s);
    DevBuf cb0, cb1;
    if (counting) {
        cb0.alloc(4 * std::max<uint64_t>(nr, 1), s);
        cb1.alloc(4 * std::max<uint64_t>(nr, 1), s);
    }
    uint64_t run_total = 0;
    for (uint32_t i = 0; i < n_runs; ++i) run_total += r_runs[i];
    if (r_runs && n_runs && run_total == nr) {   // one sorted run per source rank, read from the receive buffer
        merge_runs<KeyT>(ctx, rb0, rb1, std::vector<uint64_t>(r_runs, r_runs + n_runs), counting ? &cb0 : nullptr,
                         counting ? &cb1 : nullptr, d_r, d_rc);
    } else {
        MTG_CUDA(cudaMemcpyAsync(rb0.p, d_r, sizeof(KeyT) * nr, cudaMemcpyDeviceToDevice, s));
        if (counting) MTG_CUDA(cudaMemcpyAsync(cb0.p, d_rc, 4 * nr, cudaMemcpyDeviceToDevice, s));
        SortPlan plan;
        plan.n = 0;
        plan_add_key_bits(plan, 0, al.K * al.L);
        sort_keys<KeyT>(rb0, rb1, counting ? &cb0 : nullptr, counting ? &cb1 : nullptr, nr, plan, s, &ctx->launches);
    }
    rb1.release(); cb1.release();
    const uint64_t m2 = nx + nr;
    DevBuf merged(sizeof(KeyT) * std::max<uint64_t>(m2, 1) + 16, s), merged_counts;
    if (counting) merged_counts.alloc(4 * std::max<uint64_t>(m2, 1), s);
    {
        DevBuf part(8 * merge_partition_words(m2), s);
        merge_sorted<KeyT>((const KeyT *)d_x, nx, rb0.as<KeyT>(), nr, counting ? d_xc : nullptr,
                           counting ? cb0.as<uint32_t>() : nullptr, merged.as<KeyT>(),
                           counting ? merged_counts.as<uint32_t>() : nullptr, part.as<uint64_t>(), s, &ctx->launches);
        MTG_CUDA(cudaStreamSynchronize(s));   // d_x / d_r belong to the caller
    }
    const uint64_t n_input = ctx->mg.reals.n_input;
    ctx->mg.rkeys.release();
    ctx->mg.rcounts.release();
    ctx->mg.reals.keys = std::move(merged);
    ctx->mg.reals.counts = std::move(merged_counts);
    ctx->mg.reals.m = m2;
    ctx->mg.reals.n_input = n_input;
    *n_real = m2;
}

template <typename KeyT>
void mg_build_impl(mtg_ctx *ctx, const void *d_all_src, uint64_t n_all, uint64_t lo_bucket, uint64_t hi_bucket,
                   uint32_t pb, bool first, mtg_chunk *out, bool device_output) {
    cudaStream_t s = ctx->stream;
    const Alpha al = ctx->al;
    const uint32_t KL = al.K * al.L;
    Timer tm(s);
    SortProfile prof;
    for (int i = 0; i < 5; ++i) tm.mark();   // phases 0..4 ran in the earlier mg calls
    RealSet rs = std::move(ctx->mg.reals);
    DevBuf ctr(64 + 8 * MTG_MAX_K, s);
    unsigned long long *d_ctr = ctr.as<unsigned long long>();
    MTG_CUDA(cudaMemsetAsync(d_ctr, 0, 64 + 8 * MTG_MAX_K, s));
    // every range expands the global list of source-1 nodes (small) and keeps its own share
    DevBuf src(sizeof(KeyT) * std::max<uint64_t>(n_all, 1), s);
    MTG_CUDA(cudaMemcpyAsync(src.p, d_all_src, sizeof(KeyT) * n_all, cudaMemcpyDeviceToDevice, s));
    const uint64_t n_sink = ctx->mg.n_sink;
    // every range expands the (small, global) list of source-1 nodes but materialises only the
    // dummy sources of its own key range
    const uint64_t nb = 1ull << pb;
    const bool has_hi = pb > 0 && hi_bucket < nb;
    KeyT lo = pb ? w_shl(WordTraits<KeyT>::from_u64(lo_bucket), KL - pb) : WordTraits<KeyT>::zero();
    KeyT hi = has_hi ? w_shl(WordTraits<KeyT>::from_u64(hi_bucket), KL - pb) : WordTraits<KeyT>::zero();
    uint64_t d_mine = 0;
    DevBuf dk0, di0;
    expand_sources<KeyT>(ctx, src, n_all, ctx->mg.sinks, n_sink, d_ctr, dk0, di0, &d_mine, &lo,
                         has_hi ? &hi : nullptr);
    const uint64_t kept = d_mine - n_sink;
    DevBuf bidx;
    const bool use_index = rs.m >= (1u << 16);
    if (use_index) {
        bidx.alloc(8 * bucket_index_words(al), s);
        build_bucket_index<KeyT>(rs.keys.as<KeyT>(), rs.m, al, bidx.as<uint64_t>(), s, &ctx->launches);
    }
    const uint64_t n_input = ctx->mg.n_input;
    finish_chunk<KeyT>(ctx, rs, dk0, di0, n_sink + kept, n_sink, d_ctr, use_index ? bidx.as<uint64_t>() : nullptr,
                       tm, prof, out, device_output, first ? 2 : 1);
    out->n_input_kmers = n_input;
    ctx->mg = mtg_ctx::Mg();
}

uint32_t mg_part_bits(const mtg_ctx *ctx) {
    uint32_t lim = (ctx->al.K - 2) * ctx->al.L;
    return lim < 8 ? lim : 8;
}

}  // namespace

extern "C" {

#define MG_GUARD(ctx, body)                                                  \
    if (!ctx) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");           \
    std::lock_guard<std::mutex> lk(ctx->mu);                                 \
    try {                                                                    \
        MTG_CUDA(cudaSetDevice(ctx->device));                                \
        t_arena = ctx->arena;                                                \
        const bool wide = ctx->al.K * ctx->al.L > 64;                        \
        (void)wide;                                                          \
        body                                                                 \
    } catch (const CudaError &e) {                                           \
        int rc = fail_cuda(ctx, e);                                          \
        cudaStreamSynchronize(ctx->stream);                                  \
        return rc;                                                           \
    }                                                                        \
    return MTG_OK;

uint32_t mtg_mg_part_bits(const mtg_ctx *ctx) { return ctx ? mg_part_bits(ctx) : 0; }

int mtg_mg_extract(mtg_ctx *ctx, uint64_t *bucket_counts, void **d_keys, uint64_t *n) {
    MG_GUARD(ctx, {
        if (!ctx->runs.empty())
            return fail(ctx, MTG_ERR_STATE, "create the context with MTG_FLAG_NO_BATCHING for the mtg_mg_* phases");
        if (ctx->any_explicit_counts)
            return fail(ctx, MTG_ERR_UNSUPPORTED, "per-read counts are not supported by the range-partitioned build");
        if (!ctx->suffix.empty())
            return fail(ctx, MTG_ERR_UNSUPPORTED, "a suffix-filtered chunk is built on one GPU (mtg_build_chunk)");
        if (wide) mg_extract_impl<u128>(ctx, mg_part_bits(ctx), bucket_counts, d_keys, n);
        else mg_extract_impl<uint64_t>(ctx, mg_part_bits(ctx), bucket_counts, d_keys, n);
    })
}

int mtg_mg_count(mtg_ctx *ctx, uint64_t *bucket_counts, uint64_t *n) {
    MG_GUARD(ctx, {
        if (!ctx->runs.empty())
            return fail(ctx, MTG_ERR_STATE, "create the context with MTG_FLAG_NO_BATCHING for the mtg_mg_* phases");
        if (ctx->any_explicit_counts)
            return fail(ctx, MTG_ERR_UNSUPPORTED, "per-read counts are not supported by the range-partitioned build");
        if (!ctx->suffix.empty())
            return fail(ctx, MTG_ERR_UNSUPPORTED, "a suffix-filtered chunk is built on one GPU (mtg_build_chunk)");
        if (!bucket_counts || !n) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
        if (wide) mg_count_impl<u128>(ctx, mg_part_bits(ctx), bucket_counts, n);
        else mg_count_impl<uint64_t>(ctx, mg_part_bits(ctx), bucket_counts, n);
    })
}

int mtg_mg_route(mtg_ctx *ctx, const uint32_t *bucket_owner, uint32_t n_owners, const uint64_t *owner_dst) {
    MG_GUARD(ctx, {
        if (!bucket_owner || !owner_dst || n_owners == 0 || n_owners > 256)
            return fail(ctx, MTG_ERR_BAD_ARG, "bad owner table");
        const uint32_t w = wide ? 16 : 8;
        for (uint32_t i = 0; i < n_owners; ++i)
            if (owner_dst[i] % w) return fail(ctx, MTG_ERR_BAD_ARG, "destinations must be key-aligned");
        for (uint32_t i = 0; i < (1u << mg_part_bits(ctx)); ++i)
            if (bucket_owner[i] >= n_owners) return fail(ctx, MTG_ERR_BAD_ARG, "bucket owner out of range");
        if (wide) mg_route_impl<u128>(ctx, mg_part_bits(ctx), bucket_owner, n_owners, owner_dst);
        else mg_route_impl<uint64_t>(ctx, mg_part_bits(ctx), bucket_owner, n_owners, owner_dst);
    })
}

int mtg_mg_collect(mtg_ctx *ctx, const void *d_keys, uint64_t n, uint64_t *n_real) {
    MG_GUARD(ctx, {
        if (wide) collect_reals<u128>(ctx, &ctx->mg.reals, nullptr, nullptr, d_keys ? d_keys : ctx->mg.keys.p, n);
        else collect_reals<uint64_t>(ctx, &ctx->mg.reals, nullptr, nullptr, d_keys ? d_keys : ctx->mg.keys.p, n);
        ctx->mg.keys.release();
        *n_real = ctx->mg.reals.m;
    })
}

int mtg_mg_strands(mtg_ctx *ctx, uint64_t *x_bucket_counts, uint64_t *r_bucket_counts, void **d_x, void **d_xc,
                   uint64_t *n_x, void **d_r, void **d_rc, uint64_t *n_r) {
    MG_GUARD(ctx, {
        if (!ctx->prm.both_strands) return fail(ctx, MTG_ERR_STATE, "mtg_mg_strands is a canonical-mode phase");
        if (wide) mg_strands_impl<u128>(ctx, mg_part_bits(ctx), x_bucket_counts, r_bucket_counts, d_x, d_xc, n_x, d_r, d_rc, n_r);
        else mg_strands_impl<uint64_t>(ctx, mg_part_bits(ctx), x_bucket_counts, r_bucket_counts, d_x, d_xc, n_x, d_r, d_rc, n_r);
    })
}

int mtg_mg_merge(mtg_ctx *ctx, const void *d_x, const uint32_t *d_xc, uint64_t n_x, const void *d_r,
                 const uint32_t *d_rc, uint64_t n_r, const uint64_t *r_run_lengths, uint32_t n_runs,
                 uint64_t *n_real) {
    MG_GUARD(ctx, {
        if (wide) mg_merge_impl<u128>(ctx, d_x, d_xc, n_x, d_r, d_rc, n_r, r_run_lengths, n_runs, n_real);
        else mg_merge_impl<uint64_t>(ctx, d_x, d_xc, n_x, d_r, d_rc, n_r, r_run_lengths, n_runs, n_real);
    })
}

int mtg_mg_targets(mtg_ctx *ctx, uint64_t *bucket_counts, void **d_targets, uint64_t *n) {
    MG_GUARD(ctx, {
        if (wide) mg_targets_impl<u128>(ctx, mg_part_bits(ctx), bucket_counts, d_targets, n);
        else mg_targets_impl<uint64_t>(ctx, mg_part_bits(ctx), bucket_counts, d_targets, n);
    })
}

static int check_route_args(mtg_ctx *ctx, bool wide, const uint32_t *bucket_owner, uint32_t n_owners,
                            const uint64_t *owner_dst) {
    if (!bucket_owner || !owner_dst || n_owners == 0 || n_owners > 256) return fail(ctx, MTG_ERR_BAD_ARG, "bad owner table");
    const uint32_t w = wide ? 16 : 8;
    for (uint32_t i = 0; i < n_owners; ++i)
        if (owner_dst[i] % w) return fail(ctx, MTG_ERR_BAD_ARG, "destinations must be key-aligned");
    for (uint32_t i = 0; i < (1u << mg_part_bits(ctx)); ++i)
        if (bucket_owner[i] >= n_owners) return fail(ctx, MTG_ERR_BAD_ARG, "bucket owner out of range");
    return MTG_OK;
}

int mtg_mg_targets_count(mtg_ctx *ctx, uint64_t *bucket_counts, uint64_t *n) {
    MG_GUARD(ctx, {
        if (!bucket_counts || !n) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
        if (wide) mg_targets_count_impl<u128>(ctx, mg_part_bits(ctx), bucket_counts, n);
        else mg_targets_count_impl<uint64_t>(ctx, mg_part_bits(ctx), bucket_counts, n);
    })
}

int mtg_mg_targets_route(mtg_ctx *ctx, const uint32_t *bucket_owner, uint32_t n_owners, const uint64_t *owner_dst) {
    MG_GUARD(ctx, {
        if (int rc = check_route_args(ctx, wide, bucket_owner, n_owners, owner_dst)) return rc;
        if (wide) mg_targets_route_impl<u128>(ctx, mg_part_bits(ctx), bucket_owner, n_owners, owner_dst);
        else mg_targets_route_impl<uint64_t>(ctx, mg_part_bits(ctx), bucket_owner, n_owners, owner_dst);
    })
}

/* Super-batched range build: folds the result of the last mtg_mg_collect into the k-mers accumulated
 * by the earlier rounds of this range (merge + saturating count sums: SortedSet::shrink_data semantics).
 * last != 0: the accumulated set becomes the range's k-mer set again and the build goes on
 * (mtg_mg_strands / mtg_mg_targets...). */
int mtg_mg_accumulate(mtg_ctx *ctx, int last, uint64_t *n_real) {
    MG_GUARD(ctx, {
        if (!ctx->mg.has_acc) {
            ctx->mg.acc = std::move(ctx->mg.reals);
            ctx->mg.has_acc = true;
        } else {
            RealSet merged;
            if (wide) merge_fold_runs<u128>(ctx, ctx->mg.acc, ctx->mg.reals, &merged);
            else merge_fold_runs<uint64_t>(ctx, ctx->mg.acc, ctx->mg.reals, &merged);
            ctx->mg.acc = std::move(merged);
        }
        ctx->mg.reals = RealSet();
        if (n_real) *n_real = ctx->mg.acc.m;
        if (last) {
            ctx->mg.reals = std::move(ctx->mg.acc);
            ctx->mg.acc = RealSet();
            ctx->mg.has_acc = false;
        }
    })
}

/* verification hook of the fused exchanges: the k-mers (which = 0, after mtg_mg_count) or target
 * nodes (which = 1, after mtg_mg_targets_count) that the next route call will scatter */
int mtg_mg_debug_pending(mtg_ctx *ctx, int which, void **d_keys, uint64_t *n) {
    MG_GUARD(ctx, {
        if (!d_keys || !n) return fail(ctx, MTG_ERR_BAD_ARG, "null argument");
        *d_keys = which ? ctx->mg.targets.p : ctx->mg.keys.p;
        *n = which ? ctx->mg.n_routed_targets : ctx->mg.n_routed;
    })
}

int mtg_mg_dummies(mtg_ctx *ctx, const void *d_targets_in, uint64_t n_in, const uint64_t *run_lengths,
                   uint32_t n_runs, uint32_t runs_per_set, void **d_src_nodes, uint64_t *n_src, uint64_t *n_sink) {
    MG_GUARD(ctx, {
        if (wide) mg_dummies_impl<u128>(ctx, d_targets_in, n_in, run_lengths, n_runs, runs_per_set, d_src_nodes, n_src, n_sink);
        else mg_dummies_impl<uint64_t>(ctx, d_targets_in, n_in, run_lengths, n_runs, runs_per_set, d_src_nodes, n_src, n_sink);
    })
}

int mtg_mg_build(mtg_ctx *ctx, const void *d_all_src_nodes, uint64_t n_all, uint64_t lo_bucket,
                 uint64_t hi_bucket, int first_range, int device_output, mtg_chunk *out) {
    MG_GUARD(ctx, {
        if (wide) mg_build_impl<u128>(ctx, d_all_src_nodes, n_all, lo_bucket, hi_bucket, mg_part_bits(ctx),
                                      first_range != 0, out, device_output != 0);
        else mg_build_impl<uint64_t>(ctx, d_all_src_nodes, n_all, lo_bucket, hi_bucket, mg_part_bits(ctx),
                                     first_range != 0, out, device_output != 0);
    })
}

}  // extern "C"
