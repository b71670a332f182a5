// The reference-side binding of the CUDA path: the file a MetaGraph maintainer adds next to
// metagraph/src/graph/representation/succinct/boss_chunk_construct.cpp. It is compiled INSIDE the
// reference tree (it includes the reference's headers) and talks to the device only through the
// C ABI include/mtg_gpu.h (libmtg_gpu.so).
//
//   class BOSSChunkConstructorGPU : public IBOSSChunkConstructor
//       replaces BOSSChunkConstructor<KmerCollector> (boss_chunk_construct.cpp:948-1048) behind the
//       unchanged factory IBOSSChunkConstructor::initialize (boss_chunk_construct.hpp:22-31); the
//       factory gets ONE added line: it asks mtg_gpu_make_constructor() first.
//
// BOSS::Chunk has no public way to be filled from arrays (its constructors build from k-mers,
// boss_chunk.hpp:24-57, W_/last_/F_ are private), so the chunk crosses over the way
// BOSS::Chunk::build_boss_from_chunks moves chunks (boss_chunk.cpp:306-311): the four files of
// BOSS::Chunk::serialize (boss_chunk.cpp:372-386) are written from the library's pinned result
// buffers into the swap directory and read back with BOSS::Chunk::load (:330-370).
//
// Selection: MTG_GPU=1 in the environment (any other value: the reference's CPU constructors).
// Tight keys wider than 128 bits ((k+1)*L1 > 128: DNA k > 63, Protein k > 24) are outside the
// library (MTG_ERR_UNSUPPORTED): the factory keeps the reference's CPU constructor for them and
// says so once on stderr.
#include <atomic>
#include <cstdlib>
#include <cstring>
#include <filesystem>
#include <fstream>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "graph/representation/succinct/boss_chunk_construct.hpp"
#include "graph/representation/succinct/boss_chunk.hpp"
#include "kmer/kmer_extractor.hpp"
#include "common/logger.hpp"
#include "common/serialization.hpp"
#include "common/utils/file_utils.hpp"

#include "mtg_gpu.h"

namespace mtg {
namespace graph {
namespace boss {

using mtg::common::logger;

namespace {

std::atomic<uint64_t> g_gpu_chunks{0}, g_cpu_fallbacks{0};

struct Report {
    ~Report() {
        if (std::getenv("MTG_GPU_REPORT"))
            fprintf(stderr, "[mtg_gpu] chunks built on the GPU: %llu, constructors left to the CPU path "
                            "(keys wider than 128 bits): %llu\n",
                    (unsigned long long)g_gpu_chunks.load(), (unsigned long long)g_cpu_fallbacks.load());
    }
} g_report;

void write_int_vector(const std::string &path, const uint64_t *words, uint64_t n, uint8_t width, bool fixed_width) {
    // sdsl::int_vector<t_width>::serialize: u64 bit length, [u8 width if t_width == 0], 64-bit words
    std::ofstream out(path, std::ios::binary);
    const uint64_t bits = n * width;
    out.write(reinterpret_cast<const char *>(&bits), 8);
    if (!fixed_width)
        out.write(reinterpret_cast<const char *>(&width), 1);
    out.write(reinterpret_cast<const char *>(words), 8 * ((bits + 63) / 64));
    if (!out.good()) {
        logger->error("Cannot write {}", path);
        exit(1);
    }
}

class BOSSChunkConstructorGPU : public IBOSSChunkConstructor {
  public:
    BOSSChunkConstructorGPU(mtg_ctx *ctx, size_t k, const std::filesystem::path &swap_dir)
          : ctx_(ctx), k_(k), swap_dir_(swap_dir) {
        offsets_.push_back(0);
    }

    ~BOSSChunkConstructorGPU() { mtg_destroy(ctx_); }

    // single reads are batched like KmerCollector's BatchAccumulator (kmer_collector.hpp:69-79,
    // kLargeBufferSize kmer_collector.cpp:22) and go up in one push per batch
    void add_sequence(std::string_view sequence, uint64_t count) override {
        if (sequence.size() < k_ + 1)
            return;
        std::lock_guard<std::mutex> lock(mu_);
        bases_.append(sequence.data(), sequence.size());
        offsets_.push_back(bases_.size());
        counts_.push_back(count);
        any_count_ |= count != 1;
        if (bases_.size() >= kBatchBytes)
            flush();
    }

    void add_sequences(std::vector<std::string>&& sequences) override {
        push(sequences, [](const std::string &s) -> const std::string & { return s; },
             [](const std::string &) { return uint64_t(1); }, false);
    }

    void add_sequences(std::vector<std::pair<std::string, uint64_t>>&& sequences) override {
        push(sequences, [](const auto &p) -> const std::string & { return p.first; },
             [](const auto &p) { return p.second; }, true);
    }

    BOSS::Chunk build_chunk() override {
        {
            std::lock_guard<std::mutex> lock(mu_);
            flush();
        }
        mtg_chunk c;
        check(mtg_build_chunk(ctx_, &c));
        // the files of BOSS::Chunk::serialize, then BOSS::Chunk::load
        const std::string dir = utils::create_temp_dir(swap_dir_, "gpu_chunk");
        const std::string base = dir + "/chunk" + BOSS::Chunk::kFileExtension;
        write_int_vector(base + ".W", c.W, c.n_rows, c.W_width, false);
        write_int_vector(base + ".last", c.last, c.n_rows, 1, true);
        if (c.weights_width) {
            write_int_vector(base + ".weights", c.weights, c.n_rows, c.weights_width, false);
        } else {
            const uint64_t none = 0;
            write_int_vector(base + ".weights", &none, 0, 64, false);
        }
        {
            std::ofstream out(base, std::ios::binary);
            serialize_number_vector(out, std::vector<uint64_t>(c.F, c.F + c.sigma));
            serialize_number(out, c.sigma);
            serialize_number(out, c.k);
        }
        mtg_free_chunk(&c);
        BOSS::Chunk result(1, 0, swap_dir_.string());
        if (!result.load(base)) {
            logger->error("Cannot load the chunk built on the GPU from {}", base);
            exit(1);
        }
        utils::remove_temp_dir(dir);
        g_gpu_chunks++;
        return result;
    }

    uint64_t get_k() const override { return k_; }

  private:
    static constexpr size_t kBatchBytes = 1 << 22;

    void check(int rc) {
        if (rc == MTG_OK)
            return;
        logger->error("{}", mtg_last_error(ctx_));   // logger->error + exit(1), like the reference
        exit(1);
    }

    // caller holds mu_
    void flush() {
        if (offsets_.size() > 1)
            check(mtg_push_reads(ctx_, bases_.data(), offsets_.data(), any_count_ ? counts_.data() : nullptr,
                                 offsets_.size() - 1));
        bases_.clear();
        offsets_.assign(1, 0);
        counts_.clear();
        any_count_ = false;
    }

    // vector overloads: called concurrently (cli/build.cpp:35-52); the batch is flattened in the
    // calling thread and pushed through the thread-safe mtg_push_reads
    template <class Vec, class Seq, class Cnt>
    void push(const Vec &v, Seq seq, Cnt cnt, bool with_counts) {
        size_t total = 0;
        for (const auto &x : v) total += seq(x).size();
        std::string bases;
        bases.reserve(total);
        std::vector<uint64_t> offsets(1, 0), counts;
        offsets.reserve(v.size() + 1);
        for (const auto &x : v) {
            const std::string &s = seq(x);
            if (s.size() < k_ + 1)
                continue;
            bases += s;
            offsets.push_back(bases.size());
            if (with_counts)
                counts.push_back(cnt(x));
        }
        if (offsets.size() > 1)
            check(mtg_push_reads(ctx_, bases.data(), offsets.data(), with_counts ? counts.data() : nullptr,
                                 offsets.size() - 1));
    }

    mtg_ctx *ctx_;
    size_t k_;
    std::filesystem::path swap_dir_;
    std::mutex mu_;
    std::string bases_;
    std::vector<uint64_t> offsets_, counts_;
    bool any_count_ = false;
};

} // namespace

// Called first thing by IBOSSChunkConstructor::initialize (boss_chunk_construct.cpp:1134).
// nullptr = not selected / not supported: the reference's own constructors take over.
std::unique_ptr<IBOSSChunkConstructor>
mtg_gpu_make_constructor(size_t k,
                         bool both_strands,
                         uint8_t bits_per_count,
                         const std::string &filter_suffix,
                         size_t /* num_threads: the device does the work */,
                         double memory_preallocated,
                         mtg::kmer::ContainerType /* VECTOR and VECTOR_DISK: HBM either way */,
                         const std::filesystem::path &swap_dir,
                         size_t /* disk_cap_bytes */) {
    const char *on = std::getenv("MTG_GPU");
    if (!on || std::strcmp(on, "1"))
        return nullptr;

    if (bits_per_count > 32)   // boss_chunk_construct.cpp:1157-1159
        throw std::runtime_error("Error: trying to allocate too many bits per k-mer count");

    mtg_params prm;
    std::memset(&prm, 0, sizeof(prm));
    prm.k = k;
#if _PROTEIN_GRAPH
    prm.alphabet = MTG_ALPHABET_PROTEIN;
#elif _DNA_GRAPH
    prm.alphabet = MTG_ALPHABET_DNA;
#else
    return nullptr;   // DNA5 / DNA_CASE_SENSITIVE alphabets: CPU path
#endif
    prm.both_strands = both_strands;
    prm.bits_per_count = bits_per_count;
    prm.filter_suffix = filter_suffix.empty() ? nullptr : filter_suffix.c_str();
    prm.device = std::getenv("MTG_GPU_DEVICE") ? std::atoi(std::getenv("MTG_GPU_DEVICE")) : 0;
    prm.memory_preallocated = memory_preallocated;

    mtg_ctx *ctx = nullptr;
    const int rc = mtg_create(&prm, &ctx);
    if (rc == MTG_ERR_UNSUPPORTED) {
        static std::once_flag once;
        std::call_once(once, [&]() {
            logger->warn("GPU constructor: {} -- the CPU constructor is used for such k", mtg_last_error(nullptr));
        });
        g_cpu_fallbacks++;
        return nullptr;
    }
    if (rc != MTG_OK) {
        logger->error("{}", mtg_last_error(nullptr));   // same text as boss_chunk_construct.cpp:1057-1062
        exit(1);
    }
    return std::make_unique<BOSSChunkConstructorGPU>(ctx, k, swap_dir);
}

} // namespace boss
} // namespace graph
} // namespace mtg
