// C++ host side of the B200 BOSS chunk build: the reference's construction interface
// re-implemented on top of the C ABI (include/mtg_gpu.h). Same names, argument meaning
// and error behaviour as
//   IBOSSChunkConstructor  (graph/representation/succinct/boss_chunk_construct.hpp:18-34)
//   IGraphChunkConstructor (graph/representation/base/dbg_construct.hpp:12-22)
//   BOSSConstructor        (graph/representation/succinct/boss_construct.hpp:12-53)
//   BOSS::Chunk            (graph/representation/succinct/boss_chunk.hpp:19-104)
// so that cli/build.cpp:122-138 and tests/graph/succinct/test_boss_construct.cpp read the
// same against it. No sdsl / torch / CUDA types appear here: Chunk stores the packed
// 64-bit words of the sdsl int_vector files and writes the same bytes.
#pragma once
#include <cstdint>
#include <filesystem>
#include <memory>
#include <string>
#include <string_view>
#include <utility>
#include <vector>

struct mtg_ctx;

namespace mtg {
namespace gpu {

enum class ContainerType { VECTOR, VECTOR_DISK };   // kmer/kmer_collector.hpp (kept for signature parity)

enum class Alphabet { DNA = 0, Protein = 1 };       // compile-time macro in the reference

// BOSS::Chunk (boss_chunk.hpp:19-104). Row 0 is the placeholder.
class Chunk {
  public:
    static constexpr const char *kFileExtension = ".dbg.chunk";

    Chunk() {}
    // boss_chunk.cpp:135-144: empty chunk (only the placeholder row)
    Chunk(uint64_t alph_size, size_t k, uint8_t bits_per_count = 0);
    Chunk(Chunk &&) = default;
    Chunk &operator=(Chunk &&) = default;
    Chunk(const Chunk &) = delete;

    uint64_t size() const { return size_; }            // rows incl. the placeholder
    uint64_t alph_size() const { return alph_size_; }
    size_t get_k() const { return k_; }
    uint8_t get_W_width() const { return W_width_; }    // boss_chunk.cpp:388-390
    uint8_t bits_per_count() const { return weights_width_; }
    bool has_weights() const { return weights_width_ != 0; }

    uint64_t get_W(uint64_t i) const { return get_bits(W_, i, W_width_); }
    bool get_last(uint64_t i) const { return get_bits(last_, i, 1); }
    uint64_t get_weight(uint64_t i) const { return get_bits(weights_, i, weights_width_); }
    const std::vector<uint64_t> &get_F() const { return F_; }

    // packed words exactly as sdsl::int_vector<> stores them
    const std::vector<uint64_t> &W_words() const { return W_; }
    const std::vector<uint64_t> &last_words() const { return last_; }
    const std::vector<uint64_t> &weights_words() const { return weights_; }

    // boss_chunk.cpp:230-270
    void extend(const Chunk &other);
    // boss_chunk.cpp:372-386 / :330-370; return false / throw like the reference
    void serialize(const std::string &filename_base) const;
    bool load(const std::string &filename_base);

    // counters the reference logs (boss_chunk_construct.cpp:267-302), filled by build_chunk
    struct Stats {
        uint64_t n_input_kmers = 0, n_real = 0, n_dummy_sink = 0;
        std::vector<uint64_t> n_dummy_source;
        float device_ms = 0;
        unsigned kernel_launches = 0;
    } stats;

  private:
    friend class BOSSChunkConstructor;
    static uint64_t get_bits(const std::vector<uint64_t> &words, uint64_t i, unsigned width);
    static void append_bits(std::vector<uint64_t> *dst, uint64_t dst_n, unsigned width,
                            const std::vector<uint64_t> &src, uint64_t src_first, uint64_t src_n);

    uint64_t alph_size_ = 0;
    size_t k_ = 0;
    uint64_t size_ = 0;
    uint8_t W_width_ = 0, weights_width_ = 0;
    std::vector<uint64_t> W_, last_, weights_, F_;
};

// IBOSSChunkConstructor (boss_chunk_construct.hpp:18-34)
class IBOSSChunkConstructor {
  public:
    virtual ~IBOSSChunkConstructor() {}

    // Same parameters as the reference factory (boss_chunk_construct.cpp:1134-1178) plus the
    // alphabet (a compile-time macro there) and the CUDA device. num_threads,
    // memory_preallocated, container_type, swap_dir and disk_cap_bytes are accepted for
    // signature parity; the GPU path keeps everything in HBM.
    static std::unique_ptr<IBOSSChunkConstructor>
    initialize(size_t k, bool both_strands = false, uint8_t bits_per_count = 0,
               const std::string &filter_suffix = "", size_t num_threads = 1,
               double memory_preallocated = 0,
               ContainerType container_type = ContainerType::VECTOR,
               const std::filesystem::path &swap_dir = "/tmp/", size_t disk_cap_bytes = 1e9,
               Alphabet alphabet = Alphabet::DNA, int device = 0);

    virtual void add_sequence(std::string_view sequence, uint64_t count = 1) = 0;
    virtual void add_sequences(std::vector<std::string> &&sequences) = 0;
    virtual void add_sequences(std::vector<std::pair<std::string, uint64_t>> &&sequences) = 0;
    virtual Chunk build_chunk() = 0;
    virtual uint64_t get_k() const = 0;
    // the C ABI's native form (what host/read_feeder.hpp produces): read i = bases[offsets[i] .. offsets[i+1]).
    // Thread-safe; the buffers are free again when the call returns.
    virtual void add_packed(const char *bases, const uint64_t *offsets, size_t n_reads) = 0;
};

// BOSSConstructor (boss_construct.hpp:12-53). The reference's build_graph(BOSS*) hands the
// chunk to BOSS::initialize (sdsl wt_huff / rank-select: consumer side);
// here build_graph returns the chunk that call would consume.
class BOSSConstructor {
  public:
    BOSSConstructor(size_t k, bool both_strands = false, uint8_t bits_per_count = 0,
                    const std::string &filter_suffix = "", size_t num_threads = 1,
                    double memory_available = 0, ContainerType container_type = ContainerType::VECTOR,
                    const std::filesystem::path &swap_dir = "/tmp/", size_t disk_cap_bytes = 1e9,
                    Alphabet alphabet = Alphabet::DNA, int device = 0)
        : constructor_(IBOSSChunkConstructor::initialize(k, both_strands, bits_per_count,
                                                         filter_suffix, num_threads,
                                                         memory_available, container_type, swap_dir,
                                                         disk_cap_bytes, alphabet, device)) {}

    void add_sequence(std::string_view sequence, uint64_t count = 1) {
        constructor_->add_sequence(sequence, count);
    }
    void add_sequences(std::vector<std::string> &&sequences) {
        constructor_->add_sequences(std::move(sequences));
    }
    void add_sequences(std::vector<std::pair<std::string, uint64_t>> &&sequences) {
        constructor_->add_sequences(std::move(sequences));
    }
    void add_packed(const char *bases, const uint64_t *offsets, size_t n_reads) {
        constructor_->add_packed(bases, offsets, n_reads);
    }
    Chunk build_graph() { return constructor_->build_chunk(); }
    uint64_t get_k() const { return constructor_->get_k(); }

    // boss_construct.hpp:43-47 / boss_chunk.cpp:286-328: concatenate chunk files
    static Chunk build_graph_from_chunks(const std::vector<std::string> &chunk_filenames,
                                         bool verbose = false);

  private:
    std::unique_ptr<IBOSSChunkConstructor> constructor_;
};

}  // namespace gpu
}  // namespace mtg
