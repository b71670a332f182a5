// Implementation of the C++ host mirror (see boss_chunk_construct_gpu.hpp) on top of the C ABI.
#include "boss_chunk_construct_gpu.hpp"

#include <algorithm>
#include <cstring>
#include <fstream>
#include <mutex>
#include <stdexcept>

#include "../../include/mtg_gpu.h"

namespace mtg {
namespace gpu {

// ---------------------------------------------------------------------------
// Chunk
// ---------------------------------------------------------------------------
static uint8_t W_width_for(uint64_t alph_size) {   // boss_chunk.cpp:388-390: hi(2*sigma-1)+1
    uint8_t w = 0;
    for (uint64_t v = 2 * alph_size - 1; v; v >>= 1) ++w;
    return w;
}

Chunk::Chunk(uint64_t alph_size, size_t k, uint8_t bits_per_count)
    : alph_size_(alph_size), k_(k), size_(1), W_width_(W_width_for(alph_size)),
      weights_width_(bits_per_count), W_(1, 0), last_(1, 0), weights_(bits_per_count ? 1 : 0, 0),
      F_(alph_size, 0) {}

uint64_t Chunk::get_bits(const std::vector<uint64_t> &words, uint64_t i, unsigned width) {
    if (!width) return 0;
    uint64_t bit = i * width, w = bit >> 6, off = bit & 63;
    uint64_t v = words[w] >> off;
    if (off + width > 64) v |= words[w + 1] << (64 - off);
    return width == 64 ? v : v & ((1ull << width) - 1);
}

// appends src entries [src_first, src_first + src_n) to dst, which holds dst_n entries (its padding
// bits are zero). Word-wise: 64 source bits per step through a funnel shift (a plain copy when both
// bit offsets are word aligned), not one entry per step -- chunks hold billions of rows.
void Chunk::append_bits(std::vector<uint64_t> *dst, uint64_t dst_n, unsigned width,
                        const std::vector<uint64_t> &src, uint64_t src_first, uint64_t src_n) {
    if (!width || !src_n) return;
    const uint64_t nbits = src_n * width, sb = src_first * width, db = dst_n * width;
    dst->resize((db + nbits + 63) / 64, 0);
    auto get64 = [&](uint64_t bit) -> uint64_t {   // 64 bits of src from `bit` on (zero padded)
        const uint64_t w = bit >> 6, o = bit & 63;
        uint64_t v = w < src.size() ? src[w] >> o : 0;
        if (o && w + 1 < src.size()) v |= src[w + 1] << (64 - o);
        return v;
    };
    auto low = [](uint64_t n) -> uint64_t { return n >= 64 ? ~0ull : (1ull << n) - 1; };
    uint64_t done = 0;
    if (db & 63) {   // fill up the partial last word of dst
        const uint64_t o = db & 63, take = std::min<uint64_t>(64 - o, nbits);
        (*dst)[db >> 6] |= (get64(sb) & low(take)) << o;
        done = take;
    }
    uint64_t w = (db + done) >> 6;
    if (((sb + done) & 63) == 0 && done < nbits) {   // aligned: straight copy of the full words
        const uint64_t full = (nbits - done) / 64;
        if (full) memcpy(dst->data() + w, src.data() + ((sb + done) >> 6), 8 * full);
        w += full;
        done += full * 64;
    }
    for (; done < nbits; done += 64) {
        const uint64_t take = std::min<uint64_t>(64, nbits - done);
        (*dst)[w++] = get64(sb + done) & low(take);
    }
}

void Chunk::extend(const Chunk &other) {   // boss_chunk.cpp:230-270
    if (alph_size_ != other.alph_size_ || k_ != other.k_ || W_width_ != other.W_width_)
        throw std::runtime_error("ERROR: trying to concatenate incompatible graph chunks");
    if (other.size_ == 1) return;
    if (has_weights() != other.has_weights())
        throw std::runtime_error("ERROR: trying to concatenate weighted and unweighted blocks");
    // clear the padding bits of the last word before OR-ing new entries in
    auto trim = [](std::vector<uint64_t> *v, uint64_t n, unsigned width) {
        uint64_t bits = n * width;
        v->resize((bits + 63) / 64);
        if (bits & 63) v->back() &= (1ull << (bits & 63)) - 1;
    };
    trim(&W_, size_, W_width_);
    trim(&last_, size_, 1);
    append_bits(&W_, size_, W_width_, other.W_, 1, other.size_ - 1);
    append_bits(&last_, size_, 1, other.last_, 1, other.size_ - 1);
    if (has_weights()) {
        trim(&weights_, size_, weights_width_);
        append_bits(&weights_, size_, weights_width_, other.weights_, 1, other.size_ - 1);
    }
    for (size_t c = 0; c < F_.size(); ++c) F_[c] += other.F_[c];
    size_ += other.size_ - 1;
}

namespace {
void write_u64_be(std::ofstream &f, uint64_t v) {   // serialize_number (common/serialization.cpp:38-50)
    unsigned char b[8];
    for (int i = 0; i < 8; ++i) b[i] = (unsigned char)(v >> (56 - 8 * i));
    f.write((const char *)b, 8);
}
uint64_t read_u64_be(std::ifstream &f) {
    unsigned char b[8] = {0};
    f.read((char *)b, 8);
    uint64_t v = 0;
    for (int i = 0; i < 8; ++i) v = (v << 8) | b[i];
    return v;
}
// sdsl::int_vector<t_width>::serialize: u64 bit length LE, [u8 width if t_width == 0], words
void write_int_vector(const std::string &path, const std::vector<uint64_t> &words, uint64_t n,
                      unsigned width, bool fixed_width) {
    std::ofstream f(path, std::ios::binary);
    if (!f) throw std::runtime_error("ERROR: cannot write " + path);
    uint64_t bits = n * width;
    f.write((const char *)&bits, 8);
    if (!fixed_width) { unsigned char w = (unsigned char)width; f.write((const char *)&w, 1); }
    uint64_t nwords = (bits + 63) / 64;
    std::vector<uint64_t> tmp(words.begin(), words.begin() + std::min<uint64_t>(nwords, words.size()));
    tmp.resize(nwords, 0);
    if (bits & 63) tmp.back() &= (1ull << (bits & 63)) - 1;
    f.write((const char *)tmp.data(), 8 * nwords);
}
bool read_int_vector(const std::string &path, std::vector<uint64_t> *words, uint64_t *n, unsigned *width,
                     int fixed_width) {
    std::ifstream f(path, std::ios::binary);
    if (!f) return false;
    uint64_t bits = 0;
    f.read((char *)&bits, 8);
    unsigned w = fixed_width;
    if (fixed_width < 0) { unsigned char c = 0; f.read((char *)&c, 1); w = c; }
    words->assign((bits + 63) / 64, 0);
    f.read((char *)words->data(), 8 * words->size());
    if (!f) return false;
    *width = w;
    *n = w ? bits / w : 0;
    return true;
}
std::string with_ext(const std::string &base, const std::string &ext) {
    return base.size() >= ext.size() && base.compare(base.size() - ext.size(), ext.size(), ext) == 0
               ? base : base + ext;
}
}  // namespace

void Chunk::serialize(const std::string &outbase) const {   // boss_chunk.cpp:372-386
    const std::string fname = with_ext(outbase, kFileExtension);
    write_int_vector(fname + ".W", W_, size_, W_width_, false);
    write_int_vector(fname + ".last", last_, size_, 1, true);
    if (has_weights()) write_int_vector(fname + ".weights", weights_, size_, weights_width_, false);
    else write_int_vector(fname + ".weights", {}, 0, 64, false);   // empty int_vector_buffer<>
    std::ofstream f(fname, std::ios::binary);
    if (!f) throw std::runtime_error("ERROR: cannot write " + fname);
    // serialize_number_vector = sdsl int_vector<64>; then alph_size, k as big-endian u64
    uint64_t bits = 64 * F_.size();
    f.write((const char *)&bits, 8);
    unsigned char w = 64;
    f.write((const char *)&w, 1);
    f.write((const char *)F_.data(), 8 * F_.size());
    write_u64_be(f, alph_size_);
    write_u64_be(f, k_);
}

bool Chunk::load(const std::string &infbase) {   // boss_chunk.cpp:330-370
    const std::string fname = with_ext(infbase, kFileExtension);
    uint64_t n = 0, nl = 0, nw = 0;
    unsigned ww = 0, lw = 0, tw = 0;
    if (!read_int_vector(fname + ".W", &W_, &n, &ww, -1)) return false;
    if (!read_int_vector(fname + ".last", &last_, &nl, &lw, 1)) return false;
    if (!read_int_vector(fname + ".weights", &weights_, &nw, &tw, -1)) return false;
    std::ifstream f(fname, std::ios::binary);
    if (!f) return false;
    uint64_t bits = 0;
    f.read((char *)&bits, 8);
    unsigned char w = 0;
    f.read((char *)&w, 1);
    F_.assign(bits / 64, 0);
    f.read((char *)F_.data(), 8 * F_.size());
    alph_size_ = read_u64_be(f);
    k_ = read_u64_be(f);
    if (!f || !k_ || !alph_size_ || n != nl || F_.size() != alph_size_ || (nw && nw != n)) return false;
    size_ = n;
    W_width_ = (uint8_t)ww;
    weights_width_ = nw ? (uint8_t)tw : 0;
    if (!nw) weights_.clear();
    return true;
}

// ---------------------------------------------------------------------------
// BOSSChunkConstructor: IBOSSChunkConstructor over the C ABI
// ---------------------------------------------------------------------------
class BOSSChunkConstructor : public IBOSSChunkConstructor {
  public:
    BOSSChunkConstructor(size_t k, bool both_strands, uint8_t bits_per_count,
                         const std::string &filter_suffix, Alphabet alphabet, int device)
        : k_(k) {
        mtg_params prm;
        memset(&prm, 0, sizeof(prm));
        prm.k = (uint32_t)k;
        prm.alphabet = (uint32_t)alphabet;
        prm.both_strands = both_strands;
        prm.bits_per_count = bits_per_count;
        prm.filter_suffix = filter_suffix.empty() ? nullptr : filter_suffix.c_str();
        prm.device = device;
        if (int rc = mtg_create(&prm, &ctx_))   // same messages as boss_chunk_construct.cpp:1057-1062,1157-1159
            throw std::runtime_error(std::string(mtg_last_error(nullptr)) + " (code " + std::to_string(rc) + ")");
    }
    ~BOSSChunkConstructor() override { mtg_destroy(ctx_); }

    // kmer_collector.hpp:69-79: single sequences go through a batcher (not thread-safe there either)
    void add_sequence(std::string_view sequence, uint64_t count) override {
        std::lock_guard<std::mutex> lk(mu_);
        if (sequence.size() < k_ + 1) return;
        batch_offsets_.push_back(batch_bases_.size() + sequence.size());
        batch_bases_.append(sequence);
        batch_counts_.push_back(count);
        any_count_ |= count != 1;
        if (batch_bases_.size() > kBatchBytes) flush_locked();
    }
    void add_sequences(std::vector<std::string> &&sequences) override {
        std::string bases;
        std::vector<uint64_t> offsets(1, 0);
        for (const auto &s : sequences) {
            bases += s;
            offsets.push_back(bases.size());
        }
        push(bases, offsets, nullptr);
    }
    void add_sequences(std::vector<std::pair<std::string, uint64_t>> &&sequences) override {
        std::string bases;
        std::vector<uint64_t> offsets(1, 0), counts;
        for (const auto &s : sequences) {
            bases += s.first;
            offsets.push_back(bases.size());
            counts.push_back(s.second);
        }
        push(bases, offsets, counts.data());
    }

    void add_packed(const char *bases, const uint64_t *offsets, size_t n_reads) override {
        if (!n_reads) return;
        if (int rc = mtg_push_reads(ctx_, bases, offsets, nullptr, n_reads))
            throw std::runtime_error(std::string(mtg_last_error(ctx_)) + " (code " + std::to_string(rc) + ")");
    }

    Chunk build_chunk() override {
        {
            std::lock_guard<std::mutex> lk(mu_);
            flush_locked();
        }
        mtg_chunk c;
        if (int rc = mtg_build_chunk(ctx_, &c))
            throw std::runtime_error(std::string(mtg_last_error(ctx_)) + " (code " + std::to_string(rc) + ")");
        Chunk out;
        out.alph_size_ = c.sigma;
        out.k_ = c.k;
        out.size_ = c.n_rows;
        out.W_width_ = (uint8_t)c.W_width;
        out.weights_width_ = (uint8_t)c.weights_width;
        out.W_.assign(c.W, c.W + (c.n_rows * c.W_width + 63) / 64);
        out.last_.assign(c.last, c.last + (c.n_rows + 63) / 64);
        if (c.weights_width)
            out.weights_.assign(c.weights, c.weights + (c.n_rows * c.weights_width + 63) / 64);
        out.F_.assign(c.F, c.F + c.sigma);
        out.stats.n_input_kmers = c.n_input_kmers;
        out.stats.n_real = c.n_real;
        out.stats.n_dummy_sink = c.n_dummy_sink;
        out.stats.n_dummy_source.assign(c.n_dummy_source, c.n_dummy_source + c.k);
        out.stats.device_ms = c.ms_total;
        out.stats.kernel_launches = c.n_kernel_launches;
        mtg_free_chunk(&c);
        return out;
    }
    uint64_t get_k() const override { return k_; }

  private:
    static constexpr size_t kBatchBytes = 1000000;   // kLargeBufferSize, kmer_collector.cpp:22

    void push(const std::string &bases, const std::vector<uint64_t> &offsets, const uint64_t *counts) {
        if (offsets.size() < 2) return;
        if (int rc = mtg_push_reads(ctx_, bases.data(), offsets.data(), counts, offsets.size() - 1))
            throw std::runtime_error(std::string(mtg_last_error(ctx_)) + " (code " + std::to_string(rc) + ")");
    }
    void flush_locked() {
        if (batch_offsets_.empty()) return;
        std::vector<uint64_t> offsets(1, 0);
        offsets.insert(offsets.end(), batch_offsets_.begin(), batch_offsets_.end());
        push(batch_bases_, offsets, any_count_ ? batch_counts_.data() : nullptr);
        batch_bases_.clear();
        batch_offsets_.clear();
        batch_counts_.clear();
        any_count_ = false;
    }

    size_t k_;
    mtg_ctx *ctx_ = nullptr;
    std::mutex mu_;
    std::string batch_bases_;
    std::vector<uint64_t> batch_offsets_, batch_counts_;
    bool any_count_ = false;
};

std::unique_ptr<IBOSSChunkConstructor>
IBOSSChunkConstructor::initialize(size_t k, bool both_strands, uint8_t bits_per_count,
                                  const std::string &filter_suffix, size_t /*num_threads*/,
                                  double /*memory_preallocated*/, ContainerType /*container_type*/,
                                  const std::filesystem::path & /*swap_dir*/, size_t /*disk_cap_bytes*/,
                                  Alphabet alphabet, int device) {
    return std::make_unique<BOSSChunkConstructor>(k, both_strands, bits_per_count, filter_suffix, alphabet,
                                                  device);
}

Chunk BOSSConstructor::build_graph_from_chunks(const std::vector<std::string> &chunk_filenames,
                                               bool verbose) {   // boss_chunk.cpp:286-328
    Chunk full;
    bool first = true;
    for (const auto &f : chunk_filenames) {
        Chunk c;
        if (!c.load(f)) throw std::runtime_error("ERROR: File corrupted. Cannot load graph chunk " + f);
        if (verbose) fprintf(stderr, "Chunk %s loaded: %llu rows\n", f.c_str(), (unsigned long long)c.size());
        if (first) { full = std::move(c); first = false; }
        else full.extend(c);
    }
    return full;
}

}  // namespace gpu
}  // namespace mtg
