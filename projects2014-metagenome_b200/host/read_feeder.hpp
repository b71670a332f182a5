// Host feeder of the GPU chunk build: multi-threaded FASTA / FASTQ (plain or gzip) parsing into
// flat batches (concatenated bases + offsets, the native form of mtg_push_reads) that the workers push
// while the others keep parsing -- the counterpart of the reference's file loop + kseq reader + batcher
// (cli/build.cpp:31-56, cli/parse_sequences.hpp:103-151, seq_io/sequence_io.cpp:364-405,
// common/batch_accumulator.hpp), whose one-string-per-read batches cap it near 0.1 GB/s per thread.
//   * plain files are memory-mapped and cut into ranges at record boundaries: the ranges of ONE file
//     are parsed in parallel;
//   * gzip files are inflated by one worker each (a deflate stream is serial), files in parallel;
//   * batches live in buffers from a caller-supplied allocator (pinned host memory from
//     mtg_alloc_pinned: mtg_push_reads then copies straight out of them, and the H2D copy of one
//     worker's batch overlaps the parsing of the others).
#pragma once
#include <cstddef>
#include <cstdint>
#include <functional>
#include <string>
#include <vector>

namespace mtg {
namespace gpu {

struct FeederStats {
    uint64_t n_files = 0, n_reads = 0, n_bases = 0, n_batches = 0, input_bytes = 0;
    uint64_t checksum = 0;       // order-independent: sum over reads of (FNV-1a of the read)
    double seconds = 0;
};

class ReadFeeder {
  public:
    // called from the worker threads (concurrently): read i is bases[offsets[i] .. offsets[i+1])
    using Sink = std::function<void(const char *bases, const uint64_t *offsets, size_t n_reads)>;
    using Alloc = void *(*)(size_t);
    using Free = void (*)(void *);

    ReadFeeder(Sink sink, size_t num_threads, size_t batch_bytes = size_t(32) << 20, Alloc alloc = nullptr,
               Free free_fn = nullptr, bool checksum = false);

    // parses every file; throws std::runtime_error on unreadable input
    FeederStats run(const std::vector<std::string> &files);

  private:
    Sink sink_;
    size_t threads_, batch_bytes_;
    Alloc alloc_;
    Free free_;
    bool checksum_;
};

}  // namespace gpu
}  // namespace mtg
