// mtg_build: command-line front end with the `metagraph build` flag surface for the BOSS path
// (cli/config/config.cpp:123-195,279,339; cli/build.cpp:58-157). It runs the GPU chunk build
// and writes <out>.dbg.chunk{,.W,.last,.weights} exactly as `metagraph build --suffix ...` does
// (boss_chunk.cpp:372-386); `mtg_build concatenate` joins chunk files (build_boss_from_chunks,
// boss_chunk.cpp:286-328). The wavelet-tree / rank-select construction that turns a chunk into
// a .dbg (BOSS::initialize, boss.cpp:99-122) stays with the reference host code.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "boss_chunk_construct_gpu.hpp"
#include "read_feeder.hpp"
#include "../../include/mtg_gpu.h"

using namespace mtg::gpu;

namespace {

int usage() {
    fprintf(stderr,
            "Usage: mtg_build build -k <k> [--mode basic|canonical] [--count-kmers] [--count-width <bits>]\n"
            "                       [-p <threads>] [--mem-cap-gb <G>] [--alphabet DNA|Protein] [--device <id>]\n"
            "                       [--suffix <S>]\n"
            "                       [-v] -o <outbase> FILE1 [FILE2 ...]\n"
            "       mtg_build concatenate -o <outbase> CHUNK1 [CHUNK2 ...]\n"
            "       mtg_build parse [-p <threads>] [--checksum] FILE1 [FILE2 ...]      (feeder only: reads, bases, checksum, GB/s)\n");
    return 1;
}

}  // namespace

int main(int argc, char **argv) {
    if (argc < 2) return usage();
    const std::string cmd = argv[1];
    size_t k = 3, threads = 1;
    bool canonical = false, count_kmers = false, verbose = false, checksum = false;
    int count_width = 8, device = 0;   // config.hpp:103 default --count-width 8
    double mem_cap_gb = 1;
    Alphabet alphabet = Alphabet::DNA;
    std::string out, suffix;
    std::vector<std::string> files;
    for (int i = 2; i < argc; ++i) {
        const std::string a = argv[i];
        auto next = [&]() -> const char * {
            if (i + 1 >= argc) exit(usage());
            return argv[++i];
        };
        if (a == "-k" || a == "--kmer-length") k = atoi(next());
        else if (a == "--mode") {
            const std::string m = next();
            if (m == "canonical") canonical = true;
            else if (m != "basic") { fprintf(stderr, "Error: unsupported --mode %s\n", m.c_str()); return 1; }
        } else if (a == "--count-kmers") count_kmers = true;
        else if (a == "--count-width") count_width = atoi(next());
        else if (a == "-p" || a == "--parallel") threads = atoi(next());
        else if (a == "--mem-cap-gb") mem_cap_gb = atof(next());
        else if (a == "--alphabet") alphabet = std::string(next()) == "Protein" ? Alphabet::Protein : Alphabet::DNA;
        else if (a == "--device") device = atoi(next());
        else if (a == "--suffix") suffix = next();   // config.cpp:279: build one chunk of the graph
        else if (a == "-o" || a == "--outfile-base") out = next();
        else if (a == "-v" || a == "--verbose") verbose = true;
        else if (a == "--checksum") checksum = true;
        else if (a[0] == '-') { fprintf(stderr, "Error: unknown option %s\n", a.c_str()); return usage(); }
        else files.push_back(a);
    }
    if (cmd == "parse" && !files.empty()) {
        // the host feeder alone (no device): what it parses and how fast
        try {
            ReadFeeder feeder([](const char *, const uint64_t *, size_t) {}, threads, size_t(32) << 20, nullptr,
                              nullptr, checksum);
            FeederStats st = feeder.run(files);
            printf("files %llu reads %llu bases %llu batches %llu checksum %016llx seconds %.4f input_GBps %.3f\n",
                   (unsigned long long)st.n_files, (unsigned long long)st.n_reads, (unsigned long long)st.n_bases,
                   (unsigned long long)st.n_batches, (unsigned long long)st.checksum, st.seconds,
                   st.input_bytes / 1e9 / st.seconds);
        } catch (const std::exception &e) {
            fprintf(stderr, "%s\n", e.what());
            return 1;
        }
        return 0;
    }
    if (out.empty() || files.empty()) return usage();
    try {
        if (cmd == "concatenate") {
            Chunk c = BOSSConstructor::build_graph_from_chunks(files, verbose);
            c.serialize(out);
            printf("Graph chunks concatenated: %llu rows\n", (unsigned long long)c.size());
            return 0;
        }
        if (cmd != "build") return usage();
        // config.cpp:421-434
        if (!count_kmers) count_width = 0;
        else if (count_width < 2 || count_width > 32) {
            fprintf(stderr, "Error: bad value for --count-width, use [2..32]\n");
            return 1;
        }
        if (k < 2) { fprintf(stderr, "Error: k must be at least 2\n"); return 1; }
        auto t0 = std::chrono::steady_clock::now();
        // cli/build.cpp:100,122-134: BOSS node length = k - 1
        BOSSConstructor constructor(k - 1, canonical, (uint8_t)count_width, suffix, threads, mem_cap_gb * 1e9,
                                    ContainerType::VECTOR, "/tmp/", 1e9, alphabet, device);
        // -p threads parse (ranges of plain files / one gzip file each) into pinned batches and push them
        // while the others parse: H2D of one batch overlaps the parsing of the next (host/read_feeder.hpp)
        ReadFeeder feeder([&](const char *bases, const uint64_t *offsets, size_t n) {
                              constructor.add_packed(bases, offsets, n);
                          },
                          threads, size_t(32) << 20, mtg_alloc_pinned, mtg_free_pinned);
        FeederStats fst = feeder.run(files);
        if (verbose)
            fprintf(stderr, "Parsed %llu reads, %llu bases from %llu files with %zu threads in %.3f sec (%.2f GB/s)\n",
                    (unsigned long long)fst.n_reads, (unsigned long long)fst.n_bases,
                    (unsigned long long)fst.n_files, threads, fst.seconds, fst.input_bytes / 1e9 / fst.seconds);
        auto t1 = std::chrono::steady_clock::now();
        Chunk chunk = constructor.build_graph();
        auto t2 = std::chrono::steady_clock::now();
        if (verbose) {   // the counters boss_chunk_construct.cpp:267-302 logs
            fprintf(stderr, "Total number of real k-mers: %llu\n", (unsigned long long)chunk.stats.n_real);
            fprintf(stderr, "Added %llu dummy sink k-mers\n", (unsigned long long)chunk.stats.n_dummy_sink);
            for (size_t i = 0; i < chunk.stats.n_dummy_source.size(); ++i)
                fprintf(stderr, "Number of dummy k-mers with dummy prefix of length %zu: %llu\n", i + 1,
                        (unsigned long long)chunk.stats.n_dummy_source[i]);
        }
        printf("Graph chunk with %llu k-mers was built in %.3f sec (parse %.3f sec, device %.3f ms, %u kernels)\n",
               (unsigned long long)(chunk.size() - 1), std::chrono::duration<double>(t2 - t0).count(),
               std::chrono::duration<double>(t1 - t0).count(), chunk.stats.device_ms,
               chunk.stats.kernel_launches);
        // cli/build.cpp:139-145: a --suffix chunk is written to <outbase>.<suffix>
        chunk.serialize(suffix.empty() ? out : out + "." + suffix);
    } catch (const std::exception &e) {   // logger->error + exit(1) in the reference
        fprintf(stderr, "%s\n", e.what());
        return 1;
    }
    return 0;
}
