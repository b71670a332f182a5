// TEST INFRASTRUCTURE ONLY -- never linked into the product library.
//
// Thin driver around the UNMODIFIED reference sources under /root/reference
// (compiled where they lie by oracle/build_ref.sh, outputs into oracle/_ref/).
// It feeds reads to the reference's IBOSSChunkConstructor
// (metagraph/src/graph/representation/succinct/boss_chunk_construct.hpp:18-34),
// calls build_chunk() and serialises the chunk with the reference's own
// BOSS::Chunk::serialize (boss_chunk.cpp:372-386), i.e. it produces
//   <out>.dbg.chunk  <out>.dbg.chunk.W  <out>.dbg.chunk.last  <out>.dbg.chunk.weights
// which are the byte-compare targets for the CUDA path, and prints
//   "chunk_rows <n> read_sec <r> push_sec <p> build_sec <t>"  for the CPU baseline timing
// (build_sec = add_sequences + build_chunk with the reads already parsed in host memory).
//
// usage: ref_driver -k K [--canonical] [--count-width B] [--suffix S] [-p N]
//                   [--mem-gb G] -o OUT reads.fa|reads.txt
// Input: FASTA (">" headers, multi-line sequences) or plain text with one
// read per line (optionally "<read>\t<count>").
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <string>
#include <vector>

#include "graph/representation/succinct/boss_chunk_construct.hpp"
#include "graph/representation/succinct/boss_chunk.hpp"
#include "graph/representation/succinct/boss.hpp"
#include "common/logger.hpp"

using namespace mtg;
using namespace mtg::graph::boss;

int main(int argc, char **argv) {
    size_t K = 0, threads = 1;
    bool canonical = false, verbose = false;
    int count_width = 0;
    double mem_gb = 1;
    std::string suffix, out, in;
    for (int i = 1; i < argc; ++i) {
        std::string a = argv[i];
        if (a == "-k") K = atoi(argv[++i]);
        else if (a == "--canonical") canonical = true;
        else if (a == "--count-width") count_width = atoi(argv[++i]);
        else if (a == "--suffix") suffix = argv[++i];
        else if (a == "-p") threads = atoi(argv[++i]);
        else if (a == "--mem-gb") mem_gb = atof(argv[++i]);
        else if (a == "-o") out = argv[++i];
        else if (a == "-v") verbose = true;
        else in = a;
    }
    if (!K || out.empty() || in.empty()) {
        fprintf(stderr, "usage: ref_driver -k K [--canonical] [--count-width B] "
                        "[--suffix S] [-p N] [--mem-gb G] [-v] -o OUT reads\n");
        return 2;
    }
    if (verbose)
        common::set_verbose(true);

    // The reads are parsed into the batches of vector<pair<string, count>> the reference's interface takes
    // BEFORE the clock starts (read_sec, untimed part): the timed region then is what the GPU arm's
    // end-to-end figure covers -- reads resident in host memory -> add_sequences -> build_chunk.
    auto tr0 = std::chrono::steady_clock::now();
    std::ifstream f(in);
    if (!f) { fprintf(stderr, "cannot open %s\n", in.c_str()); return 2; }
    std::string line, seq;
    bool fasta = false, first = true;
    std::vector<std::vector<std::pair<std::string, uint64_t>>> batches(1);
    size_t batch_bytes = 0;
    auto push = [&](std::string &&s, uint64_t c) {
        batch_bytes += s.size();
        batches.back().emplace_back(std::move(s), c);
        if (batch_bytes > 1000000) {   // the CLI's batch size (kmer_collector.cpp:22)
            batches.emplace_back();
            batch_bytes = 0;
        }
    };
    while (std::getline(f, line)) {
        if (!line.empty() && line.back() == '\r') line.pop_back();
        if (first) { fasta = !line.empty() && line[0] == '>'; first = false; }
        if (fasta) {
            if (!line.empty() && line[0] == '>') {
                if (!seq.empty()) push(std::move(seq), 1);
                seq.clear();
            } else {
                seq += line;
            }
        } else {
            size_t tab = line.find('\t');
            if (tab == std::string::npos) push(std::move(line), 1);
            else push(line.substr(0, tab), strtoull(line.c_str() + tab + 1, 0, 10));
        }
    }
    if (fasta && !seq.empty()) push(std::move(seq), 1);
    auto tr1 = std::chrono::steady_clock::now();

    // same arguments as cli/build.cpp:122-134 (BOSS k = CLI k - 1)
    auto t0 = std::chrono::steady_clock::now();
    auto constructor = IBOSSChunkConstructor::initialize(
            K - 1, canonical, count_width, suffix, threads, mem_gb * 1e9,
            kmer::ContainerType::VECTOR);
    for (auto &b : batches)
        if (b.size()) constructor->add_sequences(std::move(b));
    auto t1 = std::chrono::steady_clock::now();

    BOSS::Chunk chunk = constructor->build_chunk();
    auto t2 = std::chrono::steady_clock::now();
    chunk.serialize(out);
    // read_sec: file -> batches (outside the timed region); push_sec: add_sequences of all batches
    // (extraction runs in the reference's thread pool meanwhile); build_sec: t0 -> chunk built
    printf("chunk_rows %llu read_sec %.6f push_sec %.6f build_sec %.6f\n",
           (unsigned long long)chunk.size(),
           std::chrono::duration<double>(tr1 - tr0).count(),
           std::chrono::duration<double>(t1 - t0).count(),
           std::chrono::duration<double>(t2 - t0).count());
    return 0;
}
