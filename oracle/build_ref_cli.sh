#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.
# Builds, from the UNMODIFIED reference sources where they lie under /root/reference (g++ on the
# files, the reference's cmake is not run), into oracle/_ref/ (git-ignored, shipped by gpurun):
#   metagraph_DNA          the reference's own CLI (`metagraph build ...` -> .dbg), CPU path
#   metagraph_gpu_DNA      the same objects, except that boss_chunk_construct.cpp is compiled from a
#                          COPY-ON-BUILD overlay (oracle/_ref/overlay/, never committed) whose factory
#                          IBOSSChunkConstructor::initialize asks the GPU binding first, plus the binding
#                          projects2014-metagenome_b200/host/metagraph_binding/boss_chunk_construct_gpu.cpp,
#                          linked against libmtg_gpu.so. Every other line is the reference's.
#   unit_tests_gpu_DNA     the reference's gtest file tests/graph/succinct/test_boss_construct.cpp
#                          (unmodified) linked the same way: the reference's own unit tests of this path
#                          run through the C ABI (MTG_GPU=1) or through its CPU code (MTG_GPU unset).
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
ROOT="$(dirname "$HERE")"
REF="${MTG_REFERENCE:-/root/reference}/metagraph"
OUT="$HERE/_ref"
JOBS="${JOBS:-$(nproc)}"
if [ ! -d "$REF/src" ]; then
    echo "build_ref_cli: $REF not present, keeping prebuilt oracle/_ref" >&2
    exit 0
fi
ALPH=DNA
OBJ="$OUT/obj_cli_$ALPH"
mkdir -p "$OBJ" "$OUT/overlay" "$OUT/obj_sdsl" "$OUT/test_data"
EXT="$REF/external-libraries"
SDSL="$EXT/sdsl-lite"
HTS="$EXT/nvbio/contrib/htslib"
PKG="$ROOT/projects2014-metagenome_b200"

# 6-line shim: the vendored htslib header lacks extern "C" (SURVEY.md 8c recipe)
cat > "$OUT/overlay/vcf_sweep_shim.h" <<'SHIM'
#pragma once
extern "C" {
#include <htslib/vcf_sweep.h>
}
SHIM

# the unit test's scratch/data directory (absolute: /root/repo is a symlink to the snapshot on the GPU box)
echo '#define TEST_DATA_DIR "/root/repo/oracle/_ref/test_data"' > "$OUT/overlay/test_data_dir.h"

# -O1: these binaries are the byte-parity checkers (.dbg files, gtest); the TIMED reference is
# oracle/_ref/ref_driver_* (oracle/build_ref.sh, -O3 like the reference's Release build). -O1 halves the
# ~100-file build.
CXXFLAGS="-std=gnu++17 -O1 -fopenmp -mcx16 -msse4.2 -mavx2 -mbmi -mbmi2 -mfma -D_${ALPH}_GRAPH \
 -DMODE_TI -DNDEBUG -DASIO_STANDALONE -DUSE_STANDALONE_ASIO -D_THREAD_SAFE -pthread -w \
 -include cstdint -include stdexcept"
INC="-I$REF/src -I$REF -I$SDSL/include -I$EXT/ips4o -I$EXT/spdlog/include \
 -I$EXT/hopscotch-map/include -I$EXT/ordered-map/include -I$EXT/cpp_progress_bar \
 -I$EXT/DYNAMIC/include -I$EXT/caches/include -I$EXT/binrel_wt/include \
 -I$EXT/nvbio/contrib/priority-deque -I$HTS -I$EXT/eigen -I$EXT/rollinghashcpp \
 -I$EXT/asio/asio/include -I$EXT/KMC/kmc_api -I$EXT/KMC -I$EXT/jsoncpp/include -I$EXT/tclap/include \
 -I$EXT/googletest/googletest/include -I$EXT/googletest/googletest -I$EXT/googletest/googlemock/include \
 -I$ROOT/include"

build_one() {  # $1=src $2=obj $3..=compiler + flags
    local src="$1" obj="$2"; shift 2
    if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ]; then
        "$@" -c "$src" -o "$obj" || { echo "FAILED: $src" >&2; exit 1; }
    fi
}
export -f build_one
oname() { echo "$1" | sed "s#^$REF/##; s#^$ROOT/##" | tr '/' '_' | sed 's/\.[a-z]*$/.o/'; }

# ---- the overlay: reference file + ONE hunk (the line INTEGRATION.md shows) ----
python3 - "$REF/src/graph/representation/succinct/boss_chunk_construct.cpp" "$OUT/overlay/boss_chunk_construct.cpp" <<'PY'
import sys
src = open(sys.argv[1]).read()
decl = '''
// ---- added for the GPU binding (projects2014-metagenome_b200/host/metagraph_binding) ----
std::unique_ptr<IBOSSChunkConstructor>
mtg_gpu_make_constructor(size_t, bool, uint8_t, const std::string &, size_t, double,
                         mtg::kmer::ContainerType, const std::filesystem::path &, size_t);
// -------------------------------------------------------------------------------------------

std::unique_ptr<IBOSSChunkConstructor>
IBOSSChunkConstructor::initialize('''
anchor = "\nstd::unique_ptr<IBOSSChunkConstructor>\nIBOSSChunkConstructor::initialize("
assert src.count(anchor) == 1
src = src.replace(anchor, decl)
body = "                                  size_t disk_cap_bytes) {\n#define OTHER_ARGS"
assert src.count(body) == 1
src = src.replace(body, '''                                  size_t disk_cap_bytes) {
    if (auto gpu = mtg_gpu_make_constructor(k, both_strands, bits_per_count, filter_suffix, num_threads,
                                            memory_preallocated, container_type, swap_dir, disk_cap_bytes))
        return gpu;   // MTG_GPU=1: the CUDA path behind the same interface
#define OTHER_ARGS''')
new = src
try:
    if open(sys.argv[2]).read() == new:
        sys.exit(0)
except OSError:
    pass
open(sys.argv[2], "w").write(new)
PY

{
  # every reference source except the server (its Simple-Web-Server submodule is absent)
  find "$REF/src" -name '*.cpp' | grep -v 'cli/server' | sort | while read -r s; do
    extra=""
    case "$s" in *seq_io/vcf_parser.cpp|*seq_io/sequence_io.cpp) extra="-include $OUT/overlay/vcf_sweep_shim.h";; esac
    echo "build_one $s $OBJ/$(oname "$s") g++ $CXXFLAGS $INC $extra"
  done
  echo "build_one $OUT/overlay/boss_chunk_construct.cpp $OBJ/overlay_boss_chunk_construct.o g++ $CXXFLAGS $INC -I$REF/src/graph/representation/succinct"
  echo "build_one $PKG/host/metagraph_binding/boss_chunk_construct_gpu.cpp $OBJ/binding_boss_chunk_construct_gpu.o g++ $CXXFLAGS $INC"
  for s in "$EXT"/KMC/kmc_api/{kmc_file,kmer_api,mmer}.cpp "$EXT"/jsoncpp/src/lib_json/*.cpp \
           "$EXT"/cpp_progress_bar/progress_bar.cpp "$EXT"/rollinghashcpp/mersennetwister.cpp \
           $(find "$EXT/binrel_wt/lib" -name '*.cpp'); do
    echo "build_one $s $OBJ/dep_$(oname "$s") g++ -std=gnu++17 -O2 -w -include cstdint -include stdexcept $INC"
  done
  echo "build_one $EXT/spdlog/src/fmt.cpp $OBJ/dep_spdlog_fmt.o g++ -std=gnu++17 -O2 -w -DSPDLOG_COMPILED_LIB -I$EXT/spdlog/include"
  for s in bgzf faidx hfile hfile_net hts kfunc knetfile kstring sam synced_bcf_reader tbx vcf vcf_sweep vcfutils; do
    echo "build_one $HTS/$s.c $OBJ/hts_$s.o gcc -O2 -w -DSAMTOOLS=1 -I$HTS"
  done
  for s in "$HTS"/cram/*.c; do
    echo "build_one $s $OBJ/hts_cram_$(basename "$s" .c).o gcc -O2 -w -DSAMTOOLS=1 -I$HTS"
  done
  # gtest + the reference's unit test of this path
  echo "build_one $EXT/googletest/googletest/src/gtest-all.cc $OBJ/gtest_all.o g++ -std=gnu++17 -O2 -w $INC"
  echo "build_one $EXT/googletest/googletest/src/gtest_main.cc $OBJ/gtest_main.o g++ -std=gnu++17 -O2 -w $INC"
  echo "build_one $REF/tests/utils/gtest_patch.cpp $OBJ/test_gtest_patch.o g++ $CXXFLAGS $INC"
  echo "build_one $REF/tests/graph/succinct/test_boss_construct.cpp $OBJ/test_boss_construct.o g++ $CXXFLAGS $INC -include $OUT/overlay/test_data_dir.h"
} > "$OBJ/jobs.txt"
xargs -P "$JOBS" -I{} bash -c "{}" < "$OBJ/jobs.txt"

# stub for the one excluded translation unit
cat > "$OUT/overlay/run_server_stub.cpp" <<'STUB'
namespace mtg { namespace cli { class Config; int run_server(Config *) { return 1; } } }
STUB
g++ -O2 -c "$OUT/overlay/run_server_stub.cpp" -o "$OBJ/stub_run_server.o"

# sdsl objects come from build_ref.sh (oracle/_ref/obj_sdsl)
COMMON=$(ls "$OBJ"/*.o | grep -v -e overlay_ -e binding_ -e gtest_ -e test_ -e src_main.o -e src_graph_representation_succinct_boss_chunk_construct.o)
LIBS="-lz -latomic -lpthread"
LINKGPU="-L$PKG -lmtg_gpu -Wl,-rpath,\$ORIGIN/../../projects2014-metagenome_b200"
g++ -fopenmp -pthread $COMMON "$OBJ/src_main.o" "$OBJ/src_graph_representation_succinct_boss_chunk_construct.o" \
    "$OUT"/obj_sdsl/*.o $LIBS -o "$OUT/metagraph_$ALPH"
g++ -fopenmp -pthread $COMMON "$OBJ/src_main.o" "$OBJ/overlay_boss_chunk_construct.o" "$OBJ/binding_boss_chunk_construct_gpu.o" \
    "$OUT"/obj_sdsl/*.o $LIBS $LINKGPU -o "$OUT/metagraph_gpu_$ALPH"
g++ -fopenmp -pthread $COMMON "$OBJ/overlay_boss_chunk_construct.o" "$OBJ/binding_boss_chunk_construct_gpu.o" \
    "$OBJ/test_boss_construct.o" "$OBJ/test_gtest_patch.o" "$OBJ/gtest_all.o" "$OBJ/gtest_main.o" \
    "$OUT"/obj_sdsl/*.o $LIBS $LINKGPU -o "$OUT/unit_tests_gpu_$ALPH"
echo "built $OUT/metagraph_$ALPH $OUT/metagraph_gpu_$ALPH $OUT/unit_tests_gpu_$ALPH"
