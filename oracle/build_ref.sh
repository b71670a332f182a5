#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.
# Compiles the reference's own implementation of the `metagraph build` chunk
# path (UNMODIFIED sources, where they lie under /root/reference) together with
# oracle/ref_driver.cpp into oracle/_ref/ref_driver_{DNA,Protein}.
# No reference source is copied into this repository; objects and binaries go
# to oracle/_ref/ (git-ignored, but shipped to the GPU box by gpurun).
# The reference's cmake build system is NOT run: g++ is invoked on the files.
set -euo pipefail
HERE="$(cd "$(dirname "$0")" && pwd)"
REF="${MTG_REFERENCE:-/root/reference}/metagraph"
OUT="$HERE/_ref"
JOBS="${JOBS:-$(nproc)}"
if [ ! -d "$REF/src" ]; then
    echo "build_ref: $REF not present, keeping prebuilt oracle/_ref" >&2
    exit 0
fi
mkdir -p "$OUT/obj_sdsl" "$OUT/obj_DNA" "$OUT/obj_Protein"
EXT="$REF/external-libraries"
SDSL="$EXT/sdsl-lite"

# -O3: what the reference's own Release build type uses (CMAKE_BUILD_TYPE=Release)
CXXFLAGS_COMMON="-std=gnu++17 -O3 -fopenmp -mcx16 -msse4.2 -mavx2 -mbmi -mbmi2 -mfma \
 -DMODE_TI -DNDEBUG -D_THREAD_SAFE -pthread -w -include cstdint -include stdexcept"
INC="-I$REF/src -I$SDSL/include -I$EXT/ips4o -I$EXT/spdlog/include \
 -I$EXT/hopscotch-map/include -I$EXT/ordered-map/include -I$EXT/cpp_progress_bar \
 -I$EXT/DYNAMIC/include -I$EXT/caches/include -I$EXT/binrel_wt/include \
 -I$EXT/nvbio/contrib/priority-deque -I$EXT/nvbio/contrib/htslib -I$EXT/eigen -I$EXT/rollinghashcpp"

# ---- sdsl (static part actually needed by int_vector_buffer / wt / rank-select) ----
SDSL_SRCS="bits coder_elias_delta coder_elias_gamma coder_fibonacci config io \
 memory_management ram_filebuf ram_fs sfstream uint128_t uint256_t util \
 rrr_vector_15 sd_vector wt_helper bp_support_algorithm louds_tree nn_dict_dynamic"
build_one() {  # $1=src $2=obj $3..=flags
    local src="$1" obj="$2"; shift 2
    if [ ! -f "$obj" ] || [ "$src" -nt "$obj" ]; then
        g++ "$@" -c "$src" -o "$obj"
    fi
}
export -f build_one
{
  for s in $SDSL_SRCS; do
    echo "build_one $SDSL/lib/$s.cpp $OUT/obj_sdsl/$s.o -std=gnu++17 -O2 -w -DNDEBUG -I$SDSL/include"
  done
  # structure_tree.cpp is a cmake template whose only substitution sits inside a
  # string literal; it compiles as C++ as it is.
  echo "build_one $SDSL/lib/structure_tree.cpp.cmake $OUT/obj_sdsl/structure_tree.o -x c++ -std=gnu++17 -O2 -w -DNDEBUG -I$SDSL/include"
} | xargs -P "$JOBS" -I{} bash -c "{}"

# ---- reference sources on the path ----
REF_SRCS="
 kmer/kmer_extractor.cpp kmer/kmer_collector.cpp kmer/kmer_boss.cpp kmer/kmer.cpp
 common/sorted_sets/sorted_set.cpp common/sorted_sets/sorted_multiset.cpp
 common/sorted_sets/sorted_set_disk.cpp common/sorted_sets/sorted_multiset_disk.cpp
 common/sorted_sets/sorted_set_disk_base.cpp
 common/elias_fano/elias_fano.cpp
 common/threads/threading.cpp common/logger.cpp common/unix_tools.cpp
 common/utils/file_utils.cpp common/utils/string_utils.cpp common/algorithms.cpp
 common/serialization.cpp common/seq_tools/reverse_complement.cpp
 common/vectors/bit_vector.cpp common/vectors/bit_vector_sdsl.cpp common/vectors/bit_vector_dyn.cpp
 common/vectors/bit_vector_adaptive.cpp common/vectors/bit_vector_sd.cpp
 common/vectors/wavelet_tree.cpp common/vectors/vector_algorithm.cpp common/vectors/bitmap.cpp
 graph/representation/succinct/boss_chunk.cpp graph/representation/succinct/boss_chunk_construct.cpp
 graph/representation/succinct/boss.cpp
"
for ALPH in DNA Protein; do
  DEF="-D_${ALPH^^}_GRAPH"
  {
    for s in $REF_SRCS; do
      [ -f "$REF/src/$s" ] || continue
      o="$OUT/obj_$ALPH/$(echo "$s" | tr '/' '_' | sed 's/\.cpp$/.o/')"
      echo "build_one $REF/src/$s $o $CXXFLAGS_COMMON $DEF $INC"
    done
    echo "build_one $HERE/ref_driver.cpp $OUT/obj_$ALPH/ref_driver.o $CXXFLAGS_COMMON $DEF $INC"
    echo "build_one $EXT/spdlog/src/fmt.cpp $OUT/obj_$ALPH/spdlog_fmt.o -std=gnu++17 -O2 -w -DSPDLOG_COMPILED_LIB -I$EXT/spdlog/include"
    echo "build_one $EXT/cpp_progress_bar/progress_bar.cpp $OUT/obj_$ALPH/progress_bar.o -std=gnu++17 -O2 -w -I$EXT/cpp_progress_bar"
  } | xargs -P "$JOBS" -I{} bash -c "{}"
  g++ -fopenmp -pthread "$OUT"/obj_$ALPH/*.o "$OUT"/obj_sdsl/*.o -lz -latomic \
      -o "$OUT/ref_driver_$ALPH"
  echo "built $OUT/ref_driver_$ALPH"
done
