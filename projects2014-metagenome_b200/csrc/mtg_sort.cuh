// Onesweep LSD radix sort (hand-written, sm_100a): one global histogram kernel,
// then one scatter kernel per digit with decoupled look-back across tiles.
// Replaces ips4o::parallel::sort in SortedSet / SortedMultiset
// (common/sorted_sets/sorted_set.cpp:41-48, sorted_multiset.cpp:54-84) and the
// dummy k-mer sorts (boss_chunk_construct.cpp:280-306).
#pragma once
#include "mtg_common.cuh"

namespace mtg {

constexpr int RADIX_BITS = 8;
constexpr int RADIX = 1 << RADIX_BITS;
constexpr int MAX_SORT_PASSES = 40;

// One LSD digit: `bits` bits starting at `shift` of the key (src 0) or of the
// 32-bit value payload (src 1).
struct DigitPass {
    int src, shift, bits;
};

struct SortPlan {
    int n;
    DigitPass p[MAX_SORT_PASSES];
    // the caller does not care in which order EQUAL keys (with their payloads) leave the sort: the first
    // pass may then rank unstably (one shared-memory atomic per key instead of the stable ranking)
    bool unstable_first = false;
};

// plan covering key bits [lo, hi) in 8-bit digits
inline void plan_add_key_bits(SortPlan &plan, int lo, int hi) {
    for (int b = lo; b < hi; b += RADIX_BITS) {
        int nb = hi - b < RADIX_BITS ? hi - b : RADIX_BITS;
        plan.p[plan.n++] = DigitPass{0, b, nb};
    }
}

struct SortWorkspace {
    unsigned long long *status = nullptr;   // [tiles * RADIX]
    size_t status_words = 0;
    unsigned long long *hist = nullptr;     // [MAX_SORT_PASSES * RADIX] digit offsets
    unsigned int *tile_counters = nullptr;  // [MAX_SORT_PASSES]
    // optional profiling: [2p], [2p+1] bracket pass p; [2*MAX], [2*MAX+1] the histogram
    cudaEvent_t *events = nullptr;
    // optional: the unsorted keys live here (read-only, e.g. a receive buffer); the histogram and
    // the first pass read them, the passes then alternate keys[1], keys[0], ... as usual
    const void *first_src = nullptr;
};

size_t sort_tile_keys(bool wide);
void set_sort_config(int variant, int unstable_first, int shape, int prefetch);   // tuning hook, -1 keeps a setting
// test/bench helper: pseudo-random words (top word of every key masked with hi_mask)
void fill_random_words(uint64_t *words, uint64_t nwords, uint64_t seed, uint64_t hi_mask, bool wide,
                       cudaStream_t stream);   // keys per tile of the active kernel configuration
void fill_iota_div(uint64_t *words, uint64_t n, uint32_t div, cudaStream_t stream);   // words[i] = i / div
template <typename KeyT>
size_t sort_status_words(uint64_t n) {
    return (div_up(n, sort_tile_keys(sizeof(KeyT) > 8)) + 1) * RADIX;
}

// Sorts n keys (with optional 32-bit values). keys[0]/vals[0] hold the input; the
// function ping-pongs between buffer 0 and 1 and returns the index of the buffer
// holding the sorted output. Workspace must have status_words >= sort_status_words(n).
// Returns the number of kernels launched through *launches.
template <typename KeyT>
int radix_sort(KeyT *keys[2], uint32_t *vals[2], uint64_t n, const SortPlan &plan,
               SortWorkspace &ws, cudaStream_t stream, unsigned *launches);

// raw counts of one digit: hist[0 .. 2^bits) (device)
template <typename KeyT>
void digit_histogram(const KeyT *keys, uint64_t n, DigitPass dp, unsigned long long *hist, cudaStream_t stream,
                     unsigned *launches);

// Grouping + exchange in one kernel: a scatter pass (stable unless stable = false) in which a key whose digit dp is b goes
// to owner[b] (device array of RADIX bytes) and is stored into the array at address route[owner]
// (device array of RADIX addresses, local or peer-mapped memory of another GPU).
// status: sort_status_words(n) words, tile_counter: one word.
template <typename KeyT>
void route_scatter(const KeyT *keys, uint64_t n, DigitPass dp, const unsigned char *owner,
                   const unsigned long long *route, unsigned long long *status, unsigned int *tile_counter,
                   cudaStream_t stream, unsigned *launches, bool stable = true);

}  // namespace mtg
