/*
 * mtg_gpu.h -- C ABI of the B200-native `metagraph build` hot path.
 *
 * This is the boundary a MetaGraph maintainer binds instead of the CPU
 * implementation behind IBOSSChunkConstructor
 *   (metagraph/src/graph/representation/succinct/boss_chunk_construct.hpp:18-34,
 *    factory boss_chunk_construct.cpp:1134-1178).
 * Plain pointers and sizes only; no C++/torch types cross it. All functions
 * return 0 on success or an MTG_ERR_* code; mtg_last_error() gives the text.
 * There is NO CPU fallback: every entry point fails with MTG_ERR_CUDA when no
 * sm_100 device / kernel image is available.
 *
 * Reads -> packed k-mers (tight alphabet, KMerBOSS layout kmer_boss.hpp:17-28)
 *   -> radix sort -> unique / saturating count -> [canonical] reverse-complement
 *   completion -> dummy sink / source k-mer recovery -> BOSS arrays W, last, F
 *   [, weights] in the sdsl int_vector bit layout BOSS::Chunk stores
 *   (boss_chunk.cpp:31-133, :372-386).
 */
#ifndef MTG_GPU_H
#define MTG_GPU_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MTG_OK 0
#define MTG_ERR_BAD_ARG 1
#define MTG_ERR_CUDA 2
#define MTG_ERR_OOM 3
#define MTG_ERR_STATE 4
#define MTG_ERR_UNSUPPORTED 5

#define MTG_ALPHABET_DNA 0      /* -D_DNA_GRAPH: "ACGT" / "$ACGT" (kmer/alphabets.hpp:65-77,128-141) */
#define MTG_ALPHABET_PROTEIN 1  /* -D_PROTEIN_GRAPH (kmer/alphabets.hpp:27-39,90-102) */

/* mtg_params.reserved: keep every pushed read until the build instead of folding the input into sorted
 * runs batch by batch (the range-partitioned phases mtg_mg_* consume the raw reads themselves) */
#define MTG_FLAG_NO_BATCHING 1u

#define MTG_MAX_SIGMA 32
#define MTG_MAX_K 128

typedef struct mtg_ctx mtg_ctx;

/* Mirrors the arguments of IBOSSChunkConstructor::initialize
 * (boss_chunk_construct.hpp:22-31). */
typedef struct mtg_params {
    uint32_t k;               /* BOSS node length (CLI -k minus 1); collected k-mers have k+1 chars */
    uint32_t alphabet;        /* MTG_ALPHABET_* (a compile-time macro in the reference) */
    uint32_t both_strands;    /* canonical mode: CANONICAL_ONLY collection + rc completion */
    uint32_t bits_per_count;  /* 0 = no weights, else 1..32 (--count-width) */
    const char *filter_suffix;/* NULL or "": default build. Else `--suffix S` over "$ACGT" / "$ABC..": only the
                               * framed k-mers whose characters before the last equal S are kept and the
                               * chunk is built straight from them (KmerExtractorBOSS,
                               * kmer/kmer_extractor.cpp:317-381; boss_chunk_construct.cpp:961-1013) */
    int32_t device;           /* CUDA device ordinal */
    uint32_t reserved;        /* flags: MTG_FLAG_NO_BATCHING */
    double memory_preallocated; /* bytes; hint for the initial read-staging capacity */
} mtg_params;

/* Result of mtg_build_chunk: the contents of a BOSS::Chunk (boss_chunk.hpp:19-104).
 * Pointers are library-owned pinned host memory, valid until mtg_free_chunk().
 * W / last / weights hold n_rows entries (row 0 is the placeholder) packed LSB-first
 * into little-endian 64-bit words exactly as sdsl::int_vector<> stores them. */
typedef struct mtg_chunk {
    uint64_t n_rows;
    uint32_t sigma;           /* BOSS alphabet size incl. '$' */
    uint32_t k;               /* node length */
    uint32_t W_width;         /* bits per W entry: hi(2*sigma-1)+1 (boss_chunk.cpp:388-390) */
    uint32_t weights_width;   /* bits_per_count, 0 if no weights */
    uint64_t F[MTG_MAX_SIGMA];
    const uint64_t *W;        /* ceil(n_rows*W_width/64) words */
    const uint64_t *last;     /* ceil(n_rows/64) words */
    const uint64_t *weights;  /* ceil(n_rows*weights_width/64) words or NULL */
    /* counters the reference logs (boss_chunk_construct.cpp:267-302) */
    uint64_t n_input_kmers;   /* valid (k+1)-windows extracted from the reads */
    uint64_t n_real;          /* real k-mers after sort/unique (and rc completion) */
    uint64_t n_dummy_sink;
    uint64_t n_dummy_source[MTG_MAX_K]; /* [i] = #dummy k-mers with dummy prefix length i+1 */
    /* device time (CUDA events) of the last build, ms; [0]=total, then per phase */
    float ms_total, ms_extract, ms_sort, ms_unique, ms_rc, ms_dummy, ms_emit;
    uint32_t n_kernel_launches;
    uint32_t device_resident; /* 1: W/last/weights are DEVICE pointers (mtg_build_chunk_device) */
    /* the dominant kernel (onesweep pass of the main k-mer sort), CUDA events around each launch */
    float ms_onesweep_sum;    /* sum of the pass kernels' durations, ms */
    float ms_sort_histogram;  /* the digit-histogram kernel */
    uint32_t n_onesweep;      /* number of pass launches */
    uint32_t onesweep_bytes_per_key; /* key (+payload) bytes one pass reads and writes per key */
    uint64_t onesweep_keys;   /* keys per pass launch */
    void *impl;
} mtg_chunk;

/* Creates a constructor for one chunk (= IBOSSChunkConstructor::initialize). */
int mtg_create(const mtg_params *params, mtg_ctx **out);
void mtg_destroy(mtg_ctx *ctx);
const char *mtg_last_error(const mtg_ctx *ctx); /* ctx may be NULL: last creation error */

/* add_sequences (boss_chunk_construct.cpp:988-994): `n_reads` reads stored back to back
 * in `bases`; read i is bases[offsets[i] .. offsets[i+1]). `counts` (nullable) gives the
 * per-read multiplicity (vector<pair<string,uint64_t>> overload); NULL means 1.
 * Thread-safe; the caller's buffers may be reused as soon as the call returns. */
int mtg_push_reads(mtg_ctx *ctx, const char *bases, const uint64_t *offsets,
                   const uint64_t *counts, size_t n_reads);

/* Same for reads that are already resident in device memory in the library's packed
 * layout: the reads separated (and terminated) by one 0x00 byte each, `n_bytes` total,
 * 16-byte aligned. `n_windows_max` >= number of (k+1)-windows they contain. The buffer
 * must stay valid until mtg_build_chunk returns. All counts are 1. */
int mtg_push_packed_device(mtg_ctx *ctx, const void *d_packed, size_t n_bytes,
                           uint64_t n_windows_max);

/* Page-locked host memory for the caller's read batches: mtg_push_reads copies straight out of such
 * buffers (no staging copy) and returns when the copies are done. NULL when it cannot be allocated. */
void *mtg_alloc_pinned(size_t bytes);
void mtg_free_pinned(void *p);

/* Host helper producing that packed layout (for callers that keep reads resident). */
size_t mtg_packed_size(const uint64_t *offsets, size_t n_reads);
void mtg_pack_reads(const char *bases, const uint64_t *offsets, size_t n_reads, char *dst);

/* build_chunk (boss_chunk_construct.cpp:996-1040): drains the pipeline, runs
 * sort/unique/rc/dummy/emit on the device, copies the chunk to pinned host memory.
 * The context is cleared afterwards (like kmer_collector.clear(), :353). */
int mtg_build_chunk(mtg_ctx *ctx, mtg_chunk *out);
/* Same, but the chunk stays in HBM: out->W / last / weights are device pointers (used by the
 * device-resident benchmark leg and by multi-GPU orchestration). */
int mtg_build_chunk_device(mtg_ctx *ctx, mtg_chunk *out);
void mtg_free_chunk(mtg_chunk *chunk);

/* Run all work of this context on the caller's CUDA stream (cudaStream_t) instead of the
 * library's own, so that the caller's events bracket it. Call before any push. A NULL handle -- which
 * is what the legacy default stream (and torch's default stream) looks like -- keeps the library's own
 * non-blocking stream. Ordering contract either way: every entry point that returns device pointers has
 * synchronised its stream before it returns; device memory the caller PASSES IN (mtg_push_packed_device,
 * the d_* arguments of mtg_mg_*) must be complete on the caller's side -- same stream, or the producing
 * stream synchronised -- before the call (multi_gpu.DistBuild._drain). */
int mtg_set_stream(mtg_ctx *ctx, void *cuda_stream);

/* Test hooks (parity of intermediate stages): run extraction + sort +
 * unique (+ rc completion) only and return the sorted real k-mers of the collector
 * alphabet: words[2*i] = low 64 bits, words[2*i+1] = high 64 bits; counts nullable. */
int mtg_debug_sorted_kmers(mtg_ctx *ctx, uint64_t **words, uint32_t **counts, uint64_t *n);
void mtg_debug_free(void *p);

/* Stand-alone primitive hooks (device-side parity tests of the radix sort). */
int mtg_debug_sort_u64(int device, uint64_t *keys, uint32_t *vals /*nullable*/, uint64_t n,
                       uint32_t key_bits);
int mtg_debug_sort_u128(int device, uint64_t *keys_lohi, uint32_t *vals /*nullable*/, uint64_t n,
                        uint32_t key_bits);

/* ---- Range-partitioned build across several GPUs (one context per GPU) --------------------
 * The reference shards this path by the leading characters of the sort key
 * (boss_chunk_construct.cpp:366-397 split by F/W, :462-525 and :754-792 routing, :917-931
 * ordered extend). Here the routing bucket of a k-mer is the top mtg_mg_part_bits() bits of
 * its word; bucket ranges are assigned to contexts by the caller, which moves the data
 * between contexts (NCCL all-to-all / all-gather in multi_gpu.py). All buffers are device
 * memory; element size is 8 bytes if (k+1)*L <= 64, else 16. Sequence of calls per context:
 *   mtg_push_reads* -> mtg_mg_extract -> [all-to-all by bucket]   (or: mtg_mg_count -> mtg_mg_route)
 *   -> mtg_mg_collect
 *   [canonical: -> mtg_mg_strands -> [all-to-all of both strands by final range] -> mtg_mg_merge]
 *   -> mtg_mg_targets -> [all-to-all by bucket]   (or: mtg_mg_targets_count -> mtg_mg_targets_route)
 *   -> mtg_mg_dummies -> [all-gather]
 *   -> mtg_mg_build  (chunk of this range; the host concatenates with Chunk::extend). */
uint32_t mtg_mg_part_bits(const mtg_ctx *ctx);
/* Extracts the pushed reads (the canonical k-mers in canonical mode) and groups the k-mers by bucket:
 * bucket_counts[2^part_bits] (host), *d_keys library-owned until mtg_mg_collect. */
int mtg_mg_extract(mtg_ctx *ctx, uint64_t *bucket_counts, void **d_keys, uint64_t *n);
/* The same phase with the exchange fused into the grouping pass (what multi_gpu.py uses over NVLink):
 * mtg_mg_count extracts and returns bucket_counts[2^part_bits] (the k-mers stay in the context,
 * unordered); after the ranks have agreed on the bucket ranges, mtg_mg_route scatters the k-mers in
 * ONE kernel: a k-mer of bucket b goes to owner bucket_owner[b] (< n_owners <= 256) and is stored
 * into the array at device address owner_dst[owner] -- normally this context's slice of the owner
 * GPU's peer-mapped receive buffer (room for the sum of this context's counts of the owner's
 * buckets), so the stores of the grouping pass are the all-to-all (csrc/mtg_sort.cu:
 * onesweep_pass_v3_kernel<ROUTE>). Within a slice the k-mers arrive in extraction order (not grouped
 * by bucket: mtg_mg_collect sorts them anyway). The call returns when the kernel is done; the
 * caller then synchronises the ranks. Replaces the routing of boss_chunk_construct.cpp:366-397
 * (split by F/W into per-bucket containers). */
int mtg_mg_count(mtg_ctx *ctx, uint64_t *bucket_counts, uint64_t *n);
int mtg_mg_route(mtg_ctx *ctx, const uint32_t *bucket_owner /* [2^part_bits] */, uint32_t n_owners,
                 const uint64_t *owner_dst /* [n_owners] device addresses */);
/* Sort + fold the k-mers routed to this context (d_keys NULL: the context's own extraction). */
int mtg_mg_collect(mtg_ctx *ctx, const void *d_keys, uint64_t n, uint64_t *n_real);
/* Canonical mode only (both_strands): mtg_mg_extract / mtg_mg_collect route and reduce the CANONICAL
 * k-mers; the second strand is produced from the distinct ones, as add_reverse_complements does
 * (boss_chunk_construct.cpp:179-222). mtg_mg_strands returns X = the canonical k-mers of this
 * rank's stage-1 range (sorted, [x_bucket_counts] per routing bucket) and R = the reverse
 * complements of the non-palindromes, sorted as well ([r_bucket_counts]); d_xc / d_rc are their
 * uint32 counts (NULL without --count-kmers; palindrome counts already doubled). The caller plans
 * the FINAL ranges on x + r counts, exchanges X (keeping rank order) and R, and hands the received
 * arrays to mtg_mg_merge with the lengths of the sorted runs R consists of (one per source rank;
 * NULL = unsorted, R is radix sorted); it merges the runs and X: the real k-mers of the final
 * range. Library memory stays valid until mtg_mg_merge returns. */
int mtg_mg_strands(mtg_ctx *ctx, uint64_t *x_bucket_counts, uint64_t *r_bucket_counts, void **d_x, void **d_xc,
                   uint64_t *n_x, void **d_r, void **d_rc, uint64_t *n_r);
int mtg_mg_merge(mtg_ctx *ctx, const void *d_x, const uint32_t *d_xc, uint64_t n_x, const void *d_r,
                 const uint32_t *d_rc, uint64_t n_r, const uint64_t *r_run_lengths, uint32_t n_runs,
                 uint64_t *n_real);
/* Sorted distinct target nodes (k-mer minus first char) of this range's k-mers, with the
 * per-bucket counts for routing them to the owner of the node. */
int mtg_mg_targets(mtg_ctx *ctx, uint64_t *bucket_counts, void **d_targets, uint64_t *n);
/* mtg_mg_targets with the exchange fused into its scatter pass (cf. mtg_mg_count / mtg_mg_route):
 * mtg_mg_targets_count computes the target nodes of this range's reals and their counts per routing
 * bucket (top part_bits of the (k)-char node word); mtg_mg_targets_route scatters them in ONE kernel,
 * bucket b into the array at owner_dst[bucket_owner[b]]. The caller makes one "owner" per (owner rank,
 * leading character) SEGMENT: the scatter is stable and the targets of one edge label are sorted, so
 * every (source, segment) slice is a sorted run, which mtg_mg_dummies merges (run_lengths). Duplicate
 * targets travel and are folded by mtg_mg_dummies. */
int mtg_mg_targets_count(mtg_ctx *ctx, uint64_t *bucket_counts, uint64_t *n);
int mtg_mg_targets_route(mtg_ctx *ctx, const uint32_t *bucket_owner /* [2^part_bits] */, uint32_t n_owners,
                         const uint64_t *owner_dst /* [n_owners] device addresses */);
/* Inputs larger than one round's buffers (super-batching, SortedSet::shrink_data semantics,
 * common/sorted_sets/sorted_set.hpp:69-82): the caller repeats push -> mtg_mg_count -> mtg_mg_route ->
 * mtg_mg_collect with the SAME bucket ranges for every slice of its reads and folds each result into the
 * range's accumulated k-mer set with mtg_mg_accumulate (merge + saturating count sums). last != 0 ends
 * the rounds: the accumulated set is the range's k-mer set for mtg_mg_strands / mtg_mg_targets*. */
int mtg_mg_accumulate(mtg_ctx *ctx, int last, uint64_t *n_real);
/* Verification hook of the fused exchanges (multi_gpu.py, MTG_MG_VERIFY): the device array the next
 * route call will scatter -- which = 0: the k-mers counted by mtg_mg_count, which = 1: the target nodes
 * counted by mtg_mg_targets_count. Library memory, valid until that route call. */
int mtg_mg_debug_pending(mtg_ctx *ctx, int which, void **d_keys, uint64_t *n);
/* d_targets_in: all target nodes that fall into this range. Finds the range's dummy sinks
 * (kept inside) and returns its source nodes without incoming edge (forward-lex packing).
 * run_lengths[n_runs] (host, may be NULL): d_targets_in is the concatenation of sorted runs of
 * these lengths (every rank sends its targets sorted) -- they are merged instead of re-sorted.
 * runs_per_set (0 = all runs overlap): every runs_per_set consecutive runs hold keys that sort
 * before those of the next set (mtg_mg_targets_route: one set per leading character), so only the
 * runs inside a set are merged. */
int mtg_mg_dummies(mtg_ctx *ctx, const void *d_targets_in, uint64_t n_in, const uint64_t *run_lengths,
                   uint32_t n_runs, uint32_t runs_per_set, void **d_src_nodes, uint64_t *n_src, uint64_t *n_sink);
/* d_all_src_nodes: the gathered source nodes of every range. Emits the chunk of bucket range
 * [lo_bucket, hi_bucket); first_range adds the all-'$' row (boss_chunk_construct.cpp:310-315). */
int mtg_mg_build(mtg_ctx *ctx, const void *d_all_src_nodes, uint64_t n_all, uint64_t lo_bucket,
                 uint64_t hi_bucket, int first_range, int device_output, mtg_chunk *out);
/* The exchange step as one kernel of NVLink peer stores: pushes n_dest byte ranges of this GPU's
 * memory into peer-mapped destinations (the symmetric receive buffers of the other ranks; the host
 * side maps them and orders the ranks, multi_gpu.py). Enqueued on cuda_stream of the current device. */
int mtg_peer_push(void *cuda_stream, uint32_t n_dest, void *const *dst_ptrs, const void *const *src_ptrs,
                  const uint64_t *nbytes);

/* Device-resident timing of the radix sort alone on n pseudo-random keys of key_bits bits
 * (tuning aid): ms_out[5]: [0] = digit-histogram kernel, [1] = mean onesweep pass, [2] = whole sort,
 * [3] = first pass, [4] = mean of the later passes. */
int mtg_debug_sort_bench(int device, uint64_t n, uint32_t key_bits, int with_vals, int iters,
                         float *ms_out);
/* tuning hook: which kernel runs the stable radix passes (3 = peer masks through shared memory, the
 * default; 5 = atomic slot + collision repair) and whether the first pass of an order-free sort may rank
 * unstably (default 1); shape: 0 = 256 threads x 16 keys per block, 1 = 512 x 8; prefetch: distance in
 * tiles of the L2 prefetch a block issues for a later tile (0 = off); -1 keeps a setting. Process-wide. */
void mtg_debug_sort_config(int variant, int unstable_first, int shape, int prefetch);
/* test/bench hook: times the unique (fold) kernel alone on keys[i] = i / dup; ms_out[0] = best ms,
 * ms_out[1] = number of distinct keys found */
int mtg_debug_unique_bench(int device, uint64_t n, uint32_t dup, int with_starts, int iters, float *ms_out);

#ifdef __cplusplus
}
#endif
#endif /* MTG_GPU_H */
