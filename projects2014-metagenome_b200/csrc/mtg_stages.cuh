// Device stages of the BOSS chunk build (host-callable launchers).
// Every function only enqueues work on `stream`; counters are device-resident and
// read back by the pipeline driver (mtg_api.cu).
#pragma once
#include "mtg_common.cuh"
#include "mtg_sort.cuh"

namespace mtg {

// ---- K1 extraction --------------------------------------------------------
struct ExtractLut {
    uint8_t code[256];   // ASCII -> tight code, 0xff = invalid / read separator
};
struct ExtractArgs {
    const uint8_t *packed;        // reads, 0x00 after each read, 16-byte aligned
    uint64_t nbytes;
    const ExtractLut *lut;        // host pointer, passed by value to the kernel
    uint32_t K, L;
    bool canonical;
    bool both = false;            // canonical only: emit the window AND its reverse complement
    uint32_t *out_counts;         // nullable: per-k-mer multiplicity payload
    unsigned long long *out_counter;   // device: running number of emitted k-mers
    const uint64_t *read_starts;  // device, packed start of each read (count mode)
    const uint32_t *read_counts;  // device, clipped per-read counts (count mode)
    uint64_t n_reads;
};
template <typename KeyT>
void extract_kmers(const ExtractArgs &a, KeyT *out_keys, cudaStream_t stream);
// raw concatenated reads + offsets[n_reads + 1] (device) -> packed layout; starts (nullable) = packed
// start of each read
void pack_reads_device(const uint8_t *raw, const uint64_t *offsets, uint64_t n_reads, uint8_t *packed,
                       uint64_t packed_bytes, uint64_t *starts, cudaStream_t stream);

// Row descriptor byte for dummy k-mers / merged rows:
//   bits 0..6  number of leading '$' in the node (dummy prefix length), 0 for real
//   bit  7     edge label W is '$' (dummy sink, or the all-'$' k-mer)
constexpr uint32_t INFO_W_DOLLAR = 0x80;

struct Alpha {
    uint32_t K;        // k-mer length (node length k = K-1)
    uint32_t L;        // bits per char in the collected (tight) words
    uint32_t sigma;    // BOSS alphabet size incl. '$'
    uint32_t W_width;  // bits per W entry
};

// ---- K3 unique / count-reduce --------------------------------------------
// keys sorted. Writes distinct keys to out_keys, the index of each run's first
// element to out_starts (plus n at [M]) and M to *d_num_out.
template <typename KeyT>
void unique_keys(const KeyT *keys, uint64_t n, KeyT *out_keys, uint64_t *out_starts,
                 unsigned long long *d_num_out, unsigned long long *scan_status,
                 unsigned int *tile_counter, cudaStream_t stream, unsigned *launches);
size_t unique_status_words(uint64_t n);
// counts[j] = min(cmax, vals ? sum(vals[starts[j]..starts[j+1])) : starts[j+1]-starts[j])
void run_counts(const uint64_t *starts, const uint32_t *vals, uint64_t m, uint32_t cmax,
                uint32_t *counts, cudaStream_t stream, unsigned *launches);

// ---- K1b reverse-complement completion -----------------------------------
// keys[0..m) distinct canonical k-mers. Writes rc(x) of every non-palindrome to rc_out[slot]
// (slot from *d_appended, with its count to rc_counts[slot]) and doubles (saturating at cmax)
// the counts of palindromes (boss_chunk_construct.cpp:179-222).
template <typename KeyT>
void append_reverse_complements(const KeyT *keys, KeyT *rc_out, uint32_t *counts, uint32_t *rc_counts,
                                uint64_t m, Alpha al, uint32_t cmax, unsigned long long *d_appended,
                                cudaStream_t stream, unsigned *launches);

// Merge of two sorted arrays (A first on ties), optional count payloads. `part` is scratch of
// merge_partition_words(na + nb) u64.
size_t merge_partition_words(uint64_t n_total);
template <typename KeyT>
void merge_sorted(const KeyT *A, uint64_t na, const KeyT *B, uint64_t nb, const uint32_t *Ac,
                  const uint32_t *Bc, KeyT *out, uint32_t *outc, uint64_t *part, cudaStream_t stream,
                  unsigned *launches);

// ---- K4 dummy k-mers -----------------------------------------------------
// Sinks: for every real x that is the first of its (x>>2L, W) class, emit
// [s_K..s_2 | 0] with info INFO_W_DOLLAR iff no real k-mer has that node.
// Sources-1: for every distinct node with no incoming real edge emit the node in
// forward-lexicographic packing (s_1 most significant, (K-1)*L bits).
// Both append through atomic counters; if a counter exceeds its capacity the
// entries beyond are dropped (the driver retries with the exact size).
// Bucket index over the sorted reals (top min(22, K*L) key bits -> first position), used by the
// membership probes; bucket_idx may be NULL (plain binary search).
size_t bucket_index_words(Alpha al);
template <typename KeyT>
void build_bucket_index(const KeyT *reals, uint64_t m, Alpha al, uint64_t *idx, cudaStream_t stream,
                        unsigned *launches);
// scratch: dummy_scratch_bytes(m, al) bytes of device memory (window bounds + incoming-edge bitmap)
size_t dummy_scratch_bytes(uint64_t m, Alpha al);
template <typename KeyT>
void find_dummy_candidates(const KeyT *reals, uint64_t m, Alpha al, const uint64_t *bucket_idx,
                           void *scratch, KeyT *sink_keys, uint64_t sink_cap, unsigned long long *d_num_sink,
                           KeyT *src_fwd, uint64_t src_cap, unsigned long long *d_num_src,
                           cudaStream_t stream, unsigned *launches);

// src_fwd sorted ascending (forward-lex). Node i contributes K-1-lcp(i, i-1) new distinct prefixes =
// dummy k-mers $^c s_1..s_{K-1-c} -> s_{K-c} (tight word, '$' slots zero, info = c). lo / hi
// (nullable) restrict them to the key range [lo, hi) of the chunk being built (multi-GPU):
// per_entry[i] = number of new prefixes of node i inside the range.
template <typename KeyT>
void source_prefix_counts(const KeyT *src_fwd, uint64_t d1, Alpha al, const KeyT *lo, const KeyT *hi,
                          uint32_t *per_entry, cudaStream_t stream, unsigned *launches);
// exclusive scan of uint32 -> uint64 offsets[n+1] (offsets[n] = total)
void exclusive_scan_u32(const uint32_t *in, uint64_t n, uint64_t *out,
                        unsigned long long *scan_status, unsigned int *tile_counter,
                        cudaStream_t stream, unsigned *launches);
// Writes the in-range dummies of node i from offsets[i] on; level_hist[c-1] counts ALL new
// prefixes (the log line of boss_chunk_construct.cpp:296-302 is global).
template <typename KeyT>
void generate_source_dummies(const KeyT *src_fwd, const uint64_t *offsets, uint64_t d1, Alpha al,
                             const KeyT *lo, const KeyT *hi, KeyT *out_keys, uint32_t *out_info,
                             unsigned long long *level_hist, cudaStream_t stream,
                             unsigned *launches);
void fill_u32(uint32_t *p, uint32_t v, uint64_t n, cudaStream_t stream, unsigned *launches);
// info -> sort payload ((127 - nd) | flag) and back (involution)
void flip_info(uint32_t *info, uint64_t n, cudaStream_t stream, unsigned *launches);

// ---- K5 merge ------------------------------------------------------------
// rank[j] = #reals ordered before dummy j (dummies sorted in final order).
template <typename KeyT>
void dummy_ranks(const KeyT *reals, uint64_t m, const KeyT *dummy_keys, uint64_t d, Alpha al,
                 const uint64_t *bucket_idx, uint64_t *rank, cudaStream_t stream, unsigned *launches);
// rows: [0] placeholder, [1] all-'$', then the merge of reals and dummies.
template <typename KeyT>
void merge_rows(const KeyT *reals, const uint32_t *real_counts, uint64_t m, const KeyT *dummy_keys,
                const uint32_t *dummy_info, const uint64_t *rank, uint64_t d, Alpha al,
                KeyT *row_keys, uint8_t *row_info, uint32_t *row_counts, cudaStream_t stream,
                unsigned *launches, uint32_t row_off = 2);

// ---- range-partitioned (multi-GPU) build, SURVEY.md 8e ----------------------------------------
// out[i] = target node of reals[i] (node word, (K-1)*L bits)
template <typename KeyT>
void target_nodes(const KeyT *reals, uint64_t m, Alpha al, KeyT *out, cudaStream_t stream,
                  unsigned *launches);
// sinks = targets \ source nodes of `reals`; sources-1 = source nodes \ targets (forward-lex)
// scratch: ranged_scratch_bytes(m, nt) bytes (window bounds + target hit bitmap)
size_t ranged_scratch_bytes(uint64_t m, uint64_t nt);
template <typename KeyT>
void ranged_dummy_candidates(const KeyT *reals, uint64_t m, const KeyT *targets, uint64_t nt, Alpha al,
                             void *scratch, KeyT *sink_keys, uint64_t sink_cap,
                             unsigned long long *d_num_sink, KeyT *src_fwd, uint64_t src_cap,
                             unsigned long long *d_num_src, cudaStream_t stream, unsigned *launches);
template <typename KeyT>
void bucket_bounds(const KeyT *sorted, uint64_t n, uint32_t B, uint32_t shift, uint64_t *bounds,
                   cudaStream_t stream, unsigned *launches);

// ---- K1s suffix-filtered chunks (--suffix), mtg_suffix.cu ---------------------------------
// Framed k-mers ('$'^(K-1) seg '$' per valid segment, both strands when a.both) that match the
// suffix filter, as composite sort keys  node << (L+8) | ((127-nd) << 1 | W!='$') << L | W.
// suffix_codes: BOSS codes (0 = '$', tight + 1 otherwise). Appends through a.out_counter; entries
// beyond `cap` are dropped (the driver retries with the exact size).
constexpr uint32_t MTG_SUFFIX_MAX = 32;
template <typename KeyT>
void extract_suffix_kmers(const ExtractArgs &a, Alpha al, const uint8_t *suffix_codes, uint32_t suffix_len,
                          KeyT *out_keys, uint64_t cap, cudaStream_t stream);
// keep[i] = 0 for the redundant dummy sinks initialize_chunk drops (boss_chunk.cpp:80-85)
template <typename KeyT>
void suffix_keep_flags(const KeyT *comp, uint64_t m, Alpha al, uint32_t *keep, cudaStream_t stream,
                       unsigned *launches);
// composite keys -> rows (row 0 = placeholder, row offsets[i] + 1 for every kept k-mer)
template <typename KeyT>
void suffix_rows(const KeyT *comp, const uint32_t *counts, const uint32_t *keep, const uint64_t *offsets,
                 uint64_t m, Alpha al, KeyT *row_keys, uint8_t *row_info, uint32_t *row_counts,
                 cudaStream_t stream, unsigned *launches);

// ---- K6 chunk emit -------------------------------------------------------
// Writes W / last / weights in sdsl int_vector word layout and F[sigma].
// F must be pre-set by the kernel: F[0]=0, F[c]=E-1 unless a boundary is found.
template <typename KeyT>
void emit_chunk(const KeyT *row_keys, const uint8_t *row_info, const uint32_t *row_counts,
                uint64_t n_rows, Alpha al, uint32_t weights_width, uint64_t *W, uint64_t *last,
                uint64_t *weights, unsigned long long *F, cudaStream_t stream, unsigned *launches);


// Fused merge + emit for L = 2 (DNA): W / last / F / weights straight from the sorted reals and the
// sorted dummies with their ranks; the merged rows never exist in HBM (mtg_emit.cu). `part` is
// scratch of emit_merged_scratch_words(m + d + row_off) words.
size_t emit_merged_scratch_words(uint64_t n_rows);
template <typename KeyT>
void emit_merged(const KeyT *reals, const uint32_t *real_counts, uint64_t m, const KeyT *dummy_keys,
                 const uint32_t *dummy_info, const uint64_t *rank, uint64_t d, uint32_t row_off, Alpha al,
                 uint32_t weights_width, uint64_t *part, uint64_t *W, uint64_t *last, uint64_t *weights,
                 unsigned long long *F, cudaStream_t stream, unsigned *launches);

}  // namespace mtg
