// TEST INFRASTRUCTURE ONLY -- a CPU restatement of the reference algorithm for
// the `metagraph build` (BOSS chunk construction) path. It is the checker for
// tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg; the product
// library (projects2014-metagenome_b200/csrc) never links, loads or calls it.
//
// Parity pinning: validated against the reference's own implementation compiled
// from /root/reference (oracle/_ref/ref_driver_*, see oracle/build_ref.sh) on
// tests/data/transcripts_1000.fa and synthetic reads (tests/golden/*), and
// against the golden counters in SURVEY.md 8c (538182 real k-mers, 214 sinks,
// source levels 251,251,251,251,249,242,219,157,60,16,4 for `build -k 12`).
//
// All file:line citations are relative to /root/reference/metagraph/src.
// The restatement is deliberately serial and literal: every loop follows the
// reference loop it cites, on a single fixed 256-bit word type.
#include <algorithm>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <string>
#include <utility>
#include <vector>

namespace {

// ----------------------------------------------------------------------------
// 256-bit word (stands in for uint64_t / sdsl::uint128_t / sdsl::uint256_t,
// kmer/kmer_boss.hpp:29-45). w[0] is least significant.
// ----------------------------------------------------------------------------
struct U256 {
    uint64_t w[4];
    U256() : w{0, 0, 0, 0} {}
    explicit U256(uint64_t x) : w{x, 0, 0, 0} {}
};
inline bool operator==(const U256 &a, const U256 &b) {
    return a.w[0] == b.w[0] && a.w[1] == b.w[1] && a.w[2] == b.w[2] && a.w[3] == b.w[3];
}
inline bool operator!=(const U256 &a, const U256 &b) { return !(a == b); }
inline bool operator<(const U256 &a, const U256 &b) {
    for (int i = 3; i >= 0; --i)
        if (a.w[i] != b.w[i]) return a.w[i] < b.w[i];
    return false;
}
inline bool operator<=(const U256 &a, const U256 &b) { return !(b < a); }
inline bool operator>(const U256 &a, const U256 &b) { return b < a; }
inline U256 operator|(U256 a, const U256 &b) {
    for (int i = 0; i < 4; ++i) a.w[i] |= b.w[i];
    return a;
}
inline U256 operator&(U256 a, const U256 &b) {
    for (int i = 0; i < 4; ++i) a.w[i] &= b.w[i];
    return a;
}
inline U256 operator~(U256 a) {
    for (int i = 0; i < 4; ++i) a.w[i] = ~a.w[i];
    return a;
}
inline U256 shl(const U256 &a, unsigned n) {
    U256 r;
    if (n >= 256) return r;
    unsigned ws = n / 64, bs = n % 64;
    for (int i = 3; i >= (int)ws; --i) {
        r.w[i] = a.w[i - ws] << bs;
        if (bs && i - (int)ws - 1 >= 0) r.w[i] |= a.w[i - ws - 1] >> (64 - bs);
    }
    return r;
}
inline U256 shr(const U256 &a, unsigned n) {
    U256 r;
    if (n >= 256) return r;
    unsigned ws = n / 64, bs = n % 64;
    for (unsigned i = 0; i + ws < 4; ++i) {
        r.w[i] = a.w[i + ws] >> bs;
        if (bs && i + ws + 1 < 4) r.w[i] |= a.w[i + ws + 1] << (64 - bs);
    }
    return r;
}
inline U256 add(const U256 &a, const U256 &b) {
    U256 r;
    unsigned __int128 c = 0;
    for (int i = 0; i < 4; ++i) {
        c += (unsigned __int128)a.w[i] + b.w[i];
        r.w[i] = (uint64_t)c;
        c >>= 64;
    }
    return r;
}
inline U256 low_bits(unsigned n) {  // n ones
    if (n >= 256) return ~U256();
    U256 one(1);
    U256 r = shl(one, n);
    // r - 1
    for (int i = 0; i < 4; ++i) {
        if (r.w[i]--) break;
    }
    return r;
}

// ----------------------------------------------------------------------------
// Alphabets (kmer/alphabets.hpp:27-149), rebuilt from their definitions.
// ----------------------------------------------------------------------------
struct Alphabet {
    int L1;            // bits per char, tight alphabet (no sentinel)
    int L2;            // bits per char, BOSS alphabet (with '$' = 0)
    int sigma1;        // tight alphabet size; code == sigma1 means invalid
    int sigma2;        // BOSS alphabet size incl. '$'; code == sigma2 means invalid
    uint8_t tight[128];
    uint8_t boss[128];
    std::vector<uint8_t> comp_tight;   // kComplementMapDNA (alphabets.hpp:141)
    std::vector<uint8_t> comp_boss;    // kBOSSComplementMapDNA (alphabets.hpp:78)
};

Alphabet make_dna() {
    Alphabet a;
    a.L1 = 2; a.L2 = 3; a.sigma1 = 4; a.sigma2 = 5;
    for (int i = 0; i < 128; ++i) { a.tight[i] = 4; a.boss[i] = 5; }
    const char *up = "ACGT", *lo = "acgt";
    for (int c = 0; c < 4; ++c) {
        a.tight[(int)up[c]] = c; a.tight[(int)lo[c]] = c;
        a.boss[(int)up[c]] = c + 1; a.boss[(int)lo[c]] = c + 1;
    }
    a.tight[(int)'U'] = a.tight[(int)'u'] = 3;   // RNA: U == T
    a.boss[(int)'U'] = a.boss[(int)'u'] = 4;
    a.comp_tight = {3, 2, 1, 0, 4};
    a.comp_boss = {0, 4, 3, 2, 1, 5};
    return a;
}

Alphabet make_protein() {
    Alphabet a;
    a.L1 = 5; a.L2 = 5; a.sigma1 = 26; a.sigma2 = 27;
    const std::string alph = "ABCDEFGHIJKLMNOPQRSTUVWYZX";   // alphabets.hpp:90
    for (int i = 0; i < 128; ++i) { a.tight[i] = 25; a.boss[i] = 26; }
    for (int c = 0; c < 26; ++c) {
        int up = alph[c], lo = alph[c] - 'A' + 'a';
        a.tight[up] = a.tight[lo] = c;
        a.boss[up] = a.boss[lo] = c + 1;
    }
    return a;
}

// encode() (kmer/kmer_extractor.cpp:30-34): negative chars map to table[0]
inline uint8_t encode_char(char s, const uint8_t *table) {
    return s >= 0 ? table[(size_t)s] : table[0];
}

// ----------------------------------------------------------------------------
// KMerBOSS (kmer/kmer_boss.hpp:123-212) on U256 with run-time L.
// ----------------------------------------------------------------------------
struct KB {
    int L;
    U256 mask;  // kFirstCharMask
    explicit KB(int l) : L(l), mask(low_bits(l)) {}
    // ctor kmer_boss.hpp:123-140: arr = s_1..s_k
    U256 pack(const uint8_t *arr, size_t k) const {
        U256 s;
        for (int i = (int)k - 2; i >= 0; --i) {
            s = s | U256(arr[i]);
            s = shl(s, L);
        }
        return s | U256(arr[k - 1]);
    }
    // to_next kmer_boss.hpp:147-169
    U256 to_next(U256 s, size_t k, uint8_t new_last) const {
        U256 old_last = s & mask;
        s = shr(s, L);
        s = s | shl(old_last, L * (k - 1));
        s = s & ~mask;
        return s | U256(new_last);
    }
    // to_prev kmer_boss.hpp:171-186
    U256 to_prev(U256 s, size_t k, uint8_t new_first) const {
        U256 last_char = shr(s, L * (k - 1));
        s = s & ~mask;
        s = s | U256(new_first);
        s = shl(s, L);
        s = s & low_bits(L * k);
        return s | last_char;
    }
    uint8_t at(const U256 &s, size_t i) const {   // operator[] :188-194
        return (uint8_t)(shr(s, L * i).w[0] & mask.w[0]);
    }
    bool compare_suffix(const U256 &a, const U256 &b, size_t minus = 0) const {  // :202-206
        return shr(a, (minus + 1) * L) == shr(b, (minus + 1) * L);
    }
    bool less(const U256 &a, const U256 &b, size_t minus = 0) const {  // :208-212
        return shr(a, (minus + 1) * L) < shr(b, (minus + 1) * L);
    }
};

}  // namespace

namespace {

typedef std::pair<U256, uint64_t> KC;   // (k-mer word, count)

// utils::drag_and_mark_segments (common/algorithms.hpp:50-67)
std::vector<bool> drag_and_mark_segments(const std::vector<uint8_t> &array,
                                         uint8_t label, size_t segment_length) {
    std::vector<bool> mask(array.size(), false);
    size_t last_occurrence
        = std::find(array.begin(), array.end(), label) - array.begin();
    for (size_t i = last_occurrence; i < array.size(); ++i) {
        if (array[i] == label) last_occurrence = i;
        if (i - last_occurrence < segment_length) mask[i] = true;
    }
    return mask;
}

// KMerBOSS::match_suffix (kmer/kmer_boss.hpp:106-112)
bool match_suffix(const uint8_t *kmer, size_t k, const std::vector<uint8_t> &suffix) {
    return suffix.empty()
        || std::equal(suffix.begin(), suffix.end(), kmer + k - suffix.size() - 1);
}

// ::sequence_to_kmers (kmer/kmer_extractor.cpp:203-237) with the slide variants
// (:87-108, :166-196) -- the non-slide variants (:69-84, :111-141) produce the
// same k-mers and are folded into the same loops here.
template <class Skip>
void sequence_to_kmers(const KB &kb, const uint8_t *begin, const uint8_t *end, size_t k,
                       const std::vector<uint8_t> &suffix, std::vector<U256> *out,
                       const std::vector<uint8_t> &complement_code, Skip skip) {
    if (k > (size_t)(end - begin)) return;
    size_t n = end - begin;
    if (complement_code.empty()) {
        U256 kmer = kb.pack(begin, k);
        if (!skip() && match_suffix(begin, k, suffix)) out->push_back(kmer);
        for (const uint8_t *kb_begin = begin + 1; kb_begin + k <= end; ++kb_begin) {
            kmer = kb.to_next(kmer, k, kb_begin[k - 1]);
            if (!skip() && match_suffix(kb_begin, k, suffix)) out->push_back(kmer);
        }
    } else {
        // reverse_complement of the ENCODED read (kmer_extractor.cpp:46-62)
        std::vector<uint8_t> rc(begin, end);
        std::reverse(rc.begin(), rc.end());
        for (auto &c : rc) c = complement_code[c];
        auto call_smallest = [&](const U256 &f, const uint8_t *fp, const U256 &r,
                                 const uint8_t *rp) {   // __call_smallest :143-161
            if (skip()) return;
            if (f <= r) {
                if (match_suffix(fp, k, suffix)) out->push_back(f);
            } else {
                if (match_suffix(rp, k, suffix)) out->push_back(r);
            }
        };
        U256 kmer = kb.pack(begin, k);
        U256 rev = kb.pack(&rc[n - k], k);
        call_smallest(kmer, begin, rev, &rc[n - k]);
        for (size_t forward_last = k, reverse_first = n - k - 1; forward_last < n;
             ++forward_last, --reverse_first) {
            kmer = kb.to_next(kmer, k, begin[forward_last]);
            rev = kb.to_prev(rev, k, rc[reverse_first]);
            call_smallest(kmer, begin + forward_last - (k - 1), rev, &rc[reverse_first]);
        }
    }
}

// KmerExtractorT<L>::sequence_to_kmers == KmerExtractor2Bit
// (kmer/kmer_extractor.cpp:472-507): tight alphabet, no '$' framing.
void extract_2bit(const Alphabet &al, const char *s, size_t len, size_t k,
                  bool canonical, std::vector<U256> *out) {
    if (len < k) return;
    std::vector<uint8_t> seq(len);
    for (size_t i = 0; i < len; ++i) seq[i] = encode_char(s[i], al.tight);
    std::vector<bool> invalid = drag_and_mark_segments(seq, al.sigma1, k);
    std::replace(seq.begin(), seq.end(), (uint8_t)al.sigma1, (uint8_t)0);
    size_t i = k - 1;
    KB kb(al.L1);
    sequence_to_kmers(kb, seq.data(), seq.data() + seq.size(), k, {}, out,
                      canonical ? al.comp_tight : std::vector<uint8_t>(),
                      [&]() { return (bool)invalid[i++]; });
}

// KmerExtractorBOSS::sequence_to_kmers (kmer/kmer_extractor.cpp:317-381):
// BOSS alphabet, '$'^(k-1) prefix and one '$' suffix per valid segment when a
// suffix filter is given.
void extract_boss(const Alphabet &al, const char *s, size_t len, size_t k,
                  const std::vector<uint8_t> &suffix, std::vector<U256> *out) {
    if (len < k) return;
    const size_t dummy_prefix_size = suffix.empty() ? 0 : k - 1;
    const size_t dummy_suffix_size = suffix.empty() ? 0 : 1;
    std::vector<uint8_t> seq(dummy_prefix_size + len + 1, (uint8_t)al.sigma2);
    for (size_t i = 0; i < len; ++i)
        seq[dummy_prefix_size + i] = encode_char(s[i], al.boss);
    uint8_t *end_segm = seq.data() + dummy_prefix_size;
    uint8_t *end = seq.data() + seq.size();
    KB kb(al.L2);
    do {
        uint8_t *begin_segm = end_segm - dummy_prefix_size;
        while (*end_segm < al.sigma2) end_segm++;
        if (begin_segm + dummy_prefix_size + k <= end_segm) {
            if (dummy_suffix_size) {
                *end_segm = 0;
                std::fill(begin_segm, begin_segm + dummy_prefix_size, 0);
            }
            sequence_to_kmers(kb, begin_segm, end_segm + dummy_suffix_size, k, suffix, out,
                              std::vector<uint8_t>(), []() { return false; });
        }
    } while (++end_segm + k + dummy_suffix_size <= end);
}

// string reverse complement used by Mode::BOTH (kmer/kmer_collector.cpp:47-52,
// common/seq_tools/reverse_complement.hpp:33-64). Only the DNA letters matter:
// every other character stays outside the DNA alphabet after complementing.
std::string rc_string(const char *s, size_t len) {
    std::string r(len, 'N');
    for (size_t i = 0; i < len; ++i) {
        char c = s[len - 1 - i], o;
        switch (c) {
            case 'A': o = 'T'; break; case 'C': o = 'G'; break;
            case 'G': o = 'C'; break; case 'T': o = 'A'; break; case 'U': o = 'A'; break;
            case 'a': o = 't'; break; case 'c': o = 'g'; break;
            case 'g': o = 'c'; break; case 't': o = 'a'; break; case 'u': o = 'a'; break;
            default: o = 'N';
        }
        r[i] = o;
    }
    return r;
}

}  // namespace

namespace {

struct LessFirst {
    bool operator()(const KC &a, const KC &b) const { return a.first < b.first; }
};

// SortedSet::sort_and_remove_duplicates (common/sorted_sets/sorted_set.cpp:41-48)
// and SortedMultiset::sort_and_merge_counts (sorted_multiset.cpp:54-84):
// sort by key, merge equal keys with saturation at max_count.
void sort_and_merge_counts(std::vector<KC> *data, uint64_t max_count) {
    if (data->empty()) return;
    std::sort(data->begin(), data->end(), LessFirst());
    auto first = data->begin(), last = data->end(), dest = first;
    while (++first != last) {
        if (first->first == dest->first) {
            if (first->second < max_count - dest->second) {
                dest->second += first->second;
            } else {
                dest->second = max_count;
            }
        } else {
            *++dest = *first;
        }
    }
    data->erase(++dest, data->end());
}

// kmer::reverse_complement (kmer/kmer_transform.hpp:14-35)
U256 kmer_reverse_complement(const KB &kb, size_t k, U256 kmer,
                             const std::vector<uint8_t> &complement_code) {
    const uint64_t mask = kb.mask.w[0];
    U256 last_two(complement_code[kmer.w[0] & mask]);
    kmer = shr(kmer, kb.L);
    last_two = shl(last_two, kb.L) | U256(complement_code[kmer.w[0] & mask]);
    kmer = shr(kmer, kb.L);
    U256 result;
    for (uint32_t i = 2; i < k; ++i) {
        result = shl(result, kb.L) | U256(complement_code[kmer.w[0] & mask]);
        kmer = shr(kmer, kb.L);
    }
    return shl(result, 2 * kb.L) | last_two;
}

// add_reverse_complements (boss_chunk_construct.cpp:179-222). `k` is the node
// length; count_max is the container count type's max (255 / 65535 / 2^32-1).
void add_reverse_complements(const KB &kb, size_t k, std::vector<KC> *kmers,
                             const std::vector<uint8_t> &complement_code,
                             bool with_counts, uint64_t count_max) {
    size_t size = kmers->size();
    for (size_t i = 0; i < size; ++i) {
        KC &km = (*kmers)[i];
        U256 rc = kmer_reverse_complement(kb, k + 1, km.first, complement_code);
        if (rc != km.first) {
            kmers->push_back(KC(rc, km.second));
        } else if (with_counts) {
            // count >> (bits-1) ? max : 2*count  (:205-211)
            if (km.second > count_max / 2) km.second = count_max;
            else km.second *= 2;
        }
    }
    std::sort(kmers->begin(), kmers->end(), LessFirst());
}

// get_sentinel_delta (kmer/kmer_transform.hpp:38-47)
U256 get_sentinel_delta(int char_width, size_t k) {
    U256 word;
    for (size_t i = 0; i < k; ++i) word = shl(word, char_width) | U256(1);
    return word;
}

// transform<KMER_TO>(kmer, k) (kmer/kmer_transform.hpp:75-100): re-pack every
// L1-bit char into an L2-bit slot (the LUT specialisations :102-165 are
// equivalent to this generic loop, tests/kmer/test_transform.cpp).
U256 transform(const KB &from, const KB &to, const U256 &kmer, size_t k) {
    U256 r;
    for (size_t i = 0; i < k; ++i)
        r = r | shl(U256(from.at(kmer, i)), to.L * i);
    return r;
}

// add_dummy_sink_kmers (boss_chunk_construct.cpp:54-98)
void add_dummy_sink_kmers(const Alphabet &al, size_t k, const std::vector<KC> &kmers,
                          std::vector<U256> *dummy_kmers) {
    KB real(al.L1), fin(al.L2);
    U256 kmer_delta = get_sentinel_delta(al.L2, k + 1);
    kmer_delta = kmer_delta & ~fin.mask;
    const size_t alphabet_size = al.sigma1;
    std::vector<size_t> max_it(alphabet_size), it(alphabet_size);
    for (size_t c = 1; c < alphabet_size; ++c) {
        std::vector<uint8_t> zeros(k + 1, 0);
        zeros[k - 1] = c;
        U256 probe = real.pack(zeros.data(), k + 1);
        it[c] = std::lower_bound(kmers.begin(), kmers.end(), probe,
                                 [](const KC &a, const U256 &b) { return a.first < b; })
                - kmers.begin();
        max_it[c - 1] = it[c];
    }
    max_it[alphabet_size - 1] = kmers.size();
    std::vector<U256> last_shifted(alphabet_size, U256(1));
    for (size_t i = 0; i < kmers.size(); ++i) {
        uint8_t last_char = real.at(kmers[i].first, 0);
        U256 shifted = real.to_next(kmers[i].first, k + 1, 0);
        while (it[last_char] < max_it[last_char]
               && kmers[it[last_char]].first < shifted) {
            it[last_char]++;
        }
        if (last_shifted[last_char] != shifted
            && (it[last_char] == max_it[last_char]
                || !real.compare_suffix(kmers[it[last_char]].first, shifted))) {
            dummy_kmers->push_back(add(transform(real, fin, shifted, k + 1), kmer_delta));
            last_shifted[last_char] = shifted;
        }
    }
}

// add_dummy_source_kmers (boss_chunk_construct.cpp:123-168)
void add_dummy_source_kmers(const Alphabet &al, size_t k, const std::vector<KC> &kmers,
                            std::vector<U256> *dummy_kmers) {
    KB real(al.L1), fin(al.L2);
    U256 kmer_delta = get_sentinel_delta(al.L2, k + 1);
    kmer_delta = kmer_delta & ~shl(fin.mask, al.L2);
    size_t real_it = 0;
    for (size_t i = 0; i < kmers.size(); ++i) {
        const U256 &kmer = kmers[i].first;
        if (i > 0 && real.at(kmer, k) != real.at(kmers[i - 1].first, k)) real_it = 0;
        if (i > 0 && real.compare_suffix(kmer, kmers[i - 1].first)) continue;
        U256 prev_kmer = real.to_prev(kmer, k + 1, 0);
        while (real_it < kmers.size() && real.less(kmers[real_it].first, prev_kmer, 1))
            real_it++;
        bool is_redundant = false;
        while (real_it < kmers.size()
               && real.compare_suffix(kmers[real_it].first, prev_kmer, 1)) {
            if (real.at(kmers[real_it].first, 0) == real.at(prev_kmer, 0)) {
                is_redundant = true;
                break;
            }
            real_it++;
        }
        if (!is_redundant)
            dummy_kmers->push_back(add(transform(real, fin, prev_kmer, k + 1), kmer_delta));
    }
}

}  // namespace

namespace {

struct Chunk {
    std::vector<uint8_t> W, last;
    std::vector<uint64_t> weights, F;
    uint64_t n_real = 0, n_sink = 0;
    std::vector<uint64_t> n_src;          // dummy sources per prefix length 1..k
    std::vector<KC> real_sorted;          // tight words after sort/unique(/rc)
};

// initialize_chunk (graph/representation/succinct/boss_chunk.cpp:31-133).
// `kmers` are final BOSS-alphabet words with counts, sorted; k = node length.
void initialize_chunk(const Alphabet &al, const std::vector<KC> &kmers, size_t k,
                      int bits_per_count, Chunk *out) {
    KB fin(al.L2);
    const uint64_t alph_size = al.sigma2;
    uint64_t max_count = bits_per_count ? (bits_per_count >= 64 ? ~0ull
                                           : ((1ull << bits_per_count) - 1)) : 0;
    out->W.assign(1, 0);
    out->last.assign(1, 0);
    out->weights.clear();
    if (bits_per_count) out->weights.push_back(0);
    out->F.assign(alph_size, 0);
    size_t curpos = 1;
    uint64_t lastF = 0;
    std::vector<U256> last_kmer(alph_size);
    size_t it = 0, end = kmers.size();
    while (it != end) {
        const KC &value = kmers[it];
        const U256 &kmer = value.first;
        uint64_t curW = fin.at(kmer, 0);
        uint64_t curF = fin.at(kmer, k);
        ++it;
        if (it != end && fin.compare_suffix(kmer, kmers[it].first)) {
            if (curW == 0 && curF > 0) continue;   // redundant dummy sink (:80-85)
            out->last.push_back(0);
        } else {
            out->last.push_back(1);
        }
        if (curW) {
            if (last_kmer[curW] != U256() && fin.compare_suffix(kmer, last_kmer[curW], 1)) {
                curW += alph_size;
            } else {
                last_kmer[curW] = kmer;
            }
        }
        out->W.push_back((uint8_t)curW);
        while (curF > lastF && lastF + 1 < alph_size) out->F[++lastF] = curpos - 1;
        if (bits_per_count) {
            if (value.second && curW && fin.at(kmer, 1))
                out->weights.push_back(std::min(value.second, max_count));
            else
                out->weights.push_back(0);
        }
        curpos++;
    }
    while (++lastF < alph_size) out->F[lastF] = curpos - 1;
}

// container count type max: KmerMultsetVector8/16/32
// (boss_chunk_construct.cpp:1147-1160)
uint64_t container_count_max(int bits_per_count) {
    if (bits_per_count <= 8) return 0xFFull;
    if (bits_per_count <= 16) return 0xFFFFull;
    return 0xFFFFFFFFull;
}

// The whole path: KmerCollector (kmer/kmer_collector.cpp:26-127) ->
// BOSSChunkConstructor::build_chunk (boss_chunk_construct.cpp:996-1040) ->
// construct_boss_chunk (:248-356) -> BOSS::Chunk ctor / initialize_chunk.
int build(const Alphabet &al, const char *bases, const uint64_t *offsets,
          const uint64_t *counts, uint64_t n_reads, size_t K, bool both_strands,
          int bits_per_count, const std::vector<uint8_t> &suffix, Chunk *out) {
    const size_t k = K - 1;
    const bool with_counts = bits_per_count > 0;
    const uint64_t cmax = container_count_max(bits_per_count);
    std::vector<KC> kmers;
    std::vector<U256> buf;
    KB real(al.L1), fin(al.L2);

    if (suffix.empty()) {
        // KmerExtractor2Bit collector, Mode CANONICAL_ONLY or BASIC (:961-971)
        for (uint64_t r = 0; r < n_reads; ++r) {
            const char *s = bases + offsets[r];
            size_t len = offsets[r + 1] - offsets[r];
            uint64_t c = counts ? std::min(counts[r], cmax) : 1;   // kmer_collector.cpp:92
            buf.clear();
            extract_2bit(al, s, len, K, both_strands, &buf);
            for (auto &x : buf) kmers.push_back(KC(x, c));
        }
        sort_and_merge_counts(&kmers, with_counts ? cmax : ~0ull);
        if (both_strands)
            add_reverse_complements(real, k, &kmers, al.comp_tight, with_counts, cmax);
        out->real_sorted = kmers;
        out->n_real = kmers.size();

        std::vector<U256> dummy;
        add_dummy_sink_kmers(al, k, kmers, &dummy);
        out->n_sink = dummy.size();
        size_t dummy_source_begin = dummy.size();
        add_dummy_source_kmers(al, k, kmers, &dummy);
        std::sort(dummy.begin() + dummy_source_begin, dummy.end());
        out->n_src.assign(1, dummy.size() - dummy_source_begin);
        for (size_t c = 2; c < k + 1; ++c) {   // :280-303
            size_t dummy_end = dummy.size();
            for (size_t i = dummy_source_begin; i < dummy_end; ++i) {
                U256 kmer = dummy[i];
                if (i > 0 && fin.compare_suffix(kmer, dummy[i - 1])) continue;
                dummy.push_back(fin.to_prev(kmer, k + 1, 0));
            }
            dummy_source_begin = dummy_end;
            std::sort(dummy.begin() + dummy_source_begin, dummy.end());
            out->n_src.push_back(dummy.size() - dummy_source_begin);
        }
        std::sort(dummy.begin(), dummy.end());

        // merge (:308-348)
        std::vector<KC> all;
        all.reserve(kmers.size() + dummy.size() + 1);
        all.push_back(KC(U256(), 0));
        const U256 kmer_delta = get_sentinel_delta(al.L2, k + 1);
        size_t di = 0;
        for (const KC &km : kmers) {
            U256 lifted = add(transform(real, fin, km.first, k + 1), kmer_delta);
            while (di < dummy.size() && lifted > dummy[di]) all.push_back(KC(dummy[di++], 0));
            all.push_back(KC(lifted, km.second));
        }
        for (; di < dummy.size(); ++di) all.push_back(KC(dummy[di], 0));
        initialize_chunk(al, all, k, bits_per_count, out);
    } else {
        // KmerExtractorBOSS collector, Mode BOTH or BASIC; the chunk is built
        // straight from the sorted collector contents (:1002-1013)
        bool all_dollar = true;
        for (uint8_t c : suffix) all_dollar &= (c == 0);
        if (all_dollar) kmers.push_back(KC(U256(), 1));   // add_kmer(all '$') :978-981
        for (uint64_t r = 0; r < n_reads; ++r) {
            const char *s = bases + offsets[r];
            size_t len = offsets[r + 1] - offsets[r];
            if (len < K) continue;   // KmerCollector::add_sequence kmer_collector.hpp:69-79
            uint64_t c = counts ? std::min(counts[r], cmax) : 1;
            buf.clear();
            extract_boss(al, s, len, K, suffix, &buf);
            if (both_strands) {
                std::string rc = rc_string(s, len);
                extract_boss(al, rc.data(), rc.size(), K, suffix, &buf);
            }
            for (auto &x : buf) kmers.push_back(KC(x, c));
        }
        sort_and_merge_counts(&kmers, with_counts ? cmax : ~0ull);
        out->real_sorted = kmers;
        out->n_real = kmers.size();
        initialize_chunk(al, kmers, k, bits_per_count, out);
    }
    return 0;
}

}  // namespace

// ----------------------------------------------------------------------------
// C ABI for the tests (ctypes). alphabet: 0 = DNA, 1 = Protein.
// suffix: ASCII over the BOSS alphabet ("$ACGT"), may be NULL/"".
// ----------------------------------------------------------------------------
extern "C" {

struct mtg_oracle_chunk {
    uint64_t n_rows;        // rows incl. the index-0 placeholder
    uint64_t sigma;         // BOSS alphabet size (length of F)
    uint64_t n_real, n_sink, n_src_levels;
    const uint8_t *W;       // one byte per row
    const uint8_t *last;    // one byte per row
    const uint64_t *weights;  // one u64 per row, or NULL
    const uint64_t *F;
    const uint64_t *n_src;  // per dummy prefix length
    uint64_t n_kmers;       // sorted real k-mers (collector alphabet)
    const uint64_t *kmer_words;   // 4 u64 per k-mer, least significant first
    const uint64_t *kmer_counts;
    void *impl;
};

struct OracleImpl {
    Chunk chunk;
    std::vector<uint64_t> words, counts;
};

int mtg_oracle_build(const char *bases, const uint64_t *offsets, const uint64_t *counts,
                     uint64_t n_reads, uint32_t K, int alphabet, int both_strands,
                     int bits_per_count, const char *suffix, mtg_oracle_chunk *out) {
    Alphabet al = alphabet == 1 ? make_protein() : make_dna();
    if (K < 2 || K * al.L2 > 256) return 1;
    std::vector<uint8_t> suf;
    if (suffix) {
        const std::string boss_alph = alphabet == 1 ? "$ABCDEFGHIJKLMNOPQRSTUVWYZX" : "$ACGT";
        for (const char *p = suffix; *p; ++p) {
            size_t pos = boss_alph.find(*p);
            if (pos == std::string::npos) return 2;
            suf.push_back((uint8_t)pos);
        }
    }
    OracleImpl *impl = new OracleImpl();
    build(al, bases, offsets, counts, n_reads, K, both_strands != 0, bits_per_count, suf,
          &impl->chunk);
    Chunk &c = impl->chunk;
    for (auto &kc : c.real_sorted) {
        for (int i = 0; i < 4; ++i) impl->words.push_back(kc.first.w[i]);
        impl->counts.push_back(kc.second);
    }
    out->n_rows = c.W.size();
    out->sigma = al.sigma2;
    out->n_real = c.n_real;
    out->n_sink = c.n_sink;
    out->n_src_levels = c.n_src.size();
    out->W = c.W.data();
    out->last = c.last.data();
    out->weights = bits_per_count ? c.weights.data() : nullptr;
    out->F = c.F.data();
    out->n_src = c.n_src.data();
    out->n_kmers = c.real_sorted.size();
    out->kmer_words = impl->words.data();
    out->kmer_counts = impl->counts.data();
    out->impl = impl;
    return 0;
}

void mtg_oracle_free(mtg_oracle_chunk *c) {
    delete (OracleImpl *)c->impl;
    c->impl = nullptr;
}

}  // extern "C"
