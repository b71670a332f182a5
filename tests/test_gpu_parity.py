"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same
seeded inputs. Bit-exact (integer / byte work): W, last, F, weights, counters."""
import ctypes

import numpy as np
import pytest

import mtg_b200
import oracle_lib
from mtg_b200 import synth

pytestmark = pytest.mark.gpu


def dna_reads(seed, n, length, G, e=0.01, n_frac=0.0):
    packed = synth.dna_reads_packed_np(seed, n, length, G, e)
    reads = synth.packed_to_strings(packed)
    if n_frac:
        rng = np.random.default_rng(seed)
        out = []
        for r in reads:
            b = bytearray(r.encode())
            for p in np.nonzero(rng.random(len(b)) < n_frac)[0]:
                b[p] = ord("N")
            out.append(b.decode())
        reads = out
    return reads


def gpu_build(reads, K, alphabet="DNA", canonical=False, bits_per_count=0, counts=None, suffix=""):
    ctor = mtg_b200.BOSSChunkConstructor.initialize(K - 1, canonical, bits_per_count, suffix,
                                                    alphabet=alphabet)
    if counts is not None:
        ctor.add_sequences(list(zip(reads, [int(c) for c in counts])))
    else:
        # several pushes, like the reference's 1 MB batches
        step = max(1, len(reads) // 3)
        for i in range(0, len(reads), step):
            ctor.add_sequences(reads[i:i + step])
    chunk = ctor.build_chunk()
    ctor.close()
    return chunk


def check(reads, K, alphabet="DNA", canonical=False, bits_per_count=0, counts=None, suffix=""):
    chunk = gpu_build(reads, K, alphabet, canonical, bits_per_count, counts, suffix)
    ref = oracle_lib.oracle_build(reads, K, alphabet=alphabet, canonical=canonical,
                                  bits_per_count=bits_per_count, counts=counts, suffix=suffix)
    assert chunk.n_rows == ref["n_rows"]
    assert chunk.stats["n_real"] == ref["n_real"]
    assert chunk.stats["n_dummy_sink"] == ref["n_sink"]
    assert chunk.stats["n_dummy_source"][: len(ref["n_src"])] == [int(x) for x in ref["n_src"]]
    assert np.array_equal(chunk.F, ref["F"])
    assert np.array_equal(chunk.last, ref["last"])
    assert np.array_equal(chunk.W, ref["W"])
    if bits_per_count:
        assert np.array_equal(chunk.weights, ref["weights"])
    else:
        assert chunk.weights is None
    # sdsl file bytes == what the reference's BOSS::Chunk::serialize would write
    assert (chunk._int_vector_bytes(chunk.W_words, chunk.n_rows, chunk.W_width, False)
            == oracle_lib.sdsl_int_vector_bytes(ref["W"], chunk.W_width))
    assert (chunk._int_vector_bytes(chunk.last_words, chunk.n_rows, 1, True)
            == oracle_lib.sdsl_int_vector_bytes(ref["last"], 1, True))
    return chunk


# ---- radix sort primitive ---------------------------------------------------

@pytest.mark.parametrize("n", [0, 1, 100, 4096, 4097, 50_000, 1_000_003])
@pytest.mark.parametrize("bits", [7, 24, 62, 64])
def test_radix_sort_u64(n, bits):
    rng = np.random.default_rng(n + bits)
    keys = rng.integers(0, 2 ** 63, n, dtype=np.uint64) * 2 + rng.integers(0, 2, n, dtype=np.uint64)
    if bits < 64:
        keys &= np.uint64((1 << bits) - 1)
    vals = np.arange(n, dtype=np.uint32)
    lib = mtg_b200._lib.load()
    k2 = keys.copy()
    assert lib.mtg_debug_sort_u64(0, k2.ctypes.data_as(ctypes.c_void_p), None, n, bits) == 0
    assert np.array_equal(k2, np.sort(keys))
    k3, v3 = keys.copy(), vals.copy()
    assert lib.mtg_debug_sort_u64(0, k3.ctypes.data_as(ctypes.c_void_p),
                                  v3.ctypes.data_as(ctypes.c_void_p), n, bits) == 0
    order = np.argsort(keys, kind="stable")
    assert np.array_equal(k3, keys[order])
    assert np.array_equal(v3, vals[order])   # stability


@pytest.mark.parametrize("n", [0, 5, 2048, 2049, 300_001])
@pytest.mark.parametrize("bits", [70, 126, 128])
def test_radix_sort_u128(n, bits):
    rng = np.random.default_rng(n + bits)
    lo = rng.integers(0, 2 ** 63, n, dtype=np.uint64) * 2 + rng.integers(0, 2, n, dtype=np.uint64)
    hi = rng.integers(0, 2 ** 63, n, dtype=np.uint64) * 2 + rng.integers(0, 2, n, dtype=np.uint64)
    if bits < 128:
        hi &= np.uint64((1 << (bits - 64)) - 1)
    if n > 10:
        hi[: n // 2] = hi[0]   # ties on the high word
        lo[: n // 4] = lo[0]
    keys = np.stack([lo, hi], axis=1).copy()
    vals = np.arange(n, dtype=np.uint32)
    lib = mtg_b200._lib.load()
    assert lib.mtg_debug_sort_u128(0, keys.ctypes.data_as(ctypes.c_void_p),
                                   vals.ctypes.data_as(ctypes.c_void_p), n, bits) == 0
    order = np.lexsort((lo, hi))   # stable, hi is the primary key
    assert np.array_equal(keys[:, 0], lo[order])
    assert np.array_equal(keys[:, 1], hi[order])
    assert np.array_equal(vals, np.arange(n, dtype=np.uint32)[order])


# ---- sorted k-mer set (extraction + sort + unique/count + rc) ------------------

@pytest.mark.parametrize("K,canonical,bpc", [(12, False, 0), (31, True, 0), (31, True, 8),
                                             (32, False, 16), (33, True, 8), (63, True, 0),
                                             (64, False, 8)])
def test_sorted_kmers(K, canonical, bpc):
    reads = dna_reads(3, 1500, 120, 15000, n_frac=0.003)
    ctor = mtg_b200.BOSSChunkConstructor.initialize(K - 1, canonical, bpc)
    ctor.add_sequences(reads)
    words, counts = ctor.debug_sorted_kmers()
    ctor.close()
    ref = oracle_lib.oracle_build(reads, K, canonical=canonical, bits_per_count=bpc)
    assert np.array_equal(words, ref["kmers"][:, :2])
    assert not ref["kmers"][:, 2:].any()
    if bpc:
        assert np.array_equal(counts.astype(np.uint64), ref["kmer_counts"])


# ---- whole chunk ---------------------------------------------------------------

@pytest.mark.parametrize("K", [2, 3, 4, 5, 8, 12, 20, 31, 32])
@pytest.mark.parametrize("canonical", [False, True])
def test_chunk_dna_u64(K, canonical):
    check(dna_reads(K, 800, 90, 6000, n_frac=0.002), K, canonical=canonical)


@pytest.mark.parametrize("K,canonical,bpc", [(33, False, 0), (40, True, 8), (63, True, 0),
                                             (64, False, 12), (64, True, 0)])
def test_chunk_dna_u128(K, canonical, bpc):
    check(dna_reads(K, 600, 150, 8000), K, canonical=canonical, bits_per_count=bpc)


@pytest.mark.parametrize("bpc", [1, 2, 8, 12, 16, 32])
@pytest.mark.parametrize("canonical", [False, True])
def test_chunk_weighted(bpc, canonical):
    # deep coverage of a tiny genome: counts saturate for small widths
    check(dna_reads(bpc, 3000, 80, 1200, e=0.002), 21, canonical=canonical, bits_per_count=bpc)


def test_chunk_explicit_counts():
    reads = dna_reads(9, 1000, 100, 5000)
    rng = np.random.default_rng(1)
    counts = rng.integers(1, 400, len(reads)).astype(np.uint64)
    counts[:5] = [2 ** 40, 255, 256, 65535, 65536]   # clipped to the container type first
    for bpc in (8, 12, 32):
        check(reads, 25, canonical=True, bits_per_count=bpc, counts=counts)


def test_chunk_palindromes_and_lowercase():
    reads = ["ACGT" * 20, "acgtacgtnnACGTACGTACGTTTTTTTTTTTTAAAAAAAAAAAA", "TTTTTTTTTTTTTTTTTTTT",
             "AAAAAAAAAAAAAAAAAAAAAAAA", "ACGUACGUACGUACGUUUU", "GATTACA", "", "AC", "NNNNNNNNNNNNNNNN"]
    for K in (2, 4, 6, 8):
        for canonical in (False, True):
            check(reads, K, canonical=canonical, bits_per_count=8)


def test_chunk_empty_and_short_inputs():
    for reads in ([], ["ACG"], ["ACGTN" * 3], ["", "", ""]):
        check(reads, 10)
        check(reads, 10, canonical=True, bits_per_count=8)


@pytest.mark.parametrize("K", [2, 3, 7, 12, 13, 25])
def test_chunk_protein(K):
    packed = synth.protein_reads_packed_np(4, 300, 120)
    reads = synth.packed_to_strings(packed)
    reads[0] = reads[0][:50] + "*?xbZ" + reads[0][55:]   # unknown characters are 'X'
    check(reads, K, alphabet="Protein", bits_per_count=8 if K % 2 else 0)


def test_chunk_multi_tile():
    # > 1 extraction tile, many sort tiles, several emit tiles
    reads = dna_reads(77, 20000, 150, 400_000, e=0.005)
    c = check(reads, 31, canonical=True, bits_per_count=8)
    assert c.stats["n_input_kmers"] == 20000 * 120


def test_errors():
    with pytest.raises(mtg_b200.MtgError):
        mtg_b200.BOSSChunkConstructor.initialize(0)
    with pytest.raises(mtg_b200.MtgError):
        mtg_b200.BOSSChunkConstructor.initialize(85)
    with pytest.raises(mtg_b200.MtgError):
        mtg_b200.BOSSChunkConstructor.initialize(20, bits_per_count=33)
    with pytest.raises(mtg_b200.MtgError):
        mtg_b200.BOSSChunkConstructor.initialize(20, both_strands=True, alphabet="Protein")


# ---- golden vectors of the reference's own implementation (tests/golden/*.npz) ----

from golden_cases import CASES, case_reads  # noqa: E402
from test_oracle_golden import assert_chunk_equals_golden, load_golden  # noqa: E402


@pytest.mark.parametrize("name", sorted(CASES))
def test_gpu_matches_reference_golden(name):
    case = CASES[name]
    reads, counts = case_reads(case)
    chunk = gpu_build(reads, case["K"], case.get("alphabet", "DNA"), case.get("canonical", False),
                      case.get("bpc", 0), counts, case.get("suffix", ""))
    res = {"W": chunk.W, "last": chunk.last, "F": chunk.F, "weights": chunk.weights}
    assert_chunk_equals_golden(res, load_golden(name), case.get("bpc", 0))


# ---- mid-sort folding of adjacent duplicates (large-input path, forced on small inputs) ----

@pytest.mark.parametrize("K,canonical,bpc,explicit", [(31, True, 0, False), (31, False, 8, False),
                                                       (25, True, 12, True), (63, True, 8, False),
                                                       (40, False, 0, False), (12, False, 2, False)])
def test_midsort_fold_path(K, canonical, bpc, explicit, monkeypatch):
    monkeypatch.setenv("MTG_MIDSORT_MIN", "1")
    reads = dna_reads(K + bpc, 2500, 110, 3000, e=0.004)   # ~80x coverage: many duplicates
    counts = None
    if explicit:
        counts = np.random.default_rng(3).integers(1, 300, len(reads)).astype(np.uint64)
    c = check(reads, K, canonical=canonical, bits_per_count=bpc, counts=counts)
    monkeypatch.setenv("MTG_MIDSORT_MIN", str(1 << 60))
    d = check(reads, K, canonical=canonical, bits_per_count=bpc, counts=counts)
    assert np.array_equal(c.W_words, d.W_words)


# ---- suffix-filtered chunks: `build --suffix S` (KmerExtractorBOSS) ----

@pytest.mark.parametrize("K,suffix,canonical,bpc", [
    (10, "$", False, 0), (10, "A", False, 0), (10, "T", True, 8), (10, "$$", False, 0),
    (10, "$C", True, 0), (10, "GT", True, 8), (10, "A$", False, 0), (5, "$$$$", False, 4),
    (5, "ACGT", True, 0), (31, "$", True, 8), (31, "CA", False, 0), (31, "$G", True, 0),
    (28, "TT", False, 16), (29, "G", True, 0), (45, "AC", True, 2), (60, "$", False, 0),
    (3, "$", True, 8), (2, "$", False, 0), (2, "C", True, 0)])
def test_suffix_chunk_matches_oracle(K, suffix, canonical, bpc):
    reads = dna_reads(K * 7 + len(suffix), 300, 90, 2500, e=0.01, n_frac=0.004)
    reads += ["ACGT"[i % 4] * (K - 1) for i in range(4)] + ["A" * K, "C" * (K + 1), "N" * 40, ""]
    check(reads, K, canonical=canonical, bits_per_count=bpc, suffix=suffix)


def test_suffix_chunk_explicit_counts_and_protein():
    reads = dna_reads(77, 200, 80, 1500, e=0.01, n_frac=0.005)
    counts = np.random.default_rng(5).integers(1, 300, len(reads)).astype(np.uint64)
    check(reads, 12, canonical=True, bits_per_count=8, counts=counts, suffix="$A")
    check(reads, 12, canonical=False, bits_per_count=12, counts=counts, suffix="G")
    prot = synth.packed_to_strings(synth.protein_reads_packed_np(3, 80, 70))
    prot[0] = prot[0][:10] + "*?x" + prot[0][13:]
    for suffix in ("$", "A", "$K", "XL"):
        check(prot, 8, alphabet="Protein", suffix=suffix)


def test_suffix_chunks_partition_the_kmer_set():
    """All suffixes of length 1 together hold every framed k-mer exactly once (what
    `build -l 1` + `concatenate` relies on, cli/build.cpp:100-138)."""
    reads = dna_reads(91, 250, 100, 3000, e=0.01, n_frac=0.003)
    total = 0
    for sfx in "$ACGT":
        c = check(reads, 15, suffix=sfx)
        total += c.stats["n_real"]
    ref = sum(oracle_lib.oracle_build(reads, 15, suffix=s)["n_real"] for s in "$ACGT")
    assert total == ref and total > 0


def test_suffix_rejects_bad_arguments():
    with pytest.raises(mtg_b200.MtgError):
        mtg_b200.BOSSChunkConstructor.initialize(4, filter_suffix="ACGTA")   # not shorter than k + 1
    with pytest.raises(mtg_b200.MtgError):
        mtg_b200.BOSSChunkConstructor.initialize(10, filter_suffix="AN")     # outside the alphabet


# ---- the reference binary itself, at a size well above the golden fixtures -----------------

@pytest.mark.parametrize("n_reads,K,canonical,bpc", [(200_000, 31, True, 8), (60_000, 31, False, 0),
                                                     (40_000, 63, True, 0)])
def test_chunk_files_equal_reference_binary(n_reads, K, canonical, bpc, tmp_path):
    """Files written from the GPU chunk == files the reference's own code writes for the same
    reads (oracle/_ref/ref_driver_DNA: the unmodified BOSSChunkConstructor + Chunk::serialize,
    compiled by oracle/build_ref.sh; it travels to the GPU box prebuilt)."""
    import subprocess
    drv = oracle_lib.ref_driver("DNA")
    if not drv:
        pytest.skip("oracle/_ref not built")
    packed = synth.dna_reads_packed_np(1234 + K, n_reads, 150, 2_000_000, 0.01)
    reads = synth.packed_to_strings(packed)
    src = tmp_path / "reads.txt"
    src.write_text("\n".join(reads) + "\n")
    cmd = [drv, "-k", str(K), "-p", "8", "-o", str(tmp_path / "ref"), str(src)]
    if canonical:
        cmd.append("--canonical")
    if bpc:
        cmd += ["--count-width", str(bpc)]
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    chunk = gpu_build(reads, K, "DNA", canonical, bpc)
    chunk.serialize(str(tmp_path / "gpu"))
    for ext in ("", ".W", ".last", ".weights"):
        a = (tmp_path / ("ref.dbg.chunk" + ext)).read_bytes()
        b = (tmp_path / ("gpu.dbg.chunk" + ext)).read_bytes()
        assert a == b, "chunk file %r differs from the reference's" % ext


@pytest.mark.parametrize("K,canonical,bpc", [(31, True, 8), (12, False, 0), (40, True, 12)])
def test_unfused_merge_emit_path(K, canonical, bpc, monkeypatch):
    """DNA chunks come from the fused merge + emit kernel; the row-materialising kernels (what
    Protein uses) must give the same chunk."""
    reads = dna_reads(77, 3000, 100, 20_000)
    fused = check(reads, K, canonical=canonical, bits_per_count=bpc)
    monkeypatch.setenv("MTG_EMIT_UNFUSED", "1")
    plain = check(reads, K, canonical=canonical, bits_per_count=bpc)
    assert np.array_equal(fused.W_words, plain.W_words) and np.array_equal(fused.last_words, plain.last_words)


def test_dna_character_classes():
    """Arithmetic DNA encoding (4 characters per word) against the table semantics: either case,
    U as T, every other byte (N, '@', '`', 'V', digits ...) ends the window."""
    rng = np.random.default_rng(5)
    base = dna_reads(9, 400, 120, 5000)
    odd = "NnUu@`VvTt0-*ZKk"
    reads = []
    for r in base:
        b = list(r)
        for p in rng.integers(0, len(b), 6):
            b[p] = odd[int(rng.integers(0, len(odd)))]
        if rng.random() < 0.5:
            b = [ch.lower() if rng.random() < 0.5 else ch for ch in b]
        reads.append("".join(b))
    check(reads, 9, canonical=True, bits_per_count=8)
    check(reads, 33, canonical=False)


# ---- BASELINE.json's full size (cfg #2): size-independent BOSS invariants ---------------------

def _popcount_bits(words, a, b):
    """number of set bits at positions [a, b) of a little-endian uint64 bit vector"""
    if b <= a:
        return 0
    wa, wb = a >> 6, (b - 1) >> 6
    w = words[wa:wb + 1].copy()
    w[0] &= np.uint64(0xFFFFFFFFFFFFFFFF) << np.uint64(a & 63)
    if b & 63:
        w[-1] &= (np.uint64(1) << np.uint64(b & 63)) - np.uint64(1)
    return int(np.bitwise_count(w).sum())


def test_full_size_chunk_invariants():
    """10 M x 150 bp reads, k = 31, canonical (the bench workload). The chunk must be a valid BOSS
    table: every node except the all-'$' root is entered by exactly one unflagged edge, per last
    character (the LF mapping the reference's BOSS::initialize relies on, boss.cpp:99-122):
        #nodes whose last char is c  ==  #rows with W == c (no minus flag)          c = A..T
    with the node ranges taken from F; plus F monotone and the row count = 1 + root + reals + dummies."""
    import torch
    n_reads, L, K, G = 10_000_000, 150, 31, 100_000_000
    packed = synth.dna_reads_packed_torch(1, n_reads, L, G, 0.005, device="cuda")
    ctor = mtg_b200.BOSSChunkConstructor.initialize(K - 1, True, 0)
    ctor.add_packed_device(packed.data_ptr(), packed.numel(), n_reads * (L - K + 1))
    torch.cuda.synchronize()
    chunk = ctor.build_chunk(copy=False)
    st = chunk.stats
    n = chunk.n_rows
    assert st["n_input_kmers"] == n_reads * (L - K + 1)
    # rows = placeholder + all-'$' root + reals + sinks + sources (the root may be booked as a source level)
    extra = n - (st["n_real"] + st["n_dummy_sink"] + sum(st["n_dummy_source"]))
    assert extra in (1, 2)
    F = [int(x) for x in chunk.F] + [n - 1]
    assert F[0] == 0 and all(F[i] <= F[i + 1] for i in range(5))
    last = np.asarray(chunk.last_words, dtype=np.uint64)
    Ww = np.asarray(chunk.W_words, dtype=np.uint64)[: (n * 4 + 63) // 64]
    counts = np.zeros(16, dtype=np.int64)
    for s in range(16):   # 16 nibbles per word
        counts += np.bincount(((Ww >> np.uint64(4 * s)) & np.uint64(15)).astype(np.uint8), minlength=16)
    assert counts[5] == 0 and counts[10:].sum() == 0            # codes: 0 '$', 1..4, 6..9 flagged
    nodes = _popcount_bits(last, 1, n)
    assert _popcount_bits(last, 0, 1) == 0                      # placeholder row
    assert nodes - 1 == int(counts[1:5].sum())
    assert _popcount_bits(last, 1, F[1] + 1) == 1               # the all-'$' root is the only '$'-ending node
    for c in range(1, 5):
        assert _popcount_bits(last, F[c] + 1, F[c + 1] + 1) == int(counts[c]), "LF invariant, char %d" % c
    # every real k-mer and every dummy source has an edge label; sinks and the root row carry '$'
    assert int(counts[1:5].sum() + counts[6:10].sum()) == st["n_real"] + sum(st["n_dummy_source"]) - (2 - extra)
    ctor.close()


@pytest.mark.parametrize("canonical,bpc,explicit", [(True, 8, False), (False, 16, True)])
def test_push_reads_from_many_threads(canonical, bpc, explicit):
    """mtg_push_reads is documented thread-safe because the reference calls add_sequences from an
    `omp parallel for` over files (cli/build.cpp:35-52): 8 threads push interleaved batches of pageable
    memory (pinned staging pool under contention); the chunk must equal the oracle's on the same reads."""
    import threading
    reads = dna_reads(77, 24_000, 110, 150_000, n_frac=0.002)
    counts = [1 + (i % 5) for i in range(len(reads))] if explicit else None
    ctor = mtg_b200.BOSSChunkConstructor.initialize(30, canonical, bpc)
    errors = []

    def worker(t):
        try:
            for i in range(t, 240, 8):            # 240 batches of 100 reads, round-robin over the threads
                lo, hi = i * 100, (i + 1) * 100
                if explicit:
                    ctor.add_sequences(list(zip(reads[lo:hi], counts[lo:hi])))
                else:
                    ctor.add_sequences(reads[lo:hi])
        except Exception as e:   # noqa: BLE001
            errors.append(e)

    threads = [threading.Thread(target=worker, args=(t,)) for t in range(8)]
    for th in threads:
        th.start()
    for th in threads:
        th.join()
    assert not errors, errors
    chunk = ctor.build_chunk()
    ctor.close()
    ref = oracle_lib.oracle_build(reads, 31, canonical=canonical, bits_per_count=bpc, counts=counts)
    assert chunk.n_rows == ref["n_rows"]
    assert np.array_equal(chunk.W, ref["W"]) and np.array_equal(chunk.last, ref["last"])
    assert np.array_equal(chunk.F, ref["F"]) and np.array_equal(chunk.weights, ref["weights"])


@pytest.mark.parametrize("K,alphabet,canonical,bpc,explicit", [
    (31, "DNA", True, 0, False), (31, "DNA", True, 8, False), (31, "DNA", False, 12, True),
    (63, "DNA", True, 16, False), (12, "DNA", False, 0, False), (9, "Protein", False, 8, False)])
def test_super_batched_build(K, alphabet, canonical, bpc, explicit, monkeypatch):
    """Inputs larger than one sort buffer: with MTG_BATCH_KMERS forced small the pushed reads are folded
    into sorted runs batch by batch and mtg_build_chunk merges the runs (SortedSet::shrink_data semantics,
    sorted_set.hpp:69-82; saturating count sums, sorted_multiset.cpp:70-82; palindrome doubling after the
    merge, boss_chunk_construct.cpp:205-211). The chunk must equal the oracle's single-pass chunk."""
    monkeypatch.setenv("MTG_BATCH_KMERS", "40000")
    monkeypatch.setenv("MTG_FOLD_CHUNK", "5000")     # the merged runs are folded in chunks: many chunk edges
    if alphabet == "DNA":
        reads = dna_reads(500 + K, 6000, 100, 40_000, n_frac=0.001)
        reads[7] = "ACGT" * 25          # palindromic k-mers: counts are doubled once, after the merge
        reads += reads[:800]             # the same k-mers in different batches
    else:
        rng = np.random.default_rng(5)
        reads = ["".join(rng.choice(list("ACDEFGHIKLMNPQRSTVWY"), 60)) for _ in range(3000)]
        reads += reads[:500]
    counts = [1 + (i % 300) for i in range(len(reads))] if explicit else None
    ctor = mtg_b200.BOSSChunkConstructor.initialize(K - 1, canonical, bpc, alphabet=alphabet)
    step = 500
    for i in range(0, len(reads), step):
        if explicit:
            ctor.add_sequences(list(zip(reads[i:i + step], counts[i:i + step])))
        else:
            ctor.add_sequences(reads[i:i + step])
    chunk = ctor.build_chunk()
    ctor.close()
    ref = oracle_lib.oracle_build(reads, K, alphabet=alphabet, canonical=canonical, bits_per_count=bpc, counts=counts)
    assert chunk.n_rows == ref["n_rows"] and chunk.stats["n_real"] == ref["n_real"]
    assert np.array_equal(chunk.W, ref["W"]) and np.array_equal(chunk.last, ref["last"])
    assert np.array_equal(chunk.F, ref["F"])
    if bpc:
        assert np.array_equal(chunk.weights, ref["weights"])


@pytest.mark.parametrize("pinned", [False, True])
@pytest.mark.parametrize("K,canonical,bpc", [(31, True, 8), (20, False, 0)])
def test_pipelined_host_pushes(K, canonical, bpc, pinned, monkeypatch):
    """Host pushes are pipelined: a segment's reads go up on the copy stream while the device folds the
    previous batch (MTG_PUSH_BATCH_KMERS forced small here). Several pushes from pageable and from
    pinned (torch pin_memory) host memory; the chunk must equal the oracle's."""
    import torch
    monkeypatch.setenv("MTG_PUSH_BATCH_KMERS", "30000")
    reads = dna_reads(900 + K, 5000, 110, 30_000, n_frac=0.001)
    ctor = mtg_b200.BOSSChunkConstructor.initialize(K - 1, canonical, bpc)
    for i in range(0, len(reads), 700):
        bases, offsets = oracle_lib.concat_reads(reads[i:i + 700])
        arr = np.frombuffer(bases, dtype=np.uint8)
        if pinned:
            t = torch.empty(len(arr), dtype=torch.uint8, pin_memory=True)
            t.numpy()[:] = arr
            ctor.add_packed(t.numpy(), offsets)
        else:
            ctor.add_packed(arr.copy(), offsets)
    chunk = ctor.build_chunk()
    ctor.close()
    ref = oracle_lib.oracle_build(reads, K, canonical=canonical, bits_per_count=bpc)
    assert chunk.n_rows == ref["n_rows"] and chunk.stats["n_real"] == ref["n_real"]
    assert np.array_equal(chunk.W, ref["W"]) and np.array_equal(chunk.last, ref["last"])
    assert np.array_equal(chunk.F, ref["F"])
    if bpc:
        assert np.array_equal(chunk.weights, ref["weights"])


def test_long_runs_with_explicit_counts():
    """A low-complexity k-mer with thousands of copies and per-read counts: the saturating run sums of
    long runs are computed warp-cooperatively (run_counts_kernel); 32-bit counts never saturate here, so
    the whole run is summed."""
    reads = ["A" * 150] * 300 + dna_reads(3, 400, 90, 3000) + ["ACGT" * 30] * 50
    counts = [1 + (i % 7) for i in range(len(reads))]
    for bpc in (8, 32):
        check(reads, 21, canonical=True, bits_per_count=bpc, counts=counts)
        check(reads, 21, canonical=False, bits_per_count=bpc, counts=counts)
