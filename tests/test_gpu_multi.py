"""GPU parity of the range-partitioned (multi-GPU) build, run as g virtual ranks on ONE GPU:
the same mtg_mg_* phases, the exchanges done by slicing. The concatenation of the per-range
chunks must equal the oracle's (= the reference's) single chunk bit for bit."""
import numpy as np
import pytest

import oracle_lib
from mtg_b200 import multi_gpu, synth
from golden_cases import CASES, case_reads
from test_oracle_golden import assert_chunk_equals_golden, load_golden
from test_gpu_parity import dna_reads

pytestmark = pytest.mark.gpu


def virtual_build(reads, K, world, alphabet="DNA", canonical=False, bpc=0, fused_route=True):
    vb = multi_gpu.VirtualBuild(world, K, canonical, bpc, alphabet, fused_route=fused_route)
    for r in range(world):
        vb.add_sequences(r, reads[r::world])
    chunks = vb.build_chunks()
    ranges = vb.ranges
    vb.close()
    return chunks, ranges


def check(reads, K, world, alphabet="DNA", canonical=False, bpc=0, fused_route=True):
    chunks, ranges = virtual_build(reads, K, world, alphabet, canonical, bpc, fused_route)
    stats = [c.stats for c in chunks]
    full = multi_gpu.concat_chunks(chunks)
    ref = oracle_lib.oracle_build(reads, K, alphabet=alphabet, canonical=canonical, bits_per_count=bpc)
    assert full.n_rows == ref["n_rows"]
    assert sum(s["n_real"] for s in stats) == ref["n_real"]
    assert sum(s["n_dummy_sink"] for s in stats) == ref["n_sink"]
    assert stats[0]["n_dummy_source"][: len(ref["n_src"])] == [int(x) for x in ref["n_src"]]
    assert np.array_equal(full.F, ref["F"])
    assert np.array_equal(full.last, ref["last"])
    assert np.array_equal(full.W, ref["W"])
    if bpc:
        assert np.array_equal(full.weights, ref["weights"])
    return full, ranges


@pytest.mark.parametrize("world", [1, 2, 3, 4, 8])
@pytest.mark.parametrize("K,canonical,bpc", [(31, False, 8), (31, True, 0), (12, False, 0), (63, False, 0),
                                             (40, True, 12)])
def test_partitioned_build_matches_oracle(world, K, canonical, bpc):
    reads = dna_reads(K + world, 1200, 130, 9000, e=0.01, n_frac=0.002)
    check(reads, K, world, canonical=canonical, bpc=bpc)


@pytest.mark.parametrize("K", [2, 3, 4, 5, 7])
def test_partitioned_small_k(K):
    reads = dna_reads(K, 300, 60, 500)
    for world in (2, 4):
        check(reads, K, world, canonical=True, bpc=8)
        check(reads, K, world)


def test_partitioned_protein():
    reads = synth.packed_to_strings(synth.protein_reads_packed_np(4, 200, 100))
    for world in (2, 5):
        check(reads, 12, world, alphabet="Protein")
        check(reads, 4, world, alphabet="Protein", bpc=8)


def test_partitioned_skewed_and_empty():
    # almost everything in one bucket (poly-A), ranks with empty ranges, and no input at all
    reads = ["A" * 80, "A" * 50 + "C" + "A" * 40, "ACGT" * 20] * 5
    check(reads, 21, 4, canonical=True, bpc=8)
    check(reads, 21, 4)
    check([], 15, 3)
    check(["ACGTACGTAC"], 15, 2)


@pytest.mark.parametrize("name", ["cfg2_k31_canonical", "cfg3_k31_count8", "cfg4_k63_basic",
                                  "cfg5_protein_k12", "k21_saturating_w2"])
def test_partitioned_matches_reference_golden(name):
    case = CASES[name]
    reads, _ = case_reads(case)
    chunks, _ = virtual_build(reads, case["K"], 4, case.get("alphabet", "DNA"), case.get("canonical", False),
                              case.get("bpc", 0))
    full = multi_gpu.concat_chunks(chunks)
    res = {"W": full.W, "last": full.last, "F": full.F, "weights": full.weights}
    assert_chunk_equals_golden(res, load_golden(name), case.get("bpc", 0))


@pytest.mark.parametrize("world", [2, 5])
@pytest.mark.parametrize("K,canonical,bpc", [(31, True, 8), (40, False, 0), (3, True, 0)])
def test_separate_group_and_exchange_path(world, K, canonical, bpc):
    """mtg_mg_extract + slicing (what the NCCL fallback uses) against the default, where the
    grouping pass scatters straight into the owners' buffers (mtg_mg_count + mtg_mg_route)."""
    reads = dna_reads(3 * K + world, 900, 110, 7000, e=0.01, n_frac=0.002)
    a, ra = check(reads, K, world, canonical=canonical, bpc=bpc, fused_route=False)
    b, rb = check(reads, K, world, canonical=canonical, bpc=bpc, fused_route=True)
    assert ra == rb and np.array_equal(a.W_words, b.W_words)


@pytest.mark.parametrize("world,fused", [(2, True), (4, True), (3, False)])
@pytest.mark.parametrize("K,canonical,bpc", [(31, True, 8), (31, False, 8), (63, True, 0)])
def test_partitioned_super_batched_rounds(world, fused, K, canonical, bpc):
    """cfg #3's shape: the reads of every rank arrive in three slices; each slice is routed with the
    ranges of the first round, collected, and folded into the range's accumulated k-mers
    (mtg_mg_accumulate: merge + saturating count sums). Result == the oracle's single chunk."""
    reads = dna_reads(11 * K + world, 1500, 120, 8000, e=0.01, n_frac=0.002)
    reads += reads[:400]                       # the same k-mers in different rounds (and ranks)
    vb = multi_gpu.VirtualBuild(world, K, canonical, bpc, fused_route=fused)
    per_rank = [reads[r::world] for r in range(world)]
    cut = [(0, 1 / 3), (1 / 3, 2 / 3), (2 / 3, 1)]
    slices = [[pr[int(a * len(pr)):int(b * len(pr))] for pr in per_rank] for a, b in cut]
    for r in range(world):
        vb.add_sequences(r, slices[0][r])
    chunks = vb.build_chunks(more_rounds=[{r: slices[i][r] for r in range(world)} for i in (1, 2)])
    vb.close()
    full = multi_gpu.concat_chunks(chunks)
    ref = oracle_lib.oracle_build(reads, K, canonical=canonical, bits_per_count=bpc)
    assert full.n_rows == ref["n_rows"]
    assert sum(c.stats["n_input_kmers"] for c in chunks) > 0
    assert np.array_equal(full.W, ref["W"]) and np.array_equal(full.last, ref["last"])
    assert np.array_equal(full.F, ref["F"])
    if bpc:
        assert np.array_equal(full.weights, ref["weights"])
