"""Seeded inputs of the golden fixtures (tests/golden/*.npz, made by tests/golden/make_golden.py
from the reference's own implementation). Shared by the generator, the CPU oracle tests and the
GPU parity tests, so all three see byte-identical reads."""
import numpy as np

from mtg_b200 import synth

HAND_DNA = ["ACGT" * 20, "acgtacgtnnACGTACGTACGTTTTTTTTTTTTAAAAAAAAAAAA", "TTTTTTTTTTTTTTTTTTTT",
            "AAAAAAAAAAAAAAAAAAAAAAAA", "ACGUACGUACGUACGUUUU", "GATTACA", "AC", "NNNNNNNNNNNNNNNN",
            "ACGTNACGTTGCANNACGTAGCTAGCTAGCTAGCATCGATCGATCAGCTACGACTAGCTAGCATCGAT"]

# name -> parameters. kind: "dna" (synthetic reads of a random genome), "protein", "hand".
CASES = {
    # scaled twins of the BASELINE.json configs
    "cfg1_k12_basic": dict(kind="dna", seed=21, n=400, len=100, G=12000, K=12),
    "cfg2_k31_canonical": dict(kind="dna", seed=1, n=400, len=150, G=4000, K=31, canonical=True),
    "cfg3_k31_count8": dict(kind="dna", seed=2, n=600, len=150, G=3000, K=31, bpc=8),
    "cfg4_k63_basic": dict(kind="dna", seed=3, n=300, len=150, G=3000, K=63),
    "cfg5_protein_k12": dict(kind="protein", seed=4, n=150, len=120, K=12, alphabet="Protein"),
    # edge cases the reference tests cover
    "k2_canonical_w8": dict(kind="hand", K=2, canonical=True, bpc=8),
    "k4_hand_w8": dict(kind="hand", K=4, bpc=8),
    "k8_hand_canonical": dict(kind="hand", K=8, canonical=True),
    "k21_saturating_w2": dict(kind="dna", seed=5, n=1500, len=80, G=800, e=0.002, K=21, bpc=2,
                              canonical=True),
    "k25_explicit_counts_w12": dict(kind="dna", seed=9, n=300, len=100, G=3000, K=25, bpc=12,
                                    canonical=True, counts=True),
    "k32_with_N": dict(kind="dna", seed=6, n=300, len=120, G=5000, K=32, n_frac=0.004),
    "k33_canonical_w16": dict(kind="dna", seed=7, n=300, len=120, G=4000, K=33, canonical=True, bpc=16),
    "k64_canonical": dict(kind="dna", seed=8, n=200, len=150, G=4000, K=64, canonical=True),
    "protein_k5_w8": dict(kind="protein", seed=10, n=100, len=80, K=5, alphabet="Protein", bpc=8),
    "protein_k25": dict(kind="protein", seed=11, n=60, len=100, K=25, alphabet="Protein"),
    # suffix-filtered chunks (--suffix, KmerExtractorBOSS)
    "k10_suffix_dollar": dict(kind="dna", seed=12, n=200, len=80, G=3000, K=10, suffix="$"),
    "k10_suffix_A": dict(kind="dna", seed=12, n=200, len=80, G=3000, K=10, suffix="A"),
    "k10_suffix_GT_canonical_w8": dict(kind="dna", seed=12, n=200, len=80, G=3000, K=10, suffix="GT",
                                       canonical=True, bpc=8),
    "k10_suffix_dollarC_with_N": dict(kind="dna", seed=13, n=200, len=80, G=3000, K=10, suffix="$C",
                                      n_frac=0.01),
}


def case_reads(case):
    """-> (list of str, counts list or None)"""
    kind = case["kind"]
    if kind == "hand":
        reads = list(HAND_DNA)
    elif kind == "protein":
        packed = synth.protein_reads_packed_np(case["seed"], case["n"], case["len"])
        reads = synth.packed_to_strings(packed)
        reads[0] = reads[0][:20] + "*?xbZ" + reads[0][25:]   # unknown characters are 'X'
    else:
        packed = synth.dna_reads_packed_np(case["seed"], case["n"], case["len"], case["G"],
                                           case.get("e", 0.01))
        reads = synth.packed_to_strings(packed)
        if case.get("n_frac"):
            rng = np.random.default_rng(case["seed"])
            out = []
            for r in reads:
                b = bytearray(r.encode())
                for p in np.nonzero(rng.random(len(b)) < case["n_frac"])[0]:
                    b[p] = ord("N")
                out.append(b.decode())
            reads = out
    counts = None
    if case.get("counts"):
        rng = np.random.default_rng(case["seed"] + 1000)
        counts = [int(x) for x in rng.integers(1, 400, len(reads))]
        counts[:5] = [2 ** 40, 255, 256, 65535, 65536]
    return reads, counts
