"""CPU tests pinning the oracle (oracle/mtg_oracle.cpp) against golden vectors produced by the
reference's OWN implementation (tests/golden/*.npz, generator: tests/golden/make_golden.py) and
against the goldens of the reference's tests. Bit-exact: W, last, F, weights and
the sha256 of the sdsl int_vector files BOSS::Chunk::serialize writes."""
import hashlib
import json
import os

import numpy as np
import pytest

import oracle_lib
from golden_cases import CASES, case_reads

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
TRANSCRIPTS = "/root/reference/metagraph/tests/data/transcripts_1000.fa"


def load_golden(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    g = {k: z[k] for k in z.files}
    g["sha256"] = json.loads(str(g["sha256"]))
    return g


def W_width(sigma):
    return int(2 * sigma - 1).bit_length()


def assert_chunk_equals_golden(res, g, bpc):
    """res: dict with W/last/F/weights arrays (one entry per row)."""
    assert len(res["W"]) == len(g["W"])
    assert np.array_equal(np.asarray(res["F"], dtype=np.uint64), g["F"])
    assert np.array_equal(np.asarray(res["last"], dtype=np.uint8), g["last"])
    assert np.array_equal(np.asarray(res["W"], dtype=np.uint8), g["W"])
    sigma = int(g["sigma"])
    wb = oracle_lib.sdsl_int_vector_bytes(res["W"], W_width(sigma))
    assert hashlib.sha256(wb).hexdigest() == g["sha256"][".W"]
    lb = oracle_lib.sdsl_int_vector_bytes(res["last"], 1, fixed_width=True)
    assert hashlib.sha256(lb).hexdigest() == g["sha256"][".last"]
    if bpc:
        assert np.array_equal(np.asarray(res["weights"], dtype=np.uint64), g["weights"])
        tb = oracle_lib.sdsl_int_vector_bytes(res["weights"], bpc)
        assert hashlib.sha256(tb).hexdigest() == g["sha256"][".weights"]


@pytest.mark.parametrize("name", sorted(CASES))
def test_oracle_matches_reference_golden(name):
    case = CASES[name]
    reads, counts = case_reads(case)
    res = oracle_lib.oracle_build(reads, case["K"], alphabet=case.get("alphabet", "DNA"),
                                  canonical=case.get("canonical", False),
                                  bits_per_count=case.get("bpc", 0), suffix=case.get("suffix", ""),
                                  counts=counts)
    assert_chunk_equals_golden(res, load_golden(name), case.get("bpc", 0))


@pytest.mark.skipif(not os.path.exists(TRANSCRIPTS), reason="/root/reference not on this box")
@pytest.mark.parametrize("tag", ["k12_basic", "k12_canonical_w8", "k20_basic", "k20_canonical_w8"])
def test_oracle_on_the_reference_fixture(tag):
    """BASELINE config #1 (`build -k 12` on transcripts_1000.fa) and the k=20 integration goldens."""
    gold = json.load(open(os.path.join(GOLDEN, "transcripts_1000.json")))[tag]
    reads = oracle_lib.read_fasta(TRANSCRIPTS)
    res = oracle_lib.oracle_build(reads, gold["K"], canonical=gold["canonical"],
                                  bits_per_count=gold["bpc"])
    assert res["n_rows"] == gold["n_rows"]
    assert [int(x) for x in res["F"]] == gold["F"]
    wb = oracle_lib.sdsl_int_vector_bytes(res["W"], 4)
    assert hashlib.sha256(wb).hexdigest() == gold["sha256"][".W"]
    lb = oracle_lib.sdsl_int_vector_bytes(res["last"], 1, fixed_width=True)
    assert hashlib.sha256(lb).hexdigest() == gold["sha256"][".last"]
    if gold["bpc"]:
        tb = oracle_lib.sdsl_int_vector_bytes(res["weights"], gold["bpc"])
        assert hashlib.sha256(tb).hexdigest() == gold["sha256"][".weights"]
    if tag == "k12_basic":
        # counters logged by the reference (boss_chunk_construct.cpp:267-302), SURVEY.md 8c
        assert res["n_real"] == 538182 and res["n_sink"] == 214
        assert [int(x) for x in res["n_src"]] == [251, 251, 251, 251, 249, 242, 219, 157, 60, 16, 4]
        assert res["n_rows"] - 1 == 540348          # BOSS edges
        hist = np.bincount(res["W"].astype(np.int64) % 5, minlength=5)
        assert [int(x) for x in hist[1:]] == [114392, 155313, 153984, 116444]   # `stats` W counts
        assert int(hist[0]) - 1 == 215                                        # minus row 0



def test_transcripts_goldens_are_the_reference_test_numbers():
    """k=20 node counts of metagraph/integration_tests/test_build.py:61,129 follow from the rows:
    nodes = rows with last==1 minus dummy nodes is checked by the reference's `stats`; here we pin
    what the fixture file records so that a regenerated golden cannot drift silently."""
    gold = json.load(open(os.path.join(GOLDEN, "transcripts_1000.json")))
    assert gold["k12_basic"]["n_rows"] == 540349
    assert gold["k12_basic"]["sha256"][".W"] and gold["k20_canonical_w8"]["weights_sum"] > 0
