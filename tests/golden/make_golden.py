#!/usr/bin/env python
"""TEST INFRASTRUCTURE. Generates tests/golden/*.npz and transcripts_1000.json by running the
reference's OWN implementation (oracle/_ref/ref_driver_{DNA,Protein}, compiled from the
unmodified sources under /root/reference by oracle/build_ref.sh) on small seeded inputs.

Every .npz holds the reads' generator parameters (the reads themselves are re-created from
mtg_b200.synth with the same seed, and also stored verbatim for the hand-written cases) and the
reference's output chunk: W, last, F, weights exactly as read back from the files
BOSS::Chunk::serialize wrote (boss_chunk.cpp:372-386).

Run in the build container only (needs /root/reference):  python tests/golden/make_golden.py
"""
import hashlib
import json
import os
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import oracle_lib  # noqa: E402
from golden_cases import CASES, case_reads  # noqa: E402

TRANSCRIPTS = "/root/reference/metagraph/tests/data/transcripts_1000.fa"


def run_ref(reads, counts, K, alphabet, canonical, bpc, suffix, td, threads=2):
    drv = oracle_lib.ref_driver(alphabet)
    assert drv, "build oracle/_ref first (bash oracle/build_ref.sh)"
    path = os.path.join(td, "reads.txt")
    with open(path, "w") as f:
        for i, r in enumerate(reads):
            f.write(r if counts is None else "%s\t%d" % (r, counts[i]))
            f.write("\n")
    return run_ref_file(path, K, alphabet, canonical, bpc, suffix, td, threads)


def run_ref_file(path, K, alphabet, canonical, bpc, suffix, td, threads=2):
    drv = oracle_lib.ref_driver(alphabet)
    out = os.path.join(td, "out")
    cmd = [drv, "-k", str(K), "-p", str(threads), "-o", out, path]
    if canonical:
        cmd.append("--canonical")
    if bpc:
        cmd += ["--count-width", str(bpc)]
    if suffix:
        cmd += ["--suffix", suffix]
    subprocess.check_call(cmd, stdout=subprocess.DEVNULL)
    res = oracle_lib.read_chunk_files(out)
    res["sha256"] = {ext: hashlib.sha256(open(out + ".dbg.chunk" + ext, "rb").read()).hexdigest()
                     for ext in ("", ".W", ".last", ".weights")}
    return res


def main():
    for name, case in CASES.items():
        reads, counts = case_reads(case)
        with tempfile.TemporaryDirectory() as td:
            # plain-text input cannot carry empty reads or reads with whitespace: filter like a
            # FASTA parser would (the reference never sees an empty sequence either)
            keep = [i for i, r in enumerate(reads) if r]
            r2 = [reads[i] for i in keep]
            c2 = None if counts is None else [counts[i] for i in keep]
            res = run_ref(r2, c2, case["K"], case.get("alphabet", "DNA"), case.get("canonical", False),
                          case.get("bpc", 0), case.get("suffix", ""), td)
        arrays = dict(W=res["W"].astype(np.uint8), last=res["last"].astype(np.uint8),
                      F=res["F"].astype(np.uint64),
                      sigma=np.uint64(res["sigma"]), k=np.uint64(res["k"]))
        if res["weights"] is not None:
            arrays["weights"] = res["weights"].astype(np.uint64)
        arrays["sha256"] = np.array(json.dumps(res["sha256"]))
        np.savez_compressed(os.path.join(HERE, name + ".npz"), **arrays)
        print("%-28s rows=%d" % (name, len(res["W"])))

    # BASELINE config #1 and its canonical/weighted variant on the reference's own fixture
    if os.path.exists(TRANSCRIPTS):
        gold = {}
        for tag, canonical, bpc in (("k12_basic", False, 0), ("k12_canonical_w8", True, 8),
                                    ("k20_basic", False, 0), ("k20_canonical_w8", True, 8)):
            K = int(tag[1:3])
            with tempfile.TemporaryDirectory() as td:
                res = run_ref_file(TRANSCRIPTS, K, "DNA", canonical, bpc, "", td, threads=4)
            W = res["W"]
            gold[tag] = dict(K=K, canonical=canonical, bpc=bpc, n_rows=int(len(W)),
                             F=[int(x) for x in res["F"]], sha256=res["sha256"],
                             W_hist=[int(x) for x in np.bincount(W.astype(np.int64), minlength=10)],
                             last_ones=int(res["last"].sum()),
                             weights_sum=int(res["weights"].sum()) if res["weights"] is not None else None)
            print(tag, gold[tag]["n_rows"])
        json.dump(gold, open(os.path.join(HERE, "transcripts_1000.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
