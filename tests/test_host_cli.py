"""The C++ host side (projects2014-metagenome_b200/host): mtg_build CLI + Chunk concatenation.
CPU: `concatenate` (pure host code) against the Python mirror and the golden files; the build
command must fail loudly without a GPU. GPU: `mtg_build build` output files == the reference's
chunk files (golden sha256)."""
import hashlib
import os
import subprocess

import numpy as np
import pytest

import mtg_b200
from golden_cases import CASES, case_reads
from test_host_chunk import chunk_from_arrays
from test_oracle_golden import load_golden

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "projects2014-metagenome_b200", "mtg_build")


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def sha(path):
    return hashlib.sha256(open(path, "rb").read()).hexdigest()


def test_cli_exists():
    assert os.path.exists(CLI), "build it: python -c 'import __graft_entry__ as g; g.build()'"


def test_concatenate_matches_python_extend(tmp_path):
    a = load_golden("cfg3_k31_count8")
    c1, c2 = chunk_from_arrays(a, 30, 8), chunk_from_arrays(a, 30, 8)
    c1.serialize(str(tmp_path / "a"))
    c2.serialize(str(tmp_path / "b"))
    out = subprocess.run([CLI, "concatenate", "-o", str(tmp_path / "cat"), str(tmp_path / "a.dbg.chunk"),
                          str(tmp_path / "b.dbg.chunk")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    c1.extend(c2)
    c1.serialize(str(tmp_path / "py"))
    for ext in ("", ".W", ".last", ".weights"):
        assert sha(str(tmp_path / "cat.dbg.chunk") + ext) == sha(str(tmp_path / "py.dbg.chunk") + ext), ext
    # a single chunk passes through byte-identically (== the reference's own file)
    subprocess.check_call([CLI, "concatenate", "-o", str(tmp_path / "one"), str(tmp_path / "a.dbg.chunk")],
                          stdout=subprocess.DEVNULL)
    for ext in ("", ".W", ".last", ".weights"):
        assert sha(str(tmp_path / "one.dbg.chunk") + ext) == a["sha256"][ext]


def test_cli_rejects_bad_arguments(tmp_path):
    fa = tmp_path / "r.fa"
    fa.write_text(">r\nACGTACGTACGTACGT\n")
    for args in (["--count-kmers", "--count-width", "40"], ["-k", "1"]):
        out = subprocess.run([CLI, "build", "-o", str(tmp_path / "o"), str(fa)] + args, capture_output=True,
                             text=True)
        assert out.returncode == 1


@pytest.mark.skipif(_has_gpu(), reason="checks the no-device behaviour")
def test_cli_build_fails_loudly_without_gpu(tmp_path):
    fa = tmp_path / "r.fa"
    fa.write_text(">r\nACGTACGTACGTACGT\n")
    out = subprocess.run([CLI, "build", "-k", "5", "-o", str(tmp_path / "o"), str(fa)], capture_output=True,
                         text=True)
    assert out.returncode == 1 and "no CPU fallback" in out.stderr
    assert not os.path.exists(str(tmp_path / "o.dbg.chunk"))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["cfg1_k12_basic", "cfg2_k31_canonical", "cfg3_k31_count8", "cfg4_k63_basic",
                                  "cfg5_protein_k12", "k33_canonical_w16", "k32_with_N"])
def test_cli_build_writes_the_reference_chunk_files(name, tmp_path):
    case, g = CASES[name], load_golden(name)
    reads, _ = case_reads(case)
    fa = tmp_path / "reads.fa"
    with open(fa, "w") as f:
        for i, r in enumerate(reads):
            if r:
                f.write(">r%d\n%s\n" % (i, "\n".join(r[j:j + 60] for j in range(0, len(r), 60))))
    args = [CLI, "build", "-v", "-k", str(case["K"]), "-o", str(tmp_path / "out"), str(fa),
            "--alphabet", case.get("alphabet", "DNA")]
    if case.get("canonical"):
        args += ["--mode", "canonical"]
    if case.get("bpc"):
        args += ["--count-kmers", "--count-width", str(case["bpc"])]
    out = subprocess.run(args, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "Graph chunk with %d k-mers was built" % (len(g["W"]) - 1) in out.stdout
    for ext in ("", ".W", ".last", ".weights"):
        assert sha(str(tmp_path / "out.dbg.chunk") + ext) == g["sha256"][ext], ext


@pytest.mark.gpu
def test_cli_build_parallel_feeder_plain_gzip_fastq(tmp_path):
    """`mtg_build build -p 8` over three input files of different formats (the host feeder: pinned batches
    pushed from 8 threads) == the chunk of the same reads pushed in one go."""
    import gzip
    case, g = CASES["cfg2_k31_canonical"], load_golden("cfg2_k31_canonical")
    reads, _ = case_reads(case)
    reads = [r for r in reads if r]
    third = len(reads) // 3
    with open(tmp_path / "a.fa", "w") as f:
        for r in reads[:third]:
            f.write(">r\n%s\n" % r)
    with gzip.open(tmp_path / "b.fa.gz", "wt") as f:
        for r in reads[third:2 * third]:
            f.write(">r\n%s\n" % r)
    with open(tmp_path / "c.fq", "w") as f:
        for r in reads[2 * third:]:
            f.write("@r\n%s\n+\n%s\n" % (r, "@" * len(r)))
    out = subprocess.run([CLI, "build", "-v", "-p", "8", "-k", str(case["K"]), "--mode", "canonical", "-o",
                          str(tmp_path / "out"), str(tmp_path / "a.fa"), str(tmp_path / "b.fa.gz"),
                          str(tmp_path / "c.fq")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    assert "Parsed %d reads" % len(reads) in out.stderr
    for ext in ("", ".W", ".last", ".weights"):
        assert sha(str(tmp_path / "out.dbg.chunk") + ext) == g["sha256"][ext], ext
