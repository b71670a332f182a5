"""CPU tests of the host feeder (projects2014-metagenome_b200/host/read_feeder.cpp) through `mtg_build parse`:
multi-threaded FASTA / FASTQ / gzip parsing must deliver exactly the reads a straightforward parser sees
(count, bases, order-independent checksum), whatever the thread count and however the file is cut into
ranges. Mirrors what the reference's reader hands to add_sequences (seq_io/sequence_io.cpp:364-405)."""
import gzip
import os
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CLI = os.path.join(ROOT, "projects2014-metagenome_b200", "mtg_build")


def fnv_sum(reads):
    total = 0
    for r in reads:
        h = 1469598103934665603
        for c in r.encode():
            h = ((h ^ c) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
        total = (total + h) & 0xFFFFFFFFFFFFFFFF
    return total


def parse(files, threads):
    out = subprocess.run([CLI, "parse", "--checksum", "-p", str(threads)] + [str(f) for f in files], capture_output=True,
                         text=True, check=True).stdout.split()
    return {out[i]: out[i + 1] for i in range(0, len(out), 2)}


def rand_reads(seed, n, lo, hi):
    rng = np.random.default_rng(seed)
    return ["".join(rng.choice(list("ACGTNacgt"), int(rng.integers(lo, hi)))) for _ in range(n)]


@pytest.mark.skipif(not os.path.exists(CLI), reason="host CLI not built")
def test_feeder_formats_and_thread_counts(tmp_path):
    reads = rand_reads(1, 3000, 1, 400)
    fa = tmp_path / "multi_line.fa"          # multi-line records, CRLF in places, an empty record
    with open(fa, "w") as f:
        for i, r in enumerate(reads):
            f.write(">read %d\n" % i)
            for j in range(0, len(r), 60):
                f.write(r[j:j + 60] + ("\r\n" if i % 7 == 0 else "\n"))
        f.write(">empty\n>last\nACGT")        # no trailing newline
    fq = tmp_path / "reads.fq"               # quality lines that start with '@' and '+'
    with open(fq, "w") as f:
        for i, r in enumerate(reads):
            q = ("@" if i % 3 == 0 else "+" if i % 3 == 1 else "I") + "I" * (len(r) - 1)
            f.write("@r%d\n%s\n+\n%s\n" % (i, r, q))
    gz = tmp_path / "reads.fa.gz"
    with gzip.open(gz, "wt") as f:
        for i, r in enumerate(reads):
            f.write(">g%d\n%s\n" % (i, r))
    exp_fa = {"reads": len(reads) + 1, "bases": sum(map(len, reads)) + 4, "sum": fnv_sum(reads + ["ACGT"])}
    exp = {"reads": len(reads), "bases": sum(map(len, reads)), "sum": fnv_sum(reads)}
    for threads in (1, 3, 8):
        for path, e in ((fa, exp_fa), (fq, exp), (gz, exp)):
            got = parse([path], threads)
            assert int(got["reads"]) == e["reads"] and int(got["bases"]) == e["bases"], (path, threads, got)
            assert int(got["checksum"], 16) == e["sum"], (path, threads)
        got = parse([fa, fq, gz], threads)
        assert int(got["reads"]) == exp_fa["reads"] + 2 * exp["reads"]
        assert int(got["checksum"], 16) == (exp_fa["sum"] + 2 * exp["sum"]) & 0xFFFFFFFFFFFFFFFF


@pytest.mark.skipif(not os.path.exists(CLI), reason="host CLI not built")
def test_feeder_range_cuts_of_a_large_file(tmp_path):
    """A file larger than the 64 MB parse unit: the ranges are cut at record boundaries (FASTQ: a quality
    line starting with '@' must not be taken for a header)."""
    rng = np.random.default_rng(5)
    n, L = 600_000, 150
    codes = rng.integers(0, 4, (n, L), dtype=np.uint8)
    bases = np.frombuffer(b"ACGT", dtype=np.uint8)[codes]
    rec = np.empty((n, 4 + L + 3 + L + 1), dtype=np.uint8)      # "@r\n" seq "\n+\n" qual "\n"
    rec[:, 0], rec[:, 1], rec[:, 2] = ord("@"), ord("r"), ord("\n")
    rec[:, 3:3 + L] = bases
    rec[:, 3 + L], rec[:, 4 + L], rec[:, 5 + L] = ord("\n"), ord("+"), ord("\n")
    rec[:, 6 + L:6 + 2 * L] = ord("@")                          # every quality line starts with '@'
    rec[:, 6 + 2 * L] = ord("\n")
    fq = tmp_path / "big.fq"
    rec[:, :7 + 2 * L].tofile(fq)
    assert os.path.getsize(fq) > (64 << 20) * 2
    a = parse([fq], 1)
    b = parse([fq], 8)
    assert int(a["reads"]) == n and int(a["bases"]) == n * L
    assert a["checksum"] == b["checksum"] and b["reads"] == a["reads"]
