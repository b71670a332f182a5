"""Test-side loader for the CPU oracle (oracle/mtg_oracle.cpp) and helpers that
read/write the sdsl `int_vector` byte layout the reference's BOSS::Chunk files use
(sdsl int_vector.hpp:594-603, int_vector_buffer.hpp:271).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg import this.
"""
import ctypes
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_SRC = os.path.join(ROOT, "oracle", "mtg_oracle.cpp")
ORACLE_SO = os.path.join(ROOT, "oracle", "_build", "libmtg_oracle.so")
REF_DIR = os.path.join(ROOT, "oracle", "_ref")


class _OracleChunk(ctypes.Structure):
    _fields_ = [
        ("n_rows", ctypes.c_uint64), ("sigma", ctypes.c_uint64),
        ("n_real", ctypes.c_uint64), ("n_sink", ctypes.c_uint64),
        ("n_src_levels", ctypes.c_uint64),
        ("W", ctypes.POINTER(ctypes.c_uint8)), ("last", ctypes.POINTER(ctypes.c_uint8)),
        ("weights", ctypes.POINTER(ctypes.c_uint64)), ("F", ctypes.POINTER(ctypes.c_uint64)),
        ("n_src", ctypes.POINTER(ctypes.c_uint64)),
        ("n_kmers", ctypes.c_uint64),
        ("kmer_words", ctypes.POINTER(ctypes.c_uint64)),
        ("kmer_counts", ctypes.POINTER(ctypes.c_uint64)),
        ("impl", ctypes.c_void_p),
    ]


def build_oracle(force=False):
    if force or not os.path.exists(ORACLE_SO) or \
            os.path.getmtime(ORACLE_SO) < os.path.getmtime(ORACLE_SRC):
        os.makedirs(os.path.dirname(ORACLE_SO), exist_ok=True)
        subprocess.check_call(["g++", "-std=c++17", "-O2", "-fPIC", "-shared",
                               "-o", ORACLE_SO, ORACLE_SRC])
    return ORACLE_SO


_lib = None


def load_oracle():
    global _lib
    if _lib is None:
        _lib = ctypes.CDLL(build_oracle())
        _lib.mtg_oracle_build.restype = ctypes.c_int
    return _lib


def concat_reads(reads):
    """reads: list of bytes/str -> (bases bytes, offsets uint64[n+1])."""
    bs = [r.encode() if isinstance(r, str) else bytes(r) for r in reads]
    offsets = np.zeros(len(bs) + 1, dtype=np.uint64)
    if bs:
        offsets[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
    return b"".join(bs), offsets


def oracle_build(reads, K, alphabet="DNA", canonical=False, bits_per_count=0, suffix="",
                 counts=None):
    lib = load_oracle()
    bases, offsets = concat_reads(reads)
    out = _OracleChunk()
    cnt = None
    if counts is not None:
        cnt = np.ascontiguousarray(counts, dtype=np.uint64)
    rc = lib.mtg_oracle_build(
        ctypes.c_char_p(bases), offsets.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
        cnt.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)) if cnt is not None else None,
        ctypes.c_uint64(len(reads)), ctypes.c_uint32(K),
        ctypes.c_int(1 if alphabet == "Protein" else 0), ctypes.c_int(int(canonical)),
        ctypes.c_int(bits_per_count), ctypes.c_char_p(suffix.encode()), ctypes.byref(out))
    if rc != 0:
        raise ValueError("oracle rejected arguments (rc=%d)" % rc)
    n = out.n_rows
    res = {
        "n_rows": int(n),
        "W": np.ctypeslib.as_array(out.W, (n,)).copy(),
        "last": np.ctypeslib.as_array(out.last, (n,)).copy(),
        "F": np.ctypeslib.as_array(out.F, (out.sigma,)).copy(),
        "weights": (np.ctypeslib.as_array(out.weights, (n,)).copy()
                    if bits_per_count else None),
        "n_real": int(out.n_real), "n_sink": int(out.n_sink),
        "n_src": (np.ctypeslib.as_array(out.n_src, (out.n_src_levels,)).copy()
                  if out.n_src_levels else np.zeros(0, np.uint64)),
        "kmers": (np.ctypeslib.as_array(out.kmer_words, (out.n_kmers, 4)).copy()
                  if out.n_kmers else np.zeros((0, 4), np.uint64)),
        "kmer_counts": (np.ctypeslib.as_array(out.kmer_counts, (out.n_kmers,)).copy()
                        if out.n_kmers else np.zeros(0, np.uint64)),
    }
    lib.mtg_oracle_free(ctypes.byref(out))
    return res


# ---- sdsl int_vector layout -------------------------------------------------

def pack_bits(values, width):
    """Pack integer array into little-endian 64-bit words, `width` bits each, LSB first."""
    values = np.asarray(values, dtype=np.uint64)
    n = len(values)
    nbits = n * width
    nwords = (nbits + 63) // 64
    bits = ((values[:, None] >> np.arange(width, dtype=np.uint64)[None, :]) & np.uint64(1))
    flat = np.zeros(nwords * 64, dtype=np.uint8)
    flat[:nbits] = bits.astype(np.uint8).reshape(-1)
    return np.packbits(flat, bitorder="little").view(np.uint64)


def unpack_bits(words, n, width):
    words = np.ascontiguousarray(words, dtype=np.uint64)
    flat = np.unpackbits(words.view(np.uint8), bitorder="little")[: n * width]
    bits = flat.reshape(n, width).astype(np.uint64)
    return (bits << np.arange(width, dtype=np.uint64)[None, :]).sum(axis=1).astype(np.uint64)


def sdsl_int_vector_bytes(values, width, fixed_width=False):
    """Bytes of sdsl::int_vector<t_width>::serialize: u64 bit length, [u8 width], words."""
    n = len(values)
    hdr = np.uint64(n * width).tobytes()
    if not fixed_width:
        hdr += bytes([width])
    return hdr + pack_bits(values, width).tobytes()


def read_sdsl_int_vector(path, fixed_width=None):
    b = open(path, "rb").read()
    nbits = int(np.frombuffer(b[:8], dtype=np.uint64)[0])
    if fixed_width is None:
        width, off = b[8], 9
    else:
        width, off = fixed_width, 8
    words = np.frombuffer(b[off:], dtype=np.uint64)
    n = nbits // width if width else 0
    return unpack_bits(words, n, width), width


def read_chunk_files(prefix):
    """Read <prefix>.dbg.chunk{,.W,.last,.weights} as the reference writes them
    (boss_chunk.cpp:372-386; serialize_number = big-endian u64)."""
    W, _ = read_sdsl_int_vector(prefix + ".dbg.chunk.W")
    last, _ = read_sdsl_int_vector(prefix + ".dbg.chunk.last", fixed_width=1)
    wts, wwidth = read_sdsl_int_vector(prefix + ".dbg.chunk.weights")
    b = open(prefix + ".dbg.chunk", "rb").read()
    nbits = int(np.frombuffer(b[:8], dtype=np.uint64)[0])
    nF = nbits // 64
    F = np.frombuffer(b[9:9 + 8 * nF], dtype=np.uint64).copy()
    sigma = int.from_bytes(b[9 + 8 * nF: 17 + 8 * nF], "big")
    k = int.from_bytes(b[17 + 8 * nF: 25 + 8 * nF], "big")
    return {"W": W, "last": last, "weights": wts if len(wts) else None,
            "weights_width": wwidth, "F": F, "sigma": sigma, "k": k}


def read_fasta(path):
    reads, cur = [], []
    with open(path) as f:
        for line in f:
            line = line.rstrip("\r\n")
            if line.startswith(">"):
                if cur:
                    reads.append("".join(cur))
                cur = []
            else:
                cur.append(line)
    if cur:
        reads.append("".join(cur))
    return reads


def ref_driver(alphabet="DNA"):
    p = os.path.join(REF_DIR, "ref_driver_" + alphabet)
    return p if os.path.exists(p) else None
