"""Deterministic synthetic reads for the BASELINE.json configs.

A counter-based generator (splitmix64 finaliser of seed/stream/index), so the numpy
implementation (CPU, tests, reference arm) and the torch implementation (reads
generated directly in HBM for the device-resident bench leg) produce identical bytes.

DNA: random genome of G bases; read r = genome[start_r : start_r+len] with each base
substituted with probability e by a uniform base, reverse-complemented with
probability 1/2. Protein: i.i.d. uniform over the 20 standard amino acids.

All reads have the same length, so the library's packed layout (one 0x00 after each
read) is simply an (n_reads, len+1) uint8 matrix whose last column is zero.
"""
import numpy as np

_C1 = 0x9E3779B97F4A7C15
_C2 = 0xBF58476D1CE4E5B9
_C3 = 0x94D049BB133111EB
_M64 = (1 << 64) - 1
AA = b"ACDEFGHIKLMNPQRSTVWY"


def _mix_np(x):
    x = x.astype(np.uint64, copy=True)
    with np.errstate(over="ignore"):
        x += np.uint64(_C1)
        x = (x ^ (x >> np.uint64(30))) * np.uint64(_C2)
        x = (x ^ (x >> np.uint64(27))) * np.uint64(_C3)
        x ^= x >> np.uint64(31)
    return x


def _base_np(seed, stream):
    return np.uint64(((seed * _C1) + (stream * _C2)) & _M64)


def _h_np(seed, stream, idx):
    with np.errstate(over="ignore"):
        return _mix_np(idx.astype(np.uint64) + _base_np(seed, stream))


def genome_np(seed, G, lo=0, hi=None):
    hi = G if hi is None else hi
    return (_h_np(seed, 0, np.arange(lo, hi, dtype=np.uint64)) >> np.uint64(62)).astype(np.uint8)


def dna_reads_packed_np(seed, n_reads, read_len, G, e=0.005, r0=0, r1=None, genome=None):
    """Rows r0..r1 of the packed (n_reads, read_len+1) uint8 matrix."""
    r1 = n_reads if r1 is None else r1
    if genome is None:
        genome = genome_np(seed, G)
    r = np.arange(r0, r1, dtype=np.uint64)
    start = ((_h_np(seed, 1, r) >> np.uint64(1)) % np.uint64(G - read_len + 1)).astype(np.int64)
    rc = (_h_np(seed, 2, r) >> np.uint64(63)).astype(bool)
    pos = np.arange(read_len, dtype=np.int64)
    codes = genome[start[:, None] + pos[None, :]]
    u = _h_np(seed, 3, (r[:, None] * np.uint64(read_len) + pos[None, :].astype(np.uint64)))
    sub = (u >> np.uint64(40)) < np.uint64(int(e * (1 << 24)))
    codes = np.where(sub, ((u >> np.uint64(10)) & np.uint64(3)).astype(np.uint8), codes)
    codes_rc = (3 - codes)[:, ::-1]
    codes = np.where(rc[:, None], codes_rc, codes)
    out = np.zeros((r1 - r0, read_len + 1), dtype=np.uint8)
    out[:, :read_len] = np.frombuffer(b"ACGT", dtype=np.uint8)[codes]
    return out


def protein_reads_packed_np(seed, n_reads, read_len, r0=0, r1=None):
    r1 = n_reads if r1 is None else r1
    idx = np.arange(r0 * read_len, r1 * read_len, dtype=np.uint64)
    a = ((_h_np(seed, 4, idx) >> np.uint64(32)) % np.uint64(20)).astype(np.int64)
    out = np.zeros((r1 - r0, read_len + 1), dtype=np.uint8)
    out[:, :read_len] = np.frombuffer(AA, dtype=np.uint8)[a].reshape(r1 - r0, read_len)
    return out


def packed_to_reads(packed):
    """(n, len+1) packed matrix -> (bases bytes, offsets) for the host C ABI / oracle."""
    n, lp1 = packed.shape
    bases = np.ascontiguousarray(packed[:, : lp1 - 1]).reshape(-1)
    offsets = np.arange(n + 1, dtype=np.uint64) * np.uint64(lp1 - 1)
    return bases, offsets


def packed_to_strings(packed):
    return [bytes(row[:-1]).decode() for row in packed]


# ---- torch (device) implementation: identical stream -----------------------

def _to_i64(v):
    v &= _M64
    return v - (1 << 64) if v >= (1 << 63) else v


def _lsr(x, s):
    import torch  # noqa: F401
    return (x >> s) & ((1 << (64 - s)) - 1)


def _mix_t(x):
    x = x + _to_i64(_C1)
    x = (x ^ _lsr(x, 30)) * _to_i64(_C2)
    x = (x ^ _lsr(x, 27)) * _to_i64(_C3)
    return x ^ _lsr(x, 31)


def _h_t(seed, stream, idx):
    return _mix_t(idx + _to_i64((seed * _C1) + (stream * _C2)))


def genome_torch(seed, G, device="cuda"):
    """the genome codes (uint8, 0..3) on `device`"""
    import torch
    gen = torch.empty(G, dtype=torch.uint8, device=device)
    for lo in range(0, G, 1 << 26):
        hi = min(G, lo + (1 << 26))
        j = torch.arange(lo, hi, dtype=torch.int64, device=device)
        gen[lo:hi] = _lsr(_h_t(seed, 0, j), 62).to(torch.uint8)
    return gen


def dna_reads_packed_torch(seed, n_reads, read_len, G, e=0.005, device="cuda", chunk=1 << 20, r0=0,
                           r1=None, genome=None):
    """Rows r0..r1 of the packed (n_reads, read_len+1) uint8 matrix, generated on `device`."""
    import torch
    gen = genome if genome is not None else genome_torch(seed, G, device)
    r_end = n_reads if r1 is None else r1
    r_begin = r0
    out = torch.zeros((r_end - r_begin, read_len + 1), dtype=torch.uint8, device=device)
    lut = torch.tensor(list(b"ACGT"), dtype=torch.uint8, device=device)
    pos = torch.arange(read_len, dtype=torch.int64, device=device)
    thr = int(e * (1 << 24))
    for r0 in range(r_begin, r_end, chunk):
        r1 = min(r_end, r0 + chunk)
        r = torch.arange(r0, r1, dtype=torch.int64, device=device)
        start = _lsr(_h_t(seed, 1, r), 1) % (G - read_len + 1)
        rc = _lsr(_h_t(seed, 2, r), 63).to(torch.bool)
        codes = gen[start[:, None] + pos[None, :]].to(torch.int64)
        u = _h_t(seed, 3, r[:, None] * read_len + pos[None, :])
        sub = _lsr(u, 40) < thr
        codes = torch.where(sub, _lsr(u, 10) & 3, codes)
        codes_rc = (3 - codes).flip(1)
        codes = torch.where(rc[:, None], codes_rc, codes)
        out[r0 - r_begin:r1 - r_begin, :read_len] = lut[codes]
    return out


def protein_reads_packed_torch(seed, n_reads, read_len, device="cuda", chunk=1 << 20, r0=0, r1=None):
    """Rows r0..r1 of the packed (n_reads, read_len+1) matrix of i.i.d. amino acids."""
    import torch
    r_begin, r_end = r0, (n_reads if r1 is None else r1)
    out = torch.zeros((r_end - r_begin, read_len + 1), dtype=torch.uint8, device=device)
    lut = torch.tensor(list(AA), dtype=torch.uint8, device=device)
    for a0 in range(r_begin, r_end, chunk):
        a1 = min(r_end, a0 + chunk)
        idx = torch.arange(a0 * read_len, a1 * read_len, dtype=torch.int64, device=device)
        a = _lsr(_h_t(seed, 4, idx), 32) % 20
        out[a0 - r_begin:a1 - r_begin, :read_len] = lut[a].reshape(a1 - a0, read_len)
    return out
