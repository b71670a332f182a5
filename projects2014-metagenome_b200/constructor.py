"""Python mirror of the reference's construction interface for the `metagraph build`
hot path, implemented on top of the C ABI (include/mtg_gpu.h):

  IBOSSChunkConstructor::initialize / add_sequence / add_sequences / build_chunk / get_k
      (graph/representation/succinct/boss_chunk_construct.hpp:18-34)
  BOSS::Chunk: size / extend / serialize / load
      (graph/representation/succinct/boss_chunk.hpp:19-104, boss_chunk.cpp:230-386)

Names, argument meaning and error behaviour follow the reference (errors that the
reference reports with logger->error + exit(1) are raised as MtgError).
"""
import ctypes
import weakref

import numpy as np

from . import _lib
from ._lib import MtgChunk, MtgError, MtgParams


class Chunk:
    """BOSS::Chunk: W, last, F (+ weights) of a (part of a) BOSS table.
    W / last / weights are kept as the packed little-endian 64-bit words of the sdsl
    int_vector files the reference stores (row 0 is the placeholder)."""

    kFileExtension = ".dbg.chunk"

    def __init__(self, alph_size, k, n_rows, W_words, last_words, F, W_width,
                 weights_words=None, weights_width=0, stats=None):
        self.alph_size = int(alph_size)
        self.k = int(k)
        self.n_rows = int(n_rows)
        self.W_width = int(W_width)
        self.W_words = W_words
        self.last_words = last_words
        self.F = np.asarray(F, dtype=np.uint64)
        self.weights_words = weights_words
        self.weights_width = int(weights_width)
        self.stats = stats or {}

    def size(self):
        return self.n_rows

    # ---- unpacked views -------------------------------------------------
    @staticmethod
    def _unpack(words, n, width):
        flat = np.unpackbits(np.ascontiguousarray(words).view(np.uint8), bitorder="little")
        bits = flat[: n * width].reshape(n, width).astype(np.uint64)
        return (bits << np.arange(width, dtype=np.uint64)[None, :]).sum(axis=1).astype(np.uint64)

    @staticmethod
    def _pack(values, width):
        values = np.asarray(values, dtype=np.uint64)
        n = len(values)
        nwords = (n * width + 63) // 64
        bits = (values[:, None] >> np.arange(width, dtype=np.uint64)[None, :]) & np.uint64(1)
        flat = np.zeros(nwords * 64, dtype=np.uint8)
        flat[: n * width] = bits.astype(np.uint8).reshape(-1)
        return np.packbits(flat, bitorder="little").view(np.uint64)

    @property
    def W(self):
        return self._unpack(self.W_words, self.n_rows, self.W_width)

    @property
    def last(self):
        return self._unpack(self.last_words, self.n_rows, 1)

    @property
    def weights(self):
        if self.weights_words is None:
            return None
        return self._unpack(self.weights_words, self.n_rows, self.weights_width)

    # ---- BOSS::Chunk::extend (boss_chunk.cpp:230-270) ---------------------
    @staticmethod
    def _append_bits(dst_words, dst_n, width, src_words, src_first, src_n):
        """dst (dst_n entries of `width` bits, packed LSB-first in uint64 words) + entries
        [src_first, src_first + src_n) of src -> new word array. Word-wise: every output word is a funnel
        shift of two source words (vectorised), not one entry per step -- chunks hold billions of rows."""
        nbits, sb, db = src_n * width, src_first * width, dst_n * width
        total_words = (db + nbits + 63) // 64
        out = np.zeros(total_words, dtype=np.uint64)
        keep = (db + 63) // 64
        out[:keep] = np.asarray(dst_words[:keep], dtype=np.uint64)
        if db & 63:                                   # padding bits of dst's last word must be zero
            out[keep - 1] &= np.uint64((1 << (db & 63)) - 1)
        if nbits == 0:
            return out
        src = np.concatenate([np.asarray(src_words, dtype=np.uint64), np.zeros(2, dtype=np.uint64)])
        # the appended bit stream, word aligned: stream[j] = bits [sb + 64 j, sb + 64 j + 64) of src
        nsw = (nbits + 63) // 64
        w0, o = sb >> 6, sb & 63
        lo = src[w0:w0 + nsw]
        if o:
            stream = (lo >> np.uint64(o)) | (src[w0 + 1:w0 + 1 + nsw] << np.uint64(64 - o))
        else:
            stream = lo.copy()
        if nbits & 63:
            stream[-1] &= np.uint64((1 << (nbits & 63)) - 1)
        # place it at bit offset db of out
        d0, do = db >> 6, db & 63
        if do == 0:
            out[d0:d0 + nsw] = stream
        else:
            out[d0:d0 + nsw] |= stream << np.uint64(do)
            hi = stream >> np.uint64(64 - do)
            n_hi = min(nsw, total_words - (d0 + 1))
            out[d0 + 1:d0 + 1 + n_hi] |= hi[:n_hi]
        return out

    def extend(self, other):
        if (self.alph_size != other.alph_size or self.k != other.k
                or self.W_width != other.W_width):
            raise ValueError("ERROR: trying to concatenate incompatible graph chunks")
        if other.n_rows == 1:
            return
        if (self.weights_words is not None) != (other.weights_words is not None):
            raise ValueError("ERROR: trying to concatenate weighted and unweighted blocks")
        n, m = self.n_rows, other.n_rows - 1          # other's row 0 is its placeholder
        self.W_words = self._append_bits(self.W_words, n, self.W_width, other.W_words, 1, m)
        self.last_words = self._append_bits(self.last_words, n, 1, other.last_words, 1, m)
        if self.weights_words is not None:
            self.weights_words = self._append_bits(self.weights_words, n, self.weights_width,
                                                   other.weights_words, 1, m)
        self.F = self.F + other.F
        self.n_rows = n + m

    # ---- BOSS::Chunk::serialize / load (boss_chunk.cpp:330-386) -----------
    @staticmethod
    def _int_vector_bytes(words, n, width, fixed_width):
        hdr = np.uint64(n * width).tobytes()
        if not fixed_width:
            hdr += bytes([width])
        nwords = (n * width + 63) // 64
        return hdr + np.ascontiguousarray(words[:nwords], dtype=np.uint64).tobytes()

    def serialize(self, outbase):
        fname = outbase if outbase.endswith(self.kFileExtension) else outbase + self.kFileExtension
        with open(fname + ".W", "wb") as f:
            f.write(self._int_vector_bytes(self.W_words, self.n_rows, self.W_width, False))
        with open(fname + ".last", "wb") as f:
            f.write(self._int_vector_bytes(self.last_words, self.n_rows, 1, True))
        with open(fname + ".weights", "wb") as f:
            if self.weights_words is not None:
                f.write(self._int_vector_bytes(self.weights_words, self.n_rows,
                                               self.weights_width, False))
            else:   # empty int_vector_buffer<> with the default width 64
                f.write(np.uint64(0).tobytes() + bytes([64]))
        with open(fname, "wb") as f:
            # serialize_number_vector: sdsl int_vector<64>; serialize_number: big-endian u64
            f.write(self._int_vector_bytes(self.F, len(self.F), 64, False))
            f.write(int(self.alph_size).to_bytes(8, "big"))
            f.write(int(self.k).to_bytes(8, "big"))

    @classmethod
    def load(cls, infbase):
        fname = infbase if infbase.endswith(cls.kFileExtension) else infbase + cls.kFileExtension

        def read_iv(path, fixed=None):
            b = open(path, "rb").read()
            nbits = int(np.frombuffer(b[:8], dtype=np.uint64)[0])
            width, off = (b[8], 9) if fixed is None else (fixed, 8)
            return np.frombuffer(b[off:], dtype=np.uint64).copy(), nbits // max(width, 1), width

        Ww, n, W_width = read_iv(fname + ".W")
        lw, nl, _ = read_iv(fname + ".last", 1)
        ww, nw, wwidth = read_iv(fname + ".weights")
        b = open(fname, "rb").read()
        nF = int(np.frombuffer(b[:8], dtype=np.uint64)[0]) // 64
        F = np.frombuffer(b[9:9 + 8 * nF], dtype=np.uint64).copy()
        alph = int.from_bytes(b[9 + 8 * nF:17 + 8 * nF], "big")
        k = int.from_bytes(b[17 + 8 * nF:25 + 8 * nF], "big")
        if not (k and alph and n == nl and len(F) == alph and (nw == 0 or nw == n)):
            raise ValueError("ERROR: File corrupted. Cannot load graph chunk " + fname)
        return cls(alph, k, n, Ww, lw, F, W_width, ww if nw else None, wwidth if nw else 0)


class BOSSChunkConstructor:
    """IBOSSChunkConstructor backed by the CUDA library.

    initialize(k, both_strands=False, bits_per_count=0, filter_suffix="", num_threads=1,
               memory_preallocated=0, ...)  -- boss_chunk_construct.hpp:22-31.
    `k` is the BOSS node length (CLI -k minus 1)."""

    def __init__(self, k, both_strands=False, bits_per_count=0, filter_suffix="",
                 num_threads=1, memory_preallocated=0, alphabet="DNA", device=0, no_batching=False):
        self._lib = _lib.load()
        self._ctx = ctypes.c_void_p()
        self._k = int(k)
        self._tight_bits = 2 if alphabet == "DNA" else 5   # bits per character of the sort key (alphabets.hpp)
        prm = MtgParams(k=int(k), alphabet=_lib.ALPHABETS[alphabet],
                        both_strands=int(bool(both_strands)), bits_per_count=int(bits_per_count),
                        filter_suffix=filter_suffix.encode() if filter_suffix else None,
                        device=int(device), reserved=1 if no_batching else 0,   # MTG_FLAG_NO_BATCHING
                        memory_preallocated=float(memory_preallocated))
        rc = self._lib.mtg_create(ctypes.byref(prm), ctypes.byref(self._ctx))
        if rc != _lib.MTG_OK:
            raise MtgError(rc, self._lib.mtg_last_error(None).decode())

    @classmethod
    def initialize(cls, k, both_strands=False, bits_per_count=0, filter_suffix="",
                   num_threads=1, memory_preallocated=0, **kw):
        return cls(k, both_strands, bits_per_count, filter_suffix, num_threads,
                   memory_preallocated, **kw)

    def get_k(self):
        return self._k

    def tight_bits(self):
        return self._tight_bits

    def _check(self, rc):
        if rc != _lib.MTG_OK:
            raise MtgError(rc, self._lib.mtg_last_error(self._ctx).decode())

    def add_sequence(self, sequence, count=1):
        self.add_sequences([(sequence, count)] if count != 1 else [sequence])

    def add_sequences(self, sequences):
        """sequences: list of str/bytes, or list of (str/bytes, count) pairs."""
        if not len(sequences):
            return
        counts = None
        if isinstance(sequences[0], tuple):
            counts = np.asarray([c for _, c in sequences], dtype=np.uint64)
            sequences = [s for s, _ in sequences]
        bs = [s.encode() if isinstance(s, str) else bytes(s) for s in sequences]
        offsets = np.zeros(len(bs) + 1, dtype=np.uint64)
        offsets[1:] = np.cumsum([len(b) for b in bs], dtype=np.uint64)
        self.add_packed(b"".join(bs), offsets, counts)

    def add_packed(self, bases, offsets, counts=None):
        """Concatenated reads + offsets (the C ABI's native form)."""
        offsets = np.ascontiguousarray(offsets, dtype=np.uint64)
        u64p = ctypes.POINTER(ctypes.c_uint64)
        if isinstance(bases, np.ndarray):
            bptr = bases.ctypes.data_as(ctypes.c_void_p)
        else:
            bptr = ctypes.cast(ctypes.c_char_p(bases), ctypes.c_void_p)
        cptr = None
        if counts is not None:
            counts = np.ascontiguousarray(counts, dtype=np.uint64)
            cptr = counts.ctypes.data_as(u64p)
        self._check(self._lib.mtg_push_reads(self._ctx, bptr, offsets.ctypes.data_as(u64p), cptr,
                                             len(offsets) - 1))

    def add_packed_device(self, dptr, n_bytes, n_windows_max):
        """Reads already resident in HBM in the packed layout (0x00 after each read)."""
        self._check(self._lib.mtg_push_packed_device(self._ctx, ctypes.c_void_p(dptr),
                                                     n_bytes, n_windows_max))

    def set_stream(self, cuda_stream):
        """Run the library's work on the caller's CUDA stream (an int cudaStream_t handle)."""
        self._check(self._lib.mtg_set_stream(self._ctx, ctypes.c_void_p(cuda_stream)))

    @staticmethod
    def _stats(c):
        return {
            "n_rows": int(c.n_rows),
            "n_input_kmers": int(c.n_input_kmers), "n_real": int(c.n_real),
            "n_dummy_sink": int(c.n_dummy_sink),
            "n_dummy_source": [int(c.n_dummy_source[i]) for i in range(c.k)],
            "ms_total": c.ms_total, "ms_extract": c.ms_extract, "ms_sort": c.ms_sort,
            "ms_unique": c.ms_unique, "ms_rc": c.ms_rc, "ms_dummy": c.ms_dummy,
            "ms_emit": c.ms_emit, "n_kernel_launches": int(c.n_kernel_launches),
            "ms_onesweep_sum": c.ms_onesweep_sum, "ms_sort_histogram": c.ms_sort_histogram,
            "n_onesweep": int(c.n_onesweep),
            "onesweep_bytes_per_key": int(c.onesweep_bytes_per_key),
            "onesweep_keys": int(c.onesweep_keys),
            "chunk_bytes": 8 * (((c.n_rows * c.W_width + 63) // 64) + ((c.n_rows + 63) // 64)
                                + ((c.n_rows * c.weights_width + 63) // 64)),
        }

    def build_chunk_device(self):
        """Build with the chunk left in HBM; returns the statistics only (F included)."""
        c = MtgChunk()
        self._check(self._lib.mtg_build_chunk_device(self._ctx, ctypes.byref(c)))
        try:
            st = self._stats(c)
            st["F"] = [int(c.F[i]) for i in range(c.sigma)]
            return st
        finally:
            self._lib.mtg_free_chunk(ctypes.byref(c))

    def build_chunk(self, copy=True):
        """BOSS::Chunk of everything pushed so far. copy=False leaves W / last / weights in the
        library's pinned result buffer (what the C ABI hands out): the arrays are views that stay
        valid as long as the returned Chunk object is alive."""
        c = MtgChunk()
        self._check(self._lib.mtg_build_chunk(self._ctx, ctypes.byref(c)))
        if copy:
            try:
                return self._chunk_from(c)
            finally:
                self._lib.mtg_free_chunk(ctypes.byref(c))
        chunk = self._chunk_from(c, copy=False)
        lib = self._lib
        weakref.finalize(chunk, lambda: lib.mtg_free_chunk(ctypes.byref(c)))
        return chunk

    # ---- range-partitioned (multi-GPU) phases, include/mtg_gpu.h mtg_mg_* ----------------
    def mg_part_bits(self):
        return int(self._lib.mtg_mg_part_bits(self._ctx))

    def mg_extract(self):
        """-> (bucket_counts uint64[2^part_bits], device pointer, n k-mers)"""
        counts = np.zeros(1 << self.mg_part_bits(), dtype=np.uint64)
        ptr, n = ctypes.c_void_p(), ctypes.c_uint64()
        self._check(self._lib.mtg_mg_extract(self._ctx, counts.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                                             ctypes.byref(ptr), ctypes.byref(n)))
        return counts, ptr.value or 0, n.value

    def mg_count(self):
        """extraction + per-bucket counts; the k-mers stay in the context -> (bucket_counts, n)"""
        counts = np.zeros(1 << self.mg_part_bits(), dtype=np.uint64)
        n = ctypes.c_uint64()
        self._check(self._lib.mtg_mg_count(self._ctx, counts.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                                           ctypes.byref(n)))
        return counts, n.value

    def mg_route(self, bucket_owner, owner_dst):
        """one scatter kernel: a counted k-mer of bucket b is stored into the array at device address
        owner_dst[bucket_owner[b]]"""
        own = np.ascontiguousarray(bucket_owner, dtype=np.uint32)
        dst = np.ascontiguousarray(owner_dst, dtype=np.uint64)
        assert own.shape[0] == 1 << self.mg_part_bits()
        self._check(self._lib.mtg_mg_route(self._ctx, own.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)),
                                           dst.shape[0], dst.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))))

    def mg_targets_count(self):
        """target nodes of this range's reals + their per-bucket counts -> (bucket_counts, n)"""
        counts = np.zeros(1 << self.mg_part_bits(), dtype=np.uint64)
        n = ctypes.c_uint64()
        self._check(self._lib.mtg_mg_targets_count(self._ctx, counts.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                                                   ctypes.byref(n)))
        return counts, n.value

    def mg_targets_route(self, bucket_owner, owner_dst):
        own = np.ascontiguousarray(bucket_owner, dtype=np.uint32)
        dst = np.ascontiguousarray(owner_dst, dtype=np.uint64)
        assert own.shape[0] == 1 << self.mg_part_bits()
        self._check(self._lib.mtg_mg_targets_route(self._ctx, own.ctypes.data_as(ctypes.POINTER(ctypes.c_uint32)),
                                                   dst.shape[0], dst.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64))))

    def mg_accumulate(self, last):
        """fold the last mg_collect result into the accumulated k-mers of this range (super-batched build)"""
        m = ctypes.c_uint64()
        self._check(self._lib.mtg_mg_accumulate(self._ctx, int(bool(last)), ctypes.byref(m)))
        return m.value

    def mg_debug_pending(self, which):
        """verification hook: (device pointer, n) of the array the next route call scatters
        (which = 0: k-mers after mg_count, 1: target nodes after mg_targets_count)"""
        ptr, n = ctypes.c_void_p(), ctypes.c_uint64()
        self._check(self._lib.mtg_mg_debug_pending(self._ctx, int(which), ctypes.byref(ptr), ctypes.byref(n)))
        return ptr.value or 0, n.value

    def mg_collect(self, dptr, n):
        m = ctypes.c_uint64()
        self._check(self._lib.mtg_mg_collect(self._ctx, ctypes.c_void_p(dptr), n, ctypes.byref(m)))
        return m.value

    def mg_strands(self):
        """canonical mode -> (x_counts, r_counts, (x_ptr, xc_ptr, n_x), (r_ptr, rc_ptr, n_r))"""
        nb = 1 << self.mg_part_bits()
        xc, rc = np.zeros(nb, dtype=np.uint64), np.zeros(nb, dtype=np.uint64)
        u64p = ctypes.POINTER(ctypes.c_uint64)
        px, pxc, pr, prc = ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p(), ctypes.c_void_p()
        nx, nr = ctypes.c_uint64(), ctypes.c_uint64()
        self._check(self._lib.mtg_mg_strands(self._ctx, xc.ctypes.data_as(u64p), rc.ctypes.data_as(u64p),
                                             ctypes.byref(px), ctypes.byref(pxc), ctypes.byref(nx),
                                             ctypes.byref(pr), ctypes.byref(prc), ctypes.byref(nr)))
        return xc, rc, (px.value or 0, pxc.value or 0, nx.value), (pr.value or 0, prc.value or 0, nr.value)

    def mg_merge(self, x_ptr, xc_ptr, n_x, r_ptr, rc_ptr, n_r, r_runs=None):
        m = ctypes.c_uint64()
        rl = np.ascontiguousarray(r_runs if r_runs is not None else [], dtype=np.uint64)
        self._check(self._lib.mtg_mg_merge(self._ctx, ctypes.c_void_p(x_ptr), ctypes.c_void_p(xc_ptr), n_x,
                                           ctypes.c_void_p(r_ptr), ctypes.c_void_p(rc_ptr), n_r,
                                           rl.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)) if len(rl) else None,
                                           len(rl), ctypes.byref(m)))
        return m.value

    def mg_targets(self):
        counts = np.zeros(1 << self.mg_part_bits(), dtype=np.uint64)
        ptr, n = ctypes.c_void_p(), ctypes.c_uint64()
        self._check(self._lib.mtg_mg_targets(self._ctx, counts.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)),
                                             ctypes.byref(ptr), ctypes.byref(n)))
        return counts, ptr.value or 0, n.value

    def mg_dummies(self, dptr, n_in, runs=None, runs_per_set=0):
        """runs: lengths of the sorted runs dptr consists of, or None; runs_per_set: consecutive runs
        that overlap in key space (0 = all; sets are ordered and disjoint)"""
        ptr, n_src, n_sink = ctypes.c_void_p(), ctypes.c_uint64(), ctypes.c_uint64()
        rl = np.ascontiguousarray(runs if runs is not None else [], dtype=np.uint64)
        self._check(self._lib.mtg_mg_dummies(self._ctx, ctypes.c_void_p(dptr), n_in,
                                             rl.ctypes.data_as(ctypes.POINTER(ctypes.c_uint64)) if len(rl) else None,
                                             len(rl), int(runs_per_set), ctypes.byref(ptr),
                                             ctypes.byref(n_src), ctypes.byref(n_sink)))
        return ptr.value or 0, n_src.value, n_sink.value

    def mg_build(self, dptr, n_all, lo_bucket, hi_bucket, first_range, device_output=False):
        c = MtgChunk()
        self._check(self._lib.mtg_mg_build(self._ctx, ctypes.c_void_p(dptr), n_all, lo_bucket, hi_bucket,
                                           int(first_range), int(device_output), ctypes.byref(c)))
        try:
            if device_output:
                st = self._stats(c)
                st["F"] = [int(c.F[i]) for i in range(c.sigma)]
                return st
            return self._chunk_from(c)
        finally:
            self._lib.mtg_free_chunk(ctypes.byref(c))

    def _chunk_from(self, c, copy=True):
        n = c.n_rows
        nW = (n * c.W_width + 63) // 64
        nl = (n + 63) // 64
        take = (lambda a: a.copy()) if copy else (lambda a: a)
        W = take(np.ctypeslib.as_array(c.W, (nW,)))
        last = take(np.ctypeslib.as_array(c.last, (nl,)))
        wts = None
        if c.weights_width:
            nw = (n * c.weights_width + 63) // 64
            wts = take(np.ctypeslib.as_array(c.weights, (nw,)))
        F = np.array(c.F[: c.sigma], dtype=np.uint64)
        return Chunk(c.sigma, c.k, n, W, last, F, c.W_width, wts, c.weights_width, self._stats(c))

    def debug_sorted_kmers(self):
        u64p = ctypes.POINTER(ctypes.c_uint64)
        words = u64p()
        counts = ctypes.POINTER(ctypes.c_uint32)()
        n = ctypes.c_uint64()
        self._check(self._lib.mtg_debug_sorted_kmers(self._ctx, ctypes.byref(words),
                                                     ctypes.byref(counts), ctypes.byref(n)))
        m = n.value
        w = np.ctypeslib.as_array(words, (max(m, 1), 2))[:m].copy()
        c = None
        if counts:
            c = np.ctypeslib.as_array(counts, (max(m, 1),))[:m].copy()
            self._lib.mtg_debug_free(counts)
        self._lib.mtg_debug_free(words)
        return w, c

    def close(self):
        if self._ctx:
            self._lib.mtg_destroy(self._ctx)
            self._ctx = ctypes.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


IBOSSChunkConstructor = BOSSChunkConstructor
