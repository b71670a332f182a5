"""ctypes binding of libmtg_gpu.so (include/mtg_gpu.h). No fallback: if the CUDA
library is missing or cannot be loaded, importing the product path raises."""
import ctypes
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(HERE, "libmtg_gpu.so")

MTG_OK = 0
MTG_MAX_SIGMA = 32
MTG_MAX_K = 128
ALPHABETS = {"DNA": 0, "Protein": 1}
ERR_NAMES = {1: "MTG_ERR_BAD_ARG", 2: "MTG_ERR_CUDA", 3: "MTG_ERR_OOM", 4: "MTG_ERR_STATE",
             5: "MTG_ERR_UNSUPPORTED"}


class MtgParams(ctypes.Structure):
    _fields_ = [("k", ctypes.c_uint32), ("alphabet", ctypes.c_uint32),
                ("both_strands", ctypes.c_uint32), ("bits_per_count", ctypes.c_uint32),
                ("filter_suffix", ctypes.c_char_p), ("device", ctypes.c_int32),
                ("reserved", ctypes.c_uint32), ("memory_preallocated", ctypes.c_double)]


class MtgChunk(ctypes.Structure):
    _fields_ = [("n_rows", ctypes.c_uint64), ("sigma", ctypes.c_uint32), ("k", ctypes.c_uint32),
                ("W_width", ctypes.c_uint32), ("weights_width", ctypes.c_uint32),
                ("F", ctypes.c_uint64 * MTG_MAX_SIGMA),
                ("W", ctypes.POINTER(ctypes.c_uint64)), ("last", ctypes.POINTER(ctypes.c_uint64)),
                ("weights", ctypes.POINTER(ctypes.c_uint64)),
                ("n_input_kmers", ctypes.c_uint64), ("n_real", ctypes.c_uint64),
                ("n_dummy_sink", ctypes.c_uint64), ("n_dummy_source", ctypes.c_uint64 * MTG_MAX_K),
                ("ms_total", ctypes.c_float), ("ms_extract", ctypes.c_float),
                ("ms_sort", ctypes.c_float), ("ms_unique", ctypes.c_float),
                ("ms_rc", ctypes.c_float), ("ms_dummy", ctypes.c_float), ("ms_emit", ctypes.c_float),
                ("n_kernel_launches", ctypes.c_uint32), ("device_resident", ctypes.c_uint32),
                ("ms_onesweep_sum", ctypes.c_float), ("ms_sort_histogram", ctypes.c_float),
                ("n_onesweep", ctypes.c_uint32), ("onesweep_bytes_per_key", ctypes.c_uint32),
                ("onesweep_keys", ctypes.c_uint64),
                ("impl", ctypes.c_void_p)]


EXPORTS = ["mtg_create", "mtg_destroy", "mtg_last_error", "mtg_push_reads",
           "mtg_push_packed_device", "mtg_packed_size", "mtg_pack_reads", "mtg_build_chunk",
           "mtg_build_chunk_device", "mtg_set_stream", "mtg_free_chunk", "mtg_debug_sorted_kmers", "mtg_debug_free", "mtg_debug_sort_u64",
           "mtg_debug_sort_u128", "mtg_debug_sort_bench", "mtg_debug_unique_bench", "mtg_mg_part_bits", "mtg_mg_extract", "mtg_mg_count", "mtg_mg_route", "mtg_mg_targets_count", "mtg_mg_targets_route",
           "mtg_mg_collect", "mtg_mg_strands", "mtg_mg_merge", "mtg_mg_targets", "mtg_mg_dummies", "mtg_mg_build", "mtg_peer_push", "mtg_mg_debug_pending", "mtg_mg_accumulate", "mtg_alloc_pinned", "mtg_free_pinned", "mtg_debug_sort_config"]

_lib = None


class MtgError(RuntimeError):
    def __init__(self, code, msg):
        super().__init__("%s: %s" % (ERR_NAMES.get(code, "MTG_ERR_%d" % code), msg))
        self.code = code


def load():
    """Load the CUDA library; raises if it has not been built (no CPU fallback)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError("%s is missing: build it with `python -c 'import __graft_entry__ as g; "
                          "g.build()'` (there is no CPU fallback)" % LIB_PATH)
    lib = ctypes.CDLL(LIB_PATH)
    u64p = ctypes.POINTER(ctypes.c_uint64)
    lib.mtg_create.argtypes = [ctypes.POINTER(MtgParams), ctypes.POINTER(ctypes.c_void_p)]
    lib.mtg_create.restype = ctypes.c_int
    lib.mtg_destroy.argtypes = [ctypes.c_void_p]
    lib.mtg_destroy.restype = None
    lib.mtg_last_error.argtypes = [ctypes.c_void_p]
    lib.mtg_last_error.restype = ctypes.c_char_p
    lib.mtg_push_reads.argtypes = [ctypes.c_void_p, ctypes.c_void_p, u64p, u64p, ctypes.c_size_t]
    lib.mtg_push_reads.restype = ctypes.c_int
    lib.mtg_push_packed_device.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_size_t,
                                           ctypes.c_uint64]
    lib.mtg_push_packed_device.restype = ctypes.c_int
    lib.mtg_packed_size.argtypes = [u64p, ctypes.c_size_t]
    lib.mtg_packed_size.restype = ctypes.c_size_t
    lib.mtg_pack_reads.argtypes = [ctypes.c_void_p, u64p, ctypes.c_size_t, ctypes.c_void_p]
    lib.mtg_pack_reads.restype = None
    lib.mtg_build_chunk.argtypes = [ctypes.c_void_p, ctypes.POINTER(MtgChunk)]
    lib.mtg_build_chunk.restype = ctypes.c_int
    lib.mtg_build_chunk_device.argtypes = [ctypes.c_void_p, ctypes.POINTER(MtgChunk)]
    lib.mtg_build_chunk_device.restype = ctypes.c_int
    lib.mtg_set_stream.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
    lib.mtg_set_stream.restype = ctypes.c_int
    lib.mtg_free_chunk.argtypes = [ctypes.POINTER(MtgChunk)]
    lib.mtg_free_chunk.restype = None
    lib.mtg_debug_sorted_kmers.argtypes = [ctypes.c_void_p, ctypes.POINTER(u64p),
                                           ctypes.POINTER(ctypes.POINTER(ctypes.c_uint32)), u64p]
    lib.mtg_debug_sorted_kmers.restype = ctypes.c_int
    lib.mtg_debug_free.argtypes = [ctypes.c_void_p]
    lib.mtg_debug_free.restype = None
    for name in ("mtg_debug_sort_u64", "mtg_debug_sort_u128"):
        f = getattr(lib, name)
        f.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                      ctypes.c_uint32]
        f.restype = ctypes.c_int
    lib.mtg_debug_sort_bench.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int,
                                         ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
    lib.mtg_debug_sort_bench.restype = ctypes.c_int
    lib.mtg_debug_unique_bench.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int,
                                           ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
    lib.mtg_debug_unique_bench.restype = ctypes.c_int
    vpp = ctypes.POINTER(ctypes.c_void_p)
    lib.mtg_mg_part_bits.argtypes = [ctypes.c_void_p]
    lib.mtg_mg_part_bits.restype = ctypes.c_uint32
    lib.mtg_mg_extract.argtypes = [ctypes.c_void_p, u64p, vpp, u64p]
    lib.mtg_mg_count.argtypes = [ctypes.c_void_p, u64p, u64p]
    lib.mtg_mg_route.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint32), ctypes.c_uint32, u64p]
    lib.mtg_mg_targets_count.argtypes = [ctypes.c_void_p, u64p, u64p]
    lib.mtg_mg_targets_route.argtypes = [ctypes.c_void_p, ctypes.POINTER(ctypes.c_uint32), ctypes.c_uint32, u64p]
    lib.mtg_mg_collect.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, u64p]
    lib.mtg_mg_targets.argtypes = [ctypes.c_void_p, u64p, vpp, u64p]
    lib.mtg_peer_push.argtypes = [ctypes.c_void_p, ctypes.c_uint32, ctypes.POINTER(ctypes.c_void_p),
                                  ctypes.POINTER(ctypes.c_void_p), u64p]
    lib.mtg_peer_push.restype = ctypes.c_int
    lib.mtg_debug_sort_config.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    lib.mtg_debug_sort_config.restype = None
    lib.mtg_alloc_pinned.argtypes = [ctypes.c_size_t]
    lib.mtg_alloc_pinned.restype = ctypes.c_void_p
    lib.mtg_free_pinned.argtypes = [ctypes.c_void_p]
    lib.mtg_free_pinned.restype = None
    lib.mtg_mg_accumulate.argtypes = [ctypes.c_void_p, ctypes.c_int, u64p]
    lib.mtg_mg_accumulate.restype = ctypes.c_int
    lib.mtg_mg_debug_pending.argtypes = [ctypes.c_void_p, ctypes.c_int, vpp, u64p]
    lib.mtg_mg_debug_pending.restype = ctypes.c_int
    lib.mtg_mg_strands.argtypes = [ctypes.c_void_p, u64p, u64p, vpp, vpp, u64p, vpp, vpp, u64p]
    lib.mtg_mg_merge.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64,
                                 ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, u64p, ctypes.c_uint32, u64p]
    lib.mtg_mg_dummies.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, u64p, ctypes.c_uint32, ctypes.c_uint32, vpp,
                                   u64p, u64p]
    lib.mtg_mg_build.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_uint64, ctypes.c_uint64,
                                 ctypes.c_uint64, ctypes.c_int, ctypes.c_int, ctypes.POINTER(MtgChunk)]
    for name in ("mtg_mg_extract", "mtg_mg_count", "mtg_mg_route", "mtg_mg_targets_count", "mtg_mg_targets_route", "mtg_mg_collect", "mtg_mg_strands", "mtg_mg_merge", "mtg_mg_targets", "mtg_mg_dummies", "mtg_mg_build"):
        getattr(lib, name).restype = ctypes.c_int
    _lib = lib
    return lib
