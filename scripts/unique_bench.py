"""Times the unique (fold) kernel alone: python scripts/unique_bench.py [n] [dup]"""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mtg_b200
lib = mtg_b200._lib.load()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_200_000_000
dup = int(sys.argv[2]) if len(sys.argv) > 2 else 4
ms = (ctypes.c_float * 3)()
lib.mtg_debug_unique_bench.argtypes = [ctypes.c_int, ctypes.c_uint64, ctypes.c_uint32, ctypes.c_int, ctypes.c_int, ctypes.POINTER(ctypes.c_float)]
rc = lib.mtg_debug_unique_bench(0, n, dup, 0, 5, ms)
gb = (n * 8 + ms[1] * 8) / 1e9
print("rc", rc, "env", {k: v for k, v in os.environ.items() if k.startswith("MTG_U")}, "ms %.3f" % ms[0], "distinct", int(ms[1]), "GB/s %.0f" % (gb / (ms[0] * 1e-3)))
