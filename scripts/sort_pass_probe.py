"""Pass-kernel configurations timed in ONE process (CUDA events inside mtg_debug_sort_bench; also the
program `ncu -k regex:onesweep_pass` captures): stable v3 / v5, unstable first pass, block shapes and
L2 prefetch distances. 400 M random 62-bit keys by default."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mtg_b200
lib = mtg_b200._lib.load()
ms = (ctypes.c_float * 5)()
n = int(os.environ.get("PROBE_KEYS", "400000000"))
os.environ["MTG_BENCH_UNSTABLE_FIRST"] = "1"
configs = [(3, 0, 0), (3, 0, 150), (3, 0, 300), (3, 0, 450), (3, 0, 600), (3, 0, 750), (3, 0, 900)]
if os.environ.get("PROBE_CONFIGS"):
    configs = [tuple(int(x) for x in c.split(":")) for c in os.environ["PROBE_CONFIGS"].split(",")]
for variant, shape, prefetch in configs:
    lib.mtg_debug_sort_config(variant, 1, shape, prefetch)
    for with_vals, wide in ((0, 0), (1, 0), (0, 1)):
        nn = n if not wide else n // 2
        assert lib.mtg_debug_sort_bench(0, nn, 126 if wide else 62, with_vals, 2, ms) == 0
        print("stable v%d shape %d prefetch %4d %s%s: unstable first pass %.3f ms, stable passes %.3f ms (%.0f GB/s)"
              % (variant, shape, prefetch, "u128" if wide else "u64", "+val" if with_vals else "    ", ms[3], ms[4],
                 2 * nn * ((16 if wide else 8) + 4 * with_vals) / ms[4] / 1e6), flush=True)
