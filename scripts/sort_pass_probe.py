"""One short sort per pass-kernel configuration in ONE process, for `ncu -k regex:onesweep_pass`:
v3 stable (default), v5 stable, v5 with an unstable first pass. 200 M random 62-bit keys each."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import mtg_b200
lib = mtg_b200._lib.load()
ms = (ctypes.c_float * 5)()
n = int(os.environ.get("PROBE_KEYS", "200000000"))
for variant, unstable in ((3, 0), (5, 0), (5, 1)):
    lib.mtg_debug_sort_config(variant, unstable)
    if unstable:
        os.environ["MTG_BENCH_UNSTABLE_FIRST"] = "1"
    assert lib.mtg_debug_sort_bench(0, n, 62, 0, 1, ms) == 0
    print("variant %d unstable_first %d: first pass %.3f ms, later passes %.3f ms" % (variant, unstable, ms[3], ms[4]), flush=True)
