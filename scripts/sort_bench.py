"""Times the onesweep radix sort alone (device-resident pseudo-random keys) for each kernel
configuration (MTG_SORT_VARIANT), after checking the configuration sorts correctly."""
import ctypes, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

def one(variant):
    import numpy as np
    import mtg_b200
    lib = mtg_b200._lib.load()
    # correctness (stable, ragged size, with payload)
    rng = np.random.default_rng(1)
    n = 1_000_003
    keys = rng.integers(0, 2 ** 62, n, dtype=np.uint64)
    keys[: n // 3] &= np.uint64(0xffff)
    vals = np.arange(n, dtype=np.uint32)
    k, v = keys.copy(), vals.copy()
    assert lib.mtg_debug_sort_u64(0, k.ctypes.data_as(ctypes.c_void_p), v.ctypes.data_as(ctypes.c_void_p), n, 62) == 0
    order = np.argsort(keys, kind="stable")
    ok = bool(np.array_equal(k, keys[order]) and np.array_equal(v, vals[order]))
    lo = rng.integers(0, 2 ** 62, 300_001, dtype=np.uint64); hi = rng.integers(0, 2 ** 60, 300_001, dtype=np.uint64)
    hi[:1000] = hi[0]
    kk = np.stack([lo, hi], axis=1).copy()
    assert lib.mtg_debug_sort_u128(0, kk.ctypes.data_as(ctypes.c_void_p), None, len(lo), 126) == 0
    o2 = np.lexsort((lo, hi))
    ok = ok and bool(np.array_equal(kk[:, 0], lo[o2]) and np.array_equal(kk[:, 1], hi[o2]))
    res = {"variant": variant, "correct": ok}
    ms = (ctypes.c_float * 5)()
    for name, n, bits, wv in (("u64", 400_000_000, 62, 0), ("u64+val", 300_000_000, 62, 1), ("u128", 200_000_000, 126, 0)):
        rc = lib.mtg_debug_sort_bench(0, n, bits, wv, 3, ms)
        bpk = (16 if bits > 64 else 8) + (4 if wv else 0)
        res[name] = {"rc": rc, "ms_hist": round(ms[0], 3), "ms_pass": round(ms[1], 3), "ms_sort": round(ms[2], 3), "ms_first_pass": round(ms[3], 3), "ms_later_passes": round(ms[4], 3),
                     "pass_GBps": round(2 * n * bpk / (ms[1] * 1e-3) / 1e9, 1), "n": n}
    print(json.dumps(res), flush=True)

if __name__ == "__main__":
    if len(sys.argv) > 1:
        one(int(sys.argv[1]))
    else:
        # 3 = the round-1 pass kernel (peer masks in shared memory), 5 = atomic slot + collision repair
        for v in (os.environ.get("VARIANTS", "3,5")).split(","):
            env = dict(os.environ, MTG_SORT_VARIANT=v)
            subprocess.call([sys.executable, __file__, v], env=env)
        env = dict(os.environ, MTG_SORT_VARIANT="5", MTG_BENCH_UNSTABLE_FIRST="1")
        subprocess.call([sys.executable, __file__, "5"], env=env)
