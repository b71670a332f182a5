"""dist_check.py's cases as g virtual ranks on ONE GPU: python tests/virtual_check.py [world ...]"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # this directory: oracle_lib
import mtg_b200  # noqa: E402,F401
import oracle_lib  # noqa: E402
from mtg_b200 import multi_gpu, synth  # noqa: E402

worlds = [int(x) for x in sys.argv[1:]] or [8]
cases = [(31, True, 0), (31, True, 8), (31, False, 12), (63, True, 0), (12, False, 0), (40, True, 16)]
bad = 0
for K, canonical, bpc in cases:
    reads = synth.packed_to_strings(synth.dna_reads_packed_np(100 + K, 30_000, 120, 300_000, 0.01))
    ref = oracle_lib.oracle_build(reads, K, canonical=canonical, bits_per_count=bpc)
    for world in worlds:
        for fused in (True, False):
            vb = multi_gpu.VirtualBuild(world, K, canonical, bpc, fused_route=fused)
            for r in range(world):
                vb.add_sequences(r, reads[r::world])
            full = vb.build_chunk()
            vb.close()
            good = (full.n_rows == ref["n_rows"] and np.array_equal(full.W, ref["W"])
                    and np.array_equal(full.last, ref["last"]) and np.array_equal(full.F, ref["F"])
                    and (not bpc or np.array_equal(full.weights, ref["weights"])))
            print("virtual_check K=%d canonical=%s bpc=%d world=%d fused=%s: %s (%d rows, oracle %d)"
                  % (K, canonical, bpc, world, fused, "OK" if good else "MISMATCH", full.n_rows, ref["n_rows"]), flush=True)
            bad += 0 if good else 1
sys.exit(1 if bad else 0)
