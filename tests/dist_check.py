"""Parity of the real one-process-per-GPU build (DistBuild over NCCL / NVLink peer memory) against
the CPU oracle:  torchrun --nproc-per-node N tests/dist_check.py
Every rank pushes its share of seeded reads, the per-range chunks are gathered on rank 0,
concatenated (Chunk::extend) and compared bit for bit with the oracle's single chunk.
TEST INFRASTRUCTURE (uses oracle/)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))   # this directory: oracle_lib
import mtg_b200  # noqa: E402,F401
from mtg_b200 import multi_gpu, synth  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    cases = [(31, True, 0), (31, True, 8), (31, False, 12), (63, True, 0), (12, False, 0), (40, True, 16)]
    variants = os.environ.get("DIST_CHECK_VARIANTS", "1").split(",")      # values of MTG_ROUTE
    repeats = int(os.environ.get("DIST_CHECK_REPEATS", "1"))
    ok = True
    for K, canonical, bpc in cases:
        packed = synth.dna_reads_packed_np(100 + K, 30_000, 120, 300_000, 0.01)
        reads = synth.packed_to_strings(packed)
        ref = None
        for variant in variants:
            os.environ["MTG_ROUTE"] = variant
            for rep in range(repeats):
                db = multi_gpu.DistBuild(K, canonical, bpc, device=local)
                db.rank.ctor.set_stream(torch.cuda.current_stream().cuda_stream)
                db.rank.ctor.add_sequences(reads[rank::world])
                chunk = db.build()
                st = torch.tensor([chunk.stats["n_real"], chunk.stats["n_dummy_sink"], chunk.n_rows], device="cuda")
                allst = [torch.zeros_like(st) for _ in range(world)]
                dist.all_gather(allst, st)
                full = multi_gpu.gather_chunks(chunk, dist)
                if rank == 0:
                    import oracle_lib
                    if ref is None:
                        ref = oracle_lib.oracle_build(reads, K, canonical=canonical, bits_per_count=bpc)
                    good = (full.n_rows == ref["n_rows"] and np.array_equal(full.W, ref["W"])
                            and np.array_equal(full.last, ref["last"]) and np.array_equal(full.F, ref["F"])
                            and (not bpc or np.array_equal(full.weights, ref["weights"])))
                    print("dist_check K=%d canonical=%s bpc=%d world=%d route=%s rep=%d: %s (%d rows)"
                          % (K, canonical, bpc, world, db.use_symm and db.fused_route, rep,
                             "OK" if good else "MISMATCH", full.n_rows), flush=True)
                    if not good:
                        per = [[int(v) for v in t.tolist()] for t in allst]
                        print("   per rank [n_real, n_sink, rows]:", per, "oracle n_real %d n_sink %d rows %d"
                              % (ref["n_real"], ref["n_sink"], ref["n_rows"]), flush=True)
                    ok = ok and good
                db.close()
                dist.barrier()
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    multi_gpu.release_symmetric_buffers()
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) else 1)


if __name__ == "__main__":
    main()
