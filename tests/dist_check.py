"""Parity of the real one-process-per-GPU build (DistBuild over NCCL / NVLink peer memory) against
the CPU oracle:  torchrun --nproc-per-node N tests/dist_check.py
Every rank pushes its share of seeded reads, the per-range chunks are gathered on rank 0,
concatenated (Chunk::extend) and compared bit for bit with the oracle's single chunk; every
exchange is additionally verified slice by slice (row counts + checksums, DistBuild(verify=True)).
`run_cases` is also what `bench.py --gpus N` runs (untimed) before it measures, so that every
multi-GPU bench line carries a parity verdict.
  DIST_CHECK_VARIANTS=1,0   values of MTG_ROUTE (1 = exchanges fused into the scatter kernels)
  DIST_CHECK_REPEATS=n      repeats of every case and variant
  DIST_CHECK_VERIFY=0       without the per-exchange verification (plain build timing)
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

CASES = [(31, True, 0), (31, True, 8), (31, False, 12), (63, True, 0), (12, False, 0), (40, True, 16)]


def run_cases(rank, world, local, variants=("1",), repeats=1, log=None, cases=CASES, n_reads=30_000, verify=True):
    """-> dict(cases, builds, ok, world, mismatches=[...]) on every rank (rank 0 decides, broadcast)."""
    log = log or (lambda m: print(m, flush=True))
    builds, bad = 0, []
    saved = os.environ.get("MTG_ROUTE")
    for K, canonical, bpc in cases:
        packed = synth.dna_reads_packed_np(100 + K, n_reads, 120, 10 * n_reads, 0.01)
        reads = synth.packed_to_strings(packed)
        ref = None
        for variant in variants:
            os.environ["MTG_ROUTE"] = variant
            for rep in range(repeats):
                db = multi_gpu.DistBuild(K, canonical, bpc, device=local, verify=verify)
                db.rank.ctor.add_sequences(reads[rank::world])
                err = None
                try:
                    chunk = db.build()
                except multi_gpu.ExchangeMismatch as e:     # raised on all ranks together
                    err, chunk = str(e), None
                builds += 1
                if chunk is not None:
                    st = torch.tensor([chunk.stats["n_real"], chunk.stats["n_dummy_sink"], chunk.n_rows], device="cuda")
                    allst = [torch.zeros_like(st) for _ in range(world)]
                    dist.all_gather(allst, st)
                    full = multi_gpu.gather_chunks(chunk, dist)
                tag = "K=%d canonical=%s bpc=%d world=%d route=%s rep=%d" % (
                    K, canonical, bpc, world, db.use_symm and db.fused_route, rep)
                if rank == 0:
                    import oracle_lib
                    if ref is None:
                        ref = oracle_lib.oracle_build(reads, K, canonical=canonical, bits_per_count=bpc)
                    if chunk is None:
                        good = False
                        log("dist_check %s: EXCHANGE MISMATCH (%s)" % (tag, err))
                    else:
                        good = (full.n_rows == ref["n_rows"] and np.array_equal(full.W, ref["W"])
                                and np.array_equal(full.last, ref["last"]) and np.array_equal(full.F, ref["F"])
                                and (not bpc or np.array_equal(full.weights, ref["weights"])))
                        log("dist_check %s: %s (%d rows; exchanges verified: %s)"
                            % (tag, "OK" if good else "MISMATCH", full.n_rows, ", ".join(db.verified)))
                        if not good:
                            per = [[int(v) for v in t.tolist()] for t in allst]
                            log("   per rank [n_real, n_sink, rows]: %s oracle n_real %d n_sink %d rows %d"
                                % (per, ref["n_real"], ref["n_sink"], ref["n_rows"]))
                    if not good:
                        bad.append(tag)
                db.close()
                dist.barrier()
    if saved is None:
        os.environ.pop("MTG_ROUTE", None)
    else:
        os.environ["MTG_ROUTE"] = saved
    flag = torch.tensor([len(bad)], device="cuda")
    dist.broadcast(flag, 0)
    return {"cases": len(cases), "builds": builds, "ok": int(flag.item()) == 0, "world": world,
            "variants": ["fused route" if v != "0" else "separate exchange" for v in variants],
            "repeats": repeats, "exchange_verification": bool(verify), "mismatches": bad,
            "checker": "oracle/mtg_oracle.cpp chunk (W, last, F, weights bit for bit) + per-slice exchange checksums"}


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    variants = os.environ.get("DIST_CHECK_VARIANTS", "1").split(",")      # values of MTG_ROUTE
    repeats = int(os.environ.get("DIST_CHECK_REPEATS", "1"))
    # DIST_CHECK_VERIFY=0: no per-exchange verification (it adds collectives and host synchronisation
    # after every exchange): the timing of a plain build, checked only through the final chunk
    verify = os.environ.get("DIST_CHECK_VERIFY", "1") != "0"
    res = run_cases(rank, world, local, variants, repeats, verify=verify)
    if rank == 0:
        print("dist_check summary: %s" % res, flush=True)
    multi_gpu.release_symmetric_buffers()
    dist.destroy_process_group()
    sys.exit(0 if res["ok"] else 1)


if __name__ == "__main__":
    main()
