#!/usr/bin/env python
"""Benchmark of the B200-native `metagraph build` hot path (BOSS chunk construction).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]

Metric (BASELINE.json): input k-mers/s for the BOSS build, k=31. A step is one whole pass of
the hot path (extract -> sort -> unique -> rc completion -> dummy k-mers -> W/last/F) over
one batch of synthetic reads. N=1 workload = BASELINE.json configs[1]: 10M x 150bp DNA reads,
k=31, canonical. `value` = reads resident in HBM, chunk left in HBM; `e2e` = the same through
the host C ABI (pinned host reads in, chunk in pinned host memory out).
"""
import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "input k-mers/sec for BOSS build k=31"
UNIT = "kmers/s"

WORKLOADS = {
    # name: (n_reads, read_len, genome, seed, e, K, canonical, bits_per_count)
    "cfg2": dict(name="synthetic 10Mx150bp DNA reads, k=31, canonical", n_reads=10_000_000,
                 read_len=150, G=100_000_000, seed=1, e=0.005, K=31, canonical=True, bpc=0),
    # the other BASELINE configs (named sizes; `--workload cfgN`): secondary lines under profiles/
    "cfg3": dict(name="synthetic 100Mx150bp DNA reads, k=31, --count-kmers", n_reads=100_000_000,
                 read_len=150, G=1_000_000_000, seed=2, e=0.005, K=31, canonical=False, bpc=8,
                 scaling="strong"),
    "cfg4": dict(name="synthetic 50Mx150bp DNA reads, k=63 (128-bit k-mers)", n_reads=50_000_000,
                 read_len=150, G=500_000_000, seed=3, e=0.005, K=63, canonical=False, bpc=0,
                 scaling="strong"),
    "cfg5": dict(name="synthetic 5Mx300aa protein sequences, k=12 (Protein alphabet)", n_reads=5_000_000,
                 read_len=300, G=0, seed=4, e=0.0, K=12, canonical=False, bpc=0, alphabet="Protein",
                 scaling="strong"),
}
SLICE_READS = 6_250_000      # reads are generated and pushed in slices of at most this many


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=None, help="cfg2 (default at N=1) or cfg3")
    ap.add_argument("--reads", type=int, default=0, help="override the number of reads (debug)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--cpu-sample-reads", type=int, default=0,
                    help="reads of the CPU legs (0 = the whole workload, the default: same config as the GPU arm)")
    ap.add_argument("--cpu-mem-gb", type=float, default=16.0,
                    help="--mem-cap of the reference build (16 GB: its best case, no shrink cycles on cfg #2)")
    ap.add_argument("--no-parity-check", action="store_true",
                    help="N > 1: skip the untimed DistBuild-vs-oracle verification builds")
    return ap.parse_args()


class ClockSampler:
    """SM clock / throttle reasons sampled DURING the timed region through NVML (the same
    counters `nvidia-smi --query-gpu=clocks.sm,clocks_event_reasons.*` prints), every 200 ms."""

    def __init__(self, index=0, period=0.2):
        self.index = index
        self.period = period
        self.samples = []
        self.reasons = set()
        self.stop_flag = threading.Event()
        self.thread = None
        self.max_mhz = None

    def start(self):
        if os.environ.get("BENCH_NO_CLOCKS"):
            return
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = float(pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM))
        except Exception:
            self.nv = None
            return
        self.thread = threading.Thread(target=self._run, daemon=True)
        self.thread.start()

    def _run(self):
        nv = self.nv
        names = {
            "hw_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwSlowdown", 0x8),
            "hw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonHwThermalSlowdown", 0x40),
            "sw_thermal_slowdown": getattr(nv, "nvmlClocksThrottleReasonSwThermalSlowdown", 0x20),
            "sw_power_cap": getattr(nv, "nvmlClocksThrottleReasonSwPowerCap", 0x4),
        }
        while not self.stop_flag.is_set():
            try:
                self.samples.append(float(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM)))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                for nm, bit in names.items():
                    if r & bit:
                        self.reasons.add(nm)
            except Exception:
                pass
            self.stop_flag.wait(self.period)

    def stop(self):
        if self.thread is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["clock sampling unavailable"]}
        self.stop_flag.set()
        self.thread.join(timeout=2)
        return {"sm_mhz": statistics.median(self.samples) if self.samples else None,
                "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(self.samples)}


def algorithmic_bytes(st, R, w, c, P, canonical, W_width, bpc):
    """SURVEY.md 8d byte model, evaluated on the counters of the actual build."""
    N = st["n_input_kmers"]
    Mp = st["n_real"]
    D = st["n_dummy_sink"] + sum(st["n_dummy_source"])
    E = st["n_rows"]
    M = (Mp + 1) // 2 if canonical else Mp     # distinct canonical k-mers (palindromes ~ 0)
    b = R + N * (w + c)                         # 1 extract
    b += N * (w + c) * (1 + 2 * P)              # 2 sort
    b += N * (w + c) + M * (w + c)              # 3 unique / count-reduce
    if canonical:                               # 4 rc generate, sort rc, merge
        b += 2 * M * (w + c) + M * (w + c) * (1 + 2 * P) + 2 * Mp * (w + c)
    b += 4 * Mp * w + D * w * (1 + 2 * P)       # 5 dummy recovery
    b += (Mp + D) * (w + c) + E * (W_width + 1 + bpc) / 8   # 6 merge + emit
    return b


def cpu_reference_run(wl, n_reads, threads, repeats=1, mem_gb=16.0):
    """Times the reference's own CPU implementation (oracle/_ref/ref_driver_DNA, built from the
    unmodified /root/reference sources) on a bounded sample of the workload. Returns
    (kmers_per_s, info) or (None, why)."""
    import numpy as np
    import mtg_b200
    from mtg_b200 import synth
    import oracle_lib
    alphabet = wl.get("alphabet", "DNA")
    drv = oracle_lib.ref_driver(alphabet)
    G = max(wl["read_len"] * 2, int(wl["G"] * n_reads / wl["n_reads"]))   # same coverage
    n_kmers = n_reads * (wl["read_len"] - wl["K"] + 1)
    with tempfile.TemporaryDirectory() as td:
        if drv is None:
            # oracle port (single thread) on a smaller sample
            n_reads = min(n_reads, 100_000)
            G = max(wl["read_len"] * 2, int(wl["G"] * n_reads / wl["n_reads"]))
            n_kmers = n_reads * (wl["read_len"] - wl["K"] + 1)
            packed = synth.dna_reads_packed_np(wl["seed"], n_reads, wl["read_len"], G, wl["e"])
            reads = synth.packed_to_strings(packed)
            t0 = time.perf_counter()
            oracle_lib.oracle_build(reads, wl["K"], canonical=wl["canonical"],
                                    bits_per_count=wl["bpc"])
            dt = time.perf_counter() - t0
            return n_kmers / dt, dict(kind="port", cores=1, seconds=dt, all_seconds=[dt] * repeats,
                                      n_reads=n_reads, sample="%d reads x %d bp "
                                      "(same coverage), oracle/mtg_oracle.cpp" % (n_reads, wl["read_len"]))
        shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else td
        path = os.path.join(shm, "mtg_bench_reads_%d.txt" % os.getpid())
        genome = synth.genome_np(wl["seed"], G) if alphabet == "DNA" else None
        with open(path, "wb") as f:
            for r0 in range(0, n_reads, 200_000):
                r1 = min(n_reads, r0 + 200_000)
                if alphabet == "DNA":
                    p = synth.dna_reads_packed_np(wl["seed"], n_reads, wl["read_len"], G, wl["e"], r0, r1,
                                                  genome=genome)
                else:
                    p = synth.protein_reads_packed_np(wl["seed"], n_reads, wl["read_len"], r0, r1)
                p[:, -1] = ord("\n")
                f.write(p.tobytes())
        cmd = [drv, "-k", str(wl["K"]), "-p", str(threads), "--mem-gb", str(mem_gb), "-o",
               os.path.join(td, "out"), path]
        if wl["canonical"]:
            cmd.append("--canonical")
        if wl["bpc"]:
            cmd += ["--count-width", str(wl["bpc"])]
        secs, push, read = [], [], []
        try:
            for _ in range(repeats):
                out = subprocess.run(cmd, capture_output=True, text=True, timeout=3000)
                if out.returncode != 0:
                    return None, "ref_driver failed: " + out.stderr[-200:]
                fields = out.stdout.split()
                secs.append(float(fields[fields.index("build_sec") + 1]))
                push.append(float(fields[fields.index("push_sec") + 1]))
                read.append(float(fields[fields.index("read_sec") + 1]))
        finally:
            if os.path.exists(path):
                os.remove(path)
        build_sec = secs[-1]
        whole = n_reads == wl["n_reads"]
        return n_kmers / build_sec, dict(
            kind="reference", cores=threads, seconds=build_sec, all_seconds=secs, n_reads=n_reads,
            push_seconds=push, read_seconds_untimed=read,
            sample="%s: %d reads x %d bp, genome %d bp; the reference's IBOSSChunkConstructor (unmodified sources, "
                   "g++ -O3) add_sequences + build_chunk from reads parsed into host memory beforehand, "
                   "-p %d, --mem-cap %g GB"
                   % ("the WHOLE workload" if whole else "a sample with the workload's coverage",
                      n_reads, wl["read_len"], G, threads, mem_gb))


def run_reference(args, wl):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    n_reads = min(wl["n_reads"], args.cpu_sample_reads) if args.cpu_sample_reads else wl["n_reads"]
    # a full-size reference build takes ~30 s: at most 3 of them (1 warm-up + 2 timed), said in the line
    runs = max(1, min(3, args.warmup + args.steps))
    n_warm = 1 if runs > 1 and args.warmup > 0 else 0
    v, info = cpu_reference_run(wl, n_reads, threads, repeats=runs, mem_gb=args.cpu_mem_gb)
    if v is None:
        print(json.dumps({"impl": "reference", "unavailable": info}))
        return
    secs = info["all_seconds"][n_warm:]
    per_step = info["n_reads"] * (wl["read_len"] - wl["K"] + 1)
    value = per_step * len(secs) / sum(secs)
    vals = [(value, x) for x in secs]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT,
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sum(s for _, s in vals) / len(vals), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64", "data": "synthetic",
        "config": {"workload": wl["name"], "n_reads": info["n_reads"], "read_len": wl["read_len"], "k": wl["K"],
                   "mode": "canonical" if wl["canonical"] else "basic", "count_width": wl["bpc"],
                   "sample": info["sample"],
                   "repeats": "%d full builds run (%d warm-up + %d timed): capped at 3, each takes tens of "
                              "seconds" % (runs, n_warm, len(secs))},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": info["cores"],
                         "kind": info["kind"], "sample": info["sample"],
                         "seconds_per_build": secs, "push_seconds": info.get("push_seconds"),
                         "read_seconds_untimed": info.get("read_seconds_untimed")},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line))


def run_ours(args, wl):
    import numpy as np
    import torch
    import mtg_b200
    from mtg_b200 import synth

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the hot path has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        # all-to-all over NVLink: 64 p2p channels measured 620 GB/s per GPU and direction at 2 GPUs
        # against 425 GB/s with NCCL's default (scripts/a2a_bench.py)
        # stdout carries exactly one JSON line: NCCL prints its version banner there (NCCL_DEBUG=VERSION
        # or WARN), so fd 1 points at stderr until the line is printed (multi_gpu.bench restores it)
        sys.stdout.flush()
        os.environ["MTG_BENCH_STDOUT_FD"] = str(os.dup(1))
        os.dup2(2, 1)
        os.environ.setdefault("NCCL_MIN_P2P_NCHANNELS", "64")
        os.environ.setdefault("NCCL_MAX_P2P_NCHANNELS", "64")
        dist.init_process_group("nccl", device_id=dev)

    K, L = wl["K"], wl["read_len"]
    n_reads = wl["n_reads"]
    if world > 1:
        from mtg_b200 import multi_gpu
        parity = None
        if not args.no_parity_check:
            # untimed: the six tests/dist_check.py cases through the real DistBuild, checker = oracle
            import dist_check

            def parity(rank_, world_, local_):
                return dist_check.run_cases(rank_, world_, local_, variants=("1",), repeats=1,
                                            log=lambda m: print(m, file=sys.stderr, flush=True))
        try:
            return multi_gpu.bench(args, wl, rank, world, dev, METRIC, UNIT, ClockSampler,
                                   algorithmic_bytes, parity=parity)
        finally:
            multi_gpu.release_symmetric_buffers()
            dist.destroy_process_group()

    # ---- synthetic reads, generated straight into HBM in the packed layout ----
    alphabet = wl.get("alphabet", "DNA")
    slices = []
    genome = synth.genome_torch(wl["seed"], wl["G"], dev) if alphabet == "DNA" else None
    slice_reads = n_reads if n_reads <= 12_500_000 else SLICE_READS     # the headline workload is one slice
    for r0 in range(0, n_reads, slice_reads):
        r1 = min(n_reads, r0 + slice_reads)
        if alphabet == "DNA":
            slices.append(synth.dna_reads_packed_torch(wl["seed"], n_reads, L, wl["G"], wl["e"], device=dev,
                                                       r0=r0, r1=r1, genome=genome))
        else:
            slices.append(synth.protein_reads_packed_torch(wl["seed"], n_reads, L, device=dev, r0=r0, r1=r1))
    del genome
    packed = slices[0]
    n_bytes = sum(p.numel() for p in slices)
    n_windows = n_reads * (L - K + 1)
    stream = torch.cuda.current_stream()
    ctor = mtg_b200.BOSSChunkConstructor.initialize(K - 1, wl["canonical"], wl["bpc"], device=local,
                                                    alphabet=alphabet)
    ctor.set_stream(stream.cuda_stream)

    def step_device():
        # inputs beyond one sort buffer are folded into sorted runs slice by slice inside the library
        # (super-batching) and merged by the build
        for p in slices:
            ctor.add_packed_device(p.data_ptr(), p.numel(), p.shape[0] * (L - K + 1))
        return ctor.build_chunk_device()

    for _ in range(args.warmup):
        st = step_device()
    torch.cuda.synchronize()
    sampler = ClockSampler(local)
    sampler.start()
    time.sleep(0.3)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    stats = []
    torch.cuda.synchronize()
    ev0.record(stream)
    for _ in range(args.steps):
        stats.append(step_device())
    ev1.record(stream)
    torch.cuda.synchronize()
    clocks = sampler.stop()
    ms_total = ev0.elapsed_time(ev1)
    ms_per_step = ms_total / args.steps
    st = stats[-1]
    N = st["n_input_kmers"]
    value = N * args.steps / (ms_total * 1e-3)

    # ---- roofline of the dominant kernel (onesweep pass of the main sort) ----
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "MEASURED_PEAKS.json hbm_gbs (measured copy)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    n_launch = sum(s["n_onesweep"] for s in stats)      # 0 when the build was super-batched
    ms_kernel = sum(s["ms_onesweep_sum"] for s in stats) / max(n_launch, 1)
    bytes_launch = st["onesweep_keys"] * st["onesweep_bytes_per_key"]
    achieved = bytes_launch / (ms_kernel * 1e-3) / 1e9 if ms_kernel else 0.0
    Lb = 2 if alphabet == "DNA" else 5
    w = 8 if K * Lb <= 64 else 16
    c = 0
    P = (K * Lb + 7) // 8
    alg = algorithmic_bytes(st, n_bytes, w, c, P, wl["canonical"], 4 if alphabet == "DNA" else 6, wl["bpc"])
    device_ms = sum(s["ms_total"] for s in stats) / len(stats)
    # DRAM traffic of that kernel: bytes per key of the committed ncu --set full capture of the same
    # kernel (profiles/ncu_traffic.json says which) x the keys of this launch -- not measured in this run
    traffic, traffic_src = None, None
    try:
        with open(os.path.join(ROOT, "profiles", "ncu_traffic.json")) as f:
            ent = json.load(f)["onesweep_pass_v3_kernel<u64>"]
        if w == 8 and st["onesweep_keys"]:
            traffic = float(ent["dram_bytes_per_key"]) * int(st["onesweep_keys"])
            traffic_src = "not measured in this run: " + ent["source"]
    except Exception:
        pass
    roofline = {
        "bound": "hbm",
        "kernel": "onesweep pass of the main k-mer sort over all input k-mers (%d launches/step: the first one "
                  "onesweep_pass_v5_kernel<UNSTABLE>, the others onesweep_pass_v3_kernel; mean of all)"
                  % (n_launch // max(len(stats), 1)),
        "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
        "traffic": traffic, "traffic_source": traffic_src, "peak_source": peak_src,
        "algorithmic_bytes_per_launch": bytes_launch, "ms_per_launch": ms_kernel,
        "kernel_share_of_step": sum(s["ms_onesweep_sum"] for s in stats) / max(ms_total, 1e-9),
        "whole_path": {"algorithmic_bytes_per_step": alg, "bytes_per_input_kmer": alg / max(N, 1),
                       "device_ms_per_step": device_ms,
                       "achieved": alg / (device_ms * 1e-3) / 1e9,
                       "frac": alg / (device_ms * 1e-3) / 1e9 / peak},
        "phases_ms": {k2: sum(s[k2] for s in stats) / len(stats) for k2 in
                      ("ms_extract", "ms_sort", "ms_unique", "ms_rc", "ms_dummy", "ms_emit",
                       "ms_sort_histogram")},
    }

    # ---- end to end through the host C ABI: pinned host reads in, pinned chunk out ----
    e2e = None
    if not args.no_e2e and len(slices) > 1:
        e2e = {"value": None, "unit": UNIT, "skipped": "the end-to-end leg is measured on the headline workload "
               "(cfg2); this workload's reads (%.1f GB) are kept device-resident" % (n_bytes / 1e9)}
    elif not args.no_e2e:
        host = torch.empty((n_reads, L), dtype=torch.uint8, pin_memory=True)
        host.copy_(packed[:, :L])
        torch.cuda.synchronize()
        bases = host.numpy().reshape(-1)
        offsets = np.arange(n_reads + 1, dtype=np.uint64) * np.uint64(L)
        ctor2 = mtg_b200.BOSSChunkConstructor.initialize(K - 1, wl["canonical"], wl["bpc"], device=local,
                                                         alphabet=alphabet)

        host_ms = {"push": 0.0, "build": 0.0}

        def step_host():
            ta = time.perf_counter()
            ctor2.add_packed(bases, offsets)
            tb = time.perf_counter()
            c = ctor2.build_chunk(copy=False)   # W / last / F in the pinned host buffer of the C ABI
            host_ms["push"] += 1e3 * (tb - ta)
            host_ms["build"] += 1e3 * (time.perf_counter() - tb)
            return c

        chunk = step_host()   # warm-up (pinned staging, pools)
        chunk = step_host()   # second result buffer: the previous chunk is alive while the next is built
        torch.cuda.synchronize()
        e_steps = max(1, min(args.steps, 3))
        host_ms["push"] = host_ms["build"] = 0.0
        t0 = time.perf_counter()
        for _ in range(e_steps):
            chunk = step_host()
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        e2e = {"value": N * e_steps / dt, "unit": UNIT,
               "h2d_bytes_per_step": int(n_reads * (L + 1)),
               "d2h_bytes_per_step": int(chunk.stats["chunk_bytes"]),
               "ms_per_step": 1e3 * dt / e_steps, "steps": e_steps,
               "push_ms": host_ms["push"] / e_steps, "build_ms": host_ms["build"] / e_steps,
               "build_device_ms": float(chunk.stats["ms_total"]) if "ms_total" in chunk.stats else None,
               "api": "mtg_push_reads(host bases+offsets) + mtg_build_chunk (host W/last/F)"}
        ctor2.close()
        del host

    # ---- CPU baseline: the reference's own implementation on this box's host cores ----
    cpu = None
    if not args.no_cpu_baseline and (args.workload or "cfg2") != "cfg2" and not args.cpu_sample_reads:
        cpu = {"value": None, "unit": UNIT, "kind": "reference",
               "sample": "not run by default for this workload (pass --cpu-sample-reads N)"}
    elif not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        n_cpu = min(n_reads, args.cpu_sample_reads) if args.cpu_sample_reads else n_reads
        v, info = cpu_reference_run(wl, n_cpu, threads, mem_gb=args.cpu_mem_gb)
        if v is not None:
            cpu = {"value": v, "unit": UNIT, "cores": info["cores"], "kind": info["kind"],
                   "sample": info["sample"], "seconds": info["seconds"],
                   "push_seconds": info.get("push_seconds"),
                   "read_seconds_untimed": info.get("read_seconds_untimed")}
        else:
            cpu = {"value": None, "unit": UNIT, "cores": threads, "kind": "reference", "sample": info}

    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": 1, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "u64" if w == 8 else "u128", "data": "synthetic",
        "config": {"workload": wl["name"], "alphabet": alphabet, "n_reads": n_reads, "read_len": L, "k": K,
                   "mode": "canonical" if wl["canonical"] else "basic", "count_width": wl["bpc"],
                   "genome_bp": wl["G"], "subst_rate": wl["e"], "seed": wl["seed"],
                   "input_kmers": N, "real_kmers": st["n_real"], "rows": st["n_rows"],
                   "l2_policy": "inputs larger than L2 (reads %.2f GB, keys %.2f GB per step)"
                                % (n_bytes / 1e9, N * w / 1e9)},
        "clocks": clocks, "e2e": e2e, "gpu_launches": int(sum(s["n_kernel_launches"] for s in stats)),
        "roofline": roofline, "cpu_baseline": cpu,
    }
    print(json.dumps(line))
    ctor.close()


def main():
    args = parse_args()
    wl = dict(WORKLOADS[args.workload or "cfg2"])
    if args.reads:
        wl["G"] = max(2 * wl["read_len"], int(wl["G"] * args.reads / wl["n_reads"]))
        wl["n_reads"] = args.reads
        wl["name"] += " (scaled to %d reads)" % args.reads
    if args.impl == "reference":
        run_reference(args, wl)
    else:
        run_ours(args, wl)


if __name__ == "__main__":
    main()
