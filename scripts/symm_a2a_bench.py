"""All-to-all through NVLink peer memory (torch symmetric memory: every rank maps the receive buffers
of all peers and pushes its slices with peer stores) vs NCCL all_to_all_single.
torchrun ... scripts/symm_a2a_bench.py [MB per peer]"""
import os, sys, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 1200
n = mb * (1 << 20) // 8
send = torch.arange(n * world, dtype=torch.int64, device=dev) + rank * 1000
recv = symm.empty(n * world, dtype=torch.int64, device=dev)
hdl = symm.rendezvous(recv, dist.group.WORLD.group_name)
peers = [hdl.get_buffer(r, (n * world,), torch.int64) for r in range(world)]
streams = [torch.cuda.Stream() for _ in range(world)]
for it in range(4):
    torch.cuda.synchronize(); dist.barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    cur = torch.cuda.current_stream()
    for k in range(world):
        d = (rank + k) % world          # staggered so that no peer is hit by everyone at once
        st = streams[k]
        st.wait_stream(cur)
        with torch.cuda.stream(st):
            peers[d][rank * n:(rank + 1) * n].copy_(send[d * n:(d + 1) * n], non_blocking=True)
    for st in streams:
        cur.wait_stream(st)
    hdl.barrier()
    ev1.record(); ev1.synchronize()
    ms = ev0.elapsed_time(ev1)
    if rank == 0 and it:
        print("symm world %d  %d MB/peer  %.2f ms  %.0f GB/s sent per GPU" % (world, mb, ms, n * 8 * (world - 1) / ms / 1e6), flush=True)
ok = all(int(recv[s * n]) == s * 1000 + rank * n for s in range(world))
if rank == 0:
    print("symm data ok", ok, flush=True)
dist.destroy_process_group()
