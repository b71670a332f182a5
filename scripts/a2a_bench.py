"""all_to_all_single bandwidth between the GPUs of one box: torchrun ... scripts/a2a_bench.py [MB per peer]"""
import os, sys, torch, torch.distributed as dist
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
mb = int(sys.argv[1]) if len(sys.argv) > 1 else 2400
n = mb * (1 << 20) // 8
send = torch.arange(n * world, dtype=torch.int64, device=dev)
recv = torch.empty_like(send)
sizes = [n] * world
for it in range(4):
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); dist.barrier()
    ev0.record()
    dist.all_to_all_single(recv, send, sizes, sizes)
    ev1.record(); ev1.synchronize()
    ms = ev0.elapsed_time(ev1)
    if rank == 0 and it:
        sent = n * 8 * (world - 1)
        print("world %d  %d MB/peer  %.2f ms  %.0f GB/s sent per GPU  env %s" % (
            world, mb, ms, sent / ms / 1e6, {k: v for k, v in os.environ.items() if k.startswith("NCCL_")}), flush=True)
dist.destroy_process_group()
