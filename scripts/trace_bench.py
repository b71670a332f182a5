import sys, time, os
sys.path.insert(0,'/root/repo')
import torch, mtg_b200
from mtg_b200 import synth
n=10_000_000
packed = synth.dna_reads_packed_torch(1, n, 150, 100_000_000, 0.005, device='cuda')
ctor = mtg_b200.BOSSChunkConstructor.initialize(30, True, 0)
ctor.set_stream(torch.cuda.current_stream().cuda_stream)
for i in range(4):
    torch.cuda.synchronize(); t0=time.perf_counter()
    ctor.add_packed_device(packed.data_ptr(), packed.numel(), n*120)
    t1=time.perf_counter()
    st=ctor.build_chunk_device()
    t2=time.perf_counter()
    torch.cuda.synchronize(); t3=time.perf_counter()
    print("step",i,"push %.2f build %.2f sync %.2f ms; device %.2f"%((t1-t0)*1e3,(t2-t1)*1e3,(t3-t2)*1e3, st['ms_total']), file=sys.stderr)
