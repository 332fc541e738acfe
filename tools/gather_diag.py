"""dev helper (torchrun, N >= 2): where the fused gather's time goes -- kernel alone / kernel + peer stores /
+ per-step barrier on a side stream"""
import os, sys, time
import torch, torch.distributed as dist
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from fv3net_b200 import core
from fv3net_b200.gather import CoarseGather
rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(lr); dev = torch.device("cuda", lr)
dist.init_process_group("nccl", device_id=dev)
delp, area, fields = bench.synth_device(torch, 1000 + rank, dev, "smooth")
cdim = bench.NFINE // bench.FACTOR
nbuf = 3
gath = [CoarseGather(bench.NFIELDS, bench.NTILE, bench.KM, cdim, cdim, dev) for _ in range(nbuf)]
local = torch.empty((bench.NFIELDS + 1, bench.NTILE, bench.KM, cdim, cdim), dtype=torch.float32, device=dev)
side = torch.cuda.Stream(device=dev)
def timeit(fn, n=30):
    for _ in range(5): fn(0)
    torch.cuda.synchronize(); dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n): fn(i)
    torch.cuda.current_stream().wait_stream(side)
    e1.record(); torch.cuda.synchronize()
    t = torch.tensor([e0.elapsed_time(e1) / n], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return t.item()
k = lambda **kw: core.pressure_coarsen(fields, delp, area, bench.FACTOR, bench.TOA, mask_mode="phalf", mode=core.FAST, **kw)
a = timeit(lambda i: k())
b = timeit(lambda i: k(out=local))
c = timeit(lambda i: k(out=local, peer_ptrs=gath[i % nbuf].peer_ptrs))
bar_done = [None] * nbuf
def full(i):
    b_ = i % nbuf
    comp = torch.cuda.current_stream()
    j = (b_ + 1) % nbuf
    if bar_done[j] is not None: comp.wait_event(bar_done[j])
    k(out=local, peer_ptrs=gath[b_].peer_ptrs)
    ev = torch.cuda.Event(); ev.record(comp); side.wait_event(ev)
    with torch.cuda.stream(side):
        gath[b_].barrier(); bar_done[b_] = torch.cuda.Event(); bar_done[b_].record(side)
d = timeit(full)
if rank == 0:
    print(f"world {world}: kernel {a:.4f} ms | out= {b:.4f} | + peer stores {c:.4f} | + barrier on side stream {d:.4f}")
dist.destroy_process_group()
