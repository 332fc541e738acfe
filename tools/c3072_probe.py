# one-tile C3072 -> C48 (factor 64) pressure coarsening, delp + 4 fields: time + sanity
import torch, time, sys
sys.path.insert(0, ".")
from fv3net_b200 import core
dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
nb, km, n, f = 1, 79, 3072, 64
# the bench's distribution (bench.py synth_device): smooth surface pressure + 1.5 kPa of noise
yy = torch.linspace(0, 6, n, device=dev)[:, None]
xx = torch.linspace(0, 6, n, device=dev)[None, :]
ps = (8.6e4 + 1.4e4 * torch.sin(yy * 1.3) * torch.cos(xx * 1.1) + 1500 * torch.randn((nb, n, n), device=dev, generator=g)).clamp(5.0e4, 1.04e5)[:, None]
b = torch.diff(torch.linspace(0, 1, km + 1, device=dev) ** 1.6)
delp = ((ps - 300.0) * b[None, :, None, None] * (1 + 0.02 * (torch.rand((nb, km, n, n), device=dev, generator=g) - 0.5))).float()
del ps
area = (3.6e8 + 4.7e8 * torch.rand((nb, n, n), device=dev, generator=g)).float()
fields = [(180 + 140 * torch.rand((nb, km, n, n), device=dev, generator=g)).float() for _ in range(4)]
ones = torch.ones_like(fields[0])
for _ in range(2):
    outs, dco = core.pressure_coarsen(fields, delp, area, f, 300.0, mode=core.FAST)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    outs, dco = core.pressure_coarsen(fields, delp, area, f, 300.0, mode=core.FAST)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
pts = nb * km * n * n * 4
alg = nb * n * n * (4 * km * 5 + 4)
print(f"C3072 tile f=64: {ms:.3f} ms, {pts/ms*1e-6:.1f} Gpts/s, {alg/ms*1e-6:.1f} GB/s algorithmic")
(c1,), _ = core.pressure_coarsen([ones], delp, area, f, 300.0, mode=core.FAST)
ok = ~torch.isnan(c1)
print("constant field reproduced:", float((c1[ok] - 1).abs().max()), "coverage", float(ok.float().mean()))
print("dco == wba:", torch.equal(dco, core.weighted_block_average(delp, area[:, None], f)))
print("T range", float(outs[0][~torch.isnan(outs[0])].min()), float(outs[0][~torch.isnan(outs[0])].max()))
# same tile through the regular 8 x 8 kernel: separates the cost of the wide path from the layout
for ff in (8, 16, 32):
    for _ in range(2):
        core.pressure_coarsen(fields, delp, area, ff, 300.0, mode=core.FAST)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(5):
        core.pressure_coarsen(fields, delp, area, ff, 300.0, mode=core.FAST)
    e1.record(); torch.cuda.synchronize()
    print(f"f={ff}: {e0.elapsed_time(e1)/5:.3f} ms")
e0.record()
for _ in range(5):
    core.weighted_block_average(delp, area[:, None], 64)
e1.record(); torch.cuda.synchronize()
print(f"prepass wba(delp) f=64: {e0.elapsed_time(e1)/5:.3f} ms")
