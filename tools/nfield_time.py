"""dev helper: CUDA-event time of the fused FAST kernel for 1, 2, 3, 4, 5, 9 fields at C384 (how a launch's cost
splits into geometry/control and per-field work), and agreement of the 1- and 2-field instances with the 4-field one."""
import os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fv3net_b200 import core

dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
nb, km, n, f = 6, 79, 384, 8
ps = 98000 + 1500 * torch.rand((nb, 1, n, n), device=dev, generator=g)
delp = (ps - 300.0) / km * (0.95 + 0.1 * torch.rand((nb, km, n, n), device=dev, generator=g))
area = torch.rand((nb, n, n), device=dev, generator=g) + 0.5
fields = [torch.randn((nb, km, n, n), device=dev, generator=g) for _ in range(9)]
ref, _ = core.pressure_coarsen(fields[:4], delp, area, f, 300.0, mask_mode="phalf", mode=core.FAST)
for nf in (1, 2, 3, 4, 5, 9):
    fl = fields[:nf]
    for _ in range(3):
        out, _ = core.pressure_coarsen(fl, delp, area, f, 300.0, mask_mode="phalf", mode=core.FAST)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        core.pressure_coarsen(fl, delp, area, f, 300.0, mask_mode="phalf", mode=core.FAST)
    e1.record()
    torch.cuda.synchronize()
    same = all(torch.equal(torch.nan_to_num(a), torch.nan_to_num(b)) for a, b in zip(out[:4], ref))
    print(f"{nf} field(s): {e0.elapsed_time(e1) / 10:.3f} ms   first fields identical to the 4-field launch: {same}")
