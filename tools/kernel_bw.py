"""dev helper: achieved HBM bandwidth of the thin kernels at C384 (6 x 79 x 384 x 384 fp32), algorithmic bytes
 / CUDA-event time; peak from MEASURED_PEAKS.json."""
import json, os, sys
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fv3net_b200 import core

dev = torch.device("cuda")
g = torch.Generator(device=dev).manual_seed(0)
nb, km, n, f = 6, 79, 384, 8
try:
    peak = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    peak = 6650.0
T = torch.rand((nb, km, n, n), device=dev, generator=g)
T2 = torch.rand((nb, km, n, n), device=dev, generator=g)  # second field so that inputs alternate (> L2 anyway: 280 MB each)
delp = torch.rand((nb, km, n, n), device=dev, generator=g) * 100 + 900
area = torch.rand((nb, n, n), device=dev, generator=g) + 0.5
u = torch.rand((nb, km, n + 1, n), device=dev, generator=g)
dx = torch.rand((nb, n + 1, n), device=dev, generator=g) + 0.5
coarse = torch.rand((nb, km, n // f, n // f), device=dev, generator=g)
E = 4


def timeit(name, fn, nbytes, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    gbs = nbytes / ms * 1e-6
    print(f"{name:58s} {ms:8.3f} ms  {gbs:8.1f} GB/s  {100 * gbs / peak:5.1f} % of {peak:.0f}")


fld = T.numel() * E
cf = coarse.numel() * E
timeit("K1 weighted_block_average f=8, 2-D area weights", lambda: core.weighted_block_average(T, area[:, None], f), fld + area.numel() * E + cf)
timeit("K1 mass-weighted (delp*area formed in the kernel)", lambda: core.weighted_block_average(T, area[:, None], f, weights2=delp), 2 * fld + area.numel() * E + cf)
timeit("K1 weighted_block_average f=64 (wide)", lambda: core.weighted_block_average(T, area[:, None], 64), fld + area.numel() * E)
need = nb * km * (n // f + 1) * n * E
timeit("K2 edge_weighted_block_average u/dx (needed lines only)", lambda: core.edge_weighted_block_average(u, dx[:, None], f, edge="x"), need + nb * (n // f + 1) * n * E + nb * km * (n // f + 1) * (n // f) * E)
for m in ("sum", "min", "median", "mode"):
    timeit(f"K3 block_reduce {m} f=8", lambda m=m: core.block_reduce(T, f, m), fld + cf)
timeit("K4 block_upsample f=8 (read coarse, write fine)", lambda: core.block_upsample(coarse, f), fld + cf)
timeit("K5 stagger delp (all lines)", lambda: core.stagger(delp, "y"), fld + nb * km * (n + 1) * n * E)
timeit("K5 stagger delp (every 8th line)", lambda: core.stagger(delp, "y", line_step=f), 2 * nb * km * (n // f + 1) * n * E + nb * km * (n // f + 1) * n * E)
timeit("pressure_at_interface", lambda: core.pressure_at_interface(delp, 300.0), fld + nb * (km + 1) * n * n * E)
pi = core.pressure_at_interface(delp, 300.0)
po = core.pressure_at_interface(delp * 1.01, 300.0)
timeit("K7 regrid_vertical (mappm, kord=1) [z,y,x] in place", lambda: core.regrid_vertical(pi, T, po), 2 * nb * (km + 1) * n * n * E + 2 * fld)
timeit("K7 regrid_vertical FAST (b200cs_mappm_fast)", lambda: core.regrid_vertical(pi, T, po, mode=core.FAST), 2 * nb * (km + 1) * n * n * E + 2 * fld)
# reads coarse interfaces 1..km, the fine surface level and the weights; writes km levels
timeit("mask_weights", lambda: core.mask_weights(area, po, pi, "phalf"), 2 * fld + nb * n * n * E + area.numel() * E)
