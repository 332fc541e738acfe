"""dev helper: where the drop-in call's time goes (pageable numpy in, numpy out).
(1) b200cs_host_copy pageable -> pinned at several thread counts, alone and under a concurrent DMA;
(2) the whole cubedsphere.pressure_level_coarsen call at several pool sizes."""
import os, sys, time
import numpy as np
import torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fv3net_b200 import _lib, core, cubedsphere as cs
from fv3net_b200.labelled import DataArray, Dataset

lib = _lib.load()
nb, km, n, f = 6, 79, 384, 8
rng = np.random.default_rng(0)
slab = km * n * n * 4
src = [rng.random((km, n, n), dtype=np.float32) for _ in range(8)]
pin = [torch.empty(slab, dtype=torch.uint8).pin_memory() for _ in range(4)]
dev = [torch.empty(slab, dtype=torch.uint8, device="cuda") for _ in range(4)]
print("affinity", len(os.sched_getaffinity(0)), "cpu_count", os.cpu_count())
for nt in (1, 8, 16, 32):
    for dma in (False, True):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        for i in range(32):
            lib.b200cs_host_copy(pin[i % 4].data_ptr(), src[i % 8].ctypes.data, slab, nt)
            if dma:
                dev[i % 4].copy_(pin[i % 4], non_blocking=True)
        dt = time.perf_counter() - t0
        torch.cuda.synchronize()
        print(f"host_copy nt={nt:2d} dma={int(dma)}: {32 * slab / dt * 1e-9:6.1f} GB/s ({dt / 32 * 1e3:.2f} ms/slab)")

dims = ["tile", "zaxis_1", "yaxis_2", "xaxis_1"]
names = ["W", "T", "ua", "va"]
delp = (rng.random((nb, km, n, n), dtype=np.float32) * 100 + 900)
ds = Dataset({k: DataArray(rng.random((nb, km, n, n), dtype=np.float32), dims=dims) for k in names})
d = DataArray(delp, dims=dims, name="delp")
a = DataArray(rng.random((nb, n, n), dtype=np.float32) + 0.5, dims=["tile", "yaxis_2", "xaxis_1"], name="area")
for nt in (None, 8, 16, 32):
    cs.pressure_level_coarsen(ds, d, a, 300.0, f, mode=core.FAST)
    if nt is not None:
        for r in core._STAGING.values():
            r.nthreads = nt
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(3):
        cs.pressure_level_coarsen(ds, d, a, 300.0, f, mode=core.FAST)
    torch.cuda.synchronize()
    print(f"drop-in call, pool threads {nt or 'default'}: {(time.perf_counter() - t0) / 3 * 1e3:.2f} ms")

import cProfile, pstats
pr = cProfile.Profile()
pr.enable()
for _ in range(3):
    cs.pressure_level_coarsen(ds, d, a, 300.0, f, mode=core.FAST)
torch.cuda.synchronize()
pr.disable()
pstats.Stats(pr).sort_stats("tottime").print_stats(18)
