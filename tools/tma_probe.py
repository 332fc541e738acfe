"""Dev probe for the TMA-fed fused kernel: FAST (TMA kernel) vs EXACT on small cases, then a timing loop
at the bench shape.  Run under `timeout` (a wrong barrier protocol hangs).

    python tools/tma_probe.py [--time] [--n 48] [--nf 4]
"""
import argparse
import os
import sys
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def columns(rng, nb, km, n, rough=False):
    ps = rng.uniform(5.0e4, 1.04e5, size=(nb, 1, n, n)) if rough else (
        9.0e4 + 8e3 * np.sin(np.linspace(0, 6, n))[None, None, :, None] * np.cos(np.linspace(0, 5, n))[None, None, None, :]
        + rng.normal(0, 800.0, size=(nb, 1, n, n)))
    b = np.diff(np.linspace(0, 1, km + 1) ** 1.6)[None, :, None, None]
    return ((ps - 300.0) * b * (1 + rng.uniform(-0.02, 0.02, size=(nb, km, n, n)))).astype(np.float32)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--time", action="store_true")
    ap.add_argument("--n", type=int, default=48)
    ap.add_argument("--nf", type=int, default=4)
    ap.add_argument("--km", type=int, default=79)
    ap.add_argument("--rough", action="store_true")
    a = ap.parse_args()
    import torch

    from fv3net_b200 import core

    rng = np.random.default_rng(1)
    nb, km, n, f = 2, a.km, a.n, 8
    delp = columns(rng, nb, km, n, a.rough)
    w = rng.uniform(3.6e8, 8.3e8, size=(nb, n, n)).astype(np.float32)
    fields = [rng.uniform(180, 320, size=(nb, km, n, n)).astype(np.float32)]
    fields += [(rng.normal(size=(nb, km, n, n)) * 10).astype(np.float32) for _ in range(a.nf - 1)]
    dev = [torch.as_tensor(x).cuda() for x in fields]
    dd, dw = torch.as_tensor(delp).cuda(), torch.as_tensor(w).cuda()
    t0 = time.time()
    ex, dce = core.pressure_coarsen(dev, dd, dw, f, mode=core.EXACT)
    torch.cuda.synchronize()
    print("exact done", time.time() - t0, flush=True)
    fa, dcf = core.pressure_coarsen(dev, dd, dw, f, mode=core.FAST)
    torch.cuda.synchronize()
    print("fast done", time.time() - t0, flush=True)
    print("coarse delp bit-identical:", bool(torch.equal(dce, dcf)))
    for i, (x, y, src) in enumerate(zip(ex, fa, fields)):
        x, y = x.cpu().numpy().astype(np.float64), y.cpu().numpy().astype(np.float64)
        nanpat = np.array_equal(np.isnan(x), np.isnan(y))
        err = np.nanmax(np.abs(x - y)) / np.nanmax(np.abs(src))
        print(f"field {i}: nan pattern {nanpat} max|diff|/scale {err:.3e}", flush=True)
    if a.time:
        nb, n = 6, 384
        delp = torch.as_tensor(columns(rng, nb, km, n)).cuda()
        w = torch.as_tensor(rng.uniform(3.6e8, 8.3e8, size=(nb, n, n)).astype(np.float32)).cuda()
        fs = [torch.randn(nb, km, n, n, device="cuda") * 10 + 250 for _ in range(4)]
        outs = None
        for it in range(3):
            outs = core.pressure_coarsen(fs, delp, w, f, mode=core.FAST)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for it in range(10):
            outs = core.pressure_coarsen(fs, delp, w, f, mode=core.FAST)
        e1.record()
        torch.cuda.synchronize()
        print("C384 nf=4 FAST ms/step (through core.pressure_coarsen):", e0.elapsed_time(e1) / 10, flush=True)


if __name__ == "__main__":
    main()
