"""Whole-timestep throughput of coarsen_restarts_on_pressure at C384 -> C48 with the restart
categories already resident in HBM ('Whole C384 fv_core on pressure with A-grid
winds' + fv_tracer + fv_srf_wnd).  Algorithmic bytes as SURVEY.md counts them: every A-grid field,
delp and the weights once, D-grid winds at the needed lines only, coarse outputs once.

    python tools/restart_bw.py [--n 384] [--reps 10]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fv3net_b200 import core, restarts  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--n", type=int, default=384)
ap.add_argument("--reps", type=int, default=10)
ap.add_argument("--modes", default="fast,exact")
ap.add_argument("--warm", type=int, default=3)
ap.add_argument("--rough", action="store_true",
                help="surface pressure U(5e4, 1.04e5) independently per column instead of bench.py's smooth field + "
                     "1.5 kPa noise: neighbouring columns then cross very different numbers of coarse levels per "
                     "source layer (worst case for the layer-synchronous kernel)")
args = ap.parse_args()
n, km, f, nt = args.n, 79, 8, 6
dev = core.device()
g = torch.Generator(device=dev).manual_seed(0)
U = lambda lo, hi, *s: torch.empty(s, device=dev).uniform_(lo, hi, generator=g)  # noqa: E731
N = lambda sd, *s: torch.empty(s, device=dev).normal_(0.0, sd, generator=g)  # noqa: E731

if args.rough:
    ps = U(5.0e4, 1.04e5, nt, 1, n, n)
else:  # bench.py's synth_device recipe
    yy = torch.linspace(0, 6, n, device=dev)[:, None]
    xx = torch.linspace(0, 6, n, device=dev)[None, :]
    ps = (8.6e4 + 1.4e4 * torch.sin(yy * 1.3) * torch.cos(xx * 1.1) + N(1500, nt, n, n)).clamp(5.0e4, 1.04e5)[:, None]
b = torch.diff(torch.linspace(0, 1, km + 1, device=dev) ** 1.6).view(1, km, 1, 1)
delp = ((ps - 300.0) * b * U(0.98, 1.02, nt, km, n, n)).contiguous()
fv_core = {"phis": U(0, 5e4, nt, n, n), "delp": delp, "DZ": -U(20, 800, nt, km, n, n), "W": N(10, nt, km, n, n),
           "T": U(180, 320, nt, km, n, n), "ua": N(10, nt, km, n, n), "va": N(10, nt, km, n, n),
           "u": N(10, nt, km, n + 1, n), "v": N(10, nt, km, n, n + 1)}
tracers = {k: N(1e-3, nt, km, n, n).abs_() for k in ("sphum", "liq_wat", "rainwat", "ice_wat", "snowwat", "graupel",
                                                     "o3mr", "sgs_tke", "cld_amt")}
srf = {"u_srf": N(5, nt, n, n), "v_srf": N(5, nt, n, n)}
grid = {"area": U(3.6e8, 8.3e8, nt, n, n), "dx": U(2.0e4, 2.88e4, nt, n + 1, n), "dy": U(2.0e4, 2.88e4, nt, n, n + 1)}
rs = {"fv_core.res": fv_core, "fv_tracer.res": tracers, "fv_srf_wnd.res": srf}

E, c = 4, n // f
fld, cfl = nt * km * n * n * E, nt * km * c * c * E
lines = nt * km * (c + 1) * n * E
alg = {"fv_core A-grid (delp, DZ, W, T, ua, va, phis, area)": 6 * fld + 2 * nt * n * n * E + 6 * cfl,
       "fv_core D-grid u, v (needed lines: wind + 2 delp rows + spacing)": 2 * (3 * lines + nt * (c + 1) * n * E + nt * km * (c + 1) * c * E),
       "fv_tracer (9 fields; delp, area shared with fv_core but re-read)": 10 * fld + nt * n * n * E + 9 * cfl,
       "fv_srf_wnd": 3 * nt * n * n * E}
total = sum(alg.values())


def run(mode):
    return restarts.coarsen_restarts_on_pressure(f, grid, 300.0, rs, coarsen_agrid_winds=True, mode=mode)


res = {}
for name, mode in (("fast", core.FAST), ("exact", core.EXACT)):
    if name not in args.modes.split(","):
        continue
    for _ in range(args.warm):
        run(mode)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(args.reps + 1)]
    ev[0].record()
    for i in range(args.reps):
        run(mode)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(args.reps))
    ms = ts[len(ts) // 2]
    res[name] = {"ms_per_timestep": ms, "GB_per_s": total / ms / 1e6,
                 "fine_points_per_s": (16 * nt * km * n * n + 2 * nt * km * (n + 1) * n) / ms * 1e3}
print(json.dumps({"workload": f"coarsen_restarts_on_pressure C{n}->C{c}, fv_core (A-grid winds) + 9 tracers + srf_wnd, "
                  "device resident" + (", rough surface pressure" if args.rough else ""), "algorithmic_bytes": total, "bytes_by_part": alg, **res}))
