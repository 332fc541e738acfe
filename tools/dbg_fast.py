"""debug helper (not product): FAST vs EXACT on the failing test configurations"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from fv3net_b200 import core

def _synthetic_columns(rng, nb, km, n, dt):
    yy, xx = np.meshgrid(np.linspace(0, 3, n), np.linspace(0, 3, n), indexing="ij")
    ps = 9.0e4 + 1.2e4 * np.sin(yy * 2.1)[None] * np.cos(xx * 1.7)[None] + rng.normal(0, 1500, size=(nb, n, n))
    b = np.diff(np.linspace(0, 1, km + 1) ** 1.6)
    delp = (ps[:, None] - 300.0) * b[None, :, None, None] * (1 + rng.uniform(-0.02, 0.02, size=(nb, km, n, n)))
    return delp.astype(dt)

def sprinkle_nans(rng, a, frac):
    a = a.copy(); a[rng.uniform(size=a.shape) < frac] = np.nan; return a

small = "--small" in sys.argv
cases = [(64, 2)] if small else [(64, 2), (64, 5), (64, 4)]
if "--case" in sys.argv:
    i = sys.argv.index("--case"); cases = [(int(sys.argv[i + 1]), int(sys.argv[i + 2]))]
for (n, nfields) in cases:
    rng = np.random.default_rng(99 + n + nfields)
    nb, km, f = (1 if small else 6), 79, 8
    delp = _synthetic_columns(rng, 6, km, n, np.float32)
    w = rng.uniform(3.6e8, 8.3e8, size=(6, n, n)).astype(np.float32)
    scales = [None, 10.0, 1e-3, 10.0, 3.0, 250.0, 1.0]
    fields = []
    for s in scales[:nfields]:
        fields.append(rng.uniform(180, 320, size=(6, km, n, n)).astype(np.float32) if s is None else (rng.normal(size=(6, km, n, n)) * s).astype(np.float32))
    if nfields >= 2:
        fields[1] = sprinkle_nans(rng, fields[1], 0.002)
    if small:
        delp, w, fields = delp[4:5], w[4:5], [x[4:5] for x in fields]
    for mm in ("phalf", "none"):
        ex, _ = core.pressure_coarsen(fields, delp, w, f, mode=core.EXACT, mask_mode=mm)
        for rep in range(2):
            fa, _ = core.pressure_coarsen(fields, delp, w, f, mode=core.FAST, mask_mode=mm)
            for i, (a, b, x) in enumerate(zip(ex, fa, fields)):
                err = np.abs(a.astype(np.float64) - b) / np.nanmax(np.abs(x))
                bad = np.argwhere(err > 1e-6)
                print(n, nfields, mm, "rep", rep, "field", i, "max", np.nanmax(err), "nbad", len(bad), bad[:6].tolist())
