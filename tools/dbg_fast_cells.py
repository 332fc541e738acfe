"""debug helper (not product): worst coarse cells of test_pressure_coarsen_fast_mode_tolerance[64-2], with / without NaNs"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from fv3net_b200 import core
import test_gpu_parity as G
n, nfields = int(sys.argv[1]), 2
rng = np.random.default_rng(99 + n + nfields)
nb, km, f = 6, 79, 8
delp = G._synthetic_columns(rng, nb, km, n, np.float32)
w = rng.uniform(3.6e8, 8.3e8, size=(nb, n, n)).astype(np.float32)
clean = [rng.uniform(180, 320, size=(nb, km, n, n)).astype(np.float32), (rng.normal(size=(nb, km, n, n)) * 10.0).astype(np.float32)]
withnan = [clean[0], G.sprinkle_nans(rng, clean[1], 0.002)]
import itertools
for (name, fields), mm in itertools.product((("clean", clean), ("nan", withnan)), ("phalf", "none")):
    name = name + "/" + mm
    ex, _ = core.pressure_coarsen(fields, delp, w, f, mode=core.EXACT, mask_mode=mm)
    fa, _ = core.pressure_coarsen(fields, delp, w, f, mode=core.FAST, mask_mode=mm)
    for i, (x, y) in enumerate(zip(ex, fa)):
        sc = np.nanmax(np.abs(fields[i]))
        err = np.abs(x.astype(np.float64) - y) / sc
        err[np.isnan(err)] = 0
        bad = np.argwhere(err > 1e-5)
        print(name, "field", i, "max err", err.max(), "nbad", len(bad), "nan mismatch", int((np.isnan(x) != np.isnan(y)).sum()))
        for t in bad[:8]:
            tb, k, cy, cx = (int(v) for v in t)
            blk = fields[1][tb, :, cy * 8:cy * 8 + 8, cx * 8:cx * 8 + 8]
            nanlev = sorted(set(np.argwhere(np.isnan(blk))[:, 0].tolist()))
            print("    ", (tb, k, cy, cx), "exact", x[tb, k, cy, cx], "fast", y[tb, k, cy, cx], "NaN levels in cell:", nanlev)
