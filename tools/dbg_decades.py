"""debug helper (not product): worst coarse cells of the tracer-decades parity test, FAST vs EXACT on the GPU"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from fv3net_b200 import core
import test_gpu_parity as G
rng = np.random.default_rng(2024)
nb, km, n, f = 3, 79, 64, 8
delp = G._synthetic_columns(rng, nb, km, n, np.float32)
w = rng.uniform(3.6e8, 8.3e8, size=(nb, n, n)).astype(np.float32)
z = np.linspace(0, 1, km)[None, :, None, None]
noise = lambda: rng.uniform(0.5, 1.5, size=(nb, km, n, n))
T = (210 + 80 * z ** 2 + rng.normal(0, 0.5, size=(nb, km, n, n))).astype(np.float32)
qv = (2e-2 * np.exp(-11.5 * (1 - z)) * noise()).astype(np.float32)
o3 = (1e-5 * np.exp(-((z - 0.15) / 0.08) ** 2) * noise() + 1e-9).astype(np.float32)
ql = (1e-3 * np.clip(np.sin(9 * z) - 0.7, 0, None) * noise()).astype(np.float32)
fields = [T, qv, o3, ql]
for mm in ("phalf", "none"):
    ex, _ = core.pressure_coarsen(fields, delp, w, f, mode=core.EXACT, mask_mode=mm)
    fa, _ = core.pressure_coarsen(fields, delp, w, f, mode=core.FAST, mask_mode=mm)
    sc, _ = core.pressure_coarsen([np.abs(x) for x in fields], delp, w, f, mode=core.EXACT, mask_mode=mm)
    for name, a, b, s in zip("T qv o3 ql".split(), ex, fa, sc):
        rel = np.abs(a.astype(np.float64) - b) / (s.astype(np.float64) + 1e-300)
        rel[np.isnan(rel)] = 0
        order = np.argsort(rel.ravel())[::-1][:6]
        print(mm, name, "worst", rel.max(), "n>1e-6", int((rel > 1e-6).sum()))
        for o in order:
            i = np.unravel_index(o, rel.shape)
            print("   ", tuple(int(v) for v in i), "rel %.3e exact %.9e fast %.9e scale %.9e" % (rel[i], a[i], b[i], s[i]))
