"""Host-side check of the product's streaming PPM column machine (ppm_stream.cuh, the
per-thread code of the fused pressure-coarsening kernel) against the oracle.  The header
is __host__ __device__; this compiles it with g++ (no FMA) and requires BIT-EXACT
agreement with oracle/mappm_ref.c for kord <= 6 and every iv."""
import ctypes
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))


@pytest.fixture(scope="module")
def harness():
    src = os.path.join(HERE, "harness", "ppm_stream_host.cpp")
    so = os.path.join(HERE, "harness", "libppm_stream_host.so")
    hdr = os.path.join(HERE, "..", "fv3net_b200", "csrc", "ppm_stream.cuh")
    if (not os.path.exists(so)) or max(os.path.getmtime(src), os.path.getmtime(hdr)) > os.path.getmtime(so):
        subprocess.check_call(["/usr/bin/g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                               "-x", "c++", src, "-o", so])
    lib = ctypes.CDLL(so)
    lib.ppm_stream_host.restype = ctypes.c_int
    return lib


def run_stream(lib, pe1, qs, pe2, iv, kord):
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
    rc = lib.ppm_stream_host(len(qs), ncol, km, pe1.ctypes.data_as(fp), qin, kn, pe2.ctypes.data_as(fp), qout,
                             int(iv), int(kord))
    assert rc == 0
    return outs


def columns(rng, ncol, km, kn, kind):
    top = np.full((ncol, 1), 300.0)
    if kind == "walk":  # the regression-test style: uniform thicknesses
        pe1 = np.cumsum(np.concatenate([top, rng.uniform(3, 5, (ncol, km))], 1), 1)
        pe2 = np.cumsum(np.concatenate([top, rng.uniform(3, 5, (ncol, kn))], 1), 1)
    elif kind == "sigma":  # realistic: same sigma levels, different surface pressure
        b = np.sort(rng.uniform(0, 1, km - 1))
        b = np.concatenate([[0.0], b, [1.0]])
        ps1 = rng.uniform(5e4, 1.04e5, (ncol, 1))
        ps2 = ps1 * rng.uniform(0.8, 1.2, (ncol, 1))
        pe1 = 300.0 + (ps1 - 300.0) * b[None]
        bb = np.linspace(0, 1, kn + 1) ** 1.5
        pe2 = 300.0 + (ps2 - 300.0) * bb[None]
    elif kind == "wild":  # target grid extends above and far below the source column
        pe1 = np.cumsum(np.concatenate([top, rng.uniform(1, 50, (ncol, km))], 1), 1)
        pe2 = np.cumsum(np.concatenate([top * rng.uniform(0.2, 2.0, (ncol, 1)), rng.uniform(1, 120, (ncol, kn))], 1), 1)
    return pe1, pe2


@pytest.mark.parametrize("kind", ["walk", "sigma", "wild"])
@pytest.mark.parametrize("km,kn", [(4, 4), (5, 7), (7, 7), (79, 79), (30, 12)])
def test_stream_matches_oracle_bitexact(harness, oracle, kind, km, kn):
    rng = np.random.default_rng(km * 1000 + kn)
    ncol = 200
    pe1, pe2 = columns(rng, ncol, km, kn, kind)
    q = rng.uniform(-1000, 1000, (ncol, km))
    q[::7] = np.abs(q[::7]) * 1e-3  # positive tracer-like columns
    q[3::11, km // 2:] = 0.0          # flat stretches (dm == 0 branch)
    for kord in (1, 3, 4, 5, 6):
        for iv in (-2, -1, 0, 1, 2):
            exp = oracle.mappm(pe1, q, pe2, iv=iv, kord=kord)
            got = run_stream(harness, pe1, [q], pe2, iv, kord)[0]
            assert np.array_equal(got, exp, equal_nan=True), (kord, iv)


def test_stream_multifield_and_known_answers(harness, oracle, golden):
    rng = np.random.default_rng(11)
    pe1, pe2 = columns(rng, 128, 79, 79, "sigma")
    qs = [rng.normal(size=(128, 79)) * s for s in (1.0, 300.0, 1e-3, 20.0)]
    got = run_stream(harness, pe1, qs, pe2, 1, 1)
    for g, q in zip(got, qs):
        assert np.array_equal(g, oracle.mappm(pe1, q, pe2))
    for case in ("interior", "oob", "nan"):  # external/vcm/tests/test_mappm.py:5-44
        r = run_stream(harness, golden[f"mappm_kat|{case}|p_in"], [golden[f"mappm_kat|{case}|f_in"]],
                       golden[f"mappm_kat|{case}|p_out"], 1, 1)[0]
        np.testing.assert_almost_equal(r, golden[f"mappm_kat|{case}|expected"], decimal=5)


def test_stream_poisoned_column_defined_behaviour(harness, oracle):
    """Non-monotone source pressures: reference UB; oracle and product agree on NaN."""
    pe1 = np.array([[1.0, 2.0, 3.0, 2.5, 2.6, 2.7]])
    q = np.array([[1.0, 2.0, 3.0, 4.0, 5.0]])
    pe2 = np.array([[0.5, 2.95, 2.97, 2.99, 10.0]])
    exp = oracle.mappm(pe1, q, pe2)
    got = run_stream(harness, pe1, [q], pe2, 1, 1)[0]
    assert np.array_equal(got, exp, equal_nan=True)
