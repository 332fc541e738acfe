"""Pins the CPU oracle (oracle/) against every golden vector and known answer the
reference's own tests hold for the coarse-graining hot path.  CPU only."""
import numpy as np
import pytest

PRESSURE_TAGS = [
    ("pressure-level-with-agrid-winds", True, False),
    ("pressure-level-without-agrid-winds", False, False),
    ("pressure-level-extrapolate-with-agrid-winds", False, True),
]
# The reference's own tolerance for these sets is xr.testing.assert_allclose (rtol 1e-5,
# external/vcm/tests/test_coarsen_restarts.py:119-127: "mappm does not produce
# bit-for-bit reproducible results on different platforms").  The oracle does better:
RTOL_PRESSURE = 2e-6
HYDROSTATIC = {"phis", "DZ"}  # rewritten by _impose_hydrostatic_balance ("next" row N2)


def _check(golden, tag, cat, res, rtol, skip=()):
    n = 0
    for name, val in res.items():
        if name in skip:
            continue
        exp = golden[f"{tag}|{cat}|{name}"]
        assert exp.shape == val.shape, (name, exp.shape, val.shape)
        assert np.array_equal(np.isnan(exp), np.isnan(val)), name
        if rtol == 0:
            assert np.array_equal(np.nan_to_num(exp), np.nan_to_num(val.astype(np.float64))), name
        else:
            np.testing.assert_allclose(val, exp, rtol=rtol, atol=0, err_msg=name)
        n += 1
    assert n > 0


# ---- external/vcm/tests/test_mappm.py:5-44 ------------------------------------------
@pytest.mark.parametrize("case,decimal", [("interior", 5), ("oob", 5), ("nan", 7)])
def test_mappm_known_answers(oracle, golden, case, decimal):
    res = oracle.mappm(golden[f"mappm_kat|{case}|p_in"], golden[f"mappm_kat|{case}|f_in"],
                       golden[f"mappm_kat|{case}|p_out"], 1, 1.0, 1.0, 1.0, 0.0)
    assert res.dtype == np.float32
    np.testing.assert_almost_equal(res, golden[f"mappm_kat|{case}|expected"], decimal=decimal)


def test_mappm_threads_identical(oracle):
    rng = np.random.default_rng(3)
    ncol, km = 257, 31
    pe1 = np.cumsum(np.concatenate([np.full((ncol, 1), 300.0), rng.uniform(3, 5, (ncol, km))], 1), 1)
    pe2 = np.cumsum(np.concatenate([np.full((ncol, 1), 300.0), rng.uniform(3, 5, (ncol, km))], 1), 1)
    q = rng.normal(size=(ncol, km))
    for kord in (1, 4, 7, 9, 10, 11, 16):
        for iv in (-2, -1, 0, 1, 2):
            a = oracle.mappm(pe1, q, pe2, iv=iv, kord=kord, nthreads=1)
            b = oracle.mappm(pe1, q, pe2, iv=iv, kord=kord, nthreads=4)
            assert np.array_equal(a, b, equal_nan=True)
            assert np.isfinite(a).all(), (kord, iv)


def test_mappm_identity_remap_is_conservative(oracle):
    """Remapping onto the same edges returns the input (each layer wholly inside itself)."""
    rng = np.random.default_rng(5)
    ncol, km = 64, 20
    pe = np.cumsum(np.concatenate([np.full((ncol, 1), 300.0), rng.uniform(3, 5, (ncol, km))], 1), 1)
    q = rng.uniform(200, 300, (ncol, km))
    out = oracle.mappm(pe, q, pe)
    np.testing.assert_allclose(out, q, rtol=2e-6)
    # mass conservation when remapping to a different grid with the same top/bottom
    # (layer 1 is kept: mappm.f90:62-64 assigns q1(1) to any output layer whose TOP edge is
    # at or above the old top, however deep it reaches -- not conservative by design)
    pe2 = pe.copy()
    pe2[:, 2:-1] += rng.uniform(-1, 1, (ncol, km - 2))
    out2 = oracle.mappm(pe, q, pe2).astype(np.float64)
    m1 = (q * np.diff(pe, axis=1)).sum(1)
    m2 = (out2 * np.diff(pe2.astype(np.float32).astype(np.float64), axis=1)).sum(1)
    np.testing.assert_allclose(m2, m1, rtol=1e-5)


# ---- numpy's summation order ---------------------------------------
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("fy,fx", [(2, 2), (4, 4), (8, 8), (16, 16), (64, 64), (1, 8), (8, 1), (3, 5)])
def test_blocksum_order_matches_numpy(oracle, dtype, fy, fx):
    rng = np.random.default_rng(fy * 100 + fx)
    a = (rng.normal(size=(3, 5, fy * 6, fx * 4)) * 10.0 ** rng.integers(-3, 4, size=(3, 5, 1, 1))).astype(dtype)
    assert np.array_equal(oracle.blocksum_ordered(a, fy, fx), oracle.coarsen_sum(a, fy, fx))


# ---- external/vcm/tests/test_cubedsphere.py inline known answers ---------------------
def test_weighted_block_average_kat(oracle):  # :206-236
    out = oracle.weighted_block_average(np.array([[2.0, 6.0], [6.0, 2.0]]),
                                        np.array([[6.0, 2.0], [2.0, 6.0]]), 2)
    assert out.dtype == np.float64 and np.array_equal(out, [[3.0]])


def test_edge_weighted_block_average_kat(oracle):  # :239-258 (dims (x_dim, y_dim) = axes (0,1))
    # edge='x' coarsens along x_dim (axis 0 there): our [y, x] convention is the transpose
    d, s = np.array([[2, 6, 2], [6, 2, 6]]), np.array([[6, 2, 6], [2, 6, 2]])
    out = oracle.edge_weighted_block_average(d.T, s.T, 2, edge="x").T
    assert np.array_equal(out, [[3.0, 3.0]])
    d, s = np.array([[2, 6], [6, 2], [2, 6]]), np.array([[6, 2], [2, 6], [6, 2]])
    out = oracle.edge_weighted_block_average(d.T, s.T, 2, edge="y").T
    assert np.array_equal(out, [[3.0], [3.0]])


def test_block_edge_sum_and_min_kat(oracle):  # :386-419
    d = np.array([[2, 6, 2], [6, 2, 6]])
    assert np.array_equal(oracle.block_edge_coarsen(d.T, 2, "x", "sum").T, [[8, 8]])
    assert np.array_equal(oracle.block_edge_coarsen(d.T, 2, "x", "min").T, [[2, 2]])
    d = np.array([[2, 6], [6, 2], [2, 6]])
    assert np.array_equal(oracle.block_edge_coarsen(d.T, 2, "y", "sum").T, [[8], [8]])


def test_block_reduce_regtest_checksum(oracle):  # :275-290 + _regtest_outputs
    joblib = pytest.importorskip("joblib")
    a = np.arange(32).reshape(4, 4, 2).astype(np.float32)
    for fn in (np.mean, np.median):
        res = oracle.block_reduce(a, (2, 2, 1), fn, cval=np.nan)
        assert res.dtype == np.float32
        assert joblib.hash(res) == "bc0c79ed90bacf5bf82212ab8c10303e"


def test_block_upsample_kat(oracle):  # :771-816
    assert np.array_equal(oracle.block_upsample(np.array([[1, 2], [3, 4]]), 2),
                          [[1, 1, 2, 2], [1, 1, 2, 2], [3, 3, 4, 4], [3, 3, 4, 4]])
    assert np.array_equal(oracle.block_upsample(np.array([[1, 2, 3], [4, 5, 6]]), 2),
                          [[1, 1, 2, 2, 3], [1, 1, 2, 2, 3], [4, 4, 5, 5, 6], [4, 4, 5, 5, 6]])


def test_coarsen_coords(oracle):  # :148-203, 422-614
    assert np.array_equal(oracle.coarsen_coords(np.array([5.0, 6.0, 7.0, 8.0]), 2), [3.0, 4.0])
    assert oracle.coarsen_coords(np.arange(1, 5, dtype=np.float64), 2).dtype == np.float32


def test_mode_reduce(oracle):  # :617-768 (ties -> smallest; nan omit)
    a = np.array([[1.0, 1.0, 2.0, 2.0], [3.0, np.nan, 3.0, np.nan]])
    out = oracle.mode_reduce(a.reshape(2, 2, 2), axis=(1, 2), nan_policy="omit")
    assert np.array_equal(out, [1.0, 3.0])


@pytest.mark.parametrize("extrapolate", [False, True])
def test_mask_weights_truth_table(oracle, extrapolate):  # tests/test_regridz.py:113-147
    w = np.arange(4, dtype=np.float32).reshape(2, 2)  # dims (x, y); z is LAST there
    pc = np.empty((2, 2, 3), np.float32)
    pc[..., 0], pc[..., 1], pc[..., 2] = 1.0, 2.0, 3.0
    pf = pc.copy()
    pf[0, 0, 2], pf[0, 1, 2], pf[1, :, 2] = 2.5, 2.6, 3.5
    pfull = np.diff(pc, axis=-1)
    exp = np.broadcast_to(w[..., None], (2, 2, 2)).copy()
    if not extrapolate:
        exp[0, :, 1] = 0.0
    else:
        exp[0, 0, 1] = 0.0
    # move z to axis -3 for the oracle's [z, y, x] convention
    got = oracle.mask_weights(w[None], np.moveaxis(pfull, -1, 0), np.moveaxis(pc, -1, 0),
                              np.moveaxis(pf, -1, 0), extrapolate=extrapolate)
    assert np.array_equal(np.moveaxis(got, 0, -1), exp)


def test_regrid_vertical_errors(oracle):  # tests/test_regridz.py:74-99
    z = lambda *s: np.zeros(s, np.float32)
    with pytest.raises(ValueError, match="must be same size"):
        oracle.regrid_vertical(z(6, 4, 1), z(5, 3, 1), z(3, 4, 1))
    with pytest.raises(ValueError, match="one shorter than p_in"):
        oracle.regrid_vertical(z(6, 4, 1), z(3, 4, 1), z(3, 4, 1))


# ---- external/vcm/tests/test_coarsen_restarts.py:108-127 regression sets --------------
def test_model_level_golden_bit_exact(oracle, golden):
    core, tracer, srf, grid = oracle.regression_inputs()
    tag = "mass-weighted-model-level-with-agrid-winds"
    _check(golden, tag, "fv_core.res",
           oracle.coarse_grain_fv_core(core, core["delp"], grid["area"], grid["dx"], grid["dy"], 2, True, True), 0)
    _check(golden, tag, "fv_tracer.res", oracle.coarse_grain_fv_tracer(tracer, core["delp"], grid["area"], 2, True), 0)
    _check(golden, tag, "fv_srf_wnd.res", oracle.coarse_grain_fv_srf_wnd(srf, grid["area"], 2), 0)
    tag = "area-weighted-model-level-without-agrid-winds"
    _check(golden, tag, "fv_core.res",
           oracle.coarse_grain_fv_core(core, core["delp"], grid["area"], grid["dx"], grid["dy"], 2, False, False), 0)
    _check(golden, tag, "fv_tracer.res", oracle.coarse_grain_fv_tracer(tracer, core["delp"], grid["area"], 2, False), 0)


@pytest.mark.parametrize("tag,agrid,extrapolate", PRESSURE_TAGS)
def test_pressure_level_golden(oracle, golden, tag, agrid, extrapolate):
    core, tracer, srf, grid = oracle.regression_inputs()
    res = oracle.coarse_grain_fv_core_on_pressure(core, core["delp"], grid["area"], grid["dx"], grid["dy"],
                                                  300.0, 2, agrid, extrapolate)
    rtol = 0 if extrapolate else RTOL_PRESSURE  # the extrapolate set reproduces bit-exactly
    _check(golden, tag, "fv_core.res", res, rtol, skip=HYDROSTATIC)
    res = oracle.coarse_grain_fv_tracer_on_pressure(tracer, core["delp"], grid["area"], 300.0, 2, extrapolate)
    _check(golden, tag, "fv_tracer.res", res, rtol)


# ---- N2: hydrostatic balance and the blended method (coarsen_restarts.py:559-779,990-1017) ----
@pytest.mark.parametrize("tag,agrid,extrapolate", PRESSURE_TAGS)
def test_hydrostatic_balance_golden(oracle, golden, tag, agrid, extrapolate):
    core, tracer, srf, grid = oracle.regression_inputs()
    c = oracle.coarse_grain_fv_core_on_pressure(core, core["delp"], grid["area"], grid["dx"], grid["dy"],
                                                300.0, 2, agrid, extrapolate)
    t = oracle.coarse_grain_fv_tracer_on_pressure(tracer, core["delp"], grid["area"], 300.0, 2, extrapolate)
    res = oracle.impose_hydrostatic_balance(c, t, 300.0)
    _check(golden, tag, "fv_core.res", {k: res[k] for k in HYDROSTATIC}, 5e-6)


BLENDED_TAGS = [("blended-area-weighted-without-agrid-winds", False, False),
                ("blended-mass-weighted-with-agrid-winds", True, True)]


@pytest.mark.parametrize("tag,agrid,mass", BLENDED_TAGS)
def test_blended_golden(oracle, golden, tag, agrid, mass):
    core, tracer, srf, grid = oracle.regression_inputs()
    c = oracle.coarse_grain_fv_core_via_blended_method(core, core["delp"], grid["area"], grid["dx"], grid["dy"],
                                                       300.0, 2, agrid, mass)
    t = oracle.coarse_grain_fv_tracer_via_blended_method(tracer, core["delp"], grid["area"], 300.0, 2, mass)
    c = oracle.impose_hydrostatic_balance(c, t, 300.0)
    _check(golden, tag, "fv_core.res", c, 5e-6)
    _check(golden, tag, "fv_tracer.res", t, 5e-6)


# ---- N1: sfc_data 'complex' method (coarsen_restarts.py:1111-1470) -----------------------------
def test_sfc_data_golden_bit_exact(oracle, golden):
    _, _, _, grid = oracle.regression_inputs()
    res = oracle.coarse_grain_sfc_data_complex(oracle.sfc_regression_inputs(), grid["area"], 2)
    assert len(res) == 35
    for tag in ("pressure-level-with-agrid-winds", "mass-weighted-model-level-with-agrid-winds",
                "blended-area-weighted-without-agrid-winds"):
        _check(golden, tag, "sfc_data", res, 0)
