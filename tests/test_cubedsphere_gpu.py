"""GPU tests of the reference-facing operator API (fv3net_b200/cubedsphere.py) on labelled arrays.

The cases are the reference's own (external/vcm/tests/test_cubedsphere.py:206-258, 261-419,
459-614, 731-816; test_regridz.py:24-147), restated with fv3net_b200.labelled arrays because
xarray is not installed; larger seeded cases are compared with the CPU oracle bit for bit."""
import functools

import numpy as np
import pytest

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu

from fv3net_b200.labelled import DataArray, Dataset, assert_identical_including_dtype, merge  # noqa: E402


@pytest.fixture(scope="module")
def cs():
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from fv3net_b200 import cubedsphere

    return cubedsphere


# ---- test_cubedsphere.py:206-258 ---------------------------------------------------------
@pytest.mark.parametrize("object_type", ["DataArray", "Dataset"])
def test_weighted_block_average(cs, object_type):
    dims = ["x", "y"]
    attrs = {"units": "m"}
    data = DataArray(np.array([[2.0, 6.0], [6.0, 2.0]]), dims=dims, name="foo", attrs=attrs)
    expected = DataArray(np.array([[3.0]]), dims=dims, name="foo", attrs=attrs)
    if object_type == "Dataset":
        data = data.to_dataset().assign_attrs(test="a")
        expected = expected.to_dataset().assign_attrs(test="a")
    weights = DataArray(np.array([[6.0, 2.0], [2.0, 6.0]]), dims=dims)
    result = cs.weighted_block_average(data, weights, 2, x_dim="x", y_dim="y")
    assert_identical_including_dtype(result, expected)


@pytest.mark.parametrize(
    ("data", "spacing", "factor", "edge", "expected_data"),
    [
        ([[2, 6, 2], [6, 2, 6]], [[6, 2, 6], [2, 6, 2]], 2, "x", [[3.0, 3.0]]),
        ([[2, 6], [6, 2], [2, 6]], [[6, 2], [2, 6], [6, 2]], 2, "y", [[3.0], [3.0]]),
    ],
    ids=["edge='x'", "edge='y'"],
)
def test_edge_weighted_block_average(cs, data, spacing, factor, edge, expected_data):
    dims = ["x_dim", "y_dim"]
    attrs = {"units": "m"}
    da = DataArray(data, dims=dims, attrs=attrs)
    weights = DataArray(spacing, dims=dims)
    expected = DataArray(expected_data, dims=dims, attrs=attrs)
    result = cs.edge_weighted_block_average(da, weights, factor, x_dim="x_dim", y_dim="y_dim", edge=edge)
    assert_identical_including_dtype(result, expected)


# ---- test_cubedsphere.py:261-419 ---------------------------------------------------------
@pytest.fixture()
def input_dataarray():
    shape = (4, 4, 2)
    data = np.arange(np.prod(shape)).reshape(shape).astype(np.float32)
    return DataArray(data, dims=["x", "y", "z"], name="foo", attrs={"units": "m"})


@pytest.fixture()
def input_dataset(input_dataarray):
    bar = DataArray([1, 2, 3], dims=["t"], name="bar", attrs={"units": "m"})
    return merge([input_dataarray, bar]).assign_attrs(test="a")


def _blocks(a, f):
    """[x, y, z] -> [x/f, y/f, z, f*f] windows (numpy statement of skimage's view_as_blocks)"""
    nx, ny, nz = a.shape
    return a.reshape(nx // f, f, ny // f, f, nz).transpose(0, 2, 4, 1, 3).reshape(nx // f, ny // f, nz, f * f)


@pytest.mark.parametrize("reduction_function", [np.mean, np.median])
def test_xarray_block_reduce_dataarray(cs, reduction_function, input_dataarray):
    # reference regtest output (_regtest_outputs/test_cubedsphere.test_xarray_block_reduce_dataarray*):
    # float32 a(x, y, z) with 2 x 2 x 2 values; mean and median coincide for this input
    result = cs._xarray_block_reduce_dataarray(input_dataarray, {"x": 2, "y": 2}, reduction_function)
    expected = reduction_function(_blocks(input_dataarray.values, 2), axis=-1)
    assert result.dims == ("x", "y", "z") and result.dtype == np.float32
    assert np.array_equal(result.values, expected)
    assert result.attrs == {"units": "m"} and result.name == "foo" and not result.coords


@pytest.mark.parametrize(
    "coord_func", ["mean", np.mean, {"x": np.max, "y": "median"}, {"x": np.min}],
    ids=["single str", "single function", "dict", "dict with a dimension missing"],
)
def test_block_reduce_dataarray_coordinates(cs, input_dataarray, coord_func):
    for dim, size in input_dataarray.sizes.items():
        input_dataarray[dim] = np.arange(size, dtype=np.float64)
    result = cs._xarray_block_reduce_dataarray(input_dataarray, {"x": 2, "y": 2}, np.median, coord_func=coord_func)
    # xarray's coarsen(x=2, y=2, coord_func=coord_func): windows of the coordinate through the function
    windows = np.arange(4.0).reshape(2, 2)
    fx = coord_func.get("x", "mean") if isinstance(coord_func, dict) else coord_func
    fy = coord_func.get("y", "mean") if isinstance(coord_func, dict) else coord_func
    red = lambda f: getattr(np, f)(windows, axis=1) if isinstance(f, str) else f(windows, axis=1)  # noqa: E731
    assert np.array_equal(result["x"].values, red(fx)) and np.array_equal(result["y"].values, red(fy))
    assert np.array_equal(result["z"].values, [0.0, 1.0])
    assert np.array_equal(result.values, np.median(_blocks(input_dataarray.values, 2), axis=-1))


def test_horizontal_block_reduce_dataarray(cs, input_dataarray):
    expected = cs._xarray_block_reduce_dataarray(input_dataarray, {"x": 2, "y": 2, "z": 1}, np.median)
    result = cs.horizontal_block_reduce(input_dataarray, 2, np.median, "x", "y")
    assert_identical_including_dtype(result, expected)


def test_horizontal_block_reduce_dataset(cs, input_dataset):
    expected_foo = cs._xarray_block_reduce_dataarray(input_dataset.foo, {"x": 2, "y": 2, "z": 1}, np.median)
    expected = merge([expected_foo, input_dataset.bar]).assign_attrs(test="a")
    result = cs.horizontal_block_reduce(input_dataset, 2, np.median, "x", "y")
    assert_identical_including_dtype(result, expected)  # bar has no horizontal dimensions: unchanged


def test_block_median_and_block_coarsen(cs, input_dataarray):
    expected = cs._xarray_block_reduce_dataarray(input_dataarray, {"x": 2, "y": 2, "z": 1}, np.median)
    assert_identical_including_dtype(cs.block_median(input_dataarray, 2, "x", "y"), expected)
    assert_identical_including_dtype(cs.block_coarsen(input_dataarray, 2, "x", "y", method="median"), expected)
    result = cs.block_coarsen(input_dataarray, 2, "x", "y", "min")
    assert np.array_equal(result.values, _blocks(input_dataarray.values, 2).min(axis=-1))
    assert result.dtype == np.float32 and result.name == "foo"


@pytest.mark.parametrize(
    ("data", "factor", "edge", "method", "expected_data"),
    [
        ([[2, 6, 2], [6, 2, 6]], 2, "x", "sum", [[8, 8]]),
        ([[2, 6], [6, 2], [2, 6]], 2, "y", "sum", [[8], [8]]),
        ([[2, 6, 2], [6, 2, 6]], 2, "x", "min", [[2, 2]]),
        ([[2, 6], [6, 2], [2, 6]], 2, "y", "min", [[2], [2]]),
    ],
)
def test_block_edge_sum_and_coarsen(cs, data, factor, edge, method, expected_data):
    dims = ["x_dim", "y_dim"]
    attrs = {"units": "m"}
    da = DataArray(data, dims=dims, attrs=attrs)
    expected = DataArray(expected_data, dims=dims, attrs=attrs)
    if method == "sum":
        result = cs.block_edge_sum(da, factor, x_dim="x_dim", y_dim="y_dim", edge=edge)
    else:
        result = cs.block_edge_coarsen(da, factor, x_dim="x_dim", y_dim="y_dim", edge=edge, method=method)
    assert_identical_including_dtype(result, expected)


# ---- test_cubedsphere.py:422-614: subtile coordinates --------------------------------------
@pytest.mark.parametrize("dtype", [np.float32, np.float64])
@pytest.mark.parametrize("subtile_x", [0, 1, 2])
@pytest.mark.parametrize("subtile_y", [0, 1, 2])
def test_operators_with_subtile_coordinates(cs, dtype, subtile_x, subtile_y):
    n, f = 4, 2
    x = np.arange(1.0, n + 1) + subtile_x * n
    y = np.arange(1.0, n + 1) + subtile_y * n
    ex = (np.arange(1.0, n // f + 1) + subtile_x * n // f).astype(np.float32)
    ey = (np.arange(1.0, n // f + 1) + subtile_y * n // f).astype(np.float32)
    data = np.arange(n * n).reshape(n, n).astype(dtype)
    da = DataArray(data, dims=["x", "y"], coords=[x, y], name="foo")
    weights = DataArray(np.ones((n, n), dtype), dims=["x", "y"], coords=[x, y])
    for result in (cs.weighted_block_average(da, weights, f, x_dim="x", y_dim="y"),
                   cs.horizontal_block_reduce(da, f, np.mean, x_dim="x", y_dim="y"),
                   cs.block_median(da, f, x_dim="x", y_dim="y"),
                   cs.block_coarsen(da, f, x_dim="x", y_dim="y", method="mean")):
        assert result.dtype == dtype
        assert result["x"].dtype == np.float32 and np.array_equal(result["x"].values, ex)
        assert result["y"].dtype == np.float32 and np.array_equal(result["y"].values, ey)
    # staggered: x has n + 1 interfaces
    xs = np.arange(1.0, n + 2) + subtile_x * n
    exs = (np.arange(1.0, n // f + 2) + subtile_x * n // f).astype(np.float32)
    das = DataArray(np.arange((n + 1) * n).reshape(n + 1, n).astype(dtype), dims=["x", "y"], coords=[xs, y], name="foo")
    sp = DataArray(np.ones((n + 1, n), dtype), dims=["x", "y"], coords=[xs, y])
    for result in (cs.edge_weighted_block_average(das, sp, f, x_dim="x", y_dim="y", edge="y"),
                   cs.block_edge_sum(das, f, x_dim="x", y_dim="y", edge="y")):
        assert result.shape == (n // f + 1, n // f)
        assert np.array_equal(result["x"].values, exs) and result["x"].dtype == np.float32
        assert np.array_equal(result["y"].values, ey)


# ---- test_cubedsphere.py:731-768: mode -----------------------------------------------------
def test_block_mode(cs):
    data = np.array([[0, 1, 1, 1], [0, 0, 2, 1], [3, 3, 4, 4], [3, 5, 4, 6]], dtype=np.float64)
    da = DataArray(data, dims=["x", "y"], name="foo", attrs={"units": "m"})
    expected = DataArray(np.array([[0.0, 1.0], [3.0, 4.0]]), dims=["x", "y"], name="foo", attrs={"units": "m"})
    assert_identical_including_dtype(cs._block_mode(da, 2, x_dim="x", y_dim="y"), expected)
    assert_identical_including_dtype(cs.block_coarsen(da, 2, x_dim="x", y_dim="y", method="mode"), expected)
    data[0, 0] = np.nan
    data[1, 0] = np.nan
    dan = DataArray(data, dims=["x", "y"])
    out = cs.block_coarsen(dan, 2, x_dim="x", y_dim="y", method="mode", func_kwargs={"nan_policy": "omit"})
    assert np.array_equal(out.values, [[0.0, 1.0], [3.0, 4.0]])
    r = cs.horizontal_block_reduce(dan, 2, functools.partial(cs._mode_reduce, nan_policy="omit"), "x", "y")
    assert np.array_equal(r.values, out.values)


# ---- test_cubedsphere.py:771-816: upsample -------------------------------------------------
def test_block_upsample_dataset(cs):
    x = DataArray(np.array([[1.0, 2.0], [3.0, 4.0]]), dims=["x", "y"], coords=[[1, 2], [3, 4]], name="foo",
                  attrs={"units": "m"})
    y = DataArray(np.array([1.0, 2.0]), dims=["z"], name="bar")
    ds = merge([x, y])
    result = cs.block_upsample(ds, 2, ["x", "y"])
    assert np.array_equal(result["foo"].values, np.repeat(np.repeat(x.values, 2, 0), 2, 1))
    assert np.array_equal(result["bar"].values, y.values)  # no upsampled dimension: unchanged
    assert "x" not in result.coords and "y" not in result.coords  # coordinates are dropped


@pytest.mark.parametrize(
    ("data", "expected"),
    [
        ([[1, 2], [3, 4]], [[1, 1, 2, 2], [1, 1, 2, 2], [3, 3, 4, 4], [3, 3, 4, 4]]),
        ([[1, 2, 3], [4, 5, 6]], [[1, 1, 2, 2, 3], [1, 1, 2, 2, 3], [4, 4, 5, 5, 6], [4, 4, 5, 5, 6]]),
        ([[1, 2], [3, 4], [5, 6]], [[1, 1, 2, 2], [1, 1, 2, 2], [3, 3, 4, 4], [3, 3, 4, 4], [5, 5, 6, 6]]),
    ],
    ids=["unstaggered", "y staggered", "x staggered"],
)
def test_block_upsample_dataarray(cs, data, expected):
    da = DataArray(np.array(data, np.float32), dims=["x", "y"], name="foo")
    result = cs.block_upsample(da, 2, ["x", "y"])
    assert result.dtype == np.float32 and np.array_equal(result.values, np.array(expected, np.float32))
    ref = DataArray(np.zeros(np.array(expected).shape), dims=["x", "y"],
                    coords=[np.arange(len(expected)), np.arange(len(expected[0]))])
    like = cs.block_upsample_like(da, ref, x_dim="x", y_dim="y")
    assert np.array_equal(like.values, result.values)
    assert np.array_equal(like["x"].values, ref["x"].values) and np.array_equal(like["y"].values, ref["y"].values)


# ---- larger seeded cases against the oracle, with dimension-order handling -------------------
def test_weighted_block_average_dim_order_and_mass_weights(cs, oracle):
    rng = np.random.default_rng(3)
    T = rng.uniform(180, 320, size=(2, 5, 16, 24)).astype(np.float32)  # tile, z, y, x
    area = rng.uniform(1, 2, size=(2, 16, 24)).astype(np.float32)
    T[0, 1, 3, 5] = np.nan
    da = DataArray(T, dims=["tile", "zaxis_1", "yaxis_2", "xaxis_1"], name="T", attrs={"units": "K"})
    w = DataArray(area, dims=["tile", "yaxis_2", "xaxis_1"])
    ref = oracle.weighted_block_average(T, area[:, None], 4)
    out = cs.weighted_block_average(da, w, 4)
    assert out.dims == da.dims and np.array_equal(out.values, ref, equal_nan=True)
    # x before y, z last: the result keeps obj's order
    dat = da.transpose("tile", "xaxis_1", "yaxis_2", "zaxis_1")
    out_t = cs.weighted_block_average(dat, w, 4)
    assert out_t.dims == dat.dims and np.array_equal(out_t.values, ref.transpose(0, 3, 2, 1), equal_nan=True)
    # weights with a dimension obj lacks: appended after obj's dims (xarray broadcasting)
    w3 = DataArray(rng.uniform(1, 2, size=(3, 16, 24)).astype(np.float32), dims=["t", "yaxis_2", "xaxis_1"])
    o3 = cs.weighted_block_average(da.isel(tile=0), w3, 4)
    assert o3.dims == ("zaxis_1", "yaxis_2", "xaxis_1", "t") and o3.shape == (5, 4, 6, 3)
    for t in range(3):
        assert np.array_equal(o3.values[..., t], oracle.weighted_block_average(T[0], w3.values[t][None], 4), equal_nan=True)


def test_edge_weighted_block_average_dim_order(cs, oracle):
    rng = np.random.default_rng(4)
    u = rng.normal(size=(2, 3, 17, 16)).astype(np.float32)  # tile, z, yaxis_1 (n+1), xaxis_1
    dx = rng.uniform(1, 2, size=(2, 17, 16)).astype(np.float32)
    da = DataArray(u, dims=["tile", "zaxis_1", "yaxis_1", "xaxis_1"], name="u")
    sp = DataArray(dx, dims=["tile", "yaxis_1", "xaxis_1"])
    out = cs.edge_weighted_block_average(da, sp, 4, x_dim="xaxis_1", y_dim="yaxis_1", edge="x")
    # spacing * obj: spacing's dimensions first (coarsen.py:260)
    assert out.dims == ("tile", "yaxis_1", "xaxis_1", "zaxis_1")
    ref = oracle.edge_weighted_block_average(u, dx[:, None], 4, edge="x")
    assert np.array_equal(out.values, ref.transpose(0, 2, 3, 1))


def _column_case(rng, nb=2, km=12, n=8):
    ps = 9.0e4 + 6.0e3 * rng.normal(size=(nb, 1, n, n))
    b = np.diff(np.linspace(0, 1, km + 1) ** 1.6)
    delp = ((ps - 300.0) * b[None, :, None, None] * (1 + rng.uniform(-0.02, 0.02, size=(nb, km, n, n))))
    return delp


def test_regrid_vertical_dims(cs, oracle):
    rng = np.random.default_rng(5)
    delp = _column_case(rng).astype(np.float32)
    delp2 = _column_case(rng).astype(np.float32)
    T = rng.uniform(180, 320, size=delp.shape).astype(np.float32)
    p_in = oracle.pressure_at_interface(delp, 300.0)
    p_out = oracle.pressure_at_interface(delp2, 300.0)
    dims_c = ["tile", "zaxis_1", "y", "x"]
    dims_o = ["tile", "zaxis_2", "y", "x"]
    f_in = DataArray(T, dims=dims_c, name="T", attrs={"units": "K"})
    out = cs.regrid_vertical(DataArray(p_in, dims=dims_o), f_in, DataArray(p_out, dims=dims_o))
    ref = oracle.regrid_vertical(p_in, T, p_out)
    assert out.dims == tuple(dims_c) and out.attrs == {"units": "K"} and out.dtype == np.float32
    assert np.array_equal(out.values, ref, equal_nan=True)
    # z last, pressures in another order
    f_t = f_in.transpose("tile", "y", "x", "zaxis_1")
    out_t = cs.regrid_vertical(DataArray(p_in, dims=dims_o).transpose("y", "x", "tile", "zaxis_2"), f_t,
                               DataArray(p_out, dims=dims_o))
    assert out_t.dims == f_t.dims and np.array_equal(out_t.values, ref.transpose(0, 2, 3, 1), equal_nan=True)


@pytest.mark.parametrize("extrapolate", [False, True])
def test_regrid_to_area_weighted_pressure_then_average_equals_fused(cs, oracle, extrapolate):
    rng = np.random.default_rng(6)
    nb, km, n, f = 2, 12, 8, 4
    delp = _column_case(rng, nb, km, n).astype(np.float32)
    area = rng.uniform(1, 2, size=(nb, n, n)).astype(np.float32)
    fields = {"T": rng.uniform(180, 320, size=delp.shape).astype(np.float32),
              "W": rng.normal(size=delp.shape).astype(np.float32)}
    dims = ["tile", "zaxis_1", "yaxis_2", "xaxis_1"]
    ds = Dataset({k: DataArray(v, dims=dims, attrs={"long_name": k}) for k, v in fields.items()}, attrs={"src": "t"})
    d = DataArray(delp, dims=dims, name="delp")
    a = DataArray(area, dims=["tile", "yaxis_2", "xaxis_1"], name="area")
    reg, masked = cs.regrid_to_area_weighted_pressure(ds, d, a, 300.0, f, extrapolate=extrapolate)
    oreg, omasked = oracle.regrid_to_area_weighted_pressure(fields, delp, area[:, None], 300.0, f, extrapolate=extrapolate)
    assert masked.dims == tuple(dims) and np.array_equal(masked.values, omasked)
    for k in fields:
        assert reg[k].dims == tuple(dims) and np.array_equal(reg[k].values, oreg[k], equal_nan=True)
        assert reg[k].attrs == {"long_name": k}
    coarse = cs.weighted_block_average(reg, masked, f)
    fused, dco = cs.pressure_level_coarsen(ds, d, a, 300.0, f, extrapolate=extrapolate)
    assert fused.attrs == {"src": "t"}
    for k in fields:
        assert np.array_equal(coarse[k].values, oracle.weighted_block_average(oreg[k], omasked, f), equal_nan=True)
        assert np.array_equal(fused[k].values, coarse[k].values, equal_nan=True)
        assert fused[k].dims == tuple(dims)
    assert np.array_equal(dco.values, oracle.weighted_block_average(delp, area[:, None], f))


def test_regrid_to_edge_weighted_pressure(cs, oracle):
    rng = np.random.default_rng(7)
    nb, km, n, f = 6, 10, 8, 4
    delp = _column_case(rng, nb, km, n)
    dx = rng.uniform(1, 2, size=(nb, n + 1, n))
    u = rng.normal(size=(nb, km, n + 1, n))
    d = DataArray(delp, dims=["tile", "zaxis_1", "yaxis_2", "xaxis_1"], coords={"tile": np.arange(6)})
    ds = Dataset({"u": DataArray(u, dims=["tile", "zaxis_1", "yaxis_1", "xaxis_1"])})
    length = DataArray(dx, dims=["tile", "yaxis_1", "xaxis_1"])
    reg, masked = cs.regrid_to_edge_weighted_pressure(ds, d, length, 300.0, f, x_dim="xaxis_1", y_dim="yaxis_1", edge="x")
    oreg, omasked = oracle.regrid_to_edge_weighted_pressure({"u": u}, delp, dx[:, None], 300.0, f, edge="x")
    assert np.array_equal(reg["u"].values, oreg["u"], equal_nan=True)
    assert np.array_equal(masked.values, omasked)
    out = cs.edge_weighted_block_average(reg["u"], masked, f, x_dim="xaxis_1", y_dim="yaxis_1", edge="x")
    ref = oracle.edge_weighted_block_average(oreg["u"], omasked, f, edge="x")
    assert np.array_equal(out.transpose("tile", "zaxis_1", "yaxis_1", "xaxis_1").values, ref, equal_nan=True)


def test_shift_edge_var_to_center(cs):
    # test_cubedsphere.py:57-106
    edge = DataArray(np.array([[[30.0, 40.0, 50.0], [10.0, 20.0, 30.0]]]).transpose(0, 2, 1),
                     dims=["tile", "grid_y", "grid_xt"], coords={"grid_y": [1, 2, 3], "grid_xt": [1, 2]}, name="u")
    out = cs.shift_edge_var_to_center(edge)
    assert out.dims == ("tile", "grid_yt", "grid_xt")
    assert np.array_equal(out.values[0, :, 0], [35.0, 45.0]) and np.array_equal(out.values[0, :, 1], [15.0, 25.0])
    assert np.array_equal(out["grid_yt"].values, [1, 2])
