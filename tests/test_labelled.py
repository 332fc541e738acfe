"""CPU tests of the labelled-array layer (fv3net_b200/labelled.py) and of the host-only helpers of
fv3net_b200/cubedsphere.py (coordinates, reduction-token resolution, argument errors).  The
cases follow /root/reference/external/vcm/tests/test_cubedsphere.py:148-203 (coords)."""
import functools

import numpy as np
import pytest

from fv3net_b200 import cubedsphere as cs
from fv3net_b200.labelled import DataArray, Dataset, assert_identical_including_dtype, merge


def test_dataarray_basics():
    da = DataArray(np.arange(6.0).reshape(2, 3), dims=["y", "x"], coords={"x": [1.0, 2.0, 3.0]}, name="a",
                   attrs={"units": "m"})
    assert da.sizes == {"y": 2, "x": 3}
    t = da.transpose("x", "y")
    assert t.dims == ("x", "y") and t.shape == (3, 2)
    assert np.array_equal(t.coords["x"].values, [1.0, 2.0, 3.0])
    s = da.isel(x=slice(None, None, 2))
    assert s.shape == (2, 2) and np.array_equal(s["x"].values, [1.0, 3.0])
    assert da.isel(y=0).dims == ("x",)
    r = da.rename({"x": "xx"})
    assert r.dims == ("y", "xx") and "xx" in r.coords
    assert da.rename("b").name == "b"
    assert da.identical(da.copy())
    assert not da.identical(da.assign_attrs(units="K"))
    with pytest.raises(ValueError):
        DataArray(np.zeros((2, 2)), dims=["x"])


def test_dataarray_broadcast_arithmetic():
    a = DataArray(np.ones((2, 3)), dims=["y", "x"])
    w = DataArray(np.arange(4.0), dims=["z"])
    p = a * w
    assert p.dims == ("y", "x", "z") and p.shape == (2, 3, 4)
    q = DataArray(np.arange(3.0), dims=["x"]) * a
    assert q.dims == ("x", "y")
    assert np.array_equal((2.0 * a).values, 2 * np.ones((2, 3)))


def test_dataset_basics():
    foo = DataArray(np.zeros((4, 4, 2), np.float32), dims=["x", "y", "z"], name="foo", attrs={"units": "m"})
    bar = DataArray([1, 2, 3], dims=["t"], name="bar", attrs={"units": "m"})
    ds = merge([foo, bar]).assign_attrs(test="a")
    assert list(ds) == ["foo", "bar"] and ds.attrs == {"test": "a"}
    assert ds["foo"].attrs == {"units": "m"} and ds.sizes["t"] == 3
    assert ds.identical(ds.copy())
    out = ds.map(lambda da: da, keep_attrs=True)
    assert out.identical(ds)
    with pytest.raises(ValueError):
        ds["bad"] = DataArray(np.zeros(5), dims=["x"])


@pytest.mark.parametrize(
    ("coarsening_factor", "input_coordinate", "expected_coordinate"),
    [
        (2, [1.0, 2.0, 3.0, 4.0], [1.0, 2.0]),
        (2, [5.0, 6.0, 7.0, 8.0], [3.0, 4.0]),
        (2, [1.0, 2.0, 3.0], [1.0, 2.0]),
        (2, [3.0, 4.0, 5.0], [2.0, 3.0]),
    ],
    ids=["centers; first subtile", "centers; second subtile", "interfaces; first subtile", "interfaces; second subtile"],
)
def test_coarsen_coords(coarsening_factor, input_coordinate, expected_coordinate):
    inp = DataArray(input_coordinate, dims=["x"], coords=[input_coordinate], name="x")
    exp = np.array(expected_coordinate).astype(np.float32)
    expected = DataArray(exp, dims=["x"], coords=[exp], name="x")
    result = cs.coarsen_coords(coarsening_factor, inp, ["x"])
    assert_identical_including_dtype(result["x"], expected)


@pytest.mark.parametrize("kind", ["DataArray", "Dataset"])
def test_add_coordinates(kind):
    x = np.array([1.0, 2.0], np.float32)
    y = np.array([3.0, 4.0], np.float32)
    reference = DataArray([[1, 1], [1, 1]], dims=["x", "y"], coords=[x, y], name="foo")
    coarsened = DataArray([[1]], dims=["x", "y"], name="foo")
    expected = DataArray([[1]], dims=["x", "y"], coords=[np.array([1.0], np.float32), np.array([2.0], np.float32)],
                         name="foo")
    if kind == "Dataset":
        coarsened, expected = coarsened.to_dataset(), expected.to_dataset()
    result = cs.add_coordinates(reference, coarsened, 2, ["x", "y"])
    assert_identical_including_dtype(result, expected)


def test_coord_func_forms():
    c = np.arange(1.0, 9.0)
    assert np.array_equal(cs._reduce_coord(c, 2, "mean"), [1.5, 3.5, 5.5, 7.5])
    assert np.array_equal(cs._reduce_coord(c, 4, np.max), [4.0, 8.0])
    out = cs._reduce_coord(c, 2, cs.coarsen_coords_coord_func)
    assert out.dtype == np.float32 and np.array_equal(out, [1, 2, 3, 4])
    # staggered (padded) coordinate
    assert np.array_equal(cs._reduce_coord(np.arange(1.0, 6.0), 2, cs.coarsen_coords_coord_func), [1, 2, 3])
    assert np.array_equal(cs._reduce_coord(np.arange(1.0, 6.0), 2, "mean"), [1.5, 3.5, 5.0])
    assert cs._coord_function({"x": np.min}, "y") == "mean"


def test_reduction_tokens():
    assert cs._resolve_reduction(np.median) == ("median", False)
    assert cs._resolve_reduction(np.nansum) == ("sum", True)
    assert cs._resolve_reduction(functools.partial(cs._mode_reduce, nan_policy="omit")) == ("mode", True)
    assert cs._resolve_reduction(cs._mode_reduce) == ("mode", False)
    with pytest.raises(TypeError, match="supported reductions"):
        cs._resolve_reduction(lambda a, axis: a.sum(axis))
    with pytest.raises(TypeError, match="supported reductions"):
        cs._resolve_reduction(np.std)


def test_argument_errors_raise_before_the_gpu():
    da = DataArray(np.zeros((2, 3)), dims=["x_dim", "y_dim"])
    with pytest.raises(ValueError, match="'edge' most be either 'x' or 'y'"):
        cs.edge_weighted_block_average(da, da, 2, x_dim="x_dim", y_dim="y_dim", edge="z")
    with pytest.raises(ValueError, match="'edge' most be either 'x' or 'y'"):
        cs.block_edge_coarsen(da, 2, x_dim="x_dim", y_dim="y_dim", edge="q")
    with pytest.raises(ValueError, match="Cannot repeat over"):
        cs.block_upsample(da, 2, ["nope"])
    p = DataArray(np.zeros((3, 4)), dims=["c", "zaxis_2"])
    f = DataArray(np.zeros((3, 3)), dims=["c", "zaxis_1"])
    with pytest.raises(ValueError, match="must not be equal"):
        cs.regrid_vertical(p, f, p, z_dim_center="z", z_dim_outer="z")
    with pytest.raises(ValueError, match="one shorter than p_in"):
        cs.regrid_vertical(p, DataArray(np.zeros((3, 2)), dims=["c", "zaxis_1"]), p)
    with pytest.raises(ValueError, match="All dimensions except vertical must be same size"):
        cs.regrid_vertical(DataArray(np.zeros((2, 4)), dims=["c", "zaxis_2"]), f, p)
    with pytest.raises(ValueError, match="Variable to shift to center"):
        cs.shift_edge_var_to_center(DataArray(np.zeros((2, 2)), dims=["a", "b"]))


def test_block_reduce_without_block_dims_is_identity():
    bar = DataArray([1, 2, 3], dims=["t"], name="bar", attrs={"units": "m"})
    assert cs.horizontal_block_reduce(bar, 2, np.median, "x", "y") is bar
