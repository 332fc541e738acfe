"""The reference's operator API (``vcm.cubedsphere``) on labelled arrays, backed by libb200cs.

Same function names, argument meaning, defaults and error messages as
/root/reference/external/vcm/vcm/cubedsphere/{coarsen,regridz}.py (cited per function as
``coarsen.py:LINE`` / ``regridz.py:LINE``).  Inputs are labelled arrays: ``xarray.DataArray`` /
``xarray.Dataset`` where xarray is installed, or the equivalent minimal classes of
``fv3net_b200.labelled`` (xarray is absent from this image).  Only attributes common to both are
used; results are built with ``type(obj)(...)`` so the caller gets back what it passed in.

All array arithmetic runs in the CUDA library through ``fv3net_b200.core``; this module handles
dimension names, dimension order, coordinates and attributes (SURVEY.md 8a rows A1-A5, A7, A8,
A10, A11).  There is no CPU path: without the library / a GPU every operator raises.

Documented deviation: a GPU library cannot run an arbitrary Python callable.
``horizontal_block_reduce`` / ``xarray_block_reduce`` accept the callables the reference itself
passes (``np.sum/mean/min/max/median``, their ``nan*`` forms, ``_mode_reduce`` partials) and raise
``TypeError`` naming that set for anything else.
"""
import functools

import numpy as np
import torch

from . import core
from .labelled import is_dataset

# dimension names (external/vcm/vcm/cubedsphere/constants.py)
COORD_X_CENTER, COORD_X_OUTER = "grid_xt", "grid_x"
COORD_Y_CENTER, COORD_Y_OUTER = "grid_yt", "grid_y"
COORD_Z_CENTER, COORD_Z_OUTER = "pfull", "phalf"
FV_CORE_X_CENTER, FV_CORE_X_OUTER = "xaxis_1", "xaxis_2"
FV_CORE_Y_CENTER, FV_CORE_Y_OUTER = "yaxis_2", "yaxis_1"
RESTART_Z_CENTER, RESTART_Z_OUTER = "zaxis_1", "zaxis_2"
TOA_PRESSURE = 300.0
SURFACE_LEVEL = -1


# ------------------------------------------------------------------------------------------
# labelled-array plumbing
# ------------------------------------------------------------------------------------------
def _plain_coords(obj):
    return {k: (tuple(obj.coords[k].dims), np.asarray(obj.coords[k].values)) for k in obj.coords}


def _new(like, data, dims, coords=None, name=None, attrs=None):
    sizes = dict(zip(dims, data.shape))
    keep = {}
    for k, (cd, cv) in (coords or {}).items():
        if all(d in sizes and sizes[d] == n for d, n in zip(cd, cv.shape)):
            keep[k] = (cd, cv)
    return type(like)(data, coords=keep or None, dims=tuple(dims), name=name, attrs=dict(attrs or {}))


def _map_dataset(ds, func, keep_attrs=True):
    """Applies func to every data variable; dataset attrs and per-variable attrs follow
    _propagate_attrs (coarsen.py:164-180)."""
    out, coords = {}, {}
    for k in ds.data_vars:
        r = func(ds[k])
        for ck, cv in _plain_coords(r).items():
            coords.setdefault(ck, cv)
        out[k] = (tuple(r.dims), np.asarray(r.values), dict(ds[k].attrs) if keep_attrs else dict(r.attrs))
    return type(ds)({k: v for k, v in out.items()}, coords=coords or None, attrs=dict(ds.attrs))


def _np_host(t):
    return t.cpu().numpy() if isinstance(t, torch.Tensor) else np.asarray(t)


def _aligned(obj, out_dims):
    """values of obj viewed with the axes of out_dims (size-1 where obj lacks a dimension)."""
    dims = tuple(obj.dims)
    v = np.asarray(obj.values)
    order = [dims.index(d) for d in out_dims if d in dims]
    v = v.transpose(order)
    return v.reshape(tuple(obj.sizes[d] if d in dims else 1 for d in out_dims))


def _is_dict_like(value):
    return hasattr(value, "keys") and hasattr(value, "__getitem__")


# ------------------------------------------------------------------------------------------
# coordinates (A11)
# ------------------------------------------------------------------------------------------
def coarsen_coords_coord_func(coordinate, axis=-1):
    """coarsen.py:109-132: first value of each window -> subtile-aware coarse index, float32."""
    coordinate = np.asarray(coordinate)
    return ((coordinate[:, 0] - 1) // coordinate.shape[1] + 1).astype(int).astype(np.float32)


def coarsen_coords(coarsening_factor, reference_subtile, dims):
    """coarsen.py:83-106."""
    result = {}
    for dim in dims:
        c = np.asarray(reference_subtile[dim].values)
        v = ((c[::coarsening_factor] - 1) // coarsening_factor + 1).astype(int).astype(np.float32)
        ref = reference_subtile[dim]
        result[dim] = _new(ref, v, (dim,), {dim: ((dim,), v)}, name=dim, attrs=getattr(ref, "attrs", None))
    return result


def add_coordinates(reference_obj, coarsened_obj, coarsening_factor, dims):
    """coarsen.py:135-161."""
    coarsened = coarsen_coords(coarsening_factor, reference_obj, dims)
    return coarsened_obj.assign_coords({k: np.asarray(v.values) for k, v in coarsened.items()})


_STR_FUNCS = {"mean": np.nanmean, "sum": np.nansum, "min": np.nanmin, "max": np.nanmax, "median": np.nanmedian,
              "std": np.nanstd, "var": np.nanvar, "prod": np.nanprod}


def _reduce_coord(values, f, function):
    """One 1-D coordinate reduced in windows of f; a trailing partial window is padded with NaN
    (coarsen.py:304-319: boundary='pad' for names, skimage padding for callables)."""
    values = np.asarray(values)
    n = values.shape[0]
    npad = (-n) % f
    if npad:
        values = np.concatenate([values.astype(np.result_type(values.dtype, np.float32)), np.full(npad, np.nan)])
    blocked = values.reshape(-1, f)
    if isinstance(function, str):
        if function not in _STR_FUNCS:
            raise ValueError(f"unsupported coord_func {function!r}")
        return _STR_FUNCS[function](blocked, axis=1) if npad else getattr(np, function)(blocked, axis=1)
    return np.asarray(function(blocked, axis=(1,)))


def _coord_function(coord_func, dim):
    if _is_dict_like(coord_func):
        return coord_func.get(dim, "mean")
    return coord_func


def _coarsened_coords(coords, block_sizes, coord_func):
    """coords (name -> (dims, values)) after coarsening block_sizes (dim -> f): 1-D coordinates
    along a coarsened dimension are reduced through coord_func, the others pass through."""
    out = {}
    for k, (cd, cv) in coords.items():
        hit = [d for d in cd if d in block_sizes]
        if not hit:
            out[k] = (cd, cv)
        elif len(cd) == 1:
            out[k] = (cd, _reduce_coord(cv, block_sizes[cd[0]], _coord_function(coord_func, k)))
    return out


def rename_centered_xy_coords(cell_centered_da, edge_to_center_dims=None):
    """coarsen.py:34-51."""
    edge_to_center_dims = edge_to_center_dims or {"grid_x": "grid_xt", "grid_y": "grid_yt"}
    for dim in edge_to_center_dims:
        if dim in cell_centered_da.dims:
            cell_centered_da = cell_centered_da.assign_coords({dim: np.asarray(cell_centered_da[dim].values) - 1})
            cell_centered_da = cell_centered_da.rename({dim: edge_to_center_dims[dim]})
    return cell_centered_da


def shift_edge_var_to_center(edge, edge_to_center_dims=None):
    """coarsen.py:54-80: 0.5*(e[i] + e[i-1]) along the one staggered dimension (on the device)."""
    edge_to_center_dims = edge_to_center_dims or {"grid_x": "grid_xt", "grid_y": "grid_yt"}
    for staggered_dim in [d for d in edge_to_center_dims if d in edge.dims]:
        ax = list(edge.dims).index(staggered_dim)
        t, _ = core._to_dev(np.asarray(edge.values))
        hi = t.narrow(ax, 1, t.shape[ax] - 1)
        lo = t.narrow(ax, 0, t.shape[ax] - 1)
        data = (0.5 * (hi + lo)).cpu().numpy()
        coords = {}
        for k, (cd, cv) in _plain_coords(edge).items():
            if staggered_dim in cd:
                coords[k] = (cd, cv[tuple(slice(1, None) if d == staggered_dim else slice(None) for d in cd)])
            else:
                coords[k] = (cd, cv)
        shifted = _new(edge, data, edge.dims, coords, name=edge.name, attrs=None)
        return rename_centered_xy_coords(shifted, edge_to_center_dims)
    raise ValueError(
        "Variable to shift to center must be centered on one horizontal axis and edge-valued on the other."
    )


# ------------------------------------------------------------------------------------------
# weighted block averages (A1, A2)
# ------------------------------------------------------------------------------------------
def _weighted_average_da(first, second, f, x_dim, y_dim, coord_func, edge, obj_is_first):
    """nansum(first*second)/nansum(weights) with xarray's broadcasting: output dimensions are
    first's, then second's remaining ones (coarsen.py:211 ``obj * weights``; :260 ``spacing * obj``)."""
    obj, weights = (first, second) if obj_is_first else (second, first)
    out_dims = tuple(first.dims) + tuple(d for d in second.dims if d not in first.dims)
    for d in (x_dim, y_dim):
        if d not in weights.dims:
            raise ValueError(f"weights lack dimension {d!r}")
    lead = [d for d in out_dims if d not in (x_dim, y_dim)]
    kdims = lead + [y_dim, x_dim]
    res = core.weighted_block_average(_aligned(obj, kdims), _aligned(weights, kdims), f, edge=edge)
    res = res.transpose([kdims.index(d) for d in out_dims])
    if edge is None:
        blocks = {x_dim: f, y_dim: f}
    else:
        blocks = {x_dim: f} if edge == "x" else {y_dim: f}
    coords = dict(_plain_coords(first))
    for k, v in _plain_coords(second).items():
        coords.setdefault(k, v)
    coords = _coarsened_coords(coords, blocks, coord_func)
    if edge is not None:
        down = y_dim if edge == "x" else x_dim
        coords = {k: ((cd, cv[::f]) if cd == (down,) else (cd, cv)) for k, (cd, cv) in coords.items()}
        # the downsampled dimension's coordinate goes through coord_func too (coarsen.py:282-301)
        ref = _plain_coords(obj)
        if down in ref:
            coords[down] = ((down,), _reduce_coord(ref[down][1], f, _coord_function(coord_func, down)))
    return _new(obj, res, out_dims, coords, name=obj.name, attrs=obj.attrs)


def weighted_block_average(obj, weights, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_2",
                           coord_func=coarsen_coords_coord_func):
    """coarsen.py:183-218: sum(obj*weights)/sum(weights) over f x f blocks, NaN-skipping sums,
    one fused pass on the GPU (b200cs_weighted_block_average)."""
    run = functools.partial(_weighted_average_da, second=weights, f=coarsening_factor, x_dim=x_dim, y_dim=y_dim,
                            coord_func=coord_func, edge=None, obj_is_first=True)
    if is_dataset(obj):
        return _map_dataset(obj, lambda da: run(da))
    return run(obj)


def edge_weighted_block_average(obj, spacing, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", edge="x",
                                coord_func=coarsen_coords_coord_func):
    """coarsen.py:221-273: weighted mean over f cells along the edge, every f-th line across it
    (only those lines are read)."""
    if edge not in ("x", "y"):
        raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")
    run = functools.partial(_weighted_average_da, spacing, f=coarsening_factor, x_dim=x_dim, y_dim=y_dim,
                            coord_func=coord_func, edge=edge, obj_is_first=False)
    if is_dataset(obj):
        return _map_dataset(obj, lambda da: run(da))
    return run(obj)


# ------------------------------------------------------------------------------------------
# block reductions (A3, A10)
# ------------------------------------------------------------------------------------------
def _mode(arr, axis=0, nan_policy="propagate"):
    """Identity token of coarsen.py:686-692 (scipy.stats.mode); evaluated on the GPU as op 'mode'."""
    raise TypeError("_mode is a reduction token for horizontal_block_reduce; it is evaluated on the GPU")


def _mode_reduce(arr, axis=(0,), nan_policy="propagate"):
    """Identity token of coarsen.py:743-747."""
    raise TypeError("_mode_reduce is a reduction token for horizontal_block_reduce; it is evaluated on the GPU")


_REDUCTIONS = {
    np.sum: ("sum", False), np.nansum: ("sum", True), np.mean: ("mean", False), np.nanmean: ("mean", True),
    np.min: ("min", False), np.amin: ("min", False), np.nanmin: ("min", True),
    np.max: ("max", False), np.amax: ("max", False), np.nanmax: ("max", True),
    np.median: ("median", False), np.nanmedian: ("median", True),
}


def _resolve_reduction(function):
    if isinstance(function, functools.partial) and function.func in (_mode_reduce, _mode) and not function.args:
        policy = function.keywords.get("nan_policy", "propagate")
        if policy not in ("propagate", "omit"):
            raise ValueError(f"nan_policy {policy!r} is not supported on the GPU ('propagate' or 'omit')")
        return "mode", policy == "omit"
    if function in (_mode_reduce, _mode):
        return "mode", False
    try:
        return _REDUCTIONS[function]
    except (KeyError, TypeError):
        names = "np.sum, np.mean, np.min, np.max, np.median, their nan* forms, and _mode_reduce partials"
        raise TypeError(
            f"reduction_function {function!r} cannot run on the GPU; supported reductions: {names}"
        ) from None


def _block_reduce_da(da, block_sizes, method, skipna, cval=np.nan, coord_func=coarsen_coords_coord_func,
                     down=None):
    """Reduction of da over 1-D or 2-D blocks of (at most two) named dimensions; ``down`` =
    (dim, step): that dimension is downsampled instead (the block-edge operators)."""
    red = [d for d in da.dims if d in block_sizes and block_sizes[d] != 1]
    if len(red) > 2:
        raise ValueError("block reductions over more than two dimensions are not supported")
    if not [d for d in da.dims if d in block_sizes]:
        return da  # coarsen.py:434-436
    values = np.asarray(da.values)
    int_dtype = values.dtype if values.dtype.kind in "iub" else None
    if not red and down is None:
        return da
    ydim, xdim = (red + [None, None])[:2] if len(red) == 2 else (None, red[0] if red else None)
    if down is not None and down[0] in da.dims:
        if ydim is not None:
            raise ValueError("edge reductions take one block dimension")
        ydim = down[0]
    others = [d for d in da.dims if d not in (ydim, xdim)]
    kdims = others + [d for d in (ydim, xdim) if d is not None]
    v = values.transpose([list(da.dims).index(d) for d in kdims])
    if ydim is None:
        v = v[..., None, :]
    if xdim is None:
        v = v[..., None]
    fx = block_sizes[xdim] if xdim is not None else 1
    if down is not None and ydim == down[0]:
        window = (1, fx, down[1], fx)
    else:
        fy = block_sizes[ydim] if ydim is not None else 1
        window = (fy, fx, fy, fx)
    res = core.block_reduce(v, 1, method=method, skipna=skipna, pad=(down is None), cval=cval, window=window)
    if ydim is None:
        res = res[..., 0, :]
    if xdim is None:
        res = res[..., 0]
    if int_dtype is not None and method in ("sum", "min", "max", "mode"):
        res = res.astype(np.int64 if method == "sum" and int_dtype.kind != "u" else int_dtype)
    res = res.transpose([kdims.index(d) for d in da.dims])
    coords = _coarsened_coords(_plain_coords(da), {d: block_sizes[d] for d in red}, coord_func)
    if down is not None and down[0] in da.dims:
        ref = _plain_coords(da)
        coords = {k: ((cd, cv[::down[1]]) if cd == (down[0],) else (cd, cv)) for k, (cd, cv) in coords.items()}
        if down[0] in ref:
            coords[down[0]] = ((down[0],), _reduce_coord(ref[down[0]][1], down[1], _coord_function(coord_func, down[0])))
    return _new(da, res, da.dims, coords, name=da.name, attrs=da.attrs)


def _xarray_block_reduce_dataarray(da, block_sizes, reduction_function, cval=np.nan,
                                   coord_func=coarsen_coords_coord_func):
    """coarsen.py:393-460."""
    method, skipna = _resolve_reduction(reduction_function)
    if not _is_dict_like(coord_func):
        coord_func = {d: coord_func for d in da.dims}
    return _block_reduce_da(da, dict(block_sizes), method, skipna, cval=cval, coord_func=coord_func)


def xarray_block_reduce(obj, block_sizes, reduction_function, cval=np.nan, coord_func=coarsen_coords_coord_func):
    """coarsen.py:463-517."""
    run = functools.partial(_xarray_block_reduce_dataarray, block_sizes=block_sizes,
                            reduction_function=reduction_function, cval=cval, coord_func=coord_func)
    if is_dataset(obj):
        return _map_dataset(obj, lambda da: run(da))
    return run(obj)


def horizontal_block_reduce(obj, coarsening_factor, reduction_function, x_dim="xaxis_1", y_dim="yaxis_1",
                            coord_func=coarsen_coords_coord_func):
    """coarsen.py:520-554."""
    return xarray_block_reduce(obj, {x_dim: coarsening_factor, y_dim: coarsening_factor}, reduction_function,
                               coord_func=coord_func)


def block_median(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", coord_func=coarsen_coords_coord_func):
    """coarsen.py:557-588."""
    return horizontal_block_reduce(obj, coarsening_factor, np.median, x_dim=x_dim, y_dim=y_dim, coord_func=coord_func)


def _block_mode(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", coord_func=coarsen_coords_coord_func,
                nan_policy="propagate"):
    """coarsen.py:750-786."""
    return horizontal_block_reduce(obj, coarsening_factor, functools.partial(_mode_reduce, nan_policy=nan_policy),
                                   x_dim=x_dim, y_dim=y_dim, coord_func=coord_func)


CUSTOM_COARSENING_FUNCTIONS = {"median": block_median, "mode": _block_mode}
_COARSEN_METHODS = ("sum", "mean", "min", "max")


def block_coarsen(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", method="sum",
                  coord_func=coarsen_coords_coord_func, func_kwargs=None):
    """coarsen.py:795-840: 'median' / 'mode' through the custom reducers, any other name as
    xarray's ``coarsen(...).<method>()`` (NaN-skipping)."""
    func_kwargs = func_kwargs or {}
    if method in CUSTOM_COARSENING_FUNCTIONS:
        return CUSTOM_COARSENING_FUNCTIONS[method](obj, coarsening_factor, x_dim=x_dim, y_dim=y_dim,
                                                   coord_func=coord_func, **func_kwargs)
    if method not in _COARSEN_METHODS:
        raise ValueError(f"coarsening method {method!r} is not supported on the GPU; supported: "
                         f"{_COARSEN_METHODS + tuple(CUSTOM_COARSENING_FUNCTIONS)}")
    blocks = {x_dim: coarsening_factor, y_dim: coarsening_factor}

    def run(da):
        if any(da.sizes[d] % coarsening_factor for d in blocks if d in da.dims):
            raise ValueError(f"Could not coarsen a dimension of size {da.sizes} with window {coarsening_factor}")
        out = _block_reduce_da(da, blocks, method, True, coord_func=coord_func)
        return out

    if is_dataset(obj):
        return _map_dataset(obj, run)
    return run(obj)


def block_edge_coarsen(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", edge="x",
                       coord_func=coarsen_coords_coord_func, method="sum"):
    """coarsen.py:629-683: 1-D coarsening along the edge, every f-th line across it."""
    if edge == "x":
        coarsen_dim, downsample_dim = x_dim, y_dim
    elif edge == "y":
        coarsen_dim, downsample_dim = y_dim, x_dim
    else:
        raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")
    if method not in _COARSEN_METHODS:
        raise ValueError(f"coarsening method {method!r} is not supported on the GPU; supported: {_COARSEN_METHODS}")

    def run(da):
        return _block_reduce_da(da, {coarsen_dim: coarsening_factor}, method, True, coord_func=coord_func,
                                down=(downsample_dim, coarsening_factor))

    if is_dataset(obj):
        return _map_dataset(obj, run)
    return run(obj)


def block_edge_sum(obj, coarsening_factor, x_dim="xaxis_1", y_dim="yaxis_1", edge="x",
                   coord_func=coarsen_coords_coord_func):
    """coarsen.py:591-626."""
    return block_edge_coarsen(obj, coarsening_factor, x_dim=x_dim, y_dim=y_dim, edge=edge, coord_func=coord_func,
                              method="sum")


# ------------------------------------------------------------------------------------------
# upsampling (A4)
# ------------------------------------------------------------------------------------------
def _upsample_da(da, factor, dims):
    hit = [d for d in dims if d in da.dims]
    if not hit:
        return da
    if len(hit) > 2:
        raise ValueError("block_upsample takes at most two dimensions")
    others = [d for d in da.dims if d not in hit]
    kdims = others + hit
    v = np.asarray(da.values).transpose([list(da.dims).index(d) for d in kdims])
    if len(hit) == 1:
        v = v[..., None, :]
        res = core.block_upsample(v, factor, axes=(-1,))[..., 0, :]
    else:
        res = core.block_upsample(v, factor, axes=(-2, -1))
    if res.dtype != v.dtype:
        res = res.astype(v.dtype)
    res = res.transpose([kdims.index(d) for d in da.dims])
    coords = {k: cv for k, cv in _plain_coords(da).items() if not any(d in hit for d in cv[0])}
    return _new(da, res, da.dims, coords, name=da.name, attrs=da.attrs)


def block_upsample(obj, upsampling_factor, dims):
    """coarsen.py:869-897 (+ xarray_utils.py:37-82): repeat along each named dimension; an
    odd-size dimension is an interface dimension whose last point is not repeated; coordinates
    along the upsampled dimensions are dropped."""
    for dim in dims:
        if dim not in obj.dims:
            raise ValueError("Cannot repeat over {!r}; expected one of {!r}.".format(dim, list(obj.dims)))
    if is_dataset(obj):
        return _map_dataset(obj, lambda da: _upsample_da(da, upsampling_factor, dims))
    return _upsample_da(obj, upsampling_factor, dims)


def block_upsample_like(da, reference_da, x_dim="xaxis_1", y_dim="yaxis_1"):
    """coarsen.py:900-938: factor inferred from the x dimension; the reference's horizontal
    coordinates are attached."""
    if da.sizes[x_dim] % 2 == 1:
        factor = (reference_da.sizes[x_dim] - 1) // (da.sizes[x_dim] - 1)
    else:
        factor = reference_da.sizes[x_dim] // da.sizes[x_dim]
    result = block_upsample(da, factor, [x_dim, y_dim])
    ref = _plain_coords(reference_da)
    new = {d: ref[d][1] for d in (x_dim, y_dim) if d in ref}
    return result.assign_coords(new) if new else result


# ------------------------------------------------------------------------------------------
# vertical remap (A5-A8)
# ------------------------------------------------------------------------------------------
def pressure_at_interface(delp, toa_pressure=TOA_PRESSURE, dim_center=COORD_Z_CENTER, dim_outer=COORD_Z_OUTER):
    """calc/thermo/vertically_dependent.py:41-66."""
    ax = list(delp.dims).index(dim_center)
    res = core.pressure_at_interface(np.asarray(delp.values), toa_pressure, z_axis=ax)
    dims = tuple(dim_outer if d == dim_center else d for d in delp.dims)
    coords = {k: cv for k, cv in _plain_coords(delp).items() if dim_center not in cv[0]}
    return _new(delp, res, dims, coords, name=delp.name, attrs=None)


def pressure_at_midpoint_log(delp, toa_pressure=TOA_PRESSURE, dim=COORD_Z_CENTER):
    """calc/thermo/vertically_dependent.py:153-179: delp / diff(log(p_interface))."""
    ax = list(delp.dims).index(dim)
    d, _ = core._to_dev(np.asarray(delp.values))
    pi = core.pressure_at_interface(d, toa_pressure, z_axis=ax)
    res = (d / torch.diff(torch.log(pi), dim=ax)).cpu().numpy()
    return _new(delp, res, delp.dims, _plain_coords(delp), name=delp.name, attrs=None)


def regrid_vertical(p_in, f_in, p_out, iv=1, kord=1, z_dim_center="zaxis_1", z_dim_outer="zaxis_2", mode=core.EXACT):
    """regridz.py:223-301: conservative PPM remap (mappm) of every column; float32 out like the
    Fortran; dimension order and attributes of f_in.  ``mode`` is this library's addition
    (core.FAST: 1e-6 arithmetic, kord <= 3)."""
    if z_dim_center == z_dim_outer:
        raise ValueError("'z_dim_center' and 'z_dim_outer' must not be equal.")
    dims_except_z = [d for d in f_in.dims if d != z_dim_center]
    zax = list(f_in.dims).index(z_dim_center)

    def arrange(da, zname):
        order = list(dims_except_z)
        order.insert(zax, zname)
        if set(order) != set(da.dims):
            raise ValueError("All dimensions except vertical must be same size for p_in, f_in and p_out")
        return np.asarray(da.values).transpose([list(da.dims).index(d) for d in order])

    pi, fi, po = arrange(p_in, z_dim_outer), np.asarray(f_in.values), arrange(p_out, z_dim_outer)
    cols = lambda a: a.shape[:zax] + a.shape[zax + 1:]  # noqa: E731
    if cols(pi) != cols(fi) or cols(po) != cols(fi):
        raise ValueError("All dimensions except vertical must be same size for p_in, f_in and p_out")
    if fi.shape[zax] != pi.shape[zax] - 1:
        raise ValueError("f_in must have a vertical dimension one shorter than p_in")
    res = core.regrid_vertical(pi, fi, po, iv=iv, kord=kord, z_axis=zax, mode=mode)
    coords = {k: cv for k, cv in _plain_coords(f_in).items() if z_dim_center not in cv[0]}
    return _new(f_in, res, f_in.dims, coords, name=f_in.name, attrs=f_in.attrs)


def _mask_weights(weights, pfull_coarse_on_fine, phalf_coarse_on_fine, phalf_fine, dim_center=RESTART_Z_CENTER,
                  dim_outer=RESTART_Z_OUTER, extrapolate=False):
    """regridz.py:200-220: weights where the coarse level is above the fine surface, else 0."""
    pc = pfull_coarse_on_fine if extrapolate else phalf_coarse_on_fine
    zname = dim_center if extrapolate else dim_outer
    hdims = [d for d in phalf_fine.dims if d != dim_outer]
    if len(hdims) < 2:
        raise ValueError("mask_weights needs two horizontal dimensions")
    lead, (yd, xd) = hdims[:-2], hdims[-2:]
    k_p = lead + [zname, yd, xd]
    pcv = np.asarray(pc.values).transpose([list(pc.dims).index(d) for d in k_p])
    pfv = np.asarray(phalf_fine.values).transpose([list(phalf_fine.dims).index(d) for d in lead + [dim_outer, yd, xd]])
    wv = _aligned(weights, lead + [yd, xd])
    res = core.mask_weights(wv, pcv, pfv, mask_mode="pfull" if extrapolate else "phalf")
    out_dims = lead + [dim_center, yd, xd]
    return _new(weights, res, out_dims, None, name=weights.name, attrs=None)


def _regrid_given_delp(ds, delp_fine, delp_coarse, weights, toa_pressure, x_dim=FV_CORE_X_CENTER,
                       y_dim=FV_CORE_Y_CENTER, z_dim=RESTART_Z_CENTER, extrapolate=False):
    """regridz.py:149-197."""
    delp_coarse_on_fine = block_upsample_like(delp_coarse, delp_fine, x_dim=x_dim, y_dim=y_dim)
    phalf_coarse_on_fine = pressure_at_interface(delp_coarse_on_fine, dim_center=z_dim, dim_outer=RESTART_Z_OUTER,
                                                 toa_pressure=toa_pressure)
    phalf_fine = pressure_at_interface(delp_fine, dim_center=z_dim, dim_outer=RESTART_Z_OUTER,
                                       toa_pressure=toa_pressure)
    # every operand in delp_fine's dimension order
    phalf_coarse_on_fine = phalf_coarse_on_fine.transpose(*phalf_fine.dims)
    regridded = _map_dataset(ds, lambda da: regrid_vertical(phalf_fine, da, phalf_coarse_on_fine, z_dim_center=z_dim))
    pfull_coarse_on_fine = pressure_at_midpoint_log(delp_coarse_on_fine, dim=z_dim, toa_pressure=toa_pressure)
    pfull_coarse_on_fine = pfull_coarse_on_fine.transpose(*delp_fine.dims)
    masked_weights = _mask_weights(weights, pfull_coarse_on_fine, phalf_coarse_on_fine, phalf_fine,
                                   dim_center=z_dim, extrapolate=extrapolate)
    return regridded, masked_weights


def regrid_to_area_weighted_pressure(ds, delp, area, toa_pressure, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                                     y_dim=FV_CORE_Y_CENTER, z_dim=RESTART_Z_CENTER, extrapolate=False):
    """regridz.py:31-78 -> (regridded Dataset, masked area).  The operator-by-operator form
    (five kernels, fine-size intermediates); ``pressure_level_coarsen`` is the fused equivalent
    of this call followed by ``weighted_block_average(regridded, masked_area)``."""
    delp_coarse = weighted_block_average(delp, area, coarsening_factor, x_dim=x_dim, y_dim=y_dim)
    return _regrid_given_delp(ds, delp, delp_coarse, area, toa_pressure, x_dim=x_dim, y_dim=y_dim, z_dim=z_dim,
                              extrapolate=extrapolate)


def interp_to_edges(delp, interp_dim, x_dim=FV_CORE_X_CENTER, y_dim=FV_CORE_Y_CENTER, tile_dim="tile"):
    """The ``grid.interp(delp, 'x'|'y')`` of regridz.py:118-132 / coarsen_restarts.py:825-853
    (xgcm with FV3_FACE_CONNECTIONS, cubedsphere/xgcm.py:7-34): centre -> edge along interp_dim with
    ghost lines from the neighbouring cube faces.  The interpolated dimension becomes the
    staggered one (x: xaxis_1 -> xaxis_2, y: yaxis_2 -> yaxis_1)."""
    if tile_dim not in delp.dims or delp.sizes[tile_dim] != 6:
        raise ValueError("interp_to_edges needs all six tiles")
    if tile_dim in delp.coords and list(np.asarray(delp.coords[tile_dim].values)) != list(range(6)):
        raise ValueError("tile coordinate must be 0..5")
    lead = [d for d in delp.dims if d not in (tile_dim, x_dim, y_dim)]
    kdims = [tile_dim] + lead + [y_dim, x_dim]
    v = np.asarray(delp.values).transpose([list(delp.dims).index(d) for d in kdims])
    res = core.stagger(v, interp_dim)
    new = {"x": (x_dim, FV_CORE_X_OUTER), "y": (y_dim, FV_CORE_Y_OUTER)}[interp_dim]
    out_k = [new[1] if d == new[0] else d for d in kdims]
    out_dims = [new[1] if d == new[0] else d for d in delp.dims]
    res = res.transpose([out_k.index(d) for d in out_dims])
    coords = {k: cv for k, cv in _plain_coords(delp).items() if new[0] not in cv[0]}
    return _new(delp, res, out_dims, coords, name=delp.name, attrs=None)


def regrid_to_edge_weighted_pressure(ds, delp, length, toa_pressure, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                                     y_dim=FV_CORE_Y_OUTER, z_dim=RESTART_Z_CENTER, edge="x", extrapolate=False):
    """regridz.py:81-146 -> (regridded Dataset, masked edge lengths)."""
    hor_dims = {"x": x_dim, "y": y_dim}
    interp_dim = "x" if edge == "y" else "y"
    delp_staggered = interp_to_edges(delp, interp_dim)
    stag = hor_dims[interp_dim]
    delp_staggered = delp_staggered.assign_coords(
        {stag: np.arange(1, delp.sizes[FV_CORE_X_CENTER if edge == "x" else FV_CORE_Y_CENTER] + 2, dtype=np.float32)})
    delp_staggered_coarse = edge_weighted_block_average(delp_staggered, length, coarsening_factor, x_dim=x_dim,
                                                        y_dim=y_dim, edge=edge)
    delp_staggered_coarse = delp_staggered_coarse.transpose(*delp_staggered.dims)
    return _regrid_given_delp(ds, delp_staggered, delp_staggered_coarse, length, toa_pressure, x_dim=x_dim,
                              y_dim=y_dim, z_dim=z_dim, extrapolate=extrapolate)


def pressure_level_coarsen(ds, delp, area, toa_pressure, coarsening_factor, x_dim=FV_CORE_X_CENTER,
                           y_dim=FV_CORE_Y_CENTER, z_dim=RESTART_Z_CENTER, extrapolate=False, tile_dim="tile",
                           mode=core.EXACT, coord_func=coarsen_coords_coord_func):
    """The fused hot path: ``regrid_to_area_weighted_pressure`` followed by
    ``weighted_block_average(regridded, masked_area)`` (coarsen_restarts.py:485-534,934-959) in ONE
    kernel pass (b200cs_pressure_coarsen).  Returns (coarse Dataset, coarse delp)."""
    lead = [d for d in delp.dims if d not in (z_dim, y_dim, x_dim)]
    kdims = lead + [z_dim, y_dim, x_dim]
    names = list(ds.data_vars)

    def k4(da):
        v = np.asarray(da.values).transpose([list(da.dims).index(d) for d in kdims])
        return v.reshape((-1,) + v.shape[-3:])

    lshape = tuple(delp.sizes[d] for d in lead)
    w = np.broadcast_to(_aligned(area, lead + [y_dim, x_dim]), lshape + (delp.sizes[y_dim], delp.sizes[x_dim]))
    outs, dco = core.pressure_coarsen([k4(ds[n]) for n in names], k4(delp), np.ascontiguousarray(w).reshape((-1,) + w.shape[-2:]),
                                      coarsening_factor, toa_pressure, mask_mode="pfull" if extrapolate else "phalf",
                                      mode=mode)
    blocks = {x_dim: coarsening_factor, y_dim: coarsening_factor}

    def back(v, like):
        v = _np_host(v)
        v = v.reshape(lshape + v.shape[-3:]).transpose([kdims.index(d) for d in like.dims])
        return _new(like, v, like.dims, _coarsened_coords(_plain_coords(like), blocks, coord_func), name=like.name,
                    attrs=like.attrs)

    variables = {}
    coords = {}
    for n, o in zip(names, outs):
        r = back(o, ds[n])
        variables[n] = (tuple(r.dims), np.asarray(r.values), dict(r.attrs))
        for ck, cv in _plain_coords(r).items():
            coords.setdefault(ck, cv)
    return type(ds)(variables, coords=coords or None, attrs=dict(ds.attrs)), back(dco, delp)
