"""oracle/vcm_ref.py -- TEST INFRASTRUCTURE ONLY (the checker, never the product).

numpy restatement of the array arithmetic of the reference's coarse-graining path.
Paths are relative to /root/reference/external/vcm/vcm/ unless noted.  xarray, dask
and xgcm are not installed in this image, so the labelled-array layer is dropped
and every function works on plain ndarrays with the horizontal axes LAST
(``[..., y, x]``) and, where a vertical axis exists, ``[..., z, y, x]`` -- the layout
of FV3 restart files (tests/_coarsen_restarts_regression_tests/schemas/*.json).

Third-party arithmetic restated here because it is not under /root/reference:
  * xarray 0.19.0 ``Coarsen.sum`` = reshape into windows + ``nansum`` (NaN replaced by
    0, then ``np.sum``) -- call sites cubedsphere/coarsen.py:211-212,260-261,839-840;
  * xarray ``Variable.cumsum`` = ``nancumsum`` -- calc/thermo/vertically_dependent.py:63;
  * xgcm 0.6.1 ``Grid.interp`` centre->outer with FV3_FACE_CONNECTIONS
    (cubedsphere/xgcm.py:7-34; call sites cubedsphere/regridz.py:126,
    cubedsphere/coarsen_restarts.py:847).
Pinned by the golden vectors in tests/golden/ (made by tests/golden/make_golden.py
from the reference's own regression JSONs) and the inline known answers of the
reference's unit tests (tests/test_oracle.py).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = None


def build(force=False):
    """Compile oracle/liboracle.so (gcc, -O2 -ffp-contract=off) if missing."""
    so = os.path.join(_HERE, "liboracle.so")
    srcs = [os.path.join(_HERE, s) for s in ("mappm_ref.c", "blocksum_ref.c")]
    stale = (not os.path.exists(so)) or any(
        os.path.getmtime(s) > os.path.getmtime(so) for s in srcs
    )
    if force or stale:
        subprocess.check_call(["make", "-C", _HERE, "-B", "liboracle.so"], stdout=subprocess.DEVNULL)
    return so


def lib():
    global _LIB
    if _LIB is None:
        _LIB = ctypes.CDLL(build())
        fp = ctypes.POINTER(ctypes.c_float)
        dp = ctypes.POINTER(ctypes.c_double)
        _LIB.mappm_ref.argtypes = [ctypes.c_int, ctypes.c_int, fp, fp, ctypes.c_int, fp, fp,
                                   ctypes.c_int, ctypes.c_int, ctypes.c_int]
        _LIB.mappm_ref.restype = ctypes.c_int
        _LIB.mappm_profile_ref.argtypes = [ctypes.c_int, ctypes.c_int, fp, fp, fp, fp, fp,
                                           ctypes.c_int, ctypes.c_int]
        _LIB.mappm_profile_ref.restype = ctypes.c_int
        _LIB.blocksum_ordered_f32.argtypes = [fp, ctypes.c_long] + [ctypes.c_int] * 4 + [fp]
        _LIB.blocksum_ordered_f64.argtypes = [dp, ctypes.c_long] + [ctypes.c_int] * 4 + [dp]
    return _LIB


def _fptr(a):
    return a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))


# --------------------------------------------------------------------------------------
# mappm  (external/mappm/mappm/mappm.f90:10; Python-visible f2py signature used at
# cubedsphere/regridz.py:332-334)
# --------------------------------------------------------------------------------------
def mappm(pe1, q1, pe2, i1=1, i2=None, iv=1, kord=1, ptop=0.0, nthreads=1):
    """q2[ncol,kn] float32.  f2py casts every input to float32 and coerces the scalar
    arguments with int() (external/vcm/tests/test_mappm.py:9-13 passes floats)."""
    pe1 = np.ascontiguousarray(pe1, dtype=np.float32)
    q1 = np.ascontiguousarray(q1, dtype=np.float32)
    pe2 = np.ascontiguousarray(pe2, dtype=np.float32)
    ncol, km1 = pe1.shape
    km, kn = km1 - 1, pe2.shape[1] - 1
    assert q1.shape == (ncol, km) and pe2.shape[0] == ncol
    q2 = np.empty((ncol, kn), dtype=np.float32)
    rc = lib().mappm_ref(ncol, km, _fptr(pe1), _fptr(q1), kn, _fptr(pe2), _fptr(q2),
                         int(iv), int(kord), int(nthreads))
    if rc != 0:
        raise ValueError(f"mappm_ref: unsupported km={km} for kord={kord}")
    return q2


def mappm_profile(pe1, q1, iv=1, kord=1):
    pe1 = np.ascontiguousarray(pe1, dtype=np.float32)
    q1 = np.ascontiguousarray(q1, dtype=np.float32)
    ncol, km = q1.shape
    out = [np.empty((ncol, km), dtype=np.float32) for _ in range(3)]
    rc = lib().mappm_profile_ref(ncol, km, _fptr(pe1), _fptr(q1), *[_fptr(o) for o in out],
                                 int(iv), int(kord))
    if rc != 0:
        raise ValueError("mappm_profile_ref failed")
    return out


def blocksum_ordered(a, fy, fx):
    """Explicit-order window sum (oracle/blocksum_ref.c); a is [..., ny, nx]."""
    a = np.ascontiguousarray(a)
    ny, nx = a.shape[-2:]
    out = np.empty(a.shape[:-2] + (ny // fy, nx // fx), dtype=a.dtype)
    nb = int(np.prod(a.shape[:-2], dtype=np.int64))
    if a.dtype == np.float32:
        lib().blocksum_ordered_f32(_fptr(a), nb, ny, nx, fy, fx, _fptr(out))
    elif a.dtype == np.float64:
        dp = ctypes.POINTER(ctypes.c_double)
        lib().blocksum_ordered_f64(a.ctypes.data_as(dp), nb, ny, nx, fy, fx, out.ctypes.data_as(dp))
    else:
        raise TypeError(a.dtype)
    return out


# --------------------------------------------------------------------------------------
# xarray coarsen().sum() and friends
# --------------------------------------------------------------------------------------
def coarsen_sum(a, fy, fx):
    """xarray 0.19 ``a.coarsen({y: fy, x: fx}).sum()`` on the last two axes:
    windows by reshape, NaN -> 0 (skipna default for floats), np.sum over the window
    axes.  fy or fx may be 1 (1-D coarsening as in coarsen.py:259-261)."""
    a = np.ascontiguousarray(a)
    ny, nx = a.shape[-2:]
    if ny % fy or nx % fx:
        raise ValueError("window does not divide the dimension (xarray boundary='exact')")
    r = a.reshape(a.shape[:-2] + (ny // fy, fy, nx // fx, fx))
    if np.issubdtype(r.dtype, np.floating):
        r = np.where(np.isnan(r), np.zeros((), dtype=r.dtype), r)
    return np.sum(r, axis=(-3, -1))


def coarsen_reduce(a, fy, fx, method):
    """xarray ``coarsen(...).<method>()`` for min/max/mean/sum (coarsen.py:838-840).
    NaN-skipping like xarray's defaults."""
    if method == "sum":
        return coarsen_sum(a, fy, fx)
    a = np.ascontiguousarray(a)
    ny, nx = a.shape[-2:]
    r = a.reshape(a.shape[:-2] + (ny // fy, fy, nx // fx, fx))
    fn = {"min": np.nanmin, "max": np.nanmax, "mean": np.nanmean, "median": np.nanmedian}[method]
    isfloat = np.issubdtype(r.dtype, np.floating)
    if not isfloat:
        fn = {"min": np.min, "max": np.max, "mean": np.mean, "median": np.median}[method]
    with np.errstate(all="ignore"):
        import warnings

        with warnings.catch_warnings():
            warnings.simplefilter("ignore")
            return fn(r, axis=(-3, -1))


def weighted_block_average(obj, weights, f):
    """cubedsphere/coarsen.py:183-218: nansum(obj*w)/nansum(w) over f x f windows of
    the last two axes.  ``weights`` broadcasts against ``obj`` from the right after
    optional leading-axis alignment by the caller (xarray aligns by name; here the
    caller passes arrays that numpy-broadcast, e.g. area[tile,1,1,y,x])."""
    prod = obj * weights
    num = coarsen_sum(prod, f, f)
    den = coarsen_sum(weights, f, f)
    with np.errstate(all="ignore"):
        return num / den


def edge_weighted_block_average(obj, spacing, f, edge="x"):
    """cubedsphere/coarsen.py:221-273 on [..., y, x] arrays: edge='x' sums along x and
    keeps every f-th y line; edge='y' the converse."""
    prod = spacing * obj
    if edge == "x":
        num, den = coarsen_sum(prod, 1, f), coarsen_sum(spacing, 1, f)
        with np.errstate(all="ignore"):
            res = num / den
        return res[..., ::f, :]
    elif edge == "y":
        num, den = coarsen_sum(prod, f, 1), coarsen_sum(spacing, f, 1)
        with np.errstate(all="ignore"):
            res = num / den
        return res[..., :, ::f]
    raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")


def block_edge_coarsen(obj, f, edge="x", method="sum"):
    """cubedsphere/coarsen.py:629-683."""
    if edge == "x":
        return coarsen_reduce(obj, 1, f, method)[..., ::f, :]
    elif edge == "y":
        return coarsen_reduce(obj, f, 1, method)[..., :, ::f]
    raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")


def block_reduce(image, block_size, func=np.sum, cval=0):
    """Restatement of cubedsphere/_skimage.py:125-202 (pad trailing edges with cval,
    block view, func over the block axes)."""
    image = np.asarray(image)
    if len(block_size) != image.ndim:
        raise ValueError("`block_size` must have the same length as `image.shape`.")
    pad_width = []
    for n, b in zip(image.shape, block_size):
        if b < 1:
            raise ValueError("Down-sampling factors must be >= 1.")
        pad_width.append((0, (b - n % b) % b))
    image = np.pad(image, pad_width=pad_width, mode="constant", constant_values=cval)
    shp = []
    for n, b in zip(image.shape, block_size):
        shp += [n // b, b]
    blocked = image.reshape(shp)
    # block view has all coarse axes first, then the block axes
    nd = image.ndim
    blocked = blocked.transpose([2 * i for i in range(nd)] + [2 * i + 1 for i in range(nd)])
    return func(blocked, axis=tuple(range(nd, 2 * nd)))


def horizontal_block_reduce(obj, f, func, cval=np.nan):
    """cubedsphere/coarsen.py:520-554 on the last two axes."""
    bs = (1,) * (obj.ndim - 2) + (f, f)
    return block_reduce(obj, bs, func, cval)


def mode_reduce(arr, axis, nan_policy="propagate"):
    """cubedsphere/coarsen.py:686-747 (_mode/_ureduce/_mode_reduce): scipy.stats.mode
    over a tuple of axes; ties -> smallest value."""
    import scipy.stats

    arr = np.asanyarray(arr)
    nd = arr.ndim
    keep = sorted(set(range(nd)) - set(axis))
    arr = np.moveaxis(arr, keep, range(len(keep)))
    arr = arr.reshape(arr.shape[: len(keep)] + (-1,))
    res = scipy.stats.mode(arr, axis=-1, nan_policy=nan_policy, keepdims=True).mode
    return np.squeeze(np.asarray(res), -1)


def repeat_staggered(a, f, axis):
    """cubedsphere/coarsen.py:843-866: odd-size axis = interfaces, last not repeated."""
    n = a.shape[axis]
    if n % 2 == 1:
        sl = [slice(None)] * a.ndim
        sl[axis] = slice(None, -1)
        sl2 = [slice(None)] * a.ndim
        sl2[axis] = slice(-1, None)
        return np.concatenate([np.repeat(a[tuple(sl)], f, axis=axis), a[tuple(sl2)]], axis=axis)
    return np.repeat(a, f, axis=axis)


def block_upsample(a, f, axes=(-2, -1)):
    """cubedsphere/coarsen.py:869-897."""
    for ax in axes:
        a = repeat_staggered(a, f, ax)
    return a


def block_upsample_like(a, ref_shape_yx):
    """cubedsphere/coarsen.py:900-938: factor inferred from the x axis."""
    nx_c, nx_f = a.shape[-1], ref_shape_yx[1]
    f = (nx_f - 1) // (nx_c - 1) if nx_c % 2 == 1 else nx_f // nx_c
    return block_upsample(a, f, (-1, -2))


def coarsen_coords(coord, f):
    """cubedsphere/coarsen.py:83-132: ((c[::f]-1)//f+1) as float32."""
    c = np.asarray(coord)
    return ((c[::f] - 1) // f + 1).astype(int).astype(np.float32)


# --------------------------------------------------------------------------------------
# pressure helpers (calc/thermo/vertically_dependent.py)
# --------------------------------------------------------------------------------------
def pressure_at_interface(delp, toa_pressure=300.0, z_axis=-3):
    """:41-66 -- concat(top, delp) then nancumsum along z, in delp's dtype."""
    delp = np.asarray(delp)
    top_shape = list(delp.shape)
    top_shape[z_axis] = 1
    top = np.full(top_shape, toa_pressure, dtype=delp.dtype)
    return np.nancumsum(np.concatenate([top, delp], axis=z_axis), axis=z_axis)


def pressure_at_midpoint_log(delp, toa_pressure=300.0, z_axis=-3):
    """:153-179 -- delp / diff(log(p_interface))."""
    pi = pressure_at_interface(delp, toa_pressure, z_axis)
    with np.errstate(all="ignore"):
        dlogp = np.diff(np.log(pi), axis=z_axis)
        return delp / dlogp


def surface_pressure_from_delp(delp, p_toa=300.0, z_axis=-3):
    """:189-208 -- NaN-skipping sum + p_toa."""
    return np.nansum(delp, axis=z_axis) + p_toa


# --------------------------------------------------------------------------------------
# vertical regridding (cubedsphere/regridz.py)
# --------------------------------------------------------------------------------------
def regrid_vertical(p_in, f_in, p_out, iv=1, kord=1, z_axis=-3, nthreads=1):
    """:223-338 -- columns flattened z-last, mappm (fp32), shape/dim order restored."""
    if p_in.shape[z_axis] - 1 != f_in.shape[z_axis]:
        raise ValueError("f_in must have a vertical dimension one shorter than p_in")
    pz = np.moveaxis(p_in, z_axis, -1)
    fz = np.moveaxis(f_in, z_axis, -1)
    oz = np.moveaxis(p_out, z_axis, -1)
    if pz.shape[:-1] != fz.shape[:-1] or pz.shape[:-1] != oz.shape[:-1]:
        raise ValueError("All dimensions except vertical must be same size for p_in, f_in and p_out")
    q2 = mappm(pz.reshape(-1, pz.shape[-1]), fz.reshape(-1, fz.shape[-1]),
               oz.reshape(-1, oz.shape[-1]), 1, None, iv, kord, 0.0, nthreads=nthreads)
    return np.moveaxis(q2.reshape(oz.shape[:-1] + (oz.shape[-1] - 1,)), -1, z_axis)


def mask_weights(weights, pfull_coarse_on_fine, phalf_coarse_on_fine, phalf_fine,
                 extrapolate=False, z_axis=-3):
    """:200-220.  ``weights`` must broadcast against [..., z, y, x].  The result keeps
    the weights' dtype (xarray ``where(cond, other=0.0)`` with a Python scalar)."""
    surf = np.take(phalf_fine, [-1], axis=z_axis)
    if extrapolate:
        cond = pfull_coarse_on_fine < surf
    else:
        sl = [slice(None)] * phalf_coarse_on_fine.ndim
        sl[z_axis] = slice(1, None)
        cond = phalf_coarse_on_fine[tuple(sl)] < surf
    w = np.broadcast_to(weights, cond.shape)
    return np.where(cond, w, np.zeros((), dtype=w.dtype))


def _regrid_given_delp(fields, delp_fine, delp_coarse, weights, toa_pressure, extrapolate=False,
                       iv=1, kord=1, nthreads=1):
    """:149-197.  fields: dict name -> [..., z, y, x]; weights [..., 1, y, x]-broadcastable."""
    delp_c_up = block_upsample_like(delp_coarse, delp_fine.shape[-2:])
    phalf_c_up = pressure_at_interface(delp_c_up, toa_pressure)
    phalf_fine = pressure_at_interface(delp_fine, toa_pressure)
    out = {}
    for name, f in fields.items():
        out[name] = regrid_vertical(phalf_fine, f, phalf_c_up, iv=iv, kord=kord, nthreads=nthreads)
    pfull_c_up = pressure_at_midpoint_log(delp_c_up, toa_pressure)
    masked = mask_weights(weights, pfull_c_up, phalf_c_up, phalf_fine, extrapolate=extrapolate)
    return out, masked


def regrid_to_area_weighted_pressure(fields, delp, area, toa_pressure, f, extrapolate=False,
                                     nthreads=1):
    """:31-78.  area must broadcast to delp ([tile,1,1,y,x] for [tile,Time,z,y,x])."""
    delp_coarse = weighted_block_average(delp, area, f)
    return _regrid_given_delp(fields, delp, delp_coarse, area, toa_pressure, extrapolate,
                              nthreads=nthreads)


# xgcm face connections, cubedsphere/xgcm.py:7-34: tile -> axis -> (left, right), each
# (neighbour tile, neighbour axis)
FV3_FACE_CONNECTIONS = {
    0: {"x": ((4, "y"), (1, "x")), "y": ((5, "y"), (2, "x"))},
    1: {"x": ((0, "x"), (3, "y")), "y": ((5, "x"), (2, "y"))},
    2: {"x": ((0, "y"), (3, "x")), "y": ((1, "y"), (4, "x"))},
    3: {"x": ((2, "x"), (5, "y")), "y": ((1, "x"), (4, "y"))},
    4: {"x": ((2, "y"), (5, "x")), "y": ((3, "y"), (0, "x"))},
    5: {"x": ((4, "x"), (1, "y")), "y": ((3, "x"), (0, "y"))},
}


def _edge_line(tile_arr, axis, last):
    """Boundary line of a [..., y, x] tile: along 'x' -> the first/last column (indexed
    by y); along 'y' -> the first/last row (indexed by x)."""
    idx = -1 if last else 0
    return tile_arr[..., :, idx] if axis == "x" else tile_arr[..., idx, :]


def stagger_delp(delp, axis):
    """xgcm ``grid.interp(delp, axis)`` centre -> outer for 6 tiles on axis 0
    (regridz.py:118-132).  out[j] = 0.5*(ext[j-1]+ext[j]); ghosts come from the
    neighbour tiles: left ghost = neighbour's LAST line along its connecting axis,
    right ghost = its FIRST line; reversed when the neighbour's axis differs."""
    assert delp.shape[0] == 6
    outs = []
    for t in range(6):
        (lt, lax), (rt, rax) = FV3_FACE_CONNECTIONS[t][axis]
        left = _edge_line(delp[lt], lax, last=True)
        right = _edge_line(delp[rt], rax, last=False)
        if lax != axis:
            left = left[..., ::-1]
        if rax != axis:
            right = right[..., ::-1]
        if axis == "x":
            ext = np.concatenate([left[..., :, None], delp[t], right[..., :, None]], axis=-1)
            outs.append(0.5 * (ext[..., :, :-1] + ext[..., :, 1:]))
        else:
            ext = np.concatenate([left[..., None, :], delp[t], right[..., None, :]], axis=-2)
            outs.append(0.5 * (ext[..., :-1, :] + ext[..., 1:, :]))
    return np.stack(outs, axis=0)


def regrid_to_edge_weighted_pressure(fields, delp, length, toa_pressure, f, edge="x",
                                     extrapolate=False, nthreads=1):
    """:81-146.  edge='x' (u, dx on y-interfaces) staggers delp along y."""
    interp_axis = "x" if edge == "y" else "y"
    delp_stag = stagger_delp(delp, interp_axis)
    delp_stag_coarse = edge_weighted_block_average(delp_stag, length, f, edge=edge)
    return _regrid_given_delp(fields, delp_stag, delp_stag_coarse, length, toa_pressure,
                              extrapolate, nthreads=nthreads)


# --------------------------------------------------------------------------------------
# restart recipes (cubedsphere/coarsen_restarts.py).  ds: dict name -> array laid out
# [tile, Time, z, y, x] (3-D) or [tile, Time, y, x] (2-D); area/dx/dy: [tile, y, x].
# --------------------------------------------------------------------------------------
def _w5(w):  # [tile,y,x] -> [tile,1,1,y,x]
    return w[:, None, None, :, :]


def _w4(w):  # [tile,y,x] -> [tile,1,y,x]
    return w[:, None, :, :]


def _wfor(a, w):
    return _w5(w) if a.ndim == 5 else _w4(w)


def coarse_grain_fv_core(ds, delp, area, dx, dy, f, coarsen_agrid_winds=False, mass_weighted=True):
    """:335-427."""
    if mass_weighted:
        area_vars, mass_vars = ["phis", "delp", "DZ"], ["W", "T"]
    else:
        area_vars, mass_vars = ["phis", "delp", "DZ", "W", "T"], []
    if coarsen_agrid_winds:
        if not ("ua" in ds and "va" in ds):
            raise ValueError("If 'coarsen_agrid_winds' is active, 'ua' and 'va' "
                             "must be present in the 'fv_core.res' restart files.")
        (mass_vars if mass_weighted else area_vars).extend(["ua", "va"])
    out = {}
    for v in area_vars:
        out[v] = weighted_block_average(ds[v], _wfor(ds[v], area), f)
    mass = delp * _w5(area)
    for v in mass_vars:
        out[v] = weighted_block_average(ds[v], mass, f)
    out["u"] = edge_weighted_block_average(ds["u"], _w5(dx), f, edge="x")
    out["v"] = edge_weighted_block_average(ds["v"], _w5(dy), f, edge="y")
    return out


def coarse_grain_fv_core_on_pressure(ds, delp, area, dx, dy, toa_pressure, f,
                                     coarsen_agrid_winds=False, extrapolate=False, nthreads=1):
    """:430-556."""
    area_vars, masked_vars = ["phis", "delp", "DZ"], ["W", "T"]
    if coarsen_agrid_winds:
        if not ("ua" in ds and "va" in ds):
            raise ValueError("If 'coarsen_agrid_winds' is active, 'ua' and 'va' "
                             "must be present in the 'fv_core.res' restart files.")
        masked_vars.extend(["ua", "va"])
    reg, masked_area = regrid_to_area_weighted_pressure(
        {v: ds[v] for v in masked_vars}, delp, _w5(area), toa_pressure, f, extrapolate, nthreads)
    reg_u, masked_dx = regrid_to_edge_weighted_pressure(
        {"u": ds["u"]}, delp, _w5(dx), toa_pressure, f, "x", extrapolate, nthreads)
    reg_v, masked_dy = regrid_to_edge_weighted_pressure(
        {"v": ds["v"]}, delp, _w5(dy), toa_pressure, f, "y", extrapolate, nthreads)
    out = {}
    for v in area_vars:
        out[v] = weighted_block_average(ds[v], _wfor(ds[v], area), f)
    for v in masked_vars:
        out[v] = weighted_block_average(reg[v], masked_area, f)
    out["u"] = edge_weighted_block_average(reg_u["u"], masked_dx, f, edge="x")
    out["v"] = edge_weighted_block_average(reg_v["v"], masked_dy, f, edge="y")
    return out


FRACTION_TRACERS = ["cld_amt"]
NON_FRACTION_TRACERS = ["sphum", "liq_wat", "rainwat", "ice_wat", "snowwat", "graupel", "o3mr",
                        "sgs_tke"]


def coarse_grain_fv_tracer(ds, delp, area, f, mass_weighted=True):
    """:856-900."""
    if mass_weighted:
        area_vars, mass_vars = FRACTION_TRACERS, NON_FRACTION_TRACERS
    else:
        area_vars, mass_vars = FRACTION_TRACERS + NON_FRACTION_TRACERS, []
    out = {}
    for v in area_vars:
        out[v] = weighted_block_average(ds[v], _w5(area), f)
    mass = delp * _w5(area)
    for v in mass_vars:
        out[v] = weighted_block_average(ds[v], mass, f)
    return out


def coarse_grain_fv_tracer_on_pressure(ds, delp, area, toa_pressure, f, extrapolate=False,
                                       nthreads=1):
    """:903-961."""
    reg, masked_area = regrid_to_area_weighted_pressure(ds, delp, _w5(area), toa_pressure, f,
                                                        extrapolate, nthreads)
    return {v: weighted_block_average(reg[v], masked_area, f) for v in ds}


def coarse_grain_fv_srf_wnd(ds, area, f):
    """:964-987."""
    return {v: weighted_block_average(ds[v], _w4(area), f) for v in ("u_srf", "v_srf")}


# --------------------------------------------------------------------------------------
# hydrostatic post-processing and the blended method
# --------------------------------------------------------------------------------------
GRAVITY, RDGAS, RVGAS = 9.80665, 287.05, 461.5  # calc/thermo/constants.py:2-4
SIGMA_BLEND = 0.9                               # cubedsphere/coarsen_restarts.py:54


def height_at_interface(dz, phis, z_axis=-3):
    """calc/thermo/vertically_dependent.py:69-98: reversed nancumsum of [-dz..., phis/g]."""
    bottom = np.expand_dims(phis, z_axis) / GRAVITY
    dz_with_bottom = np.concatenate([-dz, np.broadcast_to(bottom, np.take(dz, [0], axis=z_axis).shape)], axis=z_axis)
    rev = np.flip(dz_with_bottom, axis=z_axis)
    return np.flip(np.nancumsum(rev, axis=z_axis), axis=z_axis)


def hydrostatic_dz(T, q, delp, toa_pressure, z_axis=-3):
    """:210-235."""
    pi = pressure_at_interface(delp, toa_pressure, z_axis)
    tv = T * (1 + (RVGAS / RDGAS - 1) * q)
    with np.errstate(all="ignore"):
        dlogp = np.diff(np.log(pi), axis=z_axis)
    return -dlogp * RDGAS * tv / GRAVITY


def dz_and_top_to_phis(top_height, dz, z_axis=-3):
    """:182-186."""
    return GRAVITY * (top_height + np.nansum(dz, axis=z_axis))


def impose_hydrostatic_balance(core, tracer, toa_pressure, z_axis=-3):
    """cubedsphere/coarsen_restarts.py:990-1017 -> core with DZ, phis replaced."""
    height = height_at_interface(core["DZ"], core["phis"], z_axis)
    height_top = np.take(height, 0, axis=z_axis)
    dz = hydrostatic_dz(core["T"], tracer["sphum"], core["delp"], toa_pressure, z_axis)
    out = dict(core)
    out["DZ"] = dz
    out["phis"] = dz_and_top_to_phis(height_top, dz, z_axis)
    return out


def compute_blending_weights(blending_pressure, ps_coarse, pfull_coarse):
    """coarsen_restarts.py:559-577 (z axis at -3)."""
    ps = np.expand_dims(ps_coarse, -3)
    pb = np.expand_dims(blending_pressure, -3)
    with np.errstate(all="ignore"):
        w = (ps - pfull_coarse) / (ps - pb)
    return np.where(pfull_coarse > pb, w, 1.0)


def blending_weights_agrid(delp, area, toa_pressure, f):
    """:580-621."""
    delp_coarse = weighted_block_average(delp, _wfor(delp, area) if area.ndim == 3 else area, f)
    pfull_coarse = pressure_at_midpoint_log(delp_coarse, toa_pressure)
    ps = surface_pressure_from_delp(delp, toa_pressure)
    ps_coarse = surface_pressure_from_delp(delp_coarse, toa_pressure)
    pb = SIGMA_BLEND * coarsen_reduce(ps, f, f, "min")
    return compute_blending_weights(pb, ps_coarse, pfull_coarse)


def blending_weights_dgrid(delp, length, toa_pressure, f, edge):
    """:624-661."""
    delp_edge = stagger_delp(delp, "y" if edge == "x" else "x")
    delp_edge_coarse = edge_weighted_block_average(delp_edge, _wfor(delp_edge, length), f, edge=edge)
    pfull_coarse = pressure_at_midpoint_log(delp_edge_coarse, toa_pressure)
    ps = surface_pressure_from_delp(delp_edge, toa_pressure)
    ps_coarse = surface_pressure_from_delp(delp_edge_coarse, toa_pressure)
    pb = SIGMA_BLEND * block_edge_coarsen(ps, f, edge=edge, method="min")
    return compute_blending_weights(pb, ps_coarse, pfull_coarse)


def blend(weights, pressure_level, model_level):
    """:664-680."""
    return weights * pressure_level + (1 - weights) * model_level


def coarse_grain_fv_core_via_blended_method(ds, delp, area, dx, dy, toa_pressure, f, coarsen_agrid_winds=False,
                                            mass_weighted=True, nthreads=1):
    """:700-779."""
    pl = coarse_grain_fv_core_on_pressure(ds, delp, area, dx, dy, toa_pressure, f, coarsen_agrid_winds,
                                          nthreads=nthreads)
    ml = coarse_grain_fv_core(ds, delp, area, dx, dy, f, coarsen_agrid_winds, mass_weighted)
    wa = blending_weights_agrid(delp, area, toa_pressure, f)
    wu = blending_weights_dgrid(delp, dx, toa_pressure, f, "x")
    wv = blending_weights_dgrid(delp, dy, toa_pressure, f, "y")
    out = {}
    ignore = ["u", "v"] + ([] if coarsen_agrid_winds else ["ua", "va"])
    for v in ds:
        if v in ignore:
            continue
        if ds[v].ndim == delp.ndim:
            out[v] = blend(wa, pl[v], ml[v])
        else:
            out[v] = ml[v]
    out["u"] = blend(wu, pl["u"], ml["u"])
    out["v"] = blend(wv, pl["v"], ml["v"])
    return out


def coarse_grain_fv_tracer_via_blended_method(ds, delp, area, toa_pressure, f, mass_weighted=True, nthreads=1):
    """:782-822."""
    pl = coarse_grain_fv_tracer_on_pressure(ds, delp, area, toa_pressure, f, nthreads=nthreads)
    ml = coarse_grain_fv_tracer(ds, delp, area, f, mass_weighted)
    w = blending_weights_agrid(delp, area, toa_pressure, f)
    return {v: blend(w, pl[v], ml[v]) for v in ds}


# --------------------------------------------------------------------------------------
# sfc_data, 'complex' method (cubedsphere/coarsen_restarts.py:1111-1470)
# ds: name -> [tile, Time, y, x] (or [tile, Time, z, y, x] for stc/smc/slc); area [tile, y, x]
# --------------------------------------------------------------------------------------
FREEZING_TEMPERATURE, SHDMIN_THRESHOLD, STYPE_LAND_ICE, VTYPE_LAND_ICE = 273.16, 0.011, 16.0, 15.0  # :48-51


def _block_mode_omit(a, f):
    red = lambda arr, axis: mode_reduce(arr, axis, nan_policy="omit")  # noqa: E731
    return horizontal_block_reduce(a, f, red)


def _where(a, mask):
    return np.where(mask, a, np.nan)


def _wb(a, w, f):
    """weighted_block_average with weights [tile, (Time,) y, x] broadcast over a's extra z axis."""
    if a.ndim == w.ndim + 1:
        w = np.expand_dims(w, -3)
    with np.errstate(all="ignore"):
        return weighted_block_average(a, w, f)


def sfc_precompute(ds, f):
    """:1162-1206."""
    slmsk_c = _block_mode_omit(ds["slmsk"], f)
    dom_sfc = np.isclose(ds["slmsk"], block_upsample(slmsk_c, f))
    vtype_c = _block_mode_omit(_where(ds["vtype"], dom_sfc), f)
    stype_c = _block_mode_omit(_where(ds["stype"], dom_sfc), f)
    dom_v = np.isclose(ds["vtype"], block_upsample(vtype_c, f))
    dom_s = np.isclose(ds["stype"], block_upsample(stype_c, f))
    return slmsk_c, vtype_c, stype_c, dom_sfc, dom_v, dom_s


SFC_AREA = ["tsea", "alvsf", "alvwf", "alnsf", "alnwf", "facsf", "facwf", "f10m", "t2m", "q2m", "uustar", "ffmm",
            "ffhh", "tprcp", "snwdph"]
SFC_DOMINANT = ["tg3", "vfrac", "fice", "sncovr", "shdmin", "shdmax", "snoalb"]
SFC_VTYPE = ["canopy", "zorl"]
SFC_STYPE = ["smc", "slc", "stc"]


def coarse_grain_sfc_data_complex(ds, area, f):
    """:1111-1159 with the method table :1360-1393 and the chgres corrections :1396-1470."""
    slmsk_c, vtype_c, stype_c, dom_sfc, dom_v, dom_s = sfc_precompute(ds, f)
    a = np.broadcast_to(area[:, None], ds["slmsk"].shape).astype(area.dtype)
    out = {"slmsk": slmsk_c, "vtype": vtype_c, "stype": stype_c}

    def m3(mask, v):  # a 2-D mask against a possibly 3-D variable
        return np.expand_dims(mask, -3) if v.ndim == mask.ndim + 1 else mask

    for name, v in ds.items():
        if name in out:
            continue
        if name in SFC_AREA:
            r = _wb(v, a, f)
        elif name in SFC_DOMINANT:
            r = _wb(_where(v, dom_sfc), _where(a, dom_sfc), f)
        elif name in SFC_VTYPE:
            mask = dom_sfc & dom_v
            aw = _wb(_where(v, mask), _where(a, mask), f)
            avw = _wb(_where(v, mask), _where(a * ds["vfrac"], mask), f)
            tot = coarsen_sum(_where(a * ds["vfrac"], mask), f, f)
            r = np.where(tot > 0.0, avw, aw)
        elif name in SFC_STYPE:
            mask = dom_sfc & dom_s
            r = _wb(_where(v, m3(mask, v)), _where(a, mask), f)
        elif name == "srflag":
            r = _block_mode_omit(v, f)
        elif name == "slope":
            r = _block_mode_omit(_where(v, dom_sfc), f)
        elif name in ("sheleg", "hice"):
            r = _wb(v, a * ds["sncovr" if name == "sheleg" else "fice"], f)
            r = np.where(np.isnan(r), 0.0, r)
        elif name == "tisfc":
            sea_ice = _wb(_where(v, dom_sfc), _where(a * ds["fice"], dom_sfc), f)
            other = _wb(_where(v, dom_sfc), _where(a, dom_sfc), f)
            r = np.where(np.isclose(slmsk_c, 2.0), sea_ice, other)
        else:
            raise KeyError(f"no sfc_data coarsening method for {name!r}")
        out[name] = r
    # surface_chgres corrections, in order
    land_ice = np.isclose(out["vtype"], VTYPE_LAND_ICE)
    for k in ("tsea", "tg3"):
        clipped = np.where(out[k] < FREEZING_TEMPERATURE, out[k], FREEZING_TEMPERATURE)
        out[k] = np.where(land_ice, clipped, out[k])
    out["stype"] = np.where(land_ice, STYPE_LAND_ICE, out["stype"])
    out["canopy"] = np.where(out["shdmin"] < SHDMIN_THRESHOLD, 0.0, out["canopy"])
    out["shdmin"] = np.where(land_ice, 0.0, out["shdmin"])
    return {k: (v.astype(np.float32) if v.dtype == np.float64 else v) for k, v in out.items()}  # _doubles_to_floats


def sfc_regression_inputs():
    """The sfc_data inputs of test_coarsen_restarts.py (schemas/sfc_data-schema.json): float64,
    [6, 1, 4, 4] (stc/smc/slc [6, 1, 4, 4, 4]), every array drawn with np.random.seed(0)."""
    names2 = ["slmsk", "tsea", "sheleg", "tg3", "zorl", "alvsf", "alvwf", "alnsf", "alnwf", "facsf", "facwf", "vfrac",
              "canopy", "f10m", "t2m", "q2m", "vtype", "stype", "uustar", "ffmm", "ffhh", "hice", "fice", "tisfc",
              "tprcp", "srflag", "snwdph", "shdmin", "shdmax", "slope", "snoalb", "sncovr"]
    ds = {n: synth_array((6, 1, 4, 4), np.float64) for n in names2}
    for n in ("stc", "smc", "slc"):
        ds[n] = synth_array((6, 1, 4, 4, 4), np.float64)
    return ds


# --------------------------------------------------------------------------------------
# fine_res_budget variant (workflows/fine_res_budget/budget/budgets.py:49-77)
# --------------------------------------------------------------------------------------
def pressure_level_average(delp_fine, delp_coarse, area, field, f, nthreads=1):
    """budgets.py:49-57: interfaces of the COARSE delp upsampled, remap, UNMASKED
    area-weighted block average.  [..., z, y, x] arrays; area broadcastable."""
    pi = pressure_at_interface(delp_fine)
    pi_c_up = block_upsample(pressure_at_interface(delp_coarse), f, (-1, -2))
    fg = regrid_vertical(pi, field, pi_c_up, nthreads=nthreads)
    return weighted_block_average(fg, area, f)


def area_above_fine_surface(delp_fine, delp_coarse, area, f):
    """budgets.py:62-77."""
    pressure_c_up = block_upsample(pressure_at_midpoint_log(delp_coarse), f, (-1, -2))
    ps = surface_pressure_from_delp(delp_fine)
    ps = np.expand_dims(ps, -3)
    w = np.where(pressure_c_up <= ps, np.broadcast_to(area, pressure_c_up.shape),
                 np.zeros((), dtype=np.asarray(area).dtype))
    return coarsen_sum(w, f, f)


def compute_recoarsened_budget_inputs(begin, middle, end, dt=900, f=8, first_moments=("T", "sphum", "vulcan_omega_coarse"),
                                      second_moments=(("T", "vulcan_omega_coarse"), ("sphum", "vulcan_omega_coarse")),
                                      storage_terms=("T", "sphum"), nthreads=1):
    """budgets.py:217-294 on name -> [tile, z, y, x] mappings (the three ``step`` slices)."""
    area, delp = middle["area_coarse"], middle["delp"]
    delp_coarse = weighted_block_average(delp, area[:, None], f)
    raw = {v: middle[v] for v in first_moments}
    for a, b in second_moments:
        raw[f"{a}_{b}"] = middle[a] * middle[b]
    for v in storage_terms:
        raw[f"{v}_storage"] = (end[v] - begin[v]) / dt
    out = {}
    for k, v in raw.items():
        if v.ndim == delp.ndim:
            out[k] = pressure_level_average(delp, delp_coarse, area[:, None], v, f, nthreads)
        else:
            out[k] = weighted_block_average(v, area, f)
    out["delp"] = delp_coarse
    out["exposed_area"] = area_above_fine_surface(delp, delp_coarse, middle["exposed_area_coarse"][:, None], f)
    out["area"] = coarsen_sum(area, f, f)
    return out


# --------------------------------------------------------------------------------------
# synthetic inputs of the reference regression test
# (external/synth/synth/core.py:59-67,136; external/vcm/tests/test_coarsen_restarts.py:24-30)
# --------------------------------------------------------------------------------------
def synth_array(shape, dtype, lo=-1000.0, hi=1000.0):
    np.random.seed(0)
    return np.random.uniform(low=lo, high=hi, size=shape).astype(dtype)


def regression_inputs():
    """Regenerates the C4 (6 tiles, 7 levels) inputs of test_coarsen_restarts.py."""
    f8, f4 = np.float64, np.float32
    core = {
        "u": synth_array((6, 1, 7, 5, 4), f8), "v": synth_array((6, 1, 7, 4, 5), f8),
        "W": synth_array((6, 1, 7, 4, 4), f8), "DZ": synth_array((6, 1, 7, 4, 4), f8),
        "T": synth_array((6, 1, 7, 4, 4), f8), "delp": synth_array((6, 1, 7, 4, 4), f8, 3, 5),
        "phis": synth_array((6, 1, 4, 4), f8), "ua": synth_array((6, 1, 7, 4, 4), f8),
        "va": synth_array((6, 1, 7, 4, 4), f8),
    }
    tracer = {n: synth_array((6, 1, 7, 4, 4), f8) for n in NON_FRACTION_TRACERS + FRACTION_TRACERS}
    srf = {"u_srf": synth_array((6, 1, 4, 4), f8), "v_srf": synth_array((6, 1, 4, 4), f8)}
    grid = {"area": synth_array((6, 4, 4), f4, 0.5, 1), "dx": synth_array((6, 5, 4), f4, 0.5, 1),
            "dy": synth_array((6, 4, 5), f4, 0.5, 1)}
    return core, tracer, srf, grid
