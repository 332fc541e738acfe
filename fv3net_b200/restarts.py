"""Restart coarse-graining recipes on device buffers: the variable -> method tables of
vcm.cubedsphere.coarsen_restarts (external/vcm/vcm/cubedsphere/coarsen_restarts.py) re-expressed
as a few batched kernel launches instead of ~20 separate xarray calls per timestep.

Each ``ds`` is a mapping name -> array laid out like FV3 restart files:
3-D variables ``[tile, (Time,) z, y, x]``, 2-D variables ``[tile, (Time,) y, x]``; grid variables
``area[tile, y, x]``, ``dx[tile, y+1, x]``, ``dy[tile, y, x+1]``.  numpy in -> numpy out, CUDA
tensors in -> CUDA tensors out.  Outputs keep the inputs' axis order
(_sync_dimension_order, coarsen_restarts.py:1078-1081).
"""
import torch

from . import core
from .core import EXACT

FRACTION_TRACERS = ["cld_amt"]  # coarsen_restarts.py:55
NON_FRACTION_TRACERS = ["sphum", "liq_wat", "rainwat", "ice_wat", "snowwat", "graupel", "o3mr", "sgs_tke"]
_AGRID_ERR = ("If 'coarsen_agrid_winds' is active, 'ua' and 'va' "
              "must be present in the 'fv_core.res' restart files.")


def _has_time(a, ndim_without_time):
    return a.ndim == ndim_without_time + 1


def _squeeze_time(a, ndim_without_time):
    """[tile, Time, ...] -> ([tile*Time, ...] view, Time) ; arrays without Time pass through."""
    if _has_time(a, ndim_without_time):
        t = a.shape[1]
        return a.reshape((a.shape[0] * t,) + tuple(a.shape[2:])), t
    return a, None


def _restore_time(a, t, ntile):
    if t is None:
        return a
    return a.reshape((ntile, t) + tuple(a.shape[1:]))


def _rep(w, t):
    """Grid weights [tile, y, x] repeated per Time entry so they align with [tile*Time, ...]."""
    if t is None or t == 1:
        return w
    if isinstance(w, torch.Tensor):
        return w.repeat_interleave(t, dim=0)
    import numpy as np

    return np.repeat(w, t, axis=0)


def _wba(ds, names, weights, f):
    """Area-weighted block average of ds[names]; weights [tile, y, x] broadcast over the axes
    between the tile axis and (y, x)."""
    out = {}
    for v in names:
        a = ds[v]
        w = weights
        for _ in range(a.ndim - 3):
            w = w[:, None]
        out[v] = core.weighted_block_average(a, w, f)
    return out


def _upload(*objs):
    """Moves mappings / arrays to the device once; returns (converted objects, inputs_were_numpy)."""
    was_np = False
    res = []
    for o in objs:
        if isinstance(o, dict):
            d = {}
            for k, a in o.items():
                t, n = core._to_dev(a)
                was_np |= n
                d[k] = t
            res.append(d)
        else:
            t, n = core._to_dev(o)
            was_np |= n
            res.append(t)
    return res, was_np


def _download(out, was_np):
    if not was_np:
        return out
    return {k: v.cpu().numpy() for k, v in out.items()}


def coarse_grain_fv_core(ds, delp, area, dx, dy, coarsening_factor, coarsen_agrid_winds=False,
                         mass_weighted=True):
    """Model-level fv_core coarse-graining (coarsen_restarts.py:335-427)."""
    f = coarsening_factor
    (ds, delp, area, dx, dy), was_np = _upload(dict(ds), delp, area, dx, dy)
    if mass_weighted:
        area_vars, mass_vars = ["phis", "delp", "DZ"], ["W", "T"]
    else:
        area_vars, mass_vars = ["phis", "delp", "DZ", "W", "T"], []
    if coarsen_agrid_winds:
        if not ("ua" in ds and "va" in ds):
            raise ValueError(_AGRID_ERR)
        (mass_vars if mass_weighted else area_vars).extend(["ua", "va"])
    out = _wba(ds, area_vars, area, f)
    nlead = delp.ndim - 2
    a5 = area
    for _ in range(nlead - 1):
        a5 = a5[:, None]
    for v in mass_vars:  # mass = delp * area formed inside the kernel (:400)
        out[v] = core.weighted_block_average(ds[v], a5, f, weights2=delp)
    d5, e5 = dx, dy
    for _ in range(ds["u"].ndim - 3):
        d5, e5 = d5[:, None], e5[:, None]
    out["u"] = core.edge_weighted_block_average(ds["u"], d5, f, edge="x")
    out["v"] = core.edge_weighted_block_average(ds["v"], e5, f, edge="y")
    return _download(out, was_np)


def coarse_grain_fv_tracer(ds, delp, area, coarsening_factor, mass_weighted=True):
    """Model-level fv_tracer coarse-graining (coarsen_restarts.py:856-900)."""
    f = coarsening_factor
    (ds, delp, area), was_np = _upload(dict(ds), delp, area)
    if mass_weighted:
        area_vars, mass_vars = FRACTION_TRACERS, NON_FRACTION_TRACERS
    else:
        area_vars, mass_vars = FRACTION_TRACERS + NON_FRACTION_TRACERS, []
    area_vars = [v for v in area_vars if v in ds]
    mass_vars = [v for v in mass_vars if v in ds]
    out = _wba(ds, area_vars, area, f)
    a5 = area
    for _ in range(delp.ndim - 3):
        a5 = a5[:, None]
    for v in mass_vars:
        out[v] = core.weighted_block_average(ds[v], a5, f, weights2=delp)
    return _download(out, was_np)


def coarse_grain_fv_srf_wnd(ds, area, coarsening_factor):
    """coarsen_restarts.py:964-987."""
    (ds, area), was_np = _upload(dict(ds), area)
    return _download(_wba(ds, ["u_srf", "v_srf"], area, coarsening_factor), was_np)


def _pressure_fields(ds, names, delp, area, toa_pressure, f, mask_mode, mode):
    """Masked-area pressure-level average of ds[names] (A-grid); also returns coarse delp."""
    d4, t = _squeeze_time(delp, 4)
    ntile = delp.shape[0]
    fields = [_squeeze_time(ds[v], 4)[0] for v in names]
    outs, dco = core.pressure_coarsen(fields, d4, _rep(area, t), f, toa_pressure, mask_mode=mask_mode, mode=mode)
    return {v: _restore_time(o, t, ntile) for v, o in zip(names, outs)}, _restore_time(dco, t, ntile)


def _edge_pressure(u, delp, length, toa_pressure, f, edge, mask_mode, mode):
    """D-grid wind component on pressure levels (regrid_to_edge_weighted_pressure +
    edge_weighted_block_average, regridz.py:81-146, coarsen_restarts.py:496-552).  Only the lines
    that survive the reference's isel(::f) are staggered, remapped and averaged."""
    d4, t = _squeeze_time(delp, 4)
    u4, _ = _squeeze_time(u, 4)
    ntile = delp.shape[0]
    if t not in (None, 1):
        raise NotImplementedError("D-grid pressure coarsening expects a single Time entry per call")
    if edge == "x":  # u on y-interfaces: keep rows 0, f, ..
        delp_lines = core.stagger(d4, "y", line_step=f)          # [6, z, n/f+1, n]
        u_lines = u4[:, :, ::f, :]                                # strided view, x contiguous
        w_lines = length[:, ::f, :]
        outs, _ = core.pressure_coarsen([u_lines], delp_lines, w_lines, f, toa_pressure, mask_mode=mask_mode,
                                        edge_lines=True, mode=mode, return_delp_coarse=False)
        res = outs[0]                                             # [6, z, n/f+1, n/f]
    else:  # v on x-interfaces: keep columns 0, f, .. (gathered into contiguous lines)
        delp_lines = core.stagger(d4, "x", line_step=f)          # [6, z, n/f+1, n] line-major
        v_lines = core.gather_columns(u4, f)                      # [6, z, n/f+1, n]
        w_lines = core.gather_columns(length, f)                  # [6, n/f+1, n]
        outs, _ = core.pressure_coarsen([v_lines], delp_lines, w_lines, f, toa_pressure, mask_mode=mask_mode,
                                        edge_lines=True, mode=mode, return_delp_coarse=False, seq_window=True)
        res = core.transpose_last2(outs[0])                       # [6, z, n/f, n/f+1]
    return _restore_time(res, t, ntile)


def coarse_grain_fv_core_on_pressure(ds, delp, area, dx, dy, toa_pressure, coarsening_factor,
                                     coarsen_agrid_winds=False, extrapolate=False, mode=EXACT):
    """Pressure-level fv_core coarse-graining (coarsen_restarts.py:430-556), without the
    hydrostatic post-processing of :233 (a 'next' row)."""
    f = coarsening_factor
    (ds, delp, area, dx, dy), was_np = _upload(dict(ds), delp, area, dx, dy)
    masked_vars = ["W", "T"]
    if coarsen_agrid_winds:
        if not ("ua" in ds and "va" in ds):
            raise ValueError(_AGRID_ERR)
        masked_vars.extend(["ua", "va"])
    mask_mode = "pfull" if extrapolate else "phalf"
    out, delp_c = _pressure_fields(ds, masked_vars, delp, area, toa_pressure, f, mask_mode, mode)
    out["delp"] = delp_c  # the block average of delp is a by-product of the fused kernel
    out.update(_wba(ds, ["phis", "DZ"], area, f))
    out["u"] = _edge_pressure(ds["u"], delp, dx, toa_pressure, f, "x", mask_mode, mode)
    out["v"] = _edge_pressure(ds["v"], delp, dy, toa_pressure, f, "y", mask_mode, mode)
    return _download(out, was_np)


def coarse_grain_fv_tracer_on_pressure(ds, delp, area, toa_pressure, coarsening_factor, extrapolate=False,
                                       mode=EXACT):
    """coarsen_restarts.py:903-961: every tracer is a masked-area pressure-level average."""
    mask_mode = "pfull" if extrapolate else "phalf"
    (ds, delp, area), was_np = _upload(dict(ds), delp, area)
    out, _ = _pressure_fields(ds, list(ds.keys()), delp, area, toa_pressure, coarsening_factor, mask_mode, mode)
    return _download(out, was_np)


def pressure_level_average(delp_fine, area, fields, coarsening_factor, toa_pressure=300.0, mode=EXACT):
    """fine_res_budget's Grid.pressure_level_average for a list of fields at once
    (workflows/fine_res_budget/budget/budgets.py:49-57): unmasked area weights, pressure geometry
    shared by all fields.  Arrays ``[tile, z, y, x]``; returns (coarse fields, coarse delp)."""
    return core.pressure_coarsen(list(fields), delp_fine, area, coarsening_factor, toa_pressure, mask_mode="none",
                                 mode=mode)


# ------------------------------------------------------------------------------------------
# hydrostatic post-processing, blended method, whole-restart entry points
# ------------------------------------------------------------------------------------------
def impose_hydrostatic_balance(ds_fv_core, ds_fv_tracer, toa_pressure):
    """coarsen_restarts.py:990-1017: DZ from (T, sphum, delp) in hydrostatic balance, phis
    adjusted so that the model-top height is unchanged.  Returns a new fv_core mapping."""
    (core_d, tracer_d), was_np = _upload(dict(ds_fv_core), {"sphum": ds_fv_tracer["sphum"]})
    dz, phis = core.hydrostatic_balance(core_d["DZ"], core_d["phis"], core_d["T"], tracer_d["sphum"], core_d["delp"],
                                        toa_pressure)
    out = dict(core_d)
    out["DZ"], out["phis"] = dz, phis
    return _download(out, was_np)


def blending_weights_agrid(delp, area, toa_pressure, coarsening_factor, delp_coarse=None):
    """coarsen_restarts.py:580-621: 1 above 0.9 x the lowest fine surface pressure of the cell,
    falling linearly (in pressure) to 0 at the coarse surface."""
    f = coarsening_factor
    (delp, area), was_np = _upload(delp, area)
    if delp_coarse is None:
        w = area
        for _ in range(delp.ndim - 3):
            w = w[:, None]
        delp_coarse = core.weighted_block_average(delp, w, f)
    ps_min = core.block_reduce(core.surface_pressure(delp, toa_pressure), f, "min")
    out = core.blending_weights(delp_coarse, ps_min, toa_pressure)
    return out.cpu().numpy() if was_np else out


def blending_weights_dgrid(delp, length, toa_pressure, coarsening_factor, edge):
    """coarsen_restarts.py:624-661: the same along coarse cell edges (delp staggered across the
    cube faces, compute_edge_delp :825-853)."""
    f = coarsening_factor
    (delp, length), was_np = _upload(delp, length)
    delp_edge = core.stagger(delp, "y" if edge == "x" else "x")
    w = length
    for _ in range(delp.ndim - 3):
        w = w[:, None]
    delp_edge_coarse = core.edge_weighted_block_average(delp_edge, w, f, edge=edge)
    ps_min = core.block_reduce(core.surface_pressure(delp_edge, toa_pressure), f, "min", edge=edge)
    out = core.blending_weights(delp_edge_coarse, ps_min, toa_pressure)
    return out.cpu().numpy() if was_np else out


def coarse_grain_fv_core_via_blended_method(ds, delp, area, dx, dy, toa_pressure, coarsening_factor,
                                            coarsen_agrid_winds=False, mass_weighted=True, mode=EXACT):
    """coarsen_restarts.py:700-779."""
    f = coarsening_factor
    (ds, delp, area, dx, dy), was_np = _upload(dict(ds), delp, area, dx, dy)
    pl = coarse_grain_fv_core_on_pressure(ds, delp, area, dx, dy, toa_pressure, f, coarsen_agrid_winds, mode=mode)
    ml = coarse_grain_fv_core(ds, delp, area, dx, dy, f, coarsen_agrid_winds, mass_weighted)
    wa = blending_weights_agrid(delp, area, toa_pressure, f, delp_coarse=ml["delp"])
    wu = blending_weights_dgrid(delp, dx, toa_pressure, f, "x")
    wv = blending_weights_dgrid(delp, dy, toa_pressure, f, "y")
    ignore = ["u", "v"] + ([] if coarsen_agrid_winds else ["ua", "va"])
    out = {}
    for v in ds:
        if v in ignore:
            continue
        out[v] = core.blend(wa, pl[v], ml[v]) if ds[v].ndim == delp.ndim else ml[v]
    out["u"] = core.blend(wu, pl["u"], ml["u"])
    out["v"] = core.blend(wv, pl["v"], ml["v"])
    return _download(out, was_np)


def coarse_grain_fv_tracer_via_blended_method(ds, delp, area, toa_pressure, coarsening_factor, mass_weighted=True,
                                              mode=EXACT):
    """coarsen_restarts.py:782-822."""
    f = coarsening_factor
    (ds, delp, area), was_np = _upload(dict(ds), delp, area)
    pl = coarse_grain_fv_tracer_on_pressure(ds, delp, area, toa_pressure, f, mode=mode)
    ml = coarse_grain_fv_tracer(ds, delp, area, f, mass_weighted)
    w = blending_weights_agrid(delp, area, toa_pressure, f)
    return _download({v: core.blend(w, pl[v], ml[v]) for v in ds}, was_np)


def _entry(restarts, grid_spec):
    need = ("fv_core.res", "fv_srf_wnd.res", "fv_tracer.res")
    missing = [c for c in need if c not in restarts]
    if missing:
        raise KeyError(f"restarts lacks the categories {missing}")
    return restarts["fv_core.res"], restarts["fv_tracer.res"], restarts["fv_srf_wnd.res"], grid_spec["area"], \
        grid_spec["dx"], grid_spec["dy"]


def coarsen_restarts_on_sigma(coarsening_factor, grid_spec, toa_pressure, restarts, coarsen_agrid_winds=False,
                              mass_weighted=True):
    """coarsen_restarts.py:76-152 for the fv_core / fv_tracer / fv_srf_wnd categories
    (``restarts``: category -> {name: array}; ``grid_spec``: {area, dx, dy}).  sfc_data is
    coarse-grained by fv3net_b200.sfc_data when present."""
    fc, ft, fs, area, dx, dy = _entry(restarts, grid_spec)
    out = {"fv_core.res": coarse_grain_fv_core(fc, fc["delp"], area, dx, dy, coarsening_factor, coarsen_agrid_winds,
                                               mass_weighted),
           "fv_srf_wnd.res": coarse_grain_fv_srf_wnd(fs, area, coarsening_factor),
           "fv_tracer.res": coarse_grain_fv_tracer(ft, fc["delp"], area, coarsening_factor, mass_weighted)}
    return _with_sfc(out, restarts, area, coarsening_factor)


def coarsen_restarts_on_pressure(coarsening_factor, grid_spec, toa_pressure, restarts, coarsen_agrid_winds=False,
                                 extrapolate=False, mode=EXACT):
    """coarsen_restarts.py:155-240, hydrostatic balance imposed on the result (:233-235)."""
    fc, ft, fs, area, dx, dy = _entry(restarts, grid_spec)
    f = coarsening_factor
    out = {"fv_core.res": coarse_grain_fv_core_on_pressure(fc, fc["delp"], area, dx, dy, toa_pressure, f,
                                                           coarsen_agrid_winds, extrapolate=extrapolate, mode=mode),
           "fv_srf_wnd.res": coarse_grain_fv_srf_wnd(fs, area, f),
           "fv_tracer.res": coarse_grain_fv_tracer_on_pressure(ft, fc["delp"], area, toa_pressure, f,
                                                               extrapolate=extrapolate, mode=mode)}
    out["fv_core.res"] = impose_hydrostatic_balance(out["fv_core.res"], out["fv_tracer.res"], toa_pressure)
    return _with_sfc(out, restarts, area, f)


def coarsen_restarts_via_blended_method(coarsening_factor, grid_spec, toa_pressure, restarts,
                                        coarsen_agrid_winds=False, mass_weighted=True, mode=EXACT):
    """coarsen_restarts.py:243-332."""
    fc, ft, fs, area, dx, dy = _entry(restarts, grid_spec)
    f = coarsening_factor
    out = {"fv_core.res": coarse_grain_fv_core_via_blended_method(fc, fc["delp"], area, dx, dy, toa_pressure, f,
                                                                  coarsen_agrid_winds, mass_weighted, mode=mode),
           "fv_srf_wnd.res": coarse_grain_fv_srf_wnd(fs, area, f),
           "fv_tracer.res": coarse_grain_fv_tracer_via_blended_method(ft, fc["delp"], area, toa_pressure, f,
                                                                      mass_weighted, mode=mode)}
    out["fv_core.res"] = impose_hydrostatic_balance(out["fv_core.res"], out["fv_tracer.res"], toa_pressure)
    return _with_sfc(out, restarts, area, f)


def _with_sfc(out, restarts, area, f):
    if "sfc_data" in restarts:
        from . import sfc_data

        out["sfc_data"] = sfc_data.coarse_grain_sfc_data_complex(restarts["sfc_data"], area, f)
    return out
