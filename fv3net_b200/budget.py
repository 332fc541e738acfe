"""fine_res_budget re-coarse-graining on the GPU.

Mirrors ``Grid.pressure_level_average``, ``Grid.area_above_fine_surface``, ``coarsen_variables``
and the array arithmetic of ``compute_recoarsened_budget_inputs``
(/root/reference/workflows/fine_res_budget/budget/budgets.py:49-117, 217-294).  Arrays are
``[tile, z, y, x]`` (3-D) / ``[tile, y, x]`` (2-D); numpy in -> numpy out, CUDA tensors stay on the
device.  The 3-D fields of one call share ONE pressure geometry: they go through the fused
kernel (b200cs_pressure_coarsen, mask NONE) in groups of four instead of one
upsample + remap + average chain per variable.
"""
import numpy as np
import torch

from . import _lib, core
from .core import EXACT


def pressure_level_average(delp_fine, area, fields, coarsening_factor, toa_pressure=300.0, mode=EXACT):
    """budgets.py:49-57 for a list of fields: unmasked area-weighted average on the coarse
    pressure levels.  The coarse delp is the area-weighted block average of delp_fine, which is how
    every caller forms it (budgets.py:272) -- it is a by-product of the kernel and is returned
    with the coarse fields."""
    return core.pressure_coarsen(list(fields), delp_fine, area, coarsening_factor, toa_pressure, mask_mode="none",
                                 mode=mode)


def pressure_at_midpoint_log(delp, toa_pressure=300.0):
    """calc/thermo/vertically_dependent.py:153-179 for ``[..., z, y, x]``."""
    d, was_np = core._to_dev(delp)
    d, nb, km, ncol = core._zyx(d)
    out = torch.empty_like(d)
    rc = _lib.load().b200cs_pressure_at_midpoint_log(d.data_ptr(), out.data_ptr(), core._DT[d.dtype], nb, km, ncol,
                                                     float(toa_pressure), core._stream())
    _lib.check(rc, "pressure_at_midpoint_log")
    return core._ret(out, was_np)


def area_above_fine_surface(delp_fine, delp_coarse, area, toa_pressure=300.0):
    """budgets.py:62-77: block sum of the area of the fine columns whose surface pressure is at
    least the coarse mid-level pressure -> ``[tile, z, cy, cx]`` ('exposed_area')."""
    df, was_np = core._to_dev(delp_fine)
    dc, _ = core._to_dev(delp_coarse)
    a, _ = core._to_dev(area)
    f = df.shape[-1] // dc.shape[-1]
    pdt = core._promote(df, dc)
    pfull_c = pressure_at_midpoint_log(dc.to(pdt), toa_pressure)
    ps = core.surface_pressure(df.to(pdt), toa_pressure)
    nb = int(np.prod(df.shape[:-3], dtype=np.int64))
    km, ny, nx = df.shape[-3:]
    a = a.expand(tuple(df.shape[:-3]) + (ny, nx)).contiguous()
    w = torch.empty(tuple(df.shape), dtype=a.dtype, device=a.device)
    rc = _lib.load().b200cs_exposed_area_weights(pfull_c.contiguous().data_ptr(), ps.contiguous().data_ptr(),
                                                 core._DT[pdt], a.data_ptr(), core._DT[a.dtype], w.data_ptr(), nb, km,
                                                 ny, nx, f, core._stream())
    _lib.check(rc, "exposed_area_weights")
    out = core.block_reduce(w, f, "sum", skipna=True)
    return core._ret(out, was_np)


def _ew2(op, a, b, scalar=0.0):
    dt = core._promote(a, b)
    a, b = a.to(dt).contiguous(), b.to(dt).contiguous()
    out = torch.empty_like(a)
    rc = _lib.load().b200cs_elementwise(op, a.data_ptr(), b.data_ptr(), None, out.data_ptr(), core._DT[dt], a.numel(),
                                        float(scalar), core._stream())
    _lib.check(rc, "elementwise")
    return out


def coarsen_variables(fields, delp_fine, area, coarsening_factor, toa_pressure=300.0, mode=EXACT):
    """budgets.py:92-117: ``fields`` name -> array; 3-D ones are averaged on coarse pressure
    levels, 2-D ones by area.  Returns (name -> coarse array, coarse delp)."""
    dev = {k: core._to_dev(v) for k, v in fields.items()}
    was_np = any(n for _, n in dev.values())
    dev = {k: t for k, (t, _) in dev.items()}
    d, _ = core._to_dev(delp_fine)
    a, _ = core._to_dev(area)
    names3 = [k for k, t in dev.items() if t.dim() == d.dim()]
    outs3, dco = core.pressure_coarsen([dev[k] for k in names3], d, a, coarsening_factor, toa_pressure,
                                       mask_mode="none", mode=mode)
    out = dict(zip(names3, outs3))
    for k, t in dev.items():
        if k not in out:
            out[k] = core.weighted_block_average(t, a, coarsening_factor)
    if was_np:
        return {k: v.cpu().numpy() for k, v in out.items()}, dco.cpu().numpy()
    return out, dco


def compute_recoarsened_budget_inputs(begin, middle, end, dt=15 * 60, factor=8, toa_pressure=300.0,
                                      first_moments=("T", "eddy_flux_vulcan_omega_temp", "t_dt_fv_sat_adj_coarse",
                                                     "t_dt_nudge_coarse", "t_dt_phys_coarse", "sphum",
                                                     "eddy_flux_vulcan_omega_sphum", "qv_dt_fv_sat_adj_coarse",
                                                     "qv_dt_phys_coarse", "vulcan_omega_coarse"),
                                      second_moments=(("T", "vulcan_omega_coarse"), ("sphum", "vulcan_omega_coarse")),
                                      storage_terms=("T", "sphum"), mode=EXACT):
    """budgets.py:217-294.  ``begin / middle / end`` are the three ``step`` slices of the merged
    dataset as name -> array mappings (``middle`` holds delp, area_coarse, exposed_area_coarse and
    every first-moment field; begin / end the storage-term fields).  Products and storage terms
    are formed on the device, then everything is averaged on the coarse pressure levels with one
    shared geometry.  Returns name -> coarse array incl. delp, exposed_area and area."""
    up = lambda x: core._to_dev(x)[0]  # noqa: E731
    was_np = not isinstance(middle["delp"], torch.Tensor)
    area, delp = up(middle["area_coarse"]), up(middle["delp"])
    raw = {v: up(middle[v]) for v in first_moments}
    for f1, f2 in second_moments:  # budgets.py:152-156
        raw[f"{f1}_{f2}"] = _ew2(_lib.EW_MUL, up(middle[f1]), up(middle[f2]))
    for v in storage_terms:        # budgets.py:80-85: (end - begin) / time_step
        raw[f"{v}_storage"] = _ew2(_lib.EW_DIFF_OVER, up(end[v]), up(begin[v]), scalar=dt)
    out, delp_coarse = coarsen_variables(raw, delp, area, factor, toa_pressure, mode)
    out["delp"] = delp_coarse
    out["exposed_area"] = area_above_fine_surface(delp, delp_coarse, up(middle["exposed_area_coarse"]), toa_pressure)
    out["area"] = core.block_reduce(area, factor, "sum", skipna=True)
    if was_np:
        return {k: (v.cpu().numpy() if isinstance(v, torch.Tensor) else v) for k, v in out.items()}
    return out
