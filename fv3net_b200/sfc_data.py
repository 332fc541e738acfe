"""'Complex' sfc_data coarse-graining on the GPU.

Mirrors ``_coarse_grain_sfc_data_complex`` and its method table
(/root/reference/external/vcm/vcm/cubedsphere/coarsen_restarts.py:1111-1470): dominant surface /
vegetation / soil types by block mode, masked area-weighted (and area x fraction weighted) block
averages, and the surface_chgres corrections.  ``ds`` maps names to arrays ``[tile, (Time,) y, x]``
(``stc/smc/slc``: ``[tile, (Time,) z, y, x]``); ``area`` is ``[tile, y, x]``.  numpy in -> numpy out,
CUDA tensors in -> CUDA tensors out; every operation is a libb200cs kernel (block_ops.cu,
elementwise.cu), nothing is computed on the host.
"""
import numpy as np
import torch

from . import _lib, core

FREEZING_TEMPERATURE, SHDMIN_THRESHOLD, STYPE_LAND_ICE, VTYPE_LAND_ICE = 273.16, 0.011, 16.0, 15.0  # :48-51
RTOL, ATOL = 1e-5, 1e-8  # xarray_utils.isclose defaults (external/vcm/vcm/xarray_utils.py:7-35)

AREA = ["tsea", "alvsf", "alvwf", "alnsf", "alnwf", "facsf", "facwf", "f10m", "t2m", "q2m", "uustar", "ffmm", "ffhh",
        "tprcp", "snwdph"]
DOMINANT_SFC = ["tg3", "vfrac", "fice", "sncovr", "shdmin", "shdmax", "snoalb"]
DOMINANT_SFC_AND_VTYPE = ["canopy", "zorl"]
DOMINANT_SFC_AND_STYPE = ["smc", "slc", "stc"]
SUPPORTED = set(AREA + DOMINANT_SFC + DOMINANT_SFC_AND_VTYPE + DOMINANT_SFC_AND_STYPE +
                ["slmsk", "vtype", "stype", "srflag", "slope", "sheleg", "hice", "tisfc"])


def _ew(op, a, b=None, mask=None, scalar=0.0):
    if b is not None:  # numpy promotion of the reference expression (area f32 * vfrac f64 -> f64)
        dt = core._promote(a, b)
        a = a.to(dt)
        b = b.to(dt).expand_as(a).contiguous()
    a = a.contiguous()
    out = torch.empty_like(a)
    if mask is not None:
        mask = mask.expand_as(a).contiguous()
    rc = _lib.load().b200cs_elementwise(op, a.data_ptr(), b.data_ptr() if b is not None else None,
                                        mask.data_ptr() if mask is not None else None, out.data_ptr(),
                                        core._DT[a.dtype], a.numel(), float(scalar), core._stream())
    _lib.check(rc, "elementwise")
    return out


def where(a, mask, other=float("nan")):
    """``a.where(mask, other)``"""
    return _ew(_lib.EW_WHERE, a, mask=mask, scalar=other)


def where_scalar(mask, scalar, a):
    """``xr.where(mask, scalar, a)``"""
    return _ew(_lib.EW_WHERE_NOT, a, mask=mask, scalar=scalar)


def select(mask, a, b):
    """``xr.where(mask, a, b)``"""
    return _ew(_lib.EW_SELECT, a, b=b, mask=mask)


def compare_scalar(cmp, a, scalar, and_mask=None):
    a = a.contiguous()
    out = torch.empty(a.shape, dtype=torch.uint8, device=a.device)
    if and_mask is not None:
        and_mask = and_mask.expand_as(a).contiguous()
    rc = _lib.load().b200cs_compare_scalar(cmp, a.data_ptr(), core._DT[a.dtype], float(scalar), RTOL, ATOL,
                                           and_mask.data_ptr() if and_mask is not None else None, out.data_ptr(),
                                           a.numel(), core._stream())
    _lib.check(rc, "compare_scalar")
    return out


def dominant_mask(fine, coarse, factor, and_mask=None):
    """``isclose(fine, block_upsample_like(coarse, fine))`` (& and_mask) -> uint8 ``fine.shape``."""
    fine, coarse = fine.contiguous(), coarse.to(fine.dtype).contiguous()
    ny, nx = fine.shape[-2:]
    nb = fine.numel() // (ny * nx)
    out = torch.empty(fine.shape, dtype=torch.uint8, device=fine.device)
    rc = _lib.load().b200cs_dominant_mask(fine.data_ptr(), coarse.data_ptr(), core._DT[fine.dtype], nb, ny, nx,
                                          int(factor), RTOL, ATOL, and_mask.data_ptr() if and_mask is not None else None,
                                          out.data_ptr(), core._stream())
    _lib.check(rc, "dominant_mask")
    return out


def _mode(a, f):
    """block_coarsen(method='mode', func_kwargs={'nan_policy': 'omit'}) (:1165-1172)"""
    return core.block_reduce(a, f, "mode", skipna=True)


def _z(mask_or_w, v):
    """a 2-D mask / weight against a variable with an extra vertical axis before (y, x)"""
    return mask_or_w.unsqueeze(-3) if v.dim() == mask_or_w.dim() + 1 else mask_or_w


def coarse_grain_sfc_data_complex(ds, area, coarsening_factor):
    """coarsen_restarts.py:1111-1159.  Returns name -> float32 arrays (``_doubles_to_floats``)."""
    f = int(coarsening_factor)
    unknown = [k for k in ds if k not in SUPPORTED]
    if unknown:
        raise KeyError(f"no sfc_data coarsening method for {unknown}")
    dev, was_np = {}, False
    for k, v in ds.items():
        t, n = core._to_dev(v)
        dev[k] = t
        was_np |= n
    area_t, _ = core._to_dev(area)
    slmsk = dev["slmsk"]
    a = area_t
    for _ in range(slmsk.dim() - 3):
        a = a.unsqueeze(1)
    a = a.expand_as(slmsk).contiguous()          # area per [tile, (Time,) y, x]

    # :1162-1206 dominant surface type, then dominant vegetation / soil type within it
    slmsk_c = _mode(slmsk, f)
    dom_sfc = dominant_mask(slmsk, slmsk_c, f)
    vtype_c = _mode(where(dev["vtype"], dom_sfc), f)
    stype_c = _mode(where(dev["stype"], dom_sfc), f)
    dom_sfc_v = dominant_mask(dev["vtype"], vtype_c, f, and_mask=dom_sfc)
    dom_sfc_s = dominant_mask(dev["stype"], stype_c, f, and_mask=dom_sfc)

    out = {"slmsk": slmsk_c, "vtype": vtype_c, "stype": stype_c}
    a_sfc = where(a, dom_sfc)

    def wba(v, w, w2=None):
        return core.weighted_block_average(v, _z(w, v), f, weights2=_z(w2, v) if w2 is not None else None)

    for name, v in dev.items():
        if name in out:
            continue
        if name in AREA:                                             # :1209-1215
            r = wba(v, a)
        elif name in DOMINANT_SFC:                                   # :1218-1231
            r = wba(where(v, dom_sfc), a_sfc)
        elif name in DOMINANT_SFC_AND_VTYPE:                         # :1234-1267
            vm, am = where(v, dom_sfc_v), where(a, dom_sfc_v)
            area_mean = wba(vm, am)
            avf = _ew(_lib.EW_MUL, am, b=dev["vfrac"])               # (area * vfrac).where(mask)
            area_vfrac_mean = wba(vm, avf)
            total = core.block_reduce(avf, f, "sum", skipna=True)
            r = select(compare_scalar(_lib.CMP_GT, total, 0.0), area_vfrac_mean, area_mean)
        elif name in DOMINANT_SFC_AND_STYPE:                         # :1270-1285
            r = wba(where(v, _z(dom_sfc_s, v)), where(a, dom_sfc_s))
        elif name == "srflag":                                       # :1288-1299
            r = _mode(v, f)
        elif name == "slope":                                        # :1302-1316
            r = _mode(where(v, dom_sfc), f)
        elif name in ("sheleg", "hice"):                             # :1319-1340, .fillna(0.0)
            frac = dev["sncovr" if name == "sheleg" else "fice"]
            r = _ew(_lib.EW_FILLNA, wba(v, _ew(_lib.EW_MUL, a, b=frac)), scalar=0.0)
        else:                                                        # tisfc :1343-1370
            vm = where(v, dom_sfc)
            sea_ice = wba(vm, where(_ew(_lib.EW_MUL, a, b=dev["fice"]), dom_sfc))
            land_or_ocean = wba(vm, a_sfc)
            r = select(compare_scalar(_lib.CMP_ISCLOSE, slmsk_c, 2.0), sea_ice, land_or_ocean)
        out[name] = r

    # surface_chgres corrections :1396-1470, in the reference's order
    land_ice = compare_scalar(_lib.CMP_ISCLOSE, out["vtype"], VTYPE_LAND_ICE)
    for k in ("tsea", "tg3"):
        if k in out:
            clipped = _ew(_lib.EW_MIN_SCALAR, out[k], scalar=FREEZING_TEMPERATURE)
            out[k] = select(land_ice, clipped, out[k])
    out["stype"] = where_scalar(land_ice, STYPE_LAND_ICE, out["stype"])
    if "canopy" in out and "shdmin" in out:
        out["canopy"] = where_scalar(compare_scalar(_lib.CMP_LT, out["shdmin"], SHDMIN_THRESHOLD), 0.0, out["canopy"])
    if "shdmin" in out:
        out["shdmin"] = where_scalar(land_ice, 0.0, out["shdmin"])
    out = {k: (v.to(torch.float32) if v.dtype == torch.float64 else v) for k, v in out.items()}
    if was_np:
        return {k: v.cpu().numpy() for k, v in out.items()}
    return out


def coarse_grain_sfc_data(ds, area, coarsening_factor, version="simple"):
    """coarsen_restarts.py:1020-1048: 'simple' = block median with slmsk rounded, 'complex' above."""
    if version == "complex":
        return coarse_grain_sfc_data_complex(ds, area, coarsening_factor)
    if version != "simple":
        raise ValueError("Currently the only supported versions are 'simple' and 'complex'. " f"Got {version}.")
    out = {k: core.block_reduce(v, coarsening_factor, "median", skipna=False) for k, v in ds.items()}
    if "slmsk" in out:
        s = out["slmsk"]
        out["slmsk"] = torch.round(s) if isinstance(s, torch.Tensor) else np.round(s).astype(s.dtype)
    return out
