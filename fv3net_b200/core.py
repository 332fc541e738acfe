"""Array-level host API of fv3net_b200: the arithmetic of vcm.cubedsphere's coarse-graining
operators on device buffers.  Horizontal axes are the LAST two axes ``[..., y, x]`` (x contiguous,
the layout of FV3 restart files); a vertical axis, where one exists, is axis -3.

Inputs may be ``torch`` CUDA tensors (returned as CUDA tensors, nothing leaves the device) or
numpy arrays (uploaded, result downloaded: numpy in -> numpy out, the reference's contract).
Every function launches hand-written sm_100a kernels through the C ABI of ``libb200cs.so``
(include/b200cs.h); there is no CPU path.

Reference citations are relative to /root/reference/external/vcm/vcm/.
"""
import ctypes
import os

import numpy as np
import torch

from . import _lib
from ._lib import EXACT, FAST, MASK_NONE, MASK_PFULL, MASK_PHALF  # noqa: F401

_DT = {torch.float32: _lib.F32, torch.float64: _lib.F64}
_NP_OF = {torch.float32: np.float32, torch.float64: np.float64}
_MASK = {None: MASK_NONE, "none": MASK_NONE, "phalf": MASK_PHALF, "pfull": MASK_PFULL}
_OPS = {"sum": _lib.OP_SUM, "mean": _lib.OP_MEAN, "min": _lib.OP_MIN, "max": _lib.OP_MAX,
        "median": _lib.OP_MEDIAN, "mode": _lib.OP_MODE}


def device():
    if not torch.cuda.is_available():
        raise _lib.B200csError("fv3net_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def _stream():
    return torch.cuda.current_stream().cuda_stream


def _to_dev(a, floating=True):
    """-> (tensor on device, was_numpy).  Integer / bool inputs become float64 like numpy's
    true-division promotion in the reference expressions."""
    was_np = not isinstance(a, torch.Tensor)
    t = torch.as_tensor(np.ascontiguousarray(a)) if was_np else a
    if floating and t.dtype not in _DT:
        t = t.to(torch.float64)
    if t.device.type != "cuda":
        t = t.to(device(), non_blocking=True)
    return t, was_np


def _ret(t, was_np):
    return t.cpu().numpy() if was_np else t


def _promote(*ts):
    return torch.float64 if any(t is not None and t.dtype == torch.float64 for t in ts) else torch.float32


def _collapse(shape, strides, lo, hi):
    """Stride of dims[lo:hi] merged into one axis, or None if they do not merge."""
    dims = [(n, s) for n, s in zip(shape[lo:hi], strides[lo:hi]) if n != 1]
    if not dims:
        return 0
    for (n0, s0), (n1, s1) in zip(dims[:-1], dims[1:]):
        if s0 != s1 * n1:
            return None
    return dims[-1][1]


_CANON_CACHE = {}


def _canon4_plan(layouts, lead_shape):
    """(b0, b1, strides) for tensors given as (shape, stride) tuples, or None when some tensor
    needs a contiguous copy first.  Pure integer work, memoised per layout."""
    nl = len(lead_shape)
    for split in range(nl, -1, -1):
        strides = []
        for shape, stride in layouts:
            if stride[-1] != 1 and shape[-1] != 1:
                break
            s0 = _collapse(shape, stride, 0, split)
            s1 = _collapse(shape, stride, split, nl)
            if s0 is None or s1 is None:
                break
            strides.append((s0, s1, stride[-2]))
        else:
            b0 = int(np.prod(lead_shape[:split], dtype=np.int64))
            b1 = int(np.prod(lead_shape[split:], dtype=np.int64))
            return b0, b1, strides
    return None


def _canon4(ts, lead_shape):
    """Views tensors (all expanded to lead_shape + (ny, nx)) as [B0, B1, ny, nx] with common
    (B0, B1) and per-tensor element strides (s0, s1, sy).  Falls back to contiguous copies."""
    for attempt in range(2):
        key = (tuple(lead_shape), tuple((tuple(t.shape), t.stride()) for t in ts))
        plan = _CANON_CACHE.get(key, False)
        if plan is False:
            plan = _canon4_plan(key[1], lead_shape)
            if len(_CANON_CACHE) < 4096:
                _CANON_CACHE[key] = plan
        if plan is not None:
            return ts, plan[0], plan[1], plan[2]
        ts = [t.contiguous() for t in ts]
    raise AssertionError("unreachable")


def _broadcast_lead(shapes):
    """numpy broadcasting of the leading shapes (torch.broadcast_shapes costs ~20 us per call)."""
    first = shapes[0]
    if all(sh == first for sh in shapes[1:]):
        return tuple(first)
    n = max(len(sh) for sh in shapes)
    out = [1] * n
    for sh in shapes:
        for i in range(1, len(sh) + 1):
            d = sh[-i]
            if d != 1:
                if out[-i] not in (1, d):
                    raise RuntimeError(f"shapes {[tuple(x) for x in shapes]} do not broadcast")
                out[-i] = d
    return tuple(out)


def _window(f, edge):
    if edge is None:
        return f, f, f, f
    if edge == "x":
        return 1, f, f, f  # fy, fx, ystep, xstep
    if edge == "y":
        return f, 1, f, f
    raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")


def weighted_block_average(obj, weights, coarsening_factor, weights2=None, edge=None):
    """nansum(obj*W)/nansum(W) over windows of the last two axes (cubedsphere/coarsen.py:183-273).

    W = weights, or weights2*weights when ``weights2`` is given (mass = delp*area formed on the
    fly, cubedsphere/coarsen_restarts.py:400).  ``edge`` None: f x f blocks; 'x' / 'y': the
    edge-weighted average (1-D windows along that axis, every f-th line across it).
    weights (and weights2) broadcast against obj in the leading axes (numpy rules)."""
    f = int(coarsening_factor)
    fy, fx, ystep, xstep = _window(f, edge)
    o, was_np = _to_dev(obj)
    w, _ = _to_dev(weights)
    v = _to_dev(weights2)[0] if weights2 is not None else None
    ny, nx = o.shape[-2:]
    tensors = [o, w] + ([v] if v is not None else [])
    lead = _broadcast_lead([t.shape[:-2] for t in tensors])
    for t in tensors:
        if tuple(t.shape[-2:]) != (ny, nx):
            raise ValueError("obj and weights must share their horizontal extents")
    full = lead + (ny, nx)
    tensors = [t if tuple(t.shape) == full else t.expand(full) for t in tensors]
    tensors, b0, b1, st = _canon4(tensors, lead)
    # xarray's coarsen(boundary="exact") raises when the window does not divide the coarsened axis
    # (coarsen.py:210-213, 257-262); the interface axis of an edge average is subsampled, not coarsened
    if (fy == ystep and ny % fy) or (fx == xstep and nx % fx):
        raise ValueError(f"window {f} does not divide the horizontal extents {(ny, nx)}")
    cy, cx = (ny - fy) // ystep + 1, (nx - fx) // xstep + 1
    odt = _promote(*tensors)
    out = torch.empty(lead + (cy, cx), dtype=odt, device=o.device)
    lib = _lib.load()
    o, w = tensors[0], tensors[1]
    v = tensors[2] if v is not None else None
    rc = lib.b200cs_weighted_block_average(
        o.data_ptr(), _DT[o.dtype], _lib.i64arr(st[0]), w.data_ptr(), _DT[w.dtype], _lib.i64arr(st[1]),
        v.data_ptr() if v is not None else None, _DT[v.dtype] if v is not None else 0,
        _lib.i64arr(st[2]) if v is not None else None, out.data_ptr(), _DT[odt], b0, b1, ny, nx, fy, fx,
        ystep, xstep, _stream())
    _lib.check(rc, "weighted_block_average")
    return _ret(out, was_np)


def edge_weighted_block_average(obj, spacing, coarsening_factor, edge="x"):
    """cubedsphere/coarsen.py:221-273 on ``[..., y, x]`` arrays."""
    if edge not in ("x", "y"):
        raise ValueError(f"'edge' most be either 'x' or 'y'; got {edge}.")
    return weighted_block_average(obj, spacing, coarsening_factor, edge=edge)


def block_reduce(obj, coarsening_factor, method="sum", skipna=True, edge=None, pad=False, cval=np.nan, window=None):
    """Block reduction of the last two axes by ``method`` in sum/mean/min/max/median/mode.
    Replaces horizontal_block_reduce / block_coarsen / block_edge_coarsen
    (cubedsphere/coarsen.py:520-683,795-840; _skimage.py:125-202).  ``pad`` pads trailing edges
    with ``cval`` (skimage semantics); dtype is preserved.  ``window`` = (fy, fx, ystep, xstep)
    overrides the square / edge windows (1-D blocks when only one block dimension is present)."""
    if method not in _OPS:
        raise ValueError(f"unsupported reduction {method!r}; supported: {sorted(_OPS)}")
    f = int(coarsening_factor)
    fy, fx, ystep, xstep = _window(f, edge) if window is None else (int(v) for v in window)
    o, was_np = _to_dev(obj)
    o = o.contiguous()
    ny, nx = o.shape[-2:]
    lead = tuple(o.shape[:-2])
    nb = int(np.prod(lead, dtype=np.int64))
    if pad:
        cy, cx = -(-ny // fy), -(-nx // fx)
    else:
        if (fy == ystep and ny % fy) or (fx == xstep and nx % fx):
            raise ValueError(f"window {(fy, fx)} does not divide the horizontal extents {(ny, nx)}; pass pad=True "
                             "for skimage's trailing-edge padding")
        cy, cx = (ny - fy) // ystep + 1, (nx - fx) // xstep + 1
    out = torch.empty(lead + (cy, cx), dtype=o.dtype, device=o.device)
    rc = _lib.load().b200cs_block_reduce(o.data_ptr(), _DT[o.dtype], _lib.i64arr((ny * nx, nx)), out.data_ptr(),
                                         nb, ny, nx, fy, fx, ystep, xstep, _OPS[method], int(bool(skipna)),
                                         int(bool(pad)), float(cval), _stream())
    _lib.check(rc, "block_reduce")
    return _ret(out, was_np)


def block_upsample(obj, upsampling_factor, axes=(-2, -1), staggered=None):
    """Repeat each element f times along the given horizontal axes; an odd-size axis is an
    interface axis whose last element is not repeated (cubedsphere/coarsen.py:843-897).
    ``staggered`` = (y, x) booleans overrides the reference's inference from odd sizes."""
    f = int(upsampling_factor)
    o, was_np = _to_dev(obj, floating=False)
    nd = o.dim()
    axes = tuple(a % nd for a in axes)
    if not set(axes) <= {nd - 2, nd - 1}:
        raise ValueError("block_upsample works on the last two axes")
    o = o.contiguous()  # any 1/2/4/8-byte dtype keeps its type: the kernel only moves bits
    cy, cx = o.shape[-2:]
    fy = f if (nd - 2) in axes else 1
    fx = f if (nd - 1) in axes else 1
    ys, xs = int(fy > 1 and cy % 2 == 1), int(fx > 1 and cx % 2 == 1)
    if staggered is not None:
        ys, xs = int(bool(staggered[0]) and fy > 1), int(bool(staggered[1]) and fx > 1)
    ny = (cy - 1) * fy + 1 if ys else cy * fy
    nx = (cx - 1) * fx + 1 if xs else cx * fx
    out = torch.empty(tuple(o.shape[:-2]) + (ny, nx), dtype=o.dtype, device=o.device)
    nb = int(np.prod(o.shape[:-2], dtype=np.int64))
    rc = _lib.load().b200cs_block_upsample(o.data_ptr(), out.data_ptr(), o.element_size(), nb, cy, cx, fy, fx,
                                           ys, xs, _stream())
    _lib.check(rc, "block_upsample")
    return _ret(out, was_np)


def block_upsample_like(obj, reference_shape_yx):
    """cubedsphere/coarsen.py:900-938: the factor is inferred from the x axis."""
    nx_c, nx_f = obj.shape[-1], reference_shape_yx[1]
    f = (nx_f - 1) // (nx_c - 1) if nx_c % 2 == 1 else nx_f // nx_c
    return block_upsample(obj, f)


def stagger(delp, axis, line_step=1):
    """Centre -> edge interpolation across the six cube faces (xgcm ``grid.interp`` with
    FV3_FACE_CONNECTIONS, cubedsphere/xgcm.py:7-34, regridz.py:118-132).  delp ``[6, ..., n, n]``;
    axis 'y' -> ``[6, ..., n+1, n]``, 'x' -> ``[6, ..., n, n+1]``.  With line_step > 1 only lines
    0, step, .. are produced, line-major: ``[6, ..., n/step+1, n]`` for both axes."""
    d, was_np = _to_dev(delp)
    if d.shape[0] != 6 or d.shape[-1] != d.shape[-2]:
        raise ValueError("stagger needs a [6, ..., n, n] array (tile axis first)")
    d = d.contiguous()
    n = d.shape[-1]
    nb1 = int(np.prod(d.shape[1:-2], dtype=np.int64))
    ax = {"y": 0, "x": 1}[axis]
    nl = n // line_step + 1
    if ax == 0 or line_step > 1:
        oshape = tuple(d.shape[:-2]) + (nl, n)
    else:
        oshape = tuple(d.shape[:-2]) + (n, nl)
    out = torch.empty(oshape, dtype=d.dtype, device=d.device)
    rc = _lib.load().b200cs_stagger(d.data_ptr(), out.data_ptr(), _DT[d.dtype], nb1, n, ax, int(line_step), _stream())
    _lib.check(rc, "stagger")
    return _ret(out, was_np)


def pressure_at_interface(delp, toa_pressure=300.0, z_axis=-3):
    """calc/thermo/vertically_dependent.py:41-66: [toa, toa+cumsum(delp)] along z (NaN as 0)."""
    d, was_np = _to_dev(delp)
    nd = d.dim()
    z = z_axis % nd
    dz = d.movedim(z, -3) if nd >= 3 else d.unsqueeze(-1).unsqueeze(-1).movedim(z, -3)
    dz = dz.contiguous()
    km = dz.shape[-3]
    ncol = dz.shape[-2] * dz.shape[-1]
    nb0 = int(np.prod(dz.shape[:-3], dtype=np.int64))
    out = torch.empty(tuple(dz.shape[:-3]) + (km + 1,) + tuple(dz.shape[-2:]), dtype=d.dtype, device=d.device)
    rc = _lib.load().b200cs_pressure_at_interface(dz.data_ptr(), _DT[d.dtype], _lib.i64arr((km * ncol, ncol)),
                                                  out.data_ptr(), nb0, km, ncol, float(toa_pressure), _stream())
    _lib.check(rc, "pressure_at_interface")
    out = out.movedim(-3, z) if nd >= 3 else out.movedim(-3, z).squeeze(-1).squeeze(-1)
    return _ret(out, was_np)


def mappm(pe1, q1, pe2, i1=1, i2=None, iv=1, kord=1, ptop=0.0, mode=EXACT):
    """Drop-in for the f2py call ``mappm.mappm(pe1[ncol,km+1], q1[ncol,km], pe2[ncol,kn+1], i1,
    i2, iv, kord, ptop)`` -> float32 ``[ncol, kn]`` (external/mappm/mappm/mappm.f90:10;
    cubedsphere/regridz.py:329-334).  Scalars are coerced with int() like f2py does
    (tests/test_mappm.py:9-13 passes floats); ptop is unused, as in the Fortran.  ``mode=FAST``:
    b200cs_mappm_fast (kord <= 3 only), within 1e-6 relative of the default."""
    p1, was_np = _to_dev(pe1)
    q, _ = _to_dev(q1)
    p2, _ = _to_dev(pe2)
    p1, q, p2 = (t.to(torch.float32).contiguous() for t in (p1, q, p2))
    ncol, km = q.shape
    if p1.shape != (ncol, km + 1) or p2.shape[0] != ncol:
        raise ValueError("mappm: expected pe1[ncol,km+1], q1[ncol,km], pe2[ncol,kn+1]")
    kn = p2.shape[1] - 1
    out = torch.empty((ncol, kn), dtype=torch.float32, device=q.device)
    _mappm_raw(p1, (km + 1, 1), q, (km, 1), p2, (kn + 1, 1), out, (kn, 1), ncol, km, kn, int(iv), int(kord),
               mode)
    return _ret(out, was_np)


def _mappm_raw(p1, s1, q, sq, p2, s2, out, so, ncol, km, kn, iv, kord, mode=EXACT):
    lib = _lib.load()
    if mode == FAST:
        _mappm_fast_raw(p1, (0,) + tuple(s1), q, (0,) + tuple(sq), p2, (0,) + tuple(s2), out, (0,) + tuple(so), 1, ncol, km,
                        kn, iv, kord)
        return
    ws, nbytes = None, 0
    if kord > 6:
        nbytes = lib.b200cs_mappm_workspace_bytes(ncol, km)
        ws = torch.empty(max(nbytes, 1), dtype=torch.uint8, device=q.device)
    rc = lib.b200cs_mappm(p1.data_ptr(), s1[0], s1[1], q.data_ptr(), sq[0], sq[1], p2.data_ptr(), s2[0], s2[1],
                          out.data_ptr(), so[0], so[1], ncol, km, kn, iv, kord,
                          ws.data_ptr() if ws is not None else None, nbytes, _stream())
    _lib.check(rc, "mappm")


def _mappm_fast_raw(p1, s1, q, sq, p2, s2, out, so, nbatch, ncol, km, kn, iv, kord):
    arr = lambda t: (ctypes.c_int64 * 3)(*t)  # noqa: E731
    rc = _lib.load().b200cs_mappm_fast(p1.data_ptr(), arr(s1), q.data_ptr(), arr(sq), p2.data_ptr(), arr(s2),
                                       out.data_ptr(), arr(so), nbatch, ncol, km, kn, iv, kord, _stream())
    _lib.check(rc, "mappm_fast")


def regrid_vertical(p_in, f_in, p_out, iv=1, kord=1, z_axis=-3, mode=EXACT):
    """cubedsphere/regridz.py:223-338 on arrays with the vertical axis at ``z_axis`` and all other
    axes equal.  The restart layout ``[..., z, y, x]`` is consumed in place through strides (no
    transposes); output float32 like mappm's.  ``mode=FAST`` (kord <= 3, strictly increasing
    ``p_out``) runs the throughput-oriented machine: within 1e-6 relative of EXACT."""
    pi, was_np = _to_dev(p_in)
    fi, _ = _to_dev(f_in)
    po, _ = _to_dev(p_out)
    nd = fi.dim()
    z = z_axis % nd
    if pi.shape[z] - 1 != fi.shape[z]:
        raise ValueError("f_in must have a vertical dimension one shorter than p_in")
    rest = lambda t: tuple(s for i, s in enumerate(t.shape) if i != z)  # noqa: E731
    if rest(pi) != rest(fi) or rest(po) != rest(fi):
        raise ValueError("All dimensions except vertical must be same size for p_in, f_in and p_out")
    # bring to [outer, z, inner] contiguous, fp32 (f2py casts)
    outer = int(np.prod(fi.shape[:z], dtype=np.int64))
    inner = int(np.prod(fi.shape[z + 1:], dtype=np.int64))
    pi, fi, po = (t.to(torch.float32).contiguous() for t in (pi, fi, po))
    km, kn = fi.shape[z], po.shape[z] - 1
    oshape = list(fi.shape)
    oshape[z] = kn
    out = torch.empty(oshape, dtype=torch.float32, device=fi.device)
    if mode == FAST and outer <= 65535:  # all slabs in one launch
        _mappm_fast_raw(pi, ((km + 1) * inner, 1, inner), fi, (km * inner, 1, inner), po, ((kn + 1) * inner, 1, inner),
                        out, (kn * inner, 1, inner), outer, inner, km, kn, int(iv), int(kord))
        return _ret(out, was_np)
    for b in range(outer):  # each outer slab is a [z, inner] matrix: col stride 1, level stride inner
        off = lambda t, nz: t.view(outer, nz, inner)[b]  # noqa: E731
        _mappm_raw(off(pi, km + 1), (1, inner), off(fi, km), (1, inner), off(po, kn + 1), (1, inner),
                   off(out, kn), (1, inner), inner, km, kn, int(iv), int(kord), mode)
    return _ret(out, was_np)


def mask_weights(weights, p_coarse_on_fine, phalf_fine, mask_mode="phalf"):
    """cubedsphere/regridz.py:200-220 for ``[..., z, y, x]`` arrays; weights ``[..., y, x]``.
    mask_mode 'phalf' compares phalf_coarse_on_fine[k+1], 'pfull' pfull_coarse_on_fine[k]."""
    w, was_np = _to_dev(weights)
    pc, _ = _to_dev(p_coarse_on_fine)
    pf, _ = _to_dev(phalf_fine)
    dt = _promote(pc, pf)
    pc, pf, w = pc.to(dt).contiguous(), pf.to(dt).contiguous(), w.contiguous()
    km = pf.shape[-3] - 1
    ncol = pf.shape[-2] * pf.shape[-1]
    lead = tuple(pf.shape[:-3])
    want = lead + (km + 1 if mask_mode == "phalf" else km,) + tuple(pf.shape[-2:])
    if mask_mode in ("phalf", "pfull") and tuple(pc.shape) != want:
        raise ValueError(f"mask_weights: the coarse pressure must have shape {want}, got {tuple(pc.shape)}")
    if tuple(w.shape[-2:]) != tuple(pf.shape[-2:]):
        raise ValueError("mask_weights: weights and pressures must share their horizontal extents")
    nb0 = int(np.prod(lead, dtype=np.int64))
    w = w.expand(lead + tuple(w.shape[-2:])).contiguous()
    out = torch.empty(lead + (km,) + tuple(pf.shape[-2:]), dtype=w.dtype, device=w.device)
    rc = _lib.load().b200cs_mask_weights(w.data_ptr(), _DT[w.dtype], pc.data_ptr(), pf.data_ptr(), _DT[dt],
                                         out.data_ptr(), nb0, km, ncol, _MASK[mask_mode], _stream())
    _lib.check(rc, "mask_weights")
    return _ret(out, was_np)


def pressure_coarsen(fields, delp, weights, coarsening_factor, toa_pressure=300.0, mask_mode="phalf",
                     edge_lines=False, iv=1, kord=1, mode=EXACT, return_delp_coarse=True, seq_window=False,
                     out=None, peer_ptrs=None):
    """Fused pressure-level coarse-graining (one pass over the fine data).

    Equivalent to regrid_to_area_weighted_pressure + weighted_block_average with the masked
    weights (cubedsphere/regridz.py:31-220, cubedsphere/coarsen_restarts.py:485-534,934-959), or
    with mask_mode='none' to Grid.pressure_level_average of fine_res_budget
    (workflows/fine_res_budget/budget/budgets.py:49-57).

    fields: list of ``[B, km, ny, nx]`` arrays (B = tiles); delp same shape; weights ``[B, ny, nx]``.
    edge_lines: windows are 1 x f (the rows are the kept lines of a D-grid wind component).
    seq_window: the window axis was not the contiguous axis of the reference's arrays (v lines are
    gathered columns), so sums follow numpy's sequential order for strided reductions.
    out: optional preallocated device tensor ``[nfields + 1, B, km, cy, cx]`` that receives the
    coarse fields and, last, the coarse delp (the returned tensors are views of it).
    peer_ptrs: device pointers of this rank's slot in every peer's gather buffer
    (fv3net_b200.gather.CoarseGather.peer_ptrs): the kernel also stores the coarse outputs there
    over NVLink (fused all-gather; FAST mode, fp32, 8 x 8 windows only).
    Returns (list of coarse fields ``[B, km, cy, cx]``, coarse delp or None)."""
    f = int(coarsening_factor)
    was_np = not isinstance(delp, torch.Tensor)
    on_host = was_np or delp.device.type != "cuda"
    if on_host and (out is not None or peer_ptrs is not None):
        raise ValueError("pressure_coarsen: out= and peer_ptrs= name device buffers; pass device-resident inputs with them")
    if on_host and getattr(delp, "ndim", 0) == 4 and delp.shape[0] > 1 and len(fields) > 0:
        return _pressure_coarsen_from_host(fields, delp, weights, f, toa_pressure, mask_mode, edge_lines, iv, kord,
                                           mode, return_delp_coarse, seq_window, was_np)
    d, _ = _to_dev(delp)
    w, _ = _to_dev(weights)
    fs = [_to_dev(x)[0] for x in fields]
    if d.dim() != 4 or w.dim() != 3:
        raise ValueError("pressure_coarsen: delp must be [B, km, ny, nx] and weights [B, ny, nx]")
    nb0, km, ny, nx = d.shape
    fdt = _promote(*fs) if fs else torch.float32
    fs = [x.to(fdt) for x in fs]

    def strided(t):
        if t.stride(-1) != 1:
            t = t.contiguous()
        return t

    d, w = strided(d), strided(w)
    fs = [strided(x) for x in fs]
    if fs:
        s0 = tuple(fs[0].stride()[:3])
        fs = [x if tuple(x.stride()[:3]) == s0 else x.contiguous() for x in fs]
        if any(tuple(x.stride()[:3]) != tuple(fs[0].stride()[:3]) for x in fs):
            fs = [x.contiguous() for x in fs]
    for x in fs:
        if tuple(x.shape) != (nb0, km, ny, nx):
            raise ValueError("pressure_coarsen: every field must have delp's shape")
    if tuple(w.shape) != (nb0, ny, nx):
        raise ValueError("pressure_coarsen: weights must be [B, ny, nx]")
    # the reference loops over the variables of a dataset (coarsen_restarts.py:934-959): any number,
    # including none.  The C ABI takes 1..B200CS_MAX_FIELDS fields per call.
    if len(fs) == 0 or len(fs) > _lib.MAX_FIELDS:
        if out is not None or peer_ptrs:
            raise ValueError(f"pressure_coarsen: out= / peer_ptrs= need 1..{_lib.MAX_FIELDS} fields per call")
        outs, dco = [], None
        for i in range(0, len(fs), _lib.MAX_FIELDS):
            o, dc = pressure_coarsen(fs[i:i + _lib.MAX_FIELDS], d, w, f, toa_pressure, mask_mode, edge_lines, iv, kord, mode,
                                     return_delp_coarse and i == 0, seq_window)
            outs += o
            dco = dc if i == 0 else dco
        if len(fs) == 0 and return_delp_coarse:
            dco = weighted_block_average(d, w[:, None], f, edge="x" if edge_lines else None)
        if was_np:
            return [o.cpu().numpy() for o in outs], (dco.cpu().numpy() if dco is not None else None)
        return outs, dco
    fy = 1 if edge_lines else f
    cy, cx = ny // fy, nx // f
    wdt = w.dtype
    cdt = _promote(d, w)
    if f not in (2, 4, 8):  # a coarse cell wider than one CTA tile (C3072 -> C48 is factor 64)
        if edge_lines or peer_ptrs:
            raise ValueError("pressure_coarsen: edge lines / gather buffers need a factor of 2, 4 or 8")
        all_f32 = fdt == torch.float32 and d.dtype == torch.float32 and wdt == torch.float32
        wide = mode == FAST and all_f32 and f in (16, 32, 64) and int(kord) <= 3 and bool(fs)
        if out is not None and (not wide or tuple(out.shape) != (len(fs) + 1, nb0, km, cy, cx) or out.dtype != torch.float32
                                or not out.is_contiguous() or out.device != d.device):
            raise ValueError("pressure_coarsen: out= with a factor above 8 needs the FAST fp32 wide kernel and a "
                             "contiguous [nfields + 1, B, km, cy, cx] float32 device tensor")
        if wide:
            outs, dco = _pressure_coarsen_wide(fs, d, w, f, toa_pressure, mask_mode, iv, kord, return_delp_coarse, out)
        else:
            outs, dco = _pressure_coarsen_unfused(fs, d, w, f, toa_pressure, mask_mode, iv, kord,
                                                  FAST if (mode == FAST and int(kord) <= 3) else EXACT)
        if was_np:
            return [o.cpu().numpy() for o in outs], (dco.cpu().numpy() if dco is not None else None)
        return outs, dco
    if out is not None:
        if (tuple(out.shape) != (len(fs) + 1, nb0, km, cy, cx) or out.dtype != wdt or cdt != wdt
                or not out.is_contiguous() or out.device != d.device):
            raise ValueError("pressure_coarsen: out must be a contiguous [nfields + 1, B, km, cy, cx] device tensor "
                             "of the weights' dtype")
        outs, dco = [out[i] for i in range(len(fs))], out[len(fs)]
    else:
        outs = [torch.empty((nb0, km, cy, cx), dtype=wdt, device=d.device) for _ in fs]
        dco = torch.empty((nb0, km, cy, cx), dtype=cdt, device=d.device) if return_delp_coarse else None
    lib = _lib.load()
    args = (len(fs), _lib.ptrarr([x.data_ptr() for x in fs]), _DT[fdt], _lib.i64arr(fs[0].stride()[:3] if fs else (0, 0, 0)),
            d.data_ptr(), _DT[d.dtype], _lib.i64arr(d.stride()[:3]), w.data_ptr(), _DT[wdt], _lib.i64arr(w.stride()[:2]),
            _lib.ptrarr([o.data_ptr() for o in outs]), dco.data_ptr() if dco is not None else None,
            nb0, km, ny, nx, fy, f, float(toa_pressure), _MASK[mask_mode], int(iv), int(kord),
            int(mode) | (_lib.SEQ_WINDOW if seq_window else 0))
    if peer_ptrs:
        nws = lib.b200cs_pressure_coarsen_gather_workspace_bytes(nb0, ny)
        ws = torch.empty(max(nws, 4), dtype=torch.uint8, device=d.device)
        rc = lib.b200cs_pressure_coarsen_gather(*args, len(peer_ptrs), _lib.ptrarr(peer_ptrs), ws.data_ptr(), nws, _stream())
    else:
        rc = lib.b200cs_pressure_coarsen(*args, _stream())
    _lib.check(rc, "pressure_coarsen")
    if was_np:
        return [o.cpu().numpy() for o in outs], (dco.cpu().numpy() if dco is not None else None)
    return outs, dco


_WIDE_WS = {}  # (device index, stream) -> workspace tensor, grown on demand


def _pressure_coarsen_wide(fs, d, w, f, toa_pressure, mask_mode, iv, kord, return_delp_coarse, out=None):
    """b200cs_pressure_coarsen_wide: FAST arithmetic, fp32, factor 16 / 32 / 64."""
    nb0, km, ny, nx = d.shape
    cy, cx = ny // f, nx // f
    lib = _lib.load()
    if out is not None:
        outs, dco = [out[i] for i in range(len(fs))], out[len(fs)]
    else:
        outs = [torch.empty((nb0, km, cy, cx), dtype=torch.float32, device=d.device) for _ in fs]
        dco = torch.empty((nb0, km, cy, cx), dtype=torch.float32, device=d.device) if return_delp_coarse else None
    nws = lib.b200cs_pressure_coarsen_wide_workspace_bytes(nb0, km, ny, nx, f)
    # the workspace is scratch between the three launches of one call: one per (device, stream), reused
    key = (d.device.index, _stream())
    ws = _WIDE_WS.get(key)
    if ws is None or ws.numel() < nws:
        ws = _WIDE_WS[key] = torch.empty(max(nws, 4), dtype=torch.uint8, device=d.device)
    rc = lib.b200cs_pressure_coarsen_wide(
        len(fs), _lib.ptrarr([x.data_ptr() for x in fs]), _lib.i64arr(fs[0].stride()[:3]), d.data_ptr(),
        _lib.i64arr(d.stride()[:3]), w.data_ptr(), _lib.i64arr(w.stride()[:2]), _lib.ptrarr([o.data_ptr() for o in outs]),
        dco.data_ptr() if dco is not None else None, nb0, km, ny, nx, f, float(toa_pressure), _MASK[mask_mode], int(iv),
        int(kord), ws.data_ptr(), nws, _stream())
    _lib.check(rc, "pressure_coarsen_wide")
    return outs, dco


def _pressure_coarsen_unfused(fs, d, w, f, toa_pressure, mask_mode, iv, kord, remap_mode=EXACT):
    """Any factor, any dtype, every kord: the reference's own chain of operators
    (cubedsphere/regridz.py:31-220 + coarsen_restarts.py:528-534), one kernel each, bit-exact with
    the oracle.  Fine-size intermediates are materialised -- the fused kernels exist to avoid this.
    ``remap_mode=FAST`` swaps only the remap for b200cs_mappm_fast (1e-6; coarse delp and masks
    stay bit-exact)."""
    dco = weighted_block_average(d, w[:, None], f)
    dup = block_upsample(dco, f, staggered=(False, False))  # cell centres, whatever the parity of the coarse extents
    phalf_c = pressure_at_interface(dup, toa_pressure)
    phalf_f = pressure_at_interface(d, toa_pressure)
    if mask_mode == "pfull":
        nb0, km = dup.shape[:2]
        ncol = dup.shape[-2] * dup.shape[-1]
        pfull = torch.empty_like(dup)
        rc = _lib.load().b200cs_pressure_at_midpoint_log(dup.contiguous().data_ptr(), pfull.data_ptr(), _DT[dup.dtype], nb0,
                                                         km, ncol, float(toa_pressure), _stream())
        _lib.check(rc, "pressure_at_midpoint_log")
        masked = mask_weights(w, pfull, phalf_f, "pfull")
    else:
        masked = mask_weights(w, phalf_c, phalf_f, mask_mode)
    outs = []
    for x in fs:
        reg = regrid_vertical(phalf_f, x, phalf_c, iv=iv, kord=kord, mode=remap_mode)
        outs.append(weighted_block_average(reg, masked, f))
    return outs, dco


_COPY_STREAMS = {}
_STAGING = {}  # device index -> _PinnedRing


class _PinnedRing:
    """Reused pinned staging slabs for host-resident (pageable) inputs.  cudaMemcpyAsync from
    pageable memory is a synchronous staged copy on one host thread; here every array slab is
    copied into a pinned slot by all host cores (b200cs_host_copy) and leaves the slot by DMA, so
    the host copy of slab i+1 overlaps the DMA of slab i and the kernels of the previous tile."""

    NSLOT = 4

    def __init__(self, dev):
        self.dev, self.bytes, self.slots, self.free = dev, 0, [], []
        self.out = None
        self.next = 0
        self.nthreads = max(1, min(32, len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)))

    def _grow(self, nbytes):
        if nbytes <= self.bytes:
            return
        for ev in self.free:
            if ev is not None:
                ev.synchronize()
        self.bytes = int(nbytes)
        self.slots = [torch.empty(self.bytes, dtype=torch.uint8).pin_memory() for _ in range(self.NSLOT)]
        self.free = [None] * self.NSLOT

    def out_slab(self, nbytes):
        """pinned slab for results on their way to the caller's numpy arrays; its contents are consumed
        before the call that asked for it returns"""
        if self.out is None or self.out.numel() < nbytes:
            self.out = torch.empty(int(nbytes), dtype=torch.uint8).pin_memory()
        return self.out

    def upload(self, src, dst, stream):
        """src: C-contiguous host tensor (pageable or pinned); dst: device tensor of the same dtype
        and element count; the DMA is queued on ``stream``.  Returns once the host copy is done."""
        nbytes = src.numel() * src.element_size()
        if src.is_pinned():
            with torch.cuda.stream(stream):
                dst.copy_(src.view_as(dst), non_blocking=True)
            return
        self._grow(nbytes)
        i = self.next
        self.next = (i + 1) % self.NSLOT
        if self.free[i] is not None:
            self.free[i].synchronize()  # the DMA that last read this slot has finished
        slot = self.slots[i]
        _lib.check(_lib.load().b200cs_host_copy(slot.data_ptr(), src.data_ptr(), nbytes, self.nthreads), "host_copy")
        with torch.cuda.stream(stream):
            dst.view(-1).view(torch.uint8).copy_(slot[:nbytes], non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(stream)
        self.free[i] = ev


def _pressure_coarsen_from_host(fields, delp, weights, f, toa_pressure, mask_mode, edge_lines, iv, kord, mode,
                                return_delp_coarse, seq_window, was_np):
    """Host-resident inputs -- what the drop-in signatures receive: plain numpy arrays
    (regridz.py:31-78).  The upload dominates (1.4 GB over PCIe against < 1 ms of kernel at C384),
    so it is a three-stage pipeline per tile (the per-tile loop of the reference's recipes):
    all-core host copy into pinned slabs (_PinnedRing) | DMA on a copy stream | fused kernel."""
    dev = device()
    comp = torch.cuda.current_stream()
    key = dev.index
    if key not in _COPY_STREAMS:
        _COPY_STREAMS[key] = torch.cuda.Stream(device=dev)
        _STAGING[key] = _PinnedRing(dev)
    copy, ring = _COPY_STREAMS[key], _STAGING[key]

    def as_t(a):
        t = a if isinstance(a, torch.Tensor) else torch.as_tensor(np.asarray(a))
        if t.dtype not in _DT:
            t = t.to(torch.float64)
        return t

    hd, hw, hf = as_t(delp), as_t(weights), [as_t(x) for x in fields]
    nb = hd.shape[0]
    outs_t, dco_t = [], []
    copy.wait_stream(comp)
    for t in range(nb):
        dt = []
        for x in [hd, hw] + hf:
            src = x[t:t + 1]
            if not src.is_contiguous():
                src = src.contiguous()
            with torch.cuda.stream(copy):  # owned by the copy stream's pool; record_stream below covers the kernels
                dst = torch.empty(src.shape, dtype=src.dtype, device=dev)
            ring.upload(src, dst, copy)
            dt.append(dst)
        ev = torch.cuda.Event()
        ev.record(copy)
        comp.wait_event(ev)
        for x in dt:
            x.record_stream(comp)
        o, dc = pressure_coarsen(dt[2:], dt[0], dt[1], f, toa_pressure, mask_mode=mask_mode, edge_lines=edge_lines,
                                 iv=iv, kord=kord, mode=mode, return_delp_coarse=return_delp_coarse,
                                 seq_window=seq_window)
        outs_t.append(o)
        dco_t.append(dc)
    outs = [torch.cat([o[i] for o in outs_t]) for i in range(len(hf))]
    dco = torch.cat(dco_t) if return_delp_coarse else None
    if was_np:
        return _download_numpy(outs + ([dco] if dco is not None else []), ring, copy, comp, len(hf), dco is not None)
    return outs, dco


def _download_numpy(tensors, ring, copy, comp, nf, with_dco):
    """Device results -> fresh numpy arrays: ONE asynchronous DMA of all of them into a reused pinned
    slab, one synchronisation, then an all-core copy into the arrays handed to the caller
    (``Tensor.cpu()`` on each result is a synchronous staged copy through pageable memory: 2.3 ms per
    4.4 MB array at C384 against 0.4 ms for all five this way)."""
    sizes = [t.numel() * t.element_size() for t in tensors]
    offs = [0]
    for n in sizes:
        offs.append(offs[-1] + ((n + 255) & ~255))
    slab = ring.out_slab(offs[-1])
    copy.wait_stream(comp)
    with torch.cuda.stream(copy):
        for t, o, n in zip(tensors, offs, sizes):
            slab[o:o + n].view(t.dtype).view(t.shape).copy_(t, non_blocking=True)
            t.record_stream(copy)
    copy.synchronize()
    res = []
    lib = _lib.load()
    for t, o, n in zip(tensors, offs, sizes):
        a = np.empty(tuple(t.shape), dtype=_NP_OF[t.dtype])
        _lib.check(lib.b200cs_host_copy(a.ctypes.data, slab.data_ptr() + o, n, ring.nthreads), "host_copy")
        res.append(a)
    return res[:nf], (res[nf] if with_dco else None)


def gather_columns(a, step):
    """[..., ny, nx] -> [..., (nx-1)//step+1, ny]: every step-th x column as a contiguous line."""
    t, was_np = _to_dev(a, floating=False)
    t = t.contiguous()
    lead = tuple(t.shape[:-2])
    ny, nx = t.shape[-2:]
    nl = (nx - 1) // step + 1
    out = torch.empty(lead + (nl, ny), dtype=t.dtype, device=t.device)
    nb = int(np.prod(lead, dtype=np.int64))
    lib = _lib.load()
    flat_in, flat_out = t.view(nb, ny, nx), out.view(nb, nl, ny)
    for lo in range(0, nb, 65535):  # grid.z is limited to 65535 slabs per launch
        n = min(65535, nb - lo)
        rc = lib.b200cs_gather_columns(flat_in[lo:].data_ptr(), _lib.i64arr((ny * nx, nx)), flat_out[lo:].data_ptr(),
                                       t.element_size(), n, ny, nx, int(step), _stream())
        _lib.check(rc, "gather_columns")
    return _ret(out, was_np)


def transpose_last2(a):
    return gather_columns(a, 1)


# ------------------------------------------------------------------------------------------
# column-wise post-processing of the coarse state (csrc/column_ops.cu)
# ------------------------------------------------------------------------------------------
GRAVITY, RDGAS, RVGAS = 9.80665, 287.05, 461.5  # calc/thermo/constants.py:2-4
SIGMA_BLEND = 0.9                               # cubedsphere/coarsen_restarts.py:54


def _zyx(t):
    """[..., z, y, x] tensor -> contiguous, (nb, km, ncol)"""
    t = t.contiguous()
    km = t.shape[-3]
    ncol = t.shape[-2] * t.shape[-1]
    nb = int(np.prod(t.shape[:-3], dtype=np.int64))
    return t, nb, km, ncol


def surface_pressure(delp, toa_pressure=300.0):
    """calc/thermo/vertically_dependent.py:189-208 for ``[..., z, y, x]``: nansum over z + toa."""
    d, was_np = _to_dev(delp)
    d, nb, km, ncol = _zyx(d)
    out = torch.empty(tuple(d.shape[:-3]) + tuple(d.shape[-2:]), dtype=d.dtype, device=d.device)
    rc = _lib.load().b200cs_surface_pressure(d.data_ptr(), _DT[d.dtype], _lib.i64arr((km * ncol, ncol)), out.data_ptr(),
                                             nb, km, ncol, float(toa_pressure), _stream())
    _lib.check(rc, "surface_pressure")
    return _ret(out, was_np)


def blending_weights(delp_coarse, ps_min, toa_pressure=300.0, sigma_blend=SIGMA_BLEND):
    """compute_blending_weights (cubedsphere/coarsen_restarts.py:559-648) from the coarse delp
    ``[..., z, cy, cx]`` and the block minimum of the fine surface pressure ``[..., cy, cx]``."""
    d, was_np = _to_dev(delp_coarse)
    p, _ = _to_dev(ps_min)
    dt = _promote(d, p)
    d, nb, km, ncol = _zyx(d.to(dt))
    p = p.to(dt).contiguous()
    if p.numel() != nb * ncol:
        raise ValueError("blending_weights: ps_min must be delp_coarse without its vertical axis")
    out = torch.empty_like(d)
    rc = _lib.load().b200cs_blending_weights(d.data_ptr(), p.data_ptr(), out.data_ptr(), _DT[dt], nb, km, ncol,
                                             float(toa_pressure), float(sigma_blend), _stream())
    _lib.check(rc, "blending_weights")
    return _ret(out, was_np)


def blend(weights, pressure_level, model_level):
    """cubedsphere/coarsen_restarts.py:651-667: weights * pressure_level + (1 - weights) * model_level."""
    w, was_np = _to_dev(weights)
    a, _ = _to_dev(pressure_level)
    m, _ = _to_dev(model_level)
    dt = _promote(w, a, m)
    if not (w.shape == a.shape == m.shape):
        raise ValueError("blend: weights, pressure_level and model_level must have one shape")
    w, a, m = (t.to(dt).contiguous() for t in (w, a, m))
    out = torch.empty_like(a)
    rc = _lib.load().b200cs_blend(w.data_ptr(), a.data_ptr(), m.data_ptr(), out.data_ptr(), _DT[dt], a.numel(), _stream())
    _lib.check(rc, "blend")
    return _ret(out, was_np)


def hydrostatic_balance(dz, phis, temperature, sphum, delp, toa_pressure=300.0):
    """_impose_hydrostatic_balance (cubedsphere/coarsen_restarts.py:990-1017) on ``[..., z, y, x]``
    arrays (phis ``[..., y, x]``) -> (hydrostatic DZ, adjusted phis)."""
    z, was_np = _to_dev(dz)
    ph, T, q, d = (_to_dev(a)[0] for a in (phis, temperature, sphum, delp))
    dt = _promote(z, ph, T, q, d)
    if not (z.shape == T.shape == q.shape == d.shape) or tuple(ph.shape) != tuple(z.shape[:-3]) + tuple(z.shape[-2:]):
        raise ValueError("hydrostatic_balance: DZ, T, sphum, delp must share one shape; phis lacks the vertical axis")
    z, nb, km, ncol = _zyx(z.to(dt))
    ph, T, q, d = (t.to(dt).contiguous() for t in (ph, T, q, d))
    dz_out, phis_out = torch.empty_like(z), torch.empty_like(ph)
    rc = _lib.load().b200cs_hydrostatic_balance(z.data_ptr(), ph.data_ptr(), T.data_ptr(), q.data_ptr(), d.data_ptr(),
                                                dz_out.data_ptr(), phis_out.data_ptr(), _DT[dt], nb, km, ncol,
                                                float(toa_pressure), GRAVITY, RDGAS, RVGAS, _stream())
    _lib.check(rc, "hydrostatic_balance")
    return _ret(dz_out, was_np), _ret(phis_out, was_np)
