"""Sharding one timestep of pressure-level coarse-graining over the GPUs of a box.

Every coarse cell depends only on its own f x f fine columns (all levels), so the path shards with
no data-path collective: the unit of work is one coarse ROW of one tile,
``ntile * (ny / f)`` units (288 at C384 -> C48), split into contiguous ranges, one per rank.  Each
rank runs the fused kernel on its row slabs (strided views of the fine arrays, nothing is copied)
and ONE all_gather collects the small coarse outputs (NCCL over NVLink on GPUs; gloo in the CPU
tests).  The reference's Beam pipeline (workflows/dataflow/fv3net/pipelines/coarsen_restarts/
pipeline.py:42-109) fans out over timesteps instead; bench.py --gpus N measures that mode too.
"""
import torch
import torch.distributed as dist


def plan(ntile, ncy, world):
    """Contiguous unit ranges [(lo, hi)] per rank over the ntile*ncy (tile, coarse row) units."""
    units = ntile * ncy
    per = -(-units // world)
    return [(min(r * per, units), min((r + 1) * per, units)) for r in range(world)]


def segments(lo, hi, ncy):
    """Splits the unit range [lo, hi) into per-tile pieces (tile, first coarse row, end coarse row)."""
    out = []
    while lo < hi:
        tile, cy0 = divmod(lo, ncy)
        cy1 = min(ncy, cy0 + (hi - lo))
        out.append((tile, cy0, cy1))
        lo += cy1 - cy0
    return out


def pressure_coarsen_sharded(fields, delp, weights, coarsening_factor, toa_pressure=300.0, mask_mode="phalf",
                             compute=None, group=None, **kw):
    """``core.pressure_coarsen`` on this rank's share of the (tile, coarse row) units, then one
    all_gather of the coarse values.  Every rank passes the same full arrays (or arrays whose other
    ranks' rows are never touched) and receives the full coarse result:
    (list of ``[B, km, cy, cx]`` tensors, coarse delp).  ``compute`` defaults to
    ``core.pressure_coarsen`` (injectable for the CPU tests of this host logic)."""
    if compute is None:
        from . import core

        compute = core.pressure_coarsen
    f = int(coarsening_factor)
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    nb, km, ny, nx = delp.shape
    ncy, ncx = ny // f, nx // f
    ranges = plan(nb, ncy, world)
    per = max(hi - lo for lo, hi in ranges)
    nf = len(fields)
    lo, hi = ranges[rank]
    pieces = []
    for tile, cy0, cy1 in segments(lo, hi, ncy):
        rows = slice(cy0 * f, cy1 * f)
        outs, dco = compute([x[tile:tile + 1, :, rows] for x in fields], delp[tile:tile + 1, :, rows],
                            weights[tile:tile + 1, rows], f, toa_pressure, mask_mode=mask_mode, **kw)
        parts = [torch.as_tensor(o) for o in list(outs) + [dco]]
        dt = parts[0].dtype
        for q in parts[1:]:  # fp32 weights with fp64 delp: coarse fields are fp32, coarse delp fp64
            dt = torch.promote_types(dt, q.dtype)
        t = torch.stack([q.to(dt) for q in parts])                          # [nf+1, 1, km, rows, ncx]
        pieces.append(t[:, 0].permute(2, 0, 1, 3))                          # [rows, nf+1, km, ncx]
    dev = pieces[0].device if pieces else (delp.device if isinstance(delp, torch.Tensor) else torch.device("cpu"))
    # every rank derives the gathered dtype from the inputs alone (ranks without units must agree)
    gdt = torch.float32
    for q in list(fields) + [delp, weights]:
        qd = q.dtype if isinstance(q, torch.Tensor) else torch.as_tensor(q[:0]).dtype
        if qd == torch.float64:
            gdt = torch.float64
    mine = torch.zeros((per, nf + 1, km, ncx), dtype=gdt, device=dev)
    if pieces:
        mine[: hi - lo] = torch.cat(pieces).to(gdt)
    if world > 1:
        gathered = torch.empty((world * per,) + tuple(mine.shape[1:]), dtype=mine.dtype, device=dev)
        dist.all_gather_into_tensor(gathered, mine, group=group)
        gathered = gathered.view((world,) + tuple(mine.shape))
    else:
        gathered = mine[None]
    full = torch.cat([gathered[r, : ranges[r][1] - ranges[r][0]] for r in range(world)])  # [units, nf+1, km, ncx]
    full = full.reshape(nb, ncy, nf + 1, km, ncx).permute(2, 0, 3, 1, 4).contiguous()      # [nf+1, B, km, cy, cx]
    return [full[i] for i in range(nf)], full[nf]
