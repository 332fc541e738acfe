"""Host-side planning of the array layer (fv3net_b200/core.py) -- no GPU needed: how operands are
broadcast and viewed as [B0, B1, ny, nx] with element strides for the C ABI."""
import numpy as np
import pytest
import torch

from fv3net_b200 import core


def test_broadcast_lead_follows_numpy():
    cases = [[(6, 79), (6, 1)], [(79,), (6, 1)], [(), (6,)], [(6, 79), (6, 79)], [(2, 1, 5), (3, 1), (1,)]]
    for shapes in cases:
        assert core._broadcast_lead(shapes) == np.broadcast_shapes(*shapes)
    with pytest.raises(RuntimeError, match="do not broadcast"):
        core._broadcast_lead([(6, 3), (6, 2)])


def _plan(*tensors):
    lead = core._broadcast_lead([t.shape[:-2] for t in tensors])
    full = lead + tuple(tensors[0].shape[-2:])
    ts = [t if tuple(t.shape) == full else t.expand(full) for t in tensors]
    return core._canon4(ts, lead)


def test_canon4_restart_layout_and_broadcast_weights():
    o, w = torch.empty(6, 79, 16, 24), torch.empty(6, 1, 16, 24)
    ts, b0, b1, st = _plan(o, w)
    assert (b0, b1) == (6, 79)
    assert st[0] == (79 * 16 * 24, 16 * 24, 24)  # field: tile stride, level stride, row stride
    assert st[1] == (16 * 24, 0, 24)             # 2-D weights: the level stride is 0, nothing is copied
    assert ts[1].data_ptr() == w.data_ptr()
    # identical layouts collapse into one batch axis
    _, b0, b1, st = _plan(o, torch.empty(6, 79, 16, 24))
    assert (b0, b1) == (6 * 79, 1) and st[0][0] == 16 * 24


def test_canon4_time_axis_and_views():
    o = torch.empty(6, 2, 5, 8, 8)          # [tile, time, z, y, x]
    w = torch.empty(6, 1, 1, 8, 8)
    _, b0, b1, st = _plan(o, w)
    assert b0 * b1 == 60 and st[1][1] == 0
    # a strided view along x cannot be addressed with a row stride alone: planned on a contiguous copy
    v = torch.empty(4, 8, 16)[:, :, ::2]
    ts, b0, b1, st = _plan(v, torch.empty(4, 8, 8))
    assert ts[0].is_contiguous() and (b0, b1) == (4, 1) and st[0] == (64, 0, 8)
    # a view with a row pitch (columns 0..7 of 16) is used in place
    p = torch.empty(4, 8, 16)[:, :, :8]
    ts, _, _, st = _plan(p, torch.empty(4, 8, 8))
    assert ts[0].data_ptr() == p.data_ptr() and st[0][2] == 16


def test_canon4_plan_is_memoised():
    o, w = torch.empty(3, 7, 8, 8), torch.empty(3, 1, 8, 8)
    _plan(o, w)
    n = len(core._CANON_CACHE)
    for _ in range(5):
        _plan(torch.empty(3, 7, 8, 8), torch.empty(3, 1, 8, 8))
    assert len(core._CANON_CACHE) == n
