"""World-size-2 gloo test (CPU) of the multi-GPU host logic: the (tile, coarse row) sharding plan,
the strided slab views and the all_gather assembly of fv3net_b200/shard.py.  The per-slab compute is
injected (the oracle) because there is no GPU here; on GPUs the same code runs core.pressure_coarsen
over NCCL (tests/test_gpu_parity.py::test_sharded_single_rank_matches_unsharded)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

from fv3net_b200 import shard  # noqa: E402


def test_plan_and_segments():
    assert shard.plan(6, 48, 8) == [(36 * r, 36 * (r + 1)) for r in range(8)]
    assert shard.plan(6, 6, 4) == [(0, 9), (9, 18), (18, 27), (27, 36)]
    assert shard.plan(1, 3, 4) == [(0, 1), (1, 2), (2, 3), (3, 3)]  # more ranks than units
    assert shard.segments(40, 100, 48) == [(0, 40, 48), (1, 0, 48), (2, 0, 4)]
    covered = []
    for lo, hi in shard.plan(6, 5, 4):
        for tile, a, b in shard.segments(lo, hi, 5):
            covered += [(tile, r) for r in range(a, b)]
    assert covered == [(t, r) for t in range(6) for r in range(5)]


def _inputs():
    rng = np.random.default_rng(3)
    nb, km, n, f = 3, 9, 24, 4
    ps = 9.0e4 + 5.0e3 * rng.normal(size=(nb, 1, n, n))
    b = np.diff(np.linspace(0, 1, km + 1) ** 1.6)
    delp = ((ps - 300.0) * b[None, :, None, None] * (1 + rng.uniform(-0.02, 0.02, size=(nb, km, n, n)))).astype(np.float32)
    area = rng.uniform(0.5, 1.0, size=(nb, n, n)).astype(np.float32)
    fields = [rng.normal(size=(nb, km, n, n)).astype(np.float32) * s for s in (1.0, 200.0)]
    return fields, delp, area, f


def _oracle_compute(fields, delp, w, f, toa, mask_mode="phalf"):
    """unfused oracle pipeline (plain repeat for the upsample: the reference's block_upsample reads
    an odd coarse extent, which row slabs can have, as an interface axis)"""
    from oracle import vcm_ref as R

    fields = [np.asarray(x) for x in fields]
    delp, w = np.asarray(delp), np.asarray(w)
    w4 = w[:, None]
    delp_c = R.weighted_block_average(delp, w4, f)
    up = np.repeat(np.repeat(delp_c, f, axis=-1), f, axis=-2)
    phalf_c = R.pressure_at_interface(up, toa)
    phalf_f = R.pressure_at_interface(delp, toa)
    pfull = R.pressure_at_midpoint_log(up, toa)
    mw = R.mask_weights(w4, pfull, phalf_c, phalf_f, extrapolate=False)
    outs = [R.weighted_block_average(R.regrid_vertical(phalf_f, x, phalf_c, nthreads=1), mw, f) for x in fields]
    return outs, delp_c


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    fields, delp, area, f = _inputs()
    outs, dco = shard.pressure_coarsen_sharded([torch.as_tensor(x) for x in fields], torch.as_tensor(delp),
                                               torch.as_tensor(area), f, 300.0, compute=_oracle_compute)
    q.put((rank, [o.numpy() for o in outs], dco.numpy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_sharded_world2_matches_unsharded():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    fields, delp, area, f = _inputs()
    exp, edco = _oracle_compute(fields, delp, area, f, 300.0)
    for rank, outs, dco in res:
        assert np.array_equal(dco, edco)
        for o, e in zip(outs, exp):
            assert np.array_equal(o, e, equal_nan=True), rank
