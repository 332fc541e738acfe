#!/usr/bin/env python
"""bench.py -- C384->C48 pressure-level coarse-graining throughput on B200.

Workload (BASELINE.json metric, configs[2]): the A-grid part of
`_coarse_grain_fv_core_on_pressure` with coarsen_agrid_winds=True -- delp + N=4 fields
(W, T, ua, va), 6 tiles x 79 levels x 384 x 384 fp32, factor 8, extrapolate=False, kord=1, iv=1:
coarse delp -> interface pressures -> PPM remap of each field onto the coarse pressure levels ->
below-surface mask -> masked block average, ONE fused kernel launch per step.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--mode fast|exact] [--impl reference]

--mode fast (default): the B200CS_FAST kernel -- masks, level search and coarse delp bit-identical to
the reference arithmetic, PPM coefficients within 1e-6 of the field scale (the north-star tolerance,
asserted in tests/test_gpu_parity.py); --mode exact: the bit-exact kernel (fp32, no FMA, IEEE
division).  The JSON line carries the other mode's kernel time as `other_mode`.

N>1 (under torchrun): weak scaling, every rank coarsens its own synthetic timestep (the
reference's Beam pipeline fans out over timesteps the same way), then an NCCL all-gather collects
the small coarse outputs; time = max over ranks.

Prints ONE JSON line.  `value` = fine grid points/s over all ranks with inputs resident in HBM;
`e2e` = the same through the public timestep driver from pinned HOST buffers (H2D + kernel + D2H
per step, the upload of step i+1 overlapping the kernels and the download of step i);
`roofline` = algorithmic bytes / CUDA-event time of the fused kernel vs the measured HBM copy
bandwidth (MEASURED_PEAKS.json); `cpu_baseline` = the CPU oracle (numpy + C mappm restatement,
all host cores) on a bounded sample of the same workload.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

NTILE, KM, NFINE, FACTOR, NFIELDS = 6, 79, 384, 8, 4
TOA = 300.0
METRIC = "fine grid pts/s, C384->C48 pressure-level coarsening (delp + 4 fields, 6 tiles x 79 levels)"
UNIT = "fine grid points/s"
FALLBACK_HBM_GBS = 6650.0  # /opt/skills/guides/B200_PROFILING.md fallback


def algorithmic_bytes(ntile=NTILE, km=KM, n=NFINE, f=FACTOR, nf=NFIELDS, elt=4):
    """delp + nf fields + 2-D weights read once, coarse fields + coarse delp written
    once = 4*km*(nf+1)+4 B per fine column + outputs (1608.7 B/col at km=79, nf=4)."""
    cols = ntile * n * n
    read = cols * (elt * km * (nf + 1) + elt)
    write = ntile * km * (n // f) ** 2 * elt * (nf + 1)
    return read + write


def points_per_step(ntile=NTILE, km=KM, n=NFINE, nf=NFIELDS):
    return ntile * km * n * n * nf


def synth_numpy(seed, ntile=NTILE, km=KM, ny=NFINE, nx=NFINE):
    """Seeded synthetic inputs of SURVEY.md 8d (numpy, used for the CPU sample)."""
    rng = np.random.default_rng(seed)
    yy, xx = np.meshgrid(np.linspace(0, 6, ny), np.linspace(0, 6, nx), indexing="ij")
    ps = 8.6e4 + 1.4e4 * np.sin(yy * 1.3)[None] * np.cos(xx * 1.1)[None] + rng.normal(0, 1500, size=(ntile, ny, nx))
    ps = np.clip(ps, 5.0e4, 1.04e5)
    b = np.diff(np.linspace(0, 1, km + 1) ** 1.6)
    delp = ((ps[:, None] - TOA) * b[None, :, None, None] * (1 + rng.uniform(-0.02, 0.02, size=(ntile, km, ny, nx)))).astype(np.float32)
    area = (rng.uniform(3.6e8, 8.3e8, size=(ntile, ny, nx))).astype(np.float32)
    fields = [rng.uniform(180, 320, size=(ntile, km, ny, nx)).astype(np.float32)] + \
             [rng.normal(0, 10, size=(ntile, km, ny, nx)).astype(np.float32) for _ in range(NFIELDS - 1)]
    return delp, area, fields


def synth_device(torch, seed, dev):
    """Same distribution generated on the device (generating 1.4 GB with numpy would dominate the run)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    yy = torch.linspace(0, 6, NFINE, device=dev)[:, None]
    xx = torch.linspace(0, 6, NFINE, device=dev)[None, :]
    ps = 8.6e4 + 1.4e4 * torch.sin(yy * 1.3) * torch.cos(xx * 1.1) + 1500 * torch.randn((NTILE, NFINE, NFINE), device=dev, generator=g)
    ps = ps.clamp(5.0e4, 1.04e5)
    b = torch.diff(torch.linspace(0, 1, KM + 1, device=dev) ** 1.6)
    noise = 1 + 0.04 * (torch.rand((NTILE, KM, NFINE, NFINE), device=dev, generator=g) - 0.5)
    delp = ((ps[:, None] - TOA) * b[None, :, None, None] * noise).float().contiguous()
    area = (3.6e8 + 4.7e8 * torch.rand((NTILE, NFINE, NFINE), device=dev, generator=g)).float()
    fields = [(180 + 140 * torch.rand((NTILE, KM, NFINE, NFINE), device=dev, generator=g)).float()]
    fields += [(10 * torch.randn((NTILE, KM, NFINE, NFINE), device=dev, generator=g)).float() for _ in range(NFIELDS - 1)]
    return delp, area, fields


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)], stdout=subprocess.PIPE, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, smax, reasons = [], None, set()
        for ln in self.lines:
            p = [x.strip() for x in ln.split(",")]
            if len(p) < 9:
                continue
            try:
                sm.append(float(p[1]))
                smax = float(p[2])
            except ValueError:
                continue
            for name, val in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), p[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": smax, "reasons": sorted(reasons),
                "samples": len(sm)}


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    try:
        return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    except Exception:
        return FALLBACK_HBM_GBS, "fallback (B200_PROFILING.md 6.65 TB/s)"


def cpu_reference_step(sample, nthreads):
    """One pass of the reference algorithm (CPU oracle: numpy restatement of regridz.py /
    coarsen.py + the C restatement of mappm.f90) over the bounded sample."""
    from oracle import vcm_ref as R

    delp, area, fields = sample
    d = {f"f{i}": x for i, x in enumerate(fields)}
    reg, masked = R.regrid_to_area_weighted_pressure(d, delp, area[:, None], TOA, FACTOR, False, nthreads=nthreads)
    outs = [R.weighted_block_average(reg[k], masked, FACTOR) for k in d]
    outs.append(R.weighted_block_average(delp, area[:, None], FACTOR))
    return outs


SLAB_TILE, SLAB_ROWS = 2, (64, 80)  # tile 2, fine rows 64..79 = coarse rows 8, 9


def check_fast_slab(outs, dco, fields, delp, area):
    """Parity of the launch that is being timed: the FAST kernel's outputs for one slab of the bench
    input (tile 2, two coarse rows, all 79 levels, all fields) against the CPU oracle run on that
    slab.  Coarse delp must be bit-identical; fields within 1e-6 of sum|w x| / sum w of the coarse
    cell and level (the oracle run on |x|), identical NaN patterns.  Returns the worst ratio."""
    from oracle import vcm_ref as R

    R.lib()
    t, (r0, r1) = SLAB_TILE, SLAB_ROWS
    c0, c1 = r0 // FACTOR, r1 // FACTOR
    h = lambda x: x[t:t + 1, ..., r0:r1, :].cpu().numpy()
    d, a, fs = h(delp), h(area), [h(x) for x in fields]
    nthreads = os.cpu_count() or 1
    names = [f"f{i}" for i in range(len(fs))]
    reg, masked = R.regrid_to_area_weighted_pressure(dict(zip(names, fs)), d, a[:, None], TOA, FACTOR, False, nthreads=nthreads)
    exp = [R.weighted_block_average(reg[k], masked, FACTOR) for k in names]
    rega, _ = R.regrid_to_area_weighted_pressure({k: np.abs(x) for k, x in zip(names, fs)}, d, a[:, None], TOA, FACTOR, False,
                                                 nthreads=nthreads)
    scale = [R.weighted_block_average(rega[k], masked, FACTOR) for k in names]
    edco = R.weighted_block_average(d, a[:, None], FACTOR)
    got_d = dco[t:t + 1, :, c0:c1].cpu().numpy()
    assert np.array_equal(got_d, edco, equal_nan=True), "coarse delp of the timed launch differs from the oracle"
    worst = 0.0
    for o, e, s_ in zip(outs, exp, scale):
        g = o[t:t + 1, :, c0:c1].cpu().numpy().astype(np.float64)
        e = e.astype(np.float64)
        assert np.array_equal(np.isnan(g), np.isnan(e)), "NaN pattern of the timed launch differs from the oracle"
        ok = ~np.isnan(e)
        worst = max(worst, float((np.abs(g - e)[ok] / (s_.astype(np.float64)[ok] + 1e-300)).max()))
    return worst


def cpu_sample(seed):
    # one of the six tiles, all 79 levels, all 4 fields: 1/6 of the C384 workload
    delp, area, fields = synth_numpy(seed, ntile=1, km=KM, ny=NFINE, nx=NFINE)
    pts = 1 * KM * NFINE * NFINE * NFIELDS
    return (delp, area, fields), pts, "1 tile x 79 levels x 384 x 384 columns x 4 fields (1/6 of the C384 workload)"


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import vcm_ref as R

    R.lib()
    cores = os.cpu_count() or 1
    sample, pts, desc = cpu_sample(0)
    for _ in range(max(args.warmup, 1)):
        cpu_reference_step(sample, cores)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        cpu_reference_step(sample, cores)
    dt = (time.perf_counter() - t0) / args.steps
    val = pts / dt
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C384->C48 pressure-level fv_core A-grid (delp + W,T,ua,va), f=8, kord=1, iv=1, "
                               "extrapolate=False; each step = bounded sample: " + desc},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
        "note": "reference stack (xarray/dask/xgcm + Fortran mappm) cannot run in this image; this is the CPU oracle "
                "port: numpy block sums (single thread, as the reference forces dask single-threaded) + C mappm "
                "with OpenMP over columns on all host cores",
    }))


def bind_to_gpu_numa_node(index):
    """Pins this process to the CPUs NVML reports as local to its GPU, so that the pinned host
    buffers of the end-to-end leg are first-touched on that NUMA node (with 8 ranks on one box the
    uploads otherwise all pull from one node's memory)."""
    try:
        import pynvml

        pynvml.nvmlInit()
        h = pynvml.nvmlDeviceGetHandleByIndex(index)
        words = pynvml.nvmlDeviceGetCpuAffinity(h, (os.cpu_count() + 63) // 64)
        cpus = [64 * i + b for i, w in enumerate(words) for b in range(64) if (w >> b) & 1]
        cpus = [c for c in cpus if c in os.sched_getaffinity(0)]
        if cpus:
            os.sched_setaffinity(0, cpus)
    except Exception:
        pass


def run_c3072(args):
    """BASELINE configs[4] / SURVEY.md 8d config 5: one C3072 -> C48 timestep (factor 64), delp + 4
    fields, fp32, 6 tiles x 79 levels x 3072 x 3072 = 89.7 GB of input, sharded by (tile, coarse
    row) over the ranks with no data-path collective (fv3net_b200/shard.py); an asynchronous NCCL
    all_gather collects the coarse outputs (21.8 MB in total).  Strong scaling: the work is fixed."""
    import torch
    import torch.distributed as dist

    from fv3net_b200 import core, shard

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n, f, km, nf = 3072, 64, KM, NFIELDS
    ncy = n // f
    ranges = shard.plan(NTILE, ncy, world)
    per = max(hi - lo for lo, hi in ranges)
    segs = shard.segments(*ranges[rank], ncy)
    g = torch.Generator(device=dev).manual_seed(3000 + rank)
    b = torch.diff(torch.linspace(0, 1, km + 1, device=dev) ** 1.6)
    data = []
    for tile, cy0, cy1 in segs:  # this rank's row slabs only: [1, km, rows, n]
        rows = (cy1 - cy0) * f
        yy = torch.linspace(0, 6, n, device=dev)[cy0 * f:cy1 * f, None]
        xx = torch.linspace(0, 6, n, device=dev)[None, :]
        ps = (8.6e4 + 1.4e4 * torch.sin(yy * 1.3 + tile) * torch.cos(xx * 1.1)
              + 1500 * torch.randn((1, rows, n), device=dev, generator=g)).clamp(5.0e4, 1.04e5)
        delp = torch.empty((1, km, rows, n), dtype=torch.float32, device=dev)
        for k in range(km):  # level by level: no 3 GB temporaries
            delp[:, k] = (ps - TOA) * b[k] * (1 + 0.04 * (torch.rand((1, rows, n), device=dev, generator=g) - 0.5))
        area = (3.6e8 + 4.7e8 * torch.rand((1, rows, n), device=dev, generator=g)) * (384.0 / n) ** 2
        fields = [torch.empty((1, km, rows, n), dtype=torch.float32, device=dev) for _ in range(nf)]
        fields[0].uniform_(180, 320, generator=g)
        for x in fields[1:]:
            x.normal_(0, 10, generator=g)
        data.append((delp, area, fields))
    mine = torch.zeros((per, nf + 1, km, n // f), dtype=torch.float32, device=dev)
    gathered = torch.empty((world * per,) + tuple(mine.shape[1:]), dtype=torch.float32, device=dev) if world > 1 else None
    pending = [None]

    def step():
        if pending[0] is not None:
            pending[0].wait()
        row = 0
        for delp, area, fields in data:
            outs, dco = core.pressure_coarsen(fields, delp, area, f, TOA, mask_mode="phalf", mode=core.FAST)
            t = torch.stack(list(outs) + [dco])[:, 0].permute(2, 0, 1, 3)  # [rows_c, nf+1, km, cx]
            mine[row:row + t.shape[0]] = t
            row += t.shape[0]
        if world > 1:
            pending[0] = dist.all_gather_into_tensor(gathered, mine, async_op=True)

    def drain():
        if pending[0] is not None:
            pending[0].wait()
            pending[0] = None

    for _ in range(max(args.warmup, 3)):
        step()
    drain()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    steps = min(args.steps, 10)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        step()
    drain()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    ms_per_step = ms / steps
    points = NTILE * km * n * n * nf
    alg = algorithmic_bytes(NTILE, km, n, f, nf)
    peak, peak_src = measured_peak()
    if rank == 0:
        print(json.dumps({
            "metric": "fine grid pts/s, C3072->C48 pressure-level coarsening (delp + 4 fields, 6 tiles x 79 levels)",
            "value": points / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": "C3072->C48 (factor 64) pressure-level coarsening, delp + 4 fields, 89.7 GB, "
                                   "sharded by (tile, coarse row); wide-window fused kernel (FAST arithmetic)",
                       "l2": "inputs larger than L2, no flush needed",
                       "collective": "none" if world == 1 else "NCCL all_gather of the coarse outputs, asynchronous"},
            "clocks": clocks, "e2e": None, "gpu_launches": steps * len(data) * 3,
            "roofline": {"bound": "hbm", "achieved": alg / (ms_per_step * 1e-3) / 1e9 / world, "peak": peak, "unit": "GB/s",
                         "frac": alg / (ms_per_step * 1e-3) / 1e9 / world / peak, "traffic": None, "peak_source": peak_src,
                         "kernel": "wba_wide_f32_kernel<64> + pressure_coarsen_sweep_kernel<4,8,WIDE> + wide_finalize_kernel",
                         "algorithmic_bytes_per_launch": alg, "note": "per-GPU average over the three kernels of a step"},
            "cpu_baseline": None}))
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="fast", choices=["exact", "fast"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--workload", default="c384", choices=["c384", "c3072"],
                    help="c384 (default): BASELINE.json's metric, one C384 timestep per GPU (weak scaling); "
                         "c3072: BASELINE configs[4], ONE C3072 -> C48 timestep (factor 64, delp + 4 fields, 89.7 GB) "
                         "sharded by (tile, coarse row) over the GPUs (strong scaling)")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl"],
                    help="N>1: 'peer' = the coarsening kernel stores its coarse outputs into every rank's gather "
                         "buffer over NVLink peer memory (fused all-gather); 'nccl' = asynchronous NCCL all_gather")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup

    if args.impl == "reference":
        run_reference(args)
        return
    if args.workload == "c3072":
        run_c3072(args)
        return

    import torch
    import torch.distributed as dist

    from fv3net_b200 import core

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    bind_to_gpu_numa_node(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    mode = core.EXACT if args.mode == "exact" else core.FAST

    delp, area, fields = synth_device(torch, 1000 + rank, dev)
    cdim = NFINE // FACTOR
    # N > 1: each rank's coarse outputs land in one contiguous buffer (no staging copy) and are
    # all-gathered asynchronously -- the gather of step i runs under the kernel of step i + 1
    # (double-buffered); the last one is waited for inside the timed region.
    nbuf = 2
    use_peer = world > 1 and args.collective == "peer" and mode == core.FAST
    pending = [None] * nbuf
    counter = [0]
    if use_peer:
        # fused compute + all-gather: every rank's kernel writes slot `rank` of every rank's
        # symmetric-memory buffer while it computes; a device-side barrier per step orders readers
        from fv3net_b200.gather import CoarseGather

        try:
            gath = [CoarseGather(NFIELDS, NTILE, KM, cdim, cdim, dev) for _ in range(nbuf)]
            ok = 1
        except Exception as e:  # no peer mapping on this box: every rank falls back to the NCCL gather
            print(f"[bench] rank {rank}: symmetric memory unavailable ({e}); using --collective nccl", file=sys.stderr)
            ok = 0
        flag = torch.tensor([ok], device=dev)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        use_peer = bool(flag.item())
        local = torch.empty((NFIELDS + 1, NTILE, KM, cdim, cdim), dtype=torch.float32, device=dev)
    if not use_peer:
        mine = [torch.empty((NFIELDS + 1, NTILE, KM, cdim, cdim), dtype=torch.float32, device=dev) for _ in range(nbuf)]
        gathered = [torch.empty((world,) + tuple(mine[0].shape), dtype=torch.float32, device=dev) for _ in range(nbuf)] if world > 1 else None

    def step():
        i = counter[0] % nbuf
        counter[0] += 1
        if use_peer:
            outs, dco = core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=mode, out=local,
                                              peer_ptrs=gath[i].peer_ptrs)
            gath[i].barrier()  # all writers of buffer i are done: gath[i].buffer holds every rank's outputs
            return outs, dco
        if world > 1 and pending[i] is not None:
            pending[i].wait()  # the gather that last used this buffer pair
        outs, dco = core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=mode, out=mine[i])
        if world > 1:  # C1: collect only the small coarse outputs
            pending[i] = dist.all_gather_into_tensor(gathered[i], mine[i], async_op=True)
        return outs, dco

    def drain():
        for i in range(nbuf):
            if pending[i] is not None:
                pending[i].wait()
                pending[i] = None

    for _ in range(args.warmup):
        step()
    drain()
    torch.cuda.synchronize()
    if use_peer:  # once, untimed: the peer-written buffer equals an NCCL all_gather of the same outputs
        ref = torch.empty((world,) + tuple(local.shape), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(ref, local)
        got = gath[(counter[0] - 1) % nbuf].buffer
        assert torch.equal(torch.nan_to_num(ref), torch.nan_to_num(got)), "fused gather differs from NCCL all_gather"
        del ref
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    torch.cuda.synchronize()
    if world > 1:  # rank 0 slept while the sampler started: line the ranks up again before timing
        dist.barrier()
        torch.cuda.synchronize()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(args.steps):
        step()
    drain()
    ev1.record()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = ev0.elapsed_time(ev1)
    # kernel-only time (same launches, no collective) for the roofline of the fused kernel
    k0, k1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    k0.record()
    for _ in range(args.steps):
        core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=mode)
    k1.record()
    torch.cuda.synchronize()
    kernel_ms = k0.elapsed_time(k1) / args.steps
    other = core.EXACT if mode == core.FAST else core.FAST
    core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=other)
    o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    o0.record()
    for _ in range(3):
        core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=other)
    o1.record()
    torch.cuda.synchronize()
    other_ms = o0.elapsed_time(o1) / 3
    clocks = sampler.stop() if rank == 0 else None
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / args.steps
    value = world * points_per_step() / (ms_per_step * 1e-3)

    # ---- end to end through the public API from pinned host buffers -------------------------
    e2e = None
    if not args.no_e2e:
        h_delp = delp.cpu().pin_memory()
        h_area = area.cpu().pin_memory()
        h_fields = [x.cpu().pin_memory() for x in fields]
        h2d = h_delp.numel() * 4 + h_area.numel() * 4 + sum(x.numel() * 4 for x in h_fields)
        d2h = (NFIELDS + 1) * NTILE * KM * cdim * cdim * 4

        # the public timestep driver (fv3net_b200/driver.py): every step uploads its inputs from pinned
        # host memory and delivers its coarse outputs to pinned host memory; the upload of step i+1
        # overlaps the kernels and the download of step i (two buffer sets).  The result of every
        # step is read on the host (checksum) before its buffers are reused.
        from fv3net_b200.driver import TimestepPipeline

        pipe = TimestepPipeline(NFIELDS, NTILE, KM, NFINE, NFINE, FACTOR, TOA, mask_mode="phalf", mode=mode, depth=2)
        checks = []

        def e2e_run(n):
            for cf, cd in pipe.run((h_fields, h_delp, h_area) for _ in range(n)):
                checks.append(float(cd[0, 0, 0, 0]) + float(cf[0][0, 0, 0, 0]))

        n_e2e = max(3, min(args.steps, 5))
        e2e_run(2)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        e2e_run(n_e2e)
        torch.cuda.synchronize()
        e_ms = (time.perf_counter() - t0) * 1e3 / n_e2e
        assert len(set(checks)) == 1, "the timestep driver returned different results for identical inputs"
        del pipe
        if world > 1:
            t = torch.tensor([e_ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e_ms = float(t.item())
        e2e = {"value": world * points_per_step() / (e_ms * 1e-3), "unit": UNIT, "h2d_bytes_per_step": h2d,
               "d2h_bytes_per_step": d2h, "ms_per_step": e_ms, "steps": n_e2e}
        del h_delp, h_area, h_fields

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    peak, peak_src = measured_peak()
    abytes = algorithmic_bytes()
    achieved = abytes / (kernel_ms * 1e-3) / 1e9
    traffic = None
    tpath = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tpath):
        try:
            traffic = json.load(open(tpath)).get(f"pressure_coarsen_{args.mode}")
        except Exception:
            traffic = None
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C384->C48 pressure-level fv_core A-grid (delp + W,T,ua,va): 6 tiles x 79 levels x "
                               "384 x 384 fp32, f=8, kord=1, iv=1, extrapolate=False, fused kernel",
                   "arithmetic_mode": args.mode, "l2": "inputs (1.42 GB) larger than L2, no flush needed",
                   "per_rank": "one synthetic timestep per GPU (weak)", "collective": ("none" if world == 1 else "fused: the coarsening kernel stores its coarse outputs (21.8 MB/rank) into every rank's gather buffer over NVLink peer memory, one device-side barrier per step" if use_peer else "NCCL all_gather of the coarse outputs (21.8 MB/rank), asynchronous: overlaps the next step")},
        "clocks": clocks,
        "other_mode": {"arithmetic_mode": "exact" if args.mode == "fast" else "fast", "kernel_ms": other_ms},
        "e2e": e2e,
        "gpu_launches": args.steps,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src,
                     "kernel": ("pressure_coarsen_sweep_kernel<NF=4,FY=8>" if args.mode == "fast"
                                else "pressure_coarsen_kernel<EXACT,NF=4,f32,f32,FX=8>"),
                     "algorithmic_bytes_per_launch": abytes, "kernel_ms": kernel_ms},
    }
    if not args.no_cpu_baseline and world == 1:
        from oracle import vcm_ref as R

        R.lib()
        cores = os.cpu_count() or 1
        sample, pts, desc = cpu_sample(0)
        cpu_reference_step(sample, cores)
        t0 = time.perf_counter()
        reps = 3
        for _ in range(reps):
            cpu_reference_step(sample, cores)
        dt = (time.perf_counter() - t0) / reps
        line["cpu_baseline"] = {"value": pts / dt, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
