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
WORKLOAD = ("C384->C48 pressure-level fv_core A-grid (delp + W,T,ua,va): 6 tiles x 79 levels x 384 x 384 fp32, f=8, "
            "kord=1, iv=1, extrapolate=False")


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


def surface_pressure(torch, terrain, g, dev, ntile=NTILE, n=NFINE):
    """Synthetic fine surface pressure [ntile, n, n] (Pa).
    smooth:    SURVEY.md 8d -- a smooth field + 1.5 kPa of white noise (BASELINE's workload).
    orography: a smooth base with a few Gaussian mountain ranges per tile, up to 45 kPa high and
               4..10 columns wide: slopes of 3 kPa and more per column, Andes / Himalaya-like, i.e.
               neighbouring columns of one coarse cell sit on very different coarse levels.
    rough:     drawn independently per column from U(50, 104) kPa -- rougher than any terrain; the
               worst case of the layer-synchronous kernel (its lanes drift apart by tens of levels)."""
    yy = torch.linspace(0, 6, n, device=dev)[:, None]
    xx = torch.linspace(0, 6, n, device=dev)[None, :]
    if terrain == "rough":
        return 5.0e4 + 5.4e4 * torch.rand((ntile, n, n), device=dev, generator=g)
    if terrain == "orography":
        ps = 1.0e5 - 2.0e3 * torch.sin(yy * 1.3) * torch.cos(xx * 1.1) + 200 * torch.randn((ntile, n, n), device=dev, generator=g)
        iy = torch.arange(n, device=dev, dtype=torch.float32)[:, None]
        ix = torch.arange(n, device=dev, dtype=torch.float32)[None, :]
        prm = torch.rand((ntile, 6, 5), device=dev, generator=g)
        for t in range(ntile):
            for m in range(6):
                cy, cx, amp, sg, el = (float(v) for v in prm[t, m])
                sig = 4.0 + 6.0 * sg
                # an elongated range: 1 : (1 + 3 el) aspect ratio
                d2 = ((iy - cy * n) / sig) ** 2 + ((ix - cx * n) / (sig * (1 + 3 * el))) ** 2
                ps[t] -= (2.5e4 + 2.0e4 * amp) * torch.exp(-0.5 * d2)
        return ps.clamp(5.0e4, 1.04e5)
    ps = 8.6e4 + 1.4e4 * torch.sin(yy * 1.3) * torch.cos(xx * 1.1) + 1500 * torch.randn((ntile, n, n), device=dev, generator=g)
    return ps.clamp(5.0e4, 1.04e5)


def synth_delp(torch, ps, g, dev):
    b = torch.diff(torch.linspace(0, 1, KM + 1, device=dev) ** 1.6)
    noise = 1 + 0.04 * (torch.rand((ps.shape[0], KM) + tuple(ps.shape[1:]), device=dev, generator=g) - 0.5)
    return ((ps[:, None] - TOA) * b[None, :, None, None] * noise).float().contiguous()


def synth_device(torch, seed, dev, terrain="smooth"):
    """Same distribution generated on the device (generating 1.4 GB with numpy would dominate the run)."""
    g = torch.Generator(device=dev).manual_seed(seed)
    ps = surface_pressure(torch, terrain, g, dev)
    delp = synth_delp(torch, ps, g, dev)
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


def cpu_sample(seed, ntile=1, ny=NFINE):
    delp, area, fields = synth_numpy(seed, ntile=ntile, km=KM, ny=ny, nx=NFINE)
    pts = ntile * KM * ny * NFINE * NFIELDS
    frac = ntile * ny / (NTILE * NFINE)
    desc = f"{ntile} tile(s) x 79 levels x {ny} x 384 columns x 4 fields" + \
        (" (the whole C384 workload)" if frac == 1 else f" ({frac:.4f} of the C384 workload)")
    return (delp, area, fields), pts, desc


def time_cpu(sample, nthreads, reps, warm=1):
    for _ in range(warm):
        cpu_reference_step(sample, nthreads)
    t0 = time.perf_counter()
    for _ in range(reps):
        cpu_reference_step(sample, nthreads)
    return (time.perf_counter() - t0) / reps


def run_reference(args):
    """The reference arm: the reference's CPU algorithm (oracle port -- the reference stack itself
    cannot be installed here) over the WHOLE workload of our arm, all six tiles per
    step, on all host cores; plus the single-thread figure (a reference worker runs dask
    single-threaded, coarsen_restarts.py:68) on one tile."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle import vcm_ref as R

    R.lib()
    cores = os.cpu_count() or 1
    sample, pts, desc = cpu_sample(0, ntile=NTILE)
    dt = time_cpu(sample, cores, args.steps, warm=max(min(args.warmup, 2), 1))
    val = pts / dt
    one, pts1, desc1 = cpu_sample(0, ntile=1, ny=64)
    dt1 = time_cpu(one, 1, 1, warm=0)
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": WORKLOAD, "reference_step": "each step = " + desc},
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc},
        "cpu_baseline_1thread": {"value": pts1 / dt1, "unit": UNIT, "cores": 1, "kind": "port", "sample": desc1},
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


C3072_N1_CACHE = "/tmp/b200cs_c3072_n1.json"           # written by an N=1 run on this box, read by the N>1 runs after it
C3072_N1_COMMITTED = os.path.join(ROOT, "profiles", "r02_c3072_n1.json")  # the builder's own N=1 measurement


def c3072_measure(torch, dist, dev, world, rank, local_rank, steps, warmup, sample_clocks=False):
    """BASELINE configs[4] / SURVEY.md 8d config 5: one C3072 -> C48 timestep (factor 64), delp + 4
    fields, fp32, 6 tiles x 79 levels x 3072 x 3072 = 89.7 GB of input, sharded by (tile, coarse
    row) over the ranks with no data-path collective (fv3net_b200/shard.py; the reference shards
    the same way per subtile, workflows/dataflow/fv3net/pipelines/coarsen_surface_c3072.py:28-29,
    87-99); an asynchronous NCCL all_gather collects the coarse outputs (21.8 MB in total).
    Strong scaling: the work is fixed.  Returns the record on rank 0 (None elsewhere)."""
    from fv3net_b200 import core, shard

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
    # every segment's coarse outputs land, through the kernel's `out`, in their part of ONE flat
    # buffer of fixed size (the all_gather needs equal sizes on all ranks): no stack / permute /
    # copy inside the timed step; segment i = [nf + 1, 1, km, rows_i, cx] at its offset
    cx = n // f
    flat = torch.zeros(((nf + 1) * km * per * cx,), dtype=torch.float32, device=dev)
    outbuf, off = [], 0
    for _, cy0, cy1 in segs:
        cnt = (nf + 1) * km * (cy1 - cy0) * cx
        outbuf.append(flat[off:off + cnt].view(nf + 1, 1, km, cy1 - cy0, cx))
        off += cnt
    gathered = torch.empty((world,) + tuple(flat.shape), dtype=torch.float32, device=dev) if world > 1 else None
    pending = []

    def step():
        for h in pending:
            h.wait()
        pending.clear()
        for i, (delp, area, fields) in enumerate(data):
            core.pressure_coarsen(fields, delp, area, f, TOA, mask_mode="phalf", mode=core.FAST, out=outbuf[i])
        if world > 1:
            pending.append(dist.all_gather_into_tensor(gathered, flat, async_op=True))

    def drain():
        for h in pending:
            h.wait()
        pending.clear()

    for _ in range(max(warmup, 3)):
        step()
    drain()
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0 and sample_clocks:
        sampler.start()
        time.sleep(0.3)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
        torch.cuda.synchronize()
    steps = max(3, min(steps, 10))
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for _ in range(steps):
        step()
    drain()
    ev1.record()
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop() if (rank == 0 and sample_clocks) else None
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        dist.barrier()
    del data, outbuf, gathered, flat
    torch.cuda.empty_cache()
    if rank != 0:
        return None
    ms_per_step = ms / steps
    points = NTILE * km * n * n * nf
    alg = algorithmic_bytes(NTILE, km, n, f, nf)
    peak, peak_src = measured_peak()
    rec = {
        "metric": "fine grid pts/s, C3072->C48 pressure-level coarsening (delp + 4 fields, 6 tiles x 79 levels)",
        "value": points / (ms_per_step * 1e-3), "unit": UNIT, "n_gpus": world, "steps": steps,
        "warmup": max(warmup, 3), "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "C3072->C48 (factor 64) pressure-level coarsening, delp + 4 fields, 89.7 GB, "
                               "sharded by (tile, coarse row); wide-window fused kernel (FAST arithmetic)",
                   "l2": "inputs larger than L2, no flush needed",
                   "collective": "none" if world == 1 else "NCCL all_gather of the coarse outputs, asynchronous"},
        "clocks": clocks, "e2e": None, "gpu_launches": steps * len(segs) * 3,
        "roofline": {"bound": "hbm", "achieved": alg / (ms_per_step * 1e-3) / 1e9 / world, "peak": peak, "unit": "GB/s",
                     "frac": alg / (ms_per_step * 1e-3) / 1e9 / world / peak, "traffic": None, "peak_source": peak_src,
                     "kernel": "wba_wide_f32_kernel<64> + pressure_coarsen_sweep_kernel<4,8,WIDE> + wide_finalize_kernel",
                     "algorithmic_bytes_per_launch": alg, "note": "per-GPU average over the three kernels of a step"},
        "cpu_baseline": None}
    # strong-scaling reference: the N=1 time measured on THIS box by an earlier run, else the committed one
    if world == 1:
        try:
            json.dump({"ms_per_step": ms_per_step, "when": time.strftime("%Y-%m-%dT%H:%M:%SZ", time.gmtime())}, open(C3072_N1_CACHE, "w"))
        except Exception:
            pass
        rec["n1_ms_per_step"], rec["n1_source"], rec["speedup_vs_n1"] = ms_per_step, "this run", 1.0
    else:
        for path, src in ((C3072_N1_CACHE, "N=1 run on this box (" + C3072_N1_CACHE + ")"),
                          (C3072_N1_COMMITTED, "builder's N=1 measurement (profiles/r02_c3072_n1.json)")):
            try:
                n1 = float(json.load(open(path))["ms_per_step"])
                rec["n1_ms_per_step"], rec["n1_source"], rec["speedup_vs_n1"] = n1, src, n1 / ms_per_step
                break
            except Exception:
                continue
    return rec


def run_c3072(args):
    """`--workload c3072`: the C3072 strong-scaling step alone, as the JSON line."""
    import torch
    import torch.distributed as dist

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rec = c3072_measure(torch, dist, dev, world, rank, local_rank, args.steps, args.warmup, sample_clocks=True)
    if rank == 0:
        print(json.dumps(rec))
    if world > 1:
        dist.destroy_process_group()


def dropin_measure(torch, core, delp, area, fields, mode, e2e):
    """The drop-in call as a user of the reference makes it (regridz.py:31-78 + coarsen.py:183-218, fused):
    labelled arrays wrapping PLAIN numpy (pageable host memory) in, labelled numpy out, through
    fv3net_b200.cubedsphere.pressure_level_coarsen.  Every step uploads its 1.40 GB and downloads its
    coarse result; the host copy into pinned staging slabs runs on all cores under the DMA."""
    from fv3net_b200 import cubedsphere as cs
    from fv3net_b200.labelled import DataArray, Dataset

    dims = ["tile", "zaxis_1", "yaxis_2", "xaxis_1"]
    names = ["W", "T", "ua", "va"][:len(fields)]
    ds = Dataset({k: DataArray(x.cpu().numpy(), dims=dims) for k, x in zip(names, fields)})
    d = DataArray(delp.cpu().numpy(), dims=dims, name="delp")
    a = DataArray(area.cpu().numpy(), dims=["tile", "yaxis_2", "xaxis_1"], name="area")
    checks = []

    def call():
        out, dco = cs.pressure_level_coarsen(ds, d, a, TOA, FACTOR, mode=mode)
        checks.append(float(np.asarray(dco.values)[0, 0, 0, 0]) + float(np.asarray(out[names[0]].values)[0, 0, 0, 0]))

    call()
    torch.cuda.synchronize()
    n = 3
    t0 = time.perf_counter()
    for _ in range(n):
        call()
    torch.cuda.synchronize()
    ms = (time.perf_counter() - t0) * 1e3 / n
    assert len(set(checks)) == 1
    h2d = sum(int(np.asarray(ds[k].values).nbytes) for k in names) + int(np.asarray(d.values).nbytes) + int(np.asarray(a.values).nbytes)
    d2h = (len(names) + 1) * NTILE * KM * (NFINE // FACTOR) ** 2 * 4
    rec = {"value": points_per_step() / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms, "steps": n,
           "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
           "api": "fv3net_b200.cubedsphere.pressure_level_coarsen(Dataset, delp, area, toa, 8) with numpy-backed labelled arrays "
                  "(pageable memory) -> numpy-backed Dataset"}
    if e2e:
        rec["vs_pinned_pipeline"] = ms / e2e["ms_per_step"]
    return rec


def whole_timestep_measure(torch, core, delp, area, fields, mode, dev):
    """coarsen_restarts_on_pressure (coarsen_restarts.py:155-240) for ONE C384 timestep resident in HBM:
    fv_core with A-grid winds (delp, W, T, ua, va on pressure levels; phis, DZ; D-grid u, v on their
    edge lines; hydrostatic adjustment), 9 tracers on pressure levels, fv_srf_wnd.  Algorithmic
    bytes as SURVEY.md 8d counts them: every array the recipe needs read once, outputs written once."""
    from fv3net_b200 import restarts

    n, km, f, nt = NFINE, KM, FACTOR, NTILE
    g = torch.Generator(device=dev).manual_seed(5)
    U = lambda lo, hi, *sh: torch.empty(sh, device=dev).uniform_(lo, hi, generator=g)  # noqa: E731
    N = lambda sd, *sh: torch.empty(sh, device=dev).normal_(0.0, sd, generator=g)  # noqa: E731
    fv_core = {"phis": U(0, 5e4, nt, n, n), "delp": delp, "DZ": -U(20, 800, nt, km, n, n), "W": fields[1],
               "T": fields[0], "ua": fields[2], "va": fields[3],
               "u": N(10, nt, km, n + 1, n), "v": N(10, nt, km, n, n + 1)}
    tracers = {k: N(1e-3, nt, km, n, n).abs_() for k in ("sphum", "liq_wat", "rainwat", "ice_wat", "snowwat", "graupel",
                                                         "o3mr", "sgs_tke", "cld_amt")}
    srf = {"u_srf": N(5, nt, n, n), "v_srf": N(5, nt, n, n)}
    grid = {"area": area, "dx": U(2.0e4, 2.88e4, nt, n + 1, n), "dy": U(2.0e4, 2.88e4, nt, n, n + 1)}
    rs = {"fv_core.res": fv_core, "fv_tracer.res": tracers, "fv_srf_wnd.res": srf}
    E, c = 4, n // f
    fld, cfl = nt * km * n * n * E, nt * km * c * c * E
    lines = nt * km * (c + 1) * n * E
    alg = {"fv_core A-grid (delp, DZ, W, T, ua, va, phis, area)": 6 * fld + 2 * nt * n * n * E + 6 * cfl,
           "fv_core D-grid u, v (needed lines: wind + 2 delp rows + spacing)": 2 * (3 * lines + nt * (c + 1) * n * E + nt * km * (c + 1) * c * E),
           "fv_tracer (9 fields + delp, area)": 10 * fld + nt * n * n * E + 9 * cfl,
           "fv_srf_wnd": 3 * nt * n * n * E}
    total = sum(alg.values())

    def run():
        return restarts.coarsen_restarts_on_pressure(f, grid, TOA, rs, coarsen_agrid_winds=True, mode=mode)

    for _ in range(2):
        run()
    torch.cuda.synchronize()
    reps = 7
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(reps + 1)]
    ev[0].record()
    for i in range(reps):
        run()
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(reps))
    ms = ts[len(ts) // 2]
    peak, peak_src = measured_peak()
    del rs, fv_core, tracers, srf, grid
    torch.cuda.empty_cache()
    return {"ms": ms, "algorithmic_bytes": total, "bytes_by_part": alg, "achieved_gbs": total / ms / 1e6,
            "frac": total / ms / 1e6 / peak, "peak_source": peak_src,
            "fine_points_per_s": (16 * nt * km * n * n + 2 * nt * km * (n + 1) * n) / ms * 1e3,
            "what": "restarts.coarsen_restarts_on_pressure(8, grid, 300.0, {fv_core, fv_tracer, fv_srf_wnd}, "
                    "coarsen_agrid_winds=True), device resident, median of 7; mode = the headline's"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="fast", choices=["exact", "fast"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the sub-records timed after the headline (other terrains, whole timestep, C3072 strong scaling, "
                         "drop-in end-to-end): a quick kernel-only line")
    ap.add_argument("--terrain", default="smooth", choices=["smooth", "orography", "rough"],
                    help="synthetic surface pressure of the headline workload (see surface_pressure); the other two are "
                         "timed as `other_terrain`")
    ap.add_argument("--workload", default="c384", choices=["c384", "c3072"],
                    help="c384 (default): BASELINE.json's metric, one C384 timestep per GPU (weak scaling); "
                         "c3072: BASELINE configs[4], ONE C3072 -> C48 timestep (factor 64, delp + 4 fields, 89.7 GB) "
                         "sharded by (tile, coarse row) over the GPUs (strong scaling)")
    ap.add_argument("--gather", default="auto", choices=["auto", "fused", "ce"],
                    help="--collective peer: 'fused' = stores from the coarsening kernel itself, 'ce' = peer-to-peer DMA after "
                         "the kernel, under the next one (auto = ce: measured faster at every N, 0.988 vs 0.979 at N=2, 0.961 vs 0.940 at N=8)")
    ap.add_argument("--ce-streams", type=int, default=8, help="--gather ce: copy streams the peer-to-peer DMAs are spread over")
    ap.add_argument("--collective", default="peer", choices=["peer", "nccl", "none"],
                    help="N>1: 'peer' = the coarsening kernel stores its coarse outputs into every rank's gather "
                         "buffer over NVLink peer memory (fused all-gather); 'nccl' = asynchronous NCCL all_gather; "
                         "'none' = diagnosis only: no gather at all, the max over ranks of the bare kernels (how much of the "
                         "N-GPU loss is the spread between the GPUs of the box)")
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

    delp, area, fields = synth_device(torch, 1000 + rank, dev, args.terrain)
    cdim = NFINE // FACTOR
    # N > 1: each rank's coarse outputs land in one contiguous buffer (no staging copy) and are
    # all-gathered asynchronously -- the gather of step i runs under the kernel of step i + 1
    # (double-buffered); the last one is waited for inside the timed region.
    use_peer = world > 1 and args.collective == "peer" and mode == core.FAST
    # fused gather: THREE buffers, so that the kernel of step i waits for the barrier of step i-2 (a whole step old) and
    # never for the one that follows the previous kernel (with two buffers every kernel start paid the barrier latency
    # and the skew between ranks: 0.935 at N=2)
    nbuf = 3 if use_peer else 2
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

    side = torch.cuda.Stream(device=dev) if use_peer else None
    # --gather: 'fused' = the coarsening kernel itself stores into the peers (16-byte stores from its epilogue for two
    # ranks, finished rows copied by the last CTA of a tile row beyond); 'ce' = plain kernel, then one peer-to-peer DMA per
    # destination on copy streams (no SM involved), overlapped with the next step's kernel
    use_ce = use_peer and (args.gather == "ce" or (args.gather == "auto" and world > 1))
    if use_ce:
        ce_streams = [torch.cuda.Stream(device=dev) for _ in range(min(world, args.ce_streams))]
        locals_ = [local, torch.empty_like(local), torch.empty_like(local)]
    bar_done = [None] * nbuf  # event: the barrier that follows the kernel which wrote buffer i has completed

    def step():
        i = counter[0] % nbuf
        counter[0] += 1
        if use_peer:
            # Ordering, with three buffers and the barriers on a SIDE stream (off the kernels' critical path):
            #   readers of buffer i wait for the barrier after the kernel that wrote it (bar_done[i]);
            #   the kernel that REUSES buffer i (three steps later) waits for the barrier of the step after the
            #   one that wrote it -- bar_done[(i + 1) % 3], two steps old by then --, which every peer enters only
            #   after it has consumed buffer i (side-stream order).
            comp = torch.cuda.current_stream()
            j = (i + 1) % nbuf
            if bar_done[j] is not None:
                comp.wait_event(bar_done[j])
            if use_ce:
                # local buffer i was last read by the DMAs of step i-3, which finished before the barrier awaited above
                outs, dco = core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=mode, out=locals_[i])
            else:
                outs, dco = core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=mode, out=local,
                                                  peer_ptrs=gath[i].peer_ptrs)
            wrote = torch.cuda.Event()
            wrote.record(comp)
            if use_ce:
                for st in ce_streams:
                    st.wait_event(wrote)
                    if bar_done[j] is not None:
                        st.wait_event(bar_done[j])  # every peer has consumed what these DMAs overwrite
                gath[i].push(locals_[i], ce_streams)
                for st in ce_streams:
                    side.wait_stream(st)
            side.wait_event(wrote)
            with torch.cuda.stream(side):
                gath[i].barrier()  # all writers of buffer i are done: gath[i].buffer holds every rank's outputs
                bar_done[i] = torch.cuda.Event()
                bar_done[i].record(side)
            return outs, dco
        if world > 1 and pending[i] is not None:
            pending[i].wait()  # the gather that last used this buffer pair
        outs, dco = core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=mode, out=mine[i])
        if world > 1 and args.collective != "none":  # C1: collect only the small coarse outputs
            pending[i] = dist.all_gather_into_tensor(gathered[i], mine[i], async_op=True)
        return outs, dco

    def drain():
        for i in range(nbuf):
            if pending[i] is not None:
                pending[i].wait()
                pending[i] = None
        if use_peer:  # the timed region ends when the last step's outputs are complete on every rank
            torch.cuda.current_stream().wait_stream(side)

    for _ in range(args.warmup):
        step()
    drain()
    torch.cuda.synchronize()
    slab_err = None
    if rank == 0 and mode == core.FAST:
        # once, untimed: the launch that is about to be timed, checked against the CPU oracle on a slab
        # of this very input (tile 2, two coarse rows): coarse delp bit-identical, fields within 1e-6
        so, sd = core.pressure_coarsen(fields, delp, area, FACTOR, TOA, mask_mode="phalf", mode=mode)
        slab_err = check_fast_slab(so, sd, fields, delp, area)
        assert slab_err <= 1e-6, f"FAST kernel differs from the oracle on the bench slab: {slab_err:.3e} > 1e-6"
        del so, sd
    if use_peer:  # once, untimed: the peer-written buffer equals an NCCL all_gather of the same outputs
        ref = torch.empty((world,) + tuple(local.shape), dtype=torch.float32, device=dev)
        dist.all_gather_into_tensor(ref, locals_[(counter[0] - 1) % nbuf] if use_ce else local)
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

    # ---- data dependence: the same launch on the other two surface-pressure fields ----------
    other_terrain = None
    if rank == 0 and not args.no_extras:
        other_terrain = {}
        for name in ("smooth", "orography", "rough"):
            if name == args.terrain:
                continue
            g2 = torch.Generator(device=dev).manual_seed(77)
            d2 = synth_delp(torch, surface_pressure(torch, name, g2, dev), g2, dev)
            for _ in range(2):
                core.pressure_coarsen(fields, d2, area, FACTOR, TOA, mask_mode="phalf", mode=mode)
            t0_, t1_ = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0_.record()
            for _ in range(5):
                core.pressure_coarsen(fields, d2, area, FACTOR, TOA, mask_mode="phalf", mode=mode)
            t1_.record()
            torch.cuda.synchronize()
            other_terrain[name] = {"kernel_ms": t0_.elapsed_time(t1_) / 5, "vs_headline": t0_.elapsed_time(t1_) / 5 / kernel_ms}
            del d2

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

    # ---- the drop-in call as a user makes it: plain (pageable) numpy in, numpy out ------------
    e2e_dropin = None
    if world == 1 and not args.no_e2e and not args.no_extras:
        e2e_dropin = dropin_measure(torch, core, delp, area, fields, mode, e2e)

    # ---- one whole restart timestep, device resident ----------------------------------------
    whole = None
    if world == 1 and not args.no_extras:
        whole = whole_timestep_measure(torch, core, delp, area, fields, mode, dev)

    # ---- C3072 -> C48 strong scaling (BASELINE configs[4]), same process, after the headline --
    c3072 = None
    if not args.no_extras and mode == core.FAST:
        del delp, area, fields
        torch.cuda.empty_cache()
        try:
            c3072 = c3072_measure(torch, dist, dev, world, rank, local_rank, 5, 3)
        except torch.cuda.OutOfMemoryError as e:  # 89.7 GB / world per rank
            c3072 = {"unavailable": f"out of memory: {str(e)[:120]}"} if rank == 0 else None

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
        "config": {"workload": WORKLOAD, "kernel": "one fused launch per step",
                   "arithmetic_mode": args.mode, "terrain": args.terrain, "l2": "inputs (1.42 GB) larger than L2, no flush needed",
                   "per_rank": "one synthetic timestep per GPU (weak)", "collective": ("none" if world == 1 else ("peer-to-peer DMA (copy engines over NVLink, no SM) of the coarse outputs (21.8 MB/rank) into every rank's gather buffer after the kernel, under the next step's kernel; one device-side barrier per step on a side stream (three buffers)" if use_ce else "fused: the coarsening kernel stores its coarse outputs (21.8 MB/rank) into every rank's gather buffer over NVLink peer memory, one device-side barrier per step on a side stream (three buffers)") if use_peer else ("NONE (diagnosis: bare kernels, not a valid multi-GPU figure)" if args.collective == "none" else "NCCL all_gather of the coarse outputs (21.8 MB/rank), asynchronous: overlaps the next step"))},
        "clocks": clocks,
        "other_mode": {"arithmetic_mode": "exact" if args.mode == "fast" else "fast", "kernel_ms": other_ms},
        "parity_slab": None if slab_err is None else {
            "max_rel_err": slab_err, "bound": 1e-6, "what": "FAST outputs of this very launch vs the CPU oracle on tile 2, coarse "
            "rows 8-9, all levels and fields; relative to sum|w x|/sum w per coarse cell; coarse delp bit-identical"},
        "other_terrain": other_terrain,
        "whole_timestep": whole,
        "c3072_strong": c3072,
        "e2e_dropin": e2e_dropin,
        "e2e": e2e,
        "gpu_launches": args.steps,
        "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                     "traffic": traffic, "peak_source": peak_src,
                     "kernel": ("pressure_coarsen_tma_kernel<NP=2,NFL=4> (TMA tile ring, packed f32x2 PPM)" if args.mode == "fast"
                                else "pressure_coarsen_kernel<EXACT,NF=4,f32,f32,FX=8>"),
                     "algorithmic_bytes_per_launch": abytes, "kernel_ms": kernel_ms},
    }
    if not args.no_cpu_baseline and world == 1:
        from oracle import vcm_ref as R

        R.lib()
        cores = os.cpu_count() or 1
        sample, pts, desc = cpu_sample(0)
        dt = time_cpu(sample, cores, 3)
        line["cpu_baseline"] = {"value": pts / dt, "unit": UNIT, "cores": cores, "kind": "port", "sample": desc}
        # what ONE reference worker does (dask single-threaded, coarsen_restarts.py:68): BASELINE.md 3 (i)
        one, pts1, desc1 = cpu_sample(0, ntile=1, ny=64)
        dt1 = time_cpu(one, 1, 1, warm=0)
        line["cpu_baseline_1thread"] = {"value": pts1 / dt1, "unit": UNIT, "cores": 1, "kind": "port", "sample": desc1}
    print(json.dumps(line))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
