"""Timestep driver: pressure-level coarse-graining of a SEQUENCE of timesteps
whose inputs live in host memory.

The reference fans timesteps out over Beam workers, each of which opens six tile files, runs the
recipe and writes six coarse files (workflows/dataflow/fv3net/pipelines/coarsen_restarts/
pipeline.py:42-109).  On one GPU the step is bound by the PCIe upload (1.40 GB per C384 timestep
against ~1 ms of kernel), so the driver keeps that link busy: a three-stage pipeline on three
streams with ``depth`` sets of device buffers --

    copy-in   tile by tile, pinned host -> device            (stream ``h2d``)
    compute   fused kernel per tile as soon as it has landed  (the caller's current stream)
    copy-out  coarse fields + coarse delp -> pinned host     (stream ``d2h``)

-- so that the upload of timestep i+1 overlaps the kernels and the download of timestep i.
PyTorch provides pinned memory, streams and events; all arithmetic is libb200cs.
"""
import torch

from . import core


class TimestepPipeline:
    def __init__(self, nfields, nb, km, ny, nx, coarsening_factor, toa_pressure=300.0, mask_mode="phalf",
                 mode=core.FAST, depth=2, iv=1, kord=1):
        self.dev = core.device()
        self.nf, self.nb, self.km, self.ny, self.nx = int(nfields), int(nb), int(km), int(ny), int(nx)
        self.f, self.toa, self.mask_mode, self.mode, self.iv, self.kord = int(coarsening_factor), float(toa_pressure), \
            mask_mode, mode, iv, kord
        self.depth = int(depth)
        cy, cx = self.ny // self.f, self.nx // self.f
        f32 = torch.float32
        mk = lambda *s: torch.empty(s, dtype=f32, device=self.dev)  # noqa: E731
        self.d_delp = [mk(self.nb, self.km, self.ny, self.nx) for _ in range(self.depth)]
        self.d_area = [mk(self.nb, self.ny, self.nx) for _ in range(self.depth)]
        self.d_fields = [[mk(self.nb, self.km, self.ny, self.nx) for _ in range(self.nf)] for _ in range(self.depth)]
        # per tile one contiguous slot [nf + 1, 1, km, cy, cx] (what the kernel's `out` wants)
        self.d_out = [mk(self.nb, self.nf + 1, 1, self.km, cy, cx) for _ in range(self.depth)]
        self.h_out = [torch.empty((self.nb, self.nf + 1, 1, self.km, cy, cx), dtype=f32).pin_memory()
                      for _ in range(self.depth)]
        self.h2d = torch.cuda.Stream(device=self.dev)
        self.d2h = torch.cuda.Stream(device=self.dev)
        self.computed = [None] * self.depth   # event: the kernels that read slot s have run
        self.downloaded = [None] * self.depth  # event: h_out[s] is complete
        self.count = 0

    def submit(self, fields, delp, area):
        """Queues one timestep (``fields``: list of ``[nb, km, ny, nx]`` float32 host tensors,
        pinned for asynchronous copies; ``delp`` same shape; ``area`` ``[nb, ny, nx]``).  Returns a
        ticket for ``result``.  Returns as soon as the work is queued."""
        if len(fields) != self.nf:
            raise ValueError(f"expected {self.nf} fields")
        s = self.count % self.depth
        self.count += 1
        comp = torch.cuda.current_stream()
        if self.downloaded[s] is not None:
            self.downloaded[s].synchronize()  # the host buffer of this slot has been consumed by now
        if self.computed[s] is not None:
            self.h2d.wait_event(self.computed[s])  # device inputs of this slot are free again
        landed = []
        with torch.cuda.stream(self.h2d):
            for t in range(self.nb):
                self.d_delp[s][t].copy_(torch.as_tensor(delp[t]), non_blocking=True)
                self.d_area[s][t].copy_(torch.as_tensor(area[t]), non_blocking=True)
                for i in range(self.nf):
                    self.d_fields[s][i][t].copy_(torch.as_tensor(fields[i][t]), non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.h2d)
                landed.append(ev)
        for t in range(self.nb):
            comp.wait_event(landed[t])
            core.pressure_coarsen([x[t:t + 1] for x in self.d_fields[s]], self.d_delp[s][t:t + 1],
                                  self.d_area[s][t:t + 1], self.f, self.toa, mask_mode=self.mask_mode, mode=self.mode,
                                  iv=self.iv, kord=self.kord, out=self.d_out[s][t])
        ev = torch.cuda.Event()
        ev.record(comp)
        self.computed[s] = ev
        self.d2h.wait_event(ev)
        with torch.cuda.stream(self.d2h):
            self.h_out[s].copy_(self.d_out[s], non_blocking=True)
            done = torch.cuda.Event()
            done.record(self.d2h)
        self.downloaded[s] = done
        return s

    def result(self, ticket):
        """Blocks until the timestep of ``ticket`` is on the host; returns (list of coarse fields
        ``[nb, km, cy, cx]``, coarse delp) as views of the slot's pinned buffer -- valid until the
        slot is reused ``depth`` submits later."""
        self.downloaded[ticket].synchronize()
        h = self.h_out[ticket]
        return [h[:, i, 0] for i in range(self.nf)], h[:, self.nf, 0]

    def run(self, timesteps):
        """Generator over an iterable of (fields, delp, area): yields (coarse fields, coarse delp)
        in order, one timestep behind the submissions."""
        prev = None
        for fields, delp, area in timesteps:
            cur = self.submit(fields, delp, area)
            if prev is not None:
                yield self.result(prev)
            prev = cur
        if prev is not None:
            yield self.result(prev)
