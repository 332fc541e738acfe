"""Peer-mapped gather buffers for the fused coarsen + all-gather kernel
(b200cs_pressure_coarsen_gather, include/b200cs.h).

Every rank owns one buffer ``[world, nfields + 1, B, km, cy, cx]`` allocated as torch symmetric
memory (cuMem allocations exchanged over the process group, mapped into every peer's address
space over NVLink).  The coarsening kernel of rank r stores its coarse outputs straight into slot
r of every rank's buffer while it computes, so no collective kernel runs and no SM is taken from
the coarsening; ``barrier()`` (a device-side signal exchange on the current stream) orders readers
after all writers.  PyTorch is only the allocator / rendezvous carrier here.
"""
import torch
import torch.distributed as dist


class CoarseGather:
    def __init__(self, nfields, nb, km, cy, cx, device, group=None):
        import torch.distributed._symmetric_memory as symm_mem

        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        if self.world > 8:
            raise ValueError("the fused gather addresses at most 8 peers (one NVSwitch domain)")
        self.slot_shape = (nfields + 1, nb, km, cy, cx)
        self.buffer = symm_mem.empty((self.world,) + self.slot_shape, dtype=torch.float32, device=device)
        self.handle = symm_mem.rendezvous(self.buffer, self.group)
        slot_bytes = self.buffer[0].numel() * 4
        # this rank's slot inside every peer's buffer (its own included: that one is a local pointer)
        self.peer_ptrs = [int(p) + self.rank * slot_bytes for p in self.handle.buffer_ptrs]
        self._peer_slots = None

    @property
    def mine(self):
        """this rank's slot of its own buffer: [nfields + 1, B, km, cy, cx]"""
        return self.buffer[self.rank]

    def barrier(self, channel=0):
        self.handle.barrier(channel=channel)

    def push(self, local, streams):
        """Copy-engine variant: ``local`` ([nfields + 1, B, km, cy, cx], this rank's finished outputs) goes into slot
        ``rank`` of every rank's buffer with one peer-to-peer cudaMemcpyAsync per destination, round-robin over
        ``streams`` -- DMA engines over NVLink, no SM is involved, so the copies of step i run entirely under the
        coarsening kernel of step i+1."""
        if self._peer_slots is None:
            shape = tuple(self.buffer.shape)
            self._peer_slots = [self.handle.get_buffer(p, shape, torch.float32)[self.rank] for p in range(self.world)]
        for p, dst in enumerate(self._peer_slots):
            with torch.cuda.stream(streams[p % len(streams)]):
                dst.copy_(local, non_blocking=True)
