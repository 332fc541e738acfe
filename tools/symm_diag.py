"""dev helper: does torch symmetric memory (peer-mapped buffers over NVLink) work on this box?"""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as symm_mem
r = int(os.environ["RANK"]); w = int(os.environ["WORLD_SIZE"]); torch.cuda.set_device(r)
dev = torch.device("cuda", r)
dist.init_process_group("nccl", device_id=dev)
t = symm_mem.empty((w, 1 << 20), dtype=torch.float32, device=dev)
hdl = symm_mem.rendezvous(t, dist.group.WORLD)
print(r, "ptrs", [hex(p) for p in hdl.buffer_ptrs], "rank", hdl.rank, "world", hdl.world_size, flush=True)
t.zero_()
hdl.barrier()
for p in range(w):  # write my slot into every peer's buffer
    peer = hdl.get_buffer(p, (w, 1 << 20), torch.float32)
    peer[r].fill_(float(r + 1))
hdl.barrier()
torch.cuda.synchronize()
print(r, "slots", [float(t[i, 0]) for i in range(w)], flush=True)
