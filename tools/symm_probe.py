"""Probe: torch symmetric memory across ranks on this box (dev tool, run under torchrun)."""
import os, torch, torch.distributed as dist
import torch.distributed._symmetric_memory as sm
rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device("cuda", local))
g = dist.group.WORLD
try:
    sm.enable_symm_mem_for_group(g.group_name)
except Exception as e:
    print("enable:", repr(e)[:200])
t = sm.empty(1 << 20, dtype=torch.float32, device=torch.device("cuda", local))
t.fill_(float(rank + 1))
h = sm.rendezvous(t, g)
print(rank, "backend", sm.get_backend(torch.device("cuda", local)), "ptrs", [hex(p) for p in h.buffer_ptrs][:4],
      "sig", h.signal_pad_size, "mc", h.has_multicast_support, flush=True)
dist.barrier(); torch.cuda.synchronize()
peer = (rank + 1) % world
r = h.get_buffer(peer, (16,), torch.float32)
print(rank, "reads peer", peer, r[:4].tolist(), flush=True)
dist.barrier(); torch.cuda.synchronize()
os._exit(0)
