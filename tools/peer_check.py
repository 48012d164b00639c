"""Peer-memory optimiser step vs NCCL all-reduce + replicated Adam (dev check, run under torchrun)."""
import os, sys
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from monogs_dev_b200 import synthetic as syn, mapping, rasterizer as rz

class A: pass
rank, world, local = bench.dist_setup(A())
dev = torch.device("cuda", local)
name = sys.argv[1] if len(sys.argv) > 1 else "cfg1_plumbing"
graph = len(sys.argv) > 2 and sys.argv[2] == "graph"

def make(peer):
    sc = syn.make_scene(name, device=dev, n_views=4)
    kfs = bench.build_window(dev, sc)
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=sc.sh_degree)
    return mapping.MappingWindow(gm, kfs, sc.bg, cuda_graph=graph, view_streams=2, peer_adam=peer), gm

wa, ga = make(False)
wb, gb = make(True)
assert wb.peer is not None
for i in range(8):
    la, lb = float(wa.step()), float(wb.step())
torch.cuda.synchronize()
wb.peer.check()
d = (ga.flat - gb.flat).abs()
rel = float((d / (1e-6 + ga.flat.abs())).median())
bad = float((d > 1e-3 * (1 + ga.flat.abs())).float().mean())
# replicas bit-identical across ranks?
chk = gb.flat.double().sum().view(1).clone()
allc = [torch.zeros_like(chk) for _ in range(world)]
dist.all_gather(allc, chk)
same = all(float(c) == float(allc[0]) for c in allc)
gz = float(gb.flat_grad.abs().max())
gb.gather_moments(); 
dm = float((ga.flat_m - gb.flat_m).abs().max()), float(ga.flat_m.abs().max())
print(f"rank {rank}: loss nccl {la:.6f} peer {lb:.6f}  median rel diff {rel:.2e}  frac>1e-3 {bad:.2e}  replicas identical {same}  grad arena max after step {gz:.1e}  moments max diff {dm[0]:.2e} of {dm[1]:.2e}", flush=True)
# one-rank overflow: starve the binning capacity of ONE view on rank 0 only.  Every rank must
# discard that step (the barrier ORs the guard over the ranks), keep its step counter, and after the
# host's re-run the replicas must still be identical and agree with the NCCL window, which never
# overflowed (advisor finding: rank-local ticks used to diverge here).
if graph and wb._lean is not None:
    wb.flush()
    if rank == 0:
        wb._lean._alloc_binning(wb.local[0], 64)
    wb._g1 = None
    for i in range(3):
        la, lb = float(wa.step()), float(wb.step())
    wb.flush(); torch.cuda.synchronize()
    wb.peer.check()
    step_b = int(gb.optimizer._step_dev.item()) if hasattr(gb.optimizer, "_step_dev") else -1
    step_a = int(ga.optimizer._step_dev.item()) if hasattr(ga.optimizer, "_step_dev") else -2
    steps = [torch.zeros(1, dtype=torch.int64, device=dev) for _ in range(world)]
    dist.all_gather(steps, torch.tensor([step_b], dtype=torch.int64, device=dev))
    chk = gb.flat.double().sum().view(1).clone()
    allc = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk)
    d = (ga.flat - gb.flat).abs()
    bad = float((d > 1e-3 * (1 + ga.flat.abs())).float().mean())
    dn = float((ga.denom - gb.denom).abs().max()) if hasattr(ga, "denom") else -1.0
    print(f"rank {rank}: after a rank-0-only overflow: loss nccl {la:.6f} peer {lb:.6f}  step counters "
          f"{[int(x) for x in steps]} (nccl window {step_a})  replicas identical "
          f"{all(float(c) == float(allc[0]) for c in allc)}  frac>1e-3 {bad:.2e}  denom max diff {dn:.0f}  "
          f"cap {wb._lean.cap[wb.local[0]]}", flush=True)
rz.set_sync_free(False)
bench.finish(world, wa) if world > 1 else None
