"""Where a multi-rank mapping step goes: G1 (renders, losses, backward) vs the overflow check vs
G2 (all-reduce, Adam, pose updates), CUDA events per replay — dev tool, run under torchrun."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from monogs_dev_b200 import synthetic as syn, mapping, rasterizer as rz

class A: pass
rank, world, local = bench.dist_setup(A())
dev = torch.device("cuda", local)
sc = syn.make_scene(bench.WORKLOAD, device=dev, n_views=bench.N_VIEWS)
kfs, host = bench.build_window(dev, mapping.render_fused, mapping.GaussianModel, None, sc)
gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=sc.sh_degree)
win = mapping.MappingWindow(gm, kfs, sc.bg, cuda_graph=True, view_streams=4, peer_adam=True)
for _ in range(6):
    win.step()
bench.barrier_sync(world)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
t = [0.0, 0.0, 0.0]; wall = 0.0
K = 30
for _ in range(K):
    for opt in (win.g.optimizer, win.kf_optimizer):
        if opt is not None: opt.sync_lr()
    w0 = time.perf_counter()
    ev[0].record(); win._g1.replay(); ev[1].record()
    rz.check_overflow(win._gstates)
    ev[2].record(); win._g2.replay(); ev[3].record()
    torch.cuda.synchronize()
    wall += time.perf_counter() - w0
    t[0] += ev[0].elapsed_time(ev[1]); t[1] += ev[1].elapsed_time(ev[2]); t[2] += ev[2].elapsed_time(ev[3])
print(f"rank {rank}/{world}: views {len(win.local)}  G1 {t[0]/K:.3f} ms  check {t[1]/K:.3f} ms  G2 {t[2]/K:.3f} ms  wall/step {wall/K*1e3:.3f} ms", flush=True)
bench.finish(world, win)
