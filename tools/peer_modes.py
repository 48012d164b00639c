"""Data-parallel mapping window at cfg3 under torchrun: step time and end state of the three
gradient-exchange modes — NCCL all-reduce + replicated Adam, peer-memory Adam (plain P2P), peer-memory
Adam over NVLS multicast (multimem.ld_reduce / multimem.st).  Dev tool.
usage: torchrun --nproc-per-node N tools/peer_modes.py [workload] [steps]"""
import os, sys, json
import torch, torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench, synthdata
from monogs_dev_b200 import synthetic as syn, mapping

class A: pass
rank, world, local = bench.dist_setup(A())
dev = torch.device("cuda", local)
name = sys.argv[1] if len(sys.argv) > 1 else "cfg3_replica"
K = int(sys.argv[2]) if len(sys.argv) > 2 else 40
n_views = bench.WORKLOADS[name][0]
sc = syn.make_scene(name, device=dev, n_views=n_views)
kfs0 = bench.build_window(dev, sc)
res = {}
wins = []
for mode in ("nccl", "nomc", "mc"):
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=sc.sh_degree)
    kfs = []
    for v, kf in enumerate(kfs0):       # fresh cameras per mode (poses are optimised)
        cam = syn.Camera(sc.W, sc.H, kf["cam"].fx, kf["cam"].fy, kf["cam"].cx, kf["cam"].cy, w2c=synthdata.scene_tensors(name, P=8, n_views=n_views).w2c[v], device=dev)
        kfs.append(dict(kf, cam=cam))
    win = mapping.MappingWindow(gm, kfs, sc.bg, pose_lr=bench.POSE_LR, cuda_graph=True, view_streams=8,
                                peer_adam={"nccl": False, "nomc": "nomc", "mc": True}[mode])
    for _ in range(8):
        win.step()
    win.flush()
    ms = bench.max_over_ranks(bench.timed_steps(win.step, K, world, settle=win.flush), world, dev)
    if win.peer is not None:
        win.peer.check()
    ps, pa = synthdata.param_checksum([gm.flat])
    chk = torch.tensor([ps], dtype=torch.float64, device=dev)
    allc = [torch.zeros_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk)
    res[mode] = dict(ms_per_step=round(ms / K, 4), multicast=bool(win.peer.multicast) if win.peer is not None else None,
                     loss=round(bench.sum_over_ranks(float(win._loss), world, dev), 7), param_sum=round(ps, 4),
                     replicas_identical=all(float(c) == float(allc[0]) for c in allc),
                     grad_arena_max=float(gm.flat_grad.abs().max()))
    wins.append(win)
if rank == 0:
    print(json.dumps(dict(world=world, workload=name, modes=res), indent=1), flush=True)
bench.finish(world, wins[0])
