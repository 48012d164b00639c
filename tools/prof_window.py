"""One-keyframe lean mapping-window steps of BASELINE config 3 for `ncu --set full`: the BENCHED
kernel instantiations (raw parameters, fused loss epilogue, in-arena gradient accumulation, fused
Adam + zero_grad, camera step).  Dev tool.  usage: python tools/prof_window.py [steps]"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from monogs_dev_b200 import synthetic as syn, mapping

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 2
dev = torch.device("cuda", 0)
sc = syn.make_scene("cfg3_replica", device=dev, n_views=2)
kfs = bench.build_window(dev, sc)[1:2]
gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=0)
win = mapping.MappingWindow(gm, kfs, sc.bg, pose_lr=bench.POSE_LR, cuda_graph=False)
assert win._lean is not None
for _ in range(steps):
    loss = win.step()
torch.cuda.synchronize()
print("ok", float(loss))
