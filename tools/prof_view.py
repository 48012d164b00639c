"""Short single-view fwd+bwd of BASELINE config 3 for `ncu --set full` (dev tool)."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import build_case
from monogs_dev_b200 import rasterizer as rz

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3_replica"
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 2
sc, rs, inp, gc, gd = build_case(name, device="cuda")
r = rz.GaussianRasterizer(rs)
m2 = torch.zeros_like(inp["means3D"])
leaves = {k: v.clone().requires_grad_(True) for k, v in inp.items()}
for _ in range(iters):
    c, rad, d, o, nt = r(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"],
                         scales=leaves["scales"], rotations=leaves["rotations"])
    torch.autograd.backward([c, d], [gc, gd])
torch.cuda.synchronize()
print("ok", c.grad_fn.state.num_rendered)
