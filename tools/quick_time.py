"""Quick per-stage timing on the GPU box (dev tool, not the bench contract)."""
import os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import build_case, run_native
from monogs_dev_b200 import rasterizer as rz
from oracle import ref_ext

def timeit(fn, n=20, warm=5):
    for _ in range(warm): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n): fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n

name = sys.argv[1] if len(sys.argv) > 1 else "cfg3_replica"
sc, rs, inp, gc, gd = build_case(name, device="cuda")
r = rz.GaussianRasterizer(rs)
m2 = torch.zeros_like(inp["means3D"])
def fwd():
    with torch.no_grad():
        return r(means3D=inp["means3D"], means2D=m2, opacities=inp["opacities"], shs=inp["shs"], scales=inp["scales"], rotations=inp["rotations"])
leaves = {k: v.clone().requires_grad_(True) for k, v in inp.items()}
def fwdbwd():
    c, rad, d, o, nt = r(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
    torch.autograd.backward([c, d], [gc, gd])
    for v in leaves.values(): v.grad = None
out = fwd()
st = None
print(name, "P", sc.P, "WxH", sc.W, sc.H)
c, rad, d, o, nt = r(means3D=leaves["means3D"], means2D=m2, opacities=leaves["opacities"], shs=leaves["shs"], scales=leaves["scales"], rotations=leaves["rotations"])
print("num_rendered", c.grad_fn.state.num_rendered, "visible", int((rad > 0).sum()))
print("new fwd ms", timeit(fwd)); print("new fwd+bwd ms", timeit(fwdbwd))
if ref_ext.available():
    def rfwd(): return ref_ext.forward(rs, **inp)
    f = rfwd()
    def rfb():
        f = ref_ext.forward(rs, **inp)
        ref_ext.backward(rs, f, gc, gd, inp["means3D"], shs=inp["shs"], scales=inp["scales"], rotations=inp["rotations"])
    print("ref fwd ms", timeit(rfwd)); print("ref fwd+bwd ms", timeit(rfb))
from monogs_dev_b200 import _lib
_lib.lib.b2gs_profile_enable(1)
for _ in range(10): fwdbwd()
torch.cuda.synchronize()
prof = _lib.profile_read(); _lib.lib.b2gs_profile_enable(0)
tot = sum(v[1] for v in prof.values())
print("per-kernel (10 fwd+bwd): total kernel ms/iter", tot / 10)
for k, (c, ms) in sorted(prof.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:22s} launches {c:4d}  avg {ms / c * 1e3:9.1f} us  share {ms / tot:6.1%}")
