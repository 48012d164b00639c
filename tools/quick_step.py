"""Per-kernel breakdown of the native mapping step at cfg3 (eager, events per launch) — dev tool."""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from monogs_dev_b200 import synthetic as syn, mapping, _lib
dev = torch.device("cuda", 0)
sc = syn.make_scene(bench.WORKLOAD, device=dev, n_views=bench.N_VIEWS)
kfs, host = bench.build_window(dev, mapping.render_fused, mapping.GaussianModel, None, sc)
gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation)
import sys as _s
VS = int(_s.argv[1]) if len(_s.argv) > 1 else 1
win = mapping.MappingWindow(gm, kfs, sc.bg, view_streams=VS)
for _ in range(3): win.step()
torch.cuda.synchronize()
_lib.lib.b2gs_profile_enable(1)
for _ in range(3): win.step()
torch.cuda.synchronize()
prof = _lib.profile_read(); _lib.lib.b2gs_profile_enable(0)
tot = sum(v[1] for v in prof.values())
print("libb2gs kernel ms/step", tot / 3)
for k, (c, ms) in sorted(prof.items(), key=lambda kv: -kv[1][1]):
    print(f"  {k:22s} launches {c:4d}  avg {ms / c * 1e3:9.1f} us  share {ms / tot:6.1%}")
