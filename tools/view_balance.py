"""Per-keyframe cost of the cfg3 window on one GPU: a one-view lean window (graph replays) for each of
the 8 keyframes — what each rank of the 8-GPU run does before the exchange.  Dev tool."""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from monogs_dev_b200 import synthetic as syn, mapping

dev = torch.device("cuda", 0)
sc = syn.make_scene("cfg3_replica", device=dev, n_views=8)
kfs_all = bench.build_window(dev, sc)
out = {}
for v in range(8):
    for reg in ((True, False) if v == 0 else (False,)):
        gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=0)
        kf = dict(kfs_all[v])
        win = mapping.MappingWindow(gm, [kf], sc.bg, pose_lr=(0, 0, 0, 0), cuda_graph=True, regulariser=reg)
        for _ in range(6):
            win.step()
        win.flush(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(50):
            win.step()
        win.flush(); b.record(); torch.cuda.synchronize()
        info = win.work_counters(0)
        out[f"view{v}" + ("_with_iso_reg" if reg else "")] = dict(ms=round(a.elapsed_time(b) / 50, 4), instances=info["instances"],
                                                                  visible=info["visible"], bwd_blend_px=info["bwd_blend_px"])
        del win, gm
print(json.dumps(out, indent=1))
