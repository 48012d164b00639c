"""Step time of the mapping window at cfg3 on ONE GPU for k local views (k = 1 is the per-rank chain
of the 8-GPU run without the exchange), lean vs autograd route, graph replays timed with CUDA events;
then the per-kernel event profile of the lean single-view chain.  Dev tool.
usage: python tools/chain_time.py [views ...]"""
import os, sys, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench
from monogs_dev_b200 import synthetic as syn, mapping, _lib

dev = torch.device("cuda", 0)
torch.cuda.set_device(0)
views = [int(a) for a in sys.argv[1:]] or [1, 8]
sc = syn.make_scene("cfg3_replica", device=dev, n_views=8)
kfs_all = bench.build_window(dev, sc)
out = {}
for k in views:
    for lean in (True, False):
        gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=0)
        kfs = [dict(kf) for kf in (kfs_all[1:2] if k == 1 else kfs_all[:k])]
        if k == 1:
            kfs[0]["optimize_pose"] = kfs[0]["optimize_exposure"] = True
        win = mapping.MappingWindow(gm, kfs, sc.bg, pose_lr=bench.POSE_LR, cuda_graph=True, view_streams=min(8, k), lean=lean)
        for _ in range(8):
            win.step()
        win.flush(); torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        K = 50
        a.record()
        for _ in range(K):
            win.step()
        win.flush()
        b.record(); torch.cuda.synchronize()
        out[f"views{k}_{'lean' if lean else 'autograd'}"] = dict(ms_per_step=round(a.elapsed_time(b) / K, 4),
                                                                launches=win.graph_launches, loss=float(win._loss))
        if lean and k == 1:
            _lib.lib.b2gs_profile_enable(1)
            win.profile_steps(5)
            torch.cuda.synchronize()
            prof = _lib.profile_read()
            _lib.lib.b2gs_profile_enable(0)
            out["chain_us"] = {n: round(1e3 * ms / c, 2) for n, (c, ms) in sorted(prof.items(), key=lambda x: -x[1][1])}
        del win, gm
print(json.dumps(out, indent=1))
