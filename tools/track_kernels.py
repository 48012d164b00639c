"""BASELINE config 2 (tracking: 100k Gaussians, 640x480, pose-only): graph ms per iteration and the
per-kernel CUDA-event times of the eager loop (b2gs_profile_enable).  Dev tool."""
import os, sys, time, json
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from monogs_dev_b200 import synthetic as syn, mapping, rasterizer as rz, _lib

DEV = torch.device("cuda", 0)
sc = syn.make_scene("cfg2_tracking", device=DEV, n_views=1)
gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=sc.sh_degree)
cam0 = sc.cameras[0]
true_w2c = syn.se3_exp(torch.tensor([0.02, -0.015, 0.01, 0.004, -0.006, 0.003]))
cam_true = syn.Camera(sc.W, sc.H, cam0.fx, cam0.fy, cam0.cx, cam0.cy, w2c=true_w2c, device=DEV)
with torch.no_grad():
    gt = mapping.render_pose_only(cam_true, gm, None, sc.bg)
out = {}
for graph in (True, False):
    cam = syn.Camera(sc.W, sc.H, cam0.fx, cam0.fy, cam0.cx, cam0.cy, device=DEV)
    loop = mapping.TrackingLoop(gm, cam, sc.bg, iters=100, cuda_graph=graph)
    best = 1e9
    for rep in range(3):
        loop.load_frame(torch.eye(3), torch.zeros(3), gt["render"], gt["depth"])
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        loop.track()
        torch.cuda.synchronize()
        best = min(best, (time.perf_counter() - t0) * 1e3 / 100)
    out["graph_ms_per_iter" if graph else "eager_ms_per_iter"] = round(best, 4)
    if not graph:
        loop.load_frame(torch.eye(3), torch.zeros(3), gt["render"], gt["depth"])
        _lib.lib.b2gs_profile_enable(1)
        loop.track()
        prof = _lib.profile_read()
        _lib.lib.b2gs_profile_enable(0)
        out["kernels_us"] = {k: round(1e3 * ms / n, 2) for k, (n, ms) in sorted(prof.items(), key=lambda kv: -kv[1][1])}
        out["sum_us"] = round(sum(1e3 * ms / n for n, ms in prof.values()), 1)
rz.set_sync_free(False)
print(json.dumps(out))
