"""Step-by-step state difference between the product window and the clean reference loop at cfg3
(dev diagnostic): after every step, per parameter group the fraction of elements off by > 1e-5 rel and
the max |diff|; per keyframe the pose / exposure difference.  usage: python tools/loop_diff.py [steps] [freeze]
freeze = 'pose' (no camera optimisation on either side) | 'gauss' (cameras only)"""
import os, sys
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import synthdata
from monogs_dev_b200 import mapping
from monogs_dev_b200.camera import Camera
from oracle import ref_loop

steps = int(sys.argv[1]) if len(sys.argv) > 1 else 6
freeze = sys.argv[2] if len(sys.argv) > 2 else ""
DEV = "cuda"
POSE_LR = (0.0015, 0.0005, 0.01, 0.01) if freeze != "pose" else (0.0, 0.0, 0.0, 0.0)
t = synthdata.scene_tensors("cfg3_replica", n_views=8)
bg = torch.zeros(3, device=DEV)
cams_r = ref_loop.make_cameras(t, DEV)
kfs_r = ref_loop.build_window_gt(t, cams_r, bg, DEV)
gr = ref_loop.make_model(t, DEV)
win_r = ref_loop.RefMappingWindow(gr, kfs_r, bg, pose_lr=POSE_LR)
to = lambda x: x.to(DEV).contiguous()
gm = mapping.GaussianModel(to(t.xyz), to(t.f_dc), to(t.f_rest), to(t.opacity), to(t.scaling), to(t.rotation), sh_degree=0)
kfs_p = []
for v, m in enumerate(t.w2c):
    cam = Camera(t.W, t.H, t.fx, t.fy, t.cx, t.cy, w2c=m, device=DEV)
    kfs_p.append(dict(cam=cam, gt_image=kfs_r[v]["gt_image"].clone(), gt_depth=kfs_r[v]["gt_depth"].clone(),
                      optimize_pose=v > 0, optimize_exposure=v > 0))
win_p = mapping.MappingWindow(gm, kfs_p, bg, pose_lr=POSE_LR, cuda_graph=False, view_streams=1)
groups = lambda g: dict(xyz=g._xyz, f_dc=g._features_dc, opacity=g._opacity, scaling=g._scaling, rotation=g._rotation)
for s in range(steps):
    if freeze == "gauss":
        for grp in gr.optimizer.param_groups: grp["lr"] = 0.0
        for grp in gm.optimizer.param_groups: grp["lr"] = 0.0
    lr_, lp_ = float(win_r.step()), float(win_p.step())
    torch.cuda.synchronize()
    line = [f"step {s}: loss rel {abs(lp_-lr_)/abs(lr_):.2e}"]
    for k, b in groups(gr).items():
        a = gm._params[k].detach(); b = b.detach()
        d = (a - b).abs()
        line.append(f"{k} off {float((d > 1e-5*(b.abs()+b.abs().mean())).float().mean()):.1e} max {float(d.max()):.1e}")
    print(" | ".join(line))
    pl = []
    for v in range(1, 8):
        cp, cr = kfs_p[v]["cam"], cams_r[v]
        pl.append(f"kf{v} dT {float((cp.T.to(DEV)-cr.T).abs().max()):.1e} dR {float((cp.R.to(DEV)-cr.R).abs().max()):.1e} "
                  f"da {abs(float(cp.exposure_a)-float(cr.exposure_a)):.1e} db {abs(float(cp.exposure_b)-float(cr.exposure_b)):.1e}")
    print("     " + " ; ".join(pl[:4]))
    print("     ref kf1: T", [round(float(x), 6) for x in cams_r[1].T], "a", round(float(cams_r[1].exposure_a), 6), "b", round(float(cams_r[1].exposure_b), 6))
    print("     prd kf1: T", [round(float(x), 6) for x in kfs_p[1]["cam"].T], "a", round(float(kfs_p[1]["cam"].exposure_a), 6), "b", round(float(kfs_p[1]["cam"].exposure_b), 6))
