"""Generate loop-level golden vectors FROM THE REFERENCE'S OWN PYTHON MODULES.

    python tests/golden/make_loop_golden.py          (build container: /root/reference present)

utils/pose_utils.py, utils/slam_utils.py, utils/camera_utils.py and
gaussian_splatting/utils/graphics_utils.py of the reference import only torch / numpy, so they
run here on the CPU.  The only patch: `Tensor.cuda()` is made the identity (the loss functions
call `viewpoint.original_image.cuda()`, utils/slam_utils.py:65,103,117).  Outputs go to
tests/golden/loop_golden.pt and pin
  * oracle/ref_loop.py (the restated loop of the reference arm)   tests/test_ref_loop_cpu.py
  * the product's fused loss kernels and device pose update       tests/test_loop_gpu.py
"""
import os
import sys

import numpy as np
import torch

REF = "/root/reference"
sys.path.insert(0, REF)
torch.Tensor.cuda = lambda self, *a, **k: self          # CPU run of `.cuda()` call sites

from gaussian_splatting.utils.graphics_utils import getProjectionMatrix2, getWorld2View2, focal2fov  # noqa: E402
from utils import pose_utils, slam_utils  # noqa: E402
from utils.camera_utils import Camera  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "loop_golden.pt")


def camera(uid, W, H, fx, fy, cx, cy, R, T):
    proj = getProjectionMatrix2(znear=0.01, zfar=100.0, fx=fx, fy=fy, cx=cx, cy=cy, W=W, H=H).transpose(0, 1)
    cam = Camera(uid, None, None, torch.eye(4), proj, fx, fy, cx, cy, focal2fov(fx, W), focal2fov(fy, H), H, W,
                 device="cpu")
    cam.update_RT(R, T)
    return cam


def main():
    g = torch.Generator().manual_seed(20251105)
    out = {}

    # ---- pose: SE3_exp, update_pose, camera properties -----------------------------------------
    poses = []
    for scale in (0.3, 0.05, 1e-3, 1e-7, 0.0):
        tau0 = 0.3 * torch.randn(6, generator=g)
        w2c = pose_utils.SE3_exp(tau0)
        cam = camera(3, 640, 480, 517.306, 516.469, 318.643, 255.314, w2c[:3, :3].clone(), w2c[:3, 3].clone())
        tau = scale * torch.randn(6, generator=g)
        cam.cam_trans_delta.data.copy_(tau[:3])
        cam.cam_rot_delta.data.copy_(tau[3:])
        before = dict(view=cam.world_view_transform.clone(), full=cam.full_proj_transform.clone(),
                      center=cam.camera_center.clone())
        conv = pose_utils.update_pose(cam)
        poses.append(dict(w2c=w2c, tau=tau, exp_tau=pose_utils.SE3_exp(tau), converged=bool(conv),
                          R=cam.R.clone(), T=cam.T.clone(), before=before,
                          view=cam.world_view_transform.clone(), full=cam.full_proj_transform.clone(),
                          center=cam.camera_center.clone(), proj=cam.projection_matrix.clone(),
                          FoVx=cam.FoVx, FoVy=cam.FoVy))
    out["poses"] = poses
    out["intrinsics"] = dict(W=640, H=480, fx=517.306, fy=516.469, cx=318.643, cy=255.314)

    # ---- losses: values + autograd gradients ---------------------------------------------------
    H, W = 37, 53
    r = lambda *s: torch.rand(*s, generator=g)
    image, depth, opacity = r(3, H, W), 3 * r(1, H, W), r(1, H, W)
    opacity[:, : H // 2] = 0.97 + 0.03 * opacity[:, : H // 2]
    gt_image, gt_depth = r(3, H, W), 3 * r(1, H, W)
    gt_image[:, :5] = 0.0                       # below the rgb boundary threshold
    gt_depth[:, :, :7] = 0.0                    # invalid depth
    grad_mask = (r(1, H, W) > 0.3).float()
    cam = camera(1, W, H, 100.0, 100.0, W / 2, H / 2, torch.eye(3), torch.zeros(3))
    cam.original_image = gt_image
    cam.depth = gt_depth[0].numpy()
    cam.grad_mask = grad_mask
    cam.exposure_a.data.fill_(0.13)
    cam.exposure_b.data.fill_(-0.02)
    out["loss_inputs"] = dict(image=image, depth=depth, opacity=opacity, gt_image=gt_image, gt_depth=gt_depth,
                              grad_mask=grad_mask, exposure_a=0.13, exposure_b=-0.02)
    cases = {}
    for name, mono, fn, init in (("mapping_rgbd", False, "map", False), ("mapping_rgb", True, "map", False),
                                 ("mapping_rgbd_init", False, "map", True),
                                 ("tracking_rgbd", False, "track", False), ("tracking_rgb", True, "track", False)):
        cfg = {"Training": {"monocular": mono, "rgb_boundary_threshold": 0.01, "alpha": 0.95}}
        im, de, op = (t.clone().requires_grad_(True) for t in (image, depth, opacity))
        cam.exposure_a.grad = cam.exposure_b.grad = None
        if fn == "map":
            loss = slam_utils.get_loss_mapping(cfg, im, de, cam, op, initialization=init)
        else:
            loss = slam_utils.get_loss_tracking(cfg, im, de, op, cam)
        (2.5 * loss).backward()
        z = lambda t, like: t.clone() if t is not None else torch.zeros_like(like)
        cases[name] = dict(loss=loss.detach().clone(), d_image=z(im.grad, image), d_depth=z(de.grad, depth),
                           d_opacity=z(op.grad, opacity), upstream=2.5,
                           d_exposure_a=z(cam.exposure_a.grad, cam.exposure_a),
                           d_exposure_b=z(cam.exposure_b.grad, cam.exposure_b))
    out["loss_cases"] = cases

    # ---- median depth (utils/slam_utils.py:131-142) --------------------------------------------
    out["median_depth"] = slam_utils.get_median_depth(depth, opacity).clone()
    torch.save(out, OUT)
    print("wrote", OUT, os.path.getsize(OUT), "bytes")


if __name__ == "__main__":
    main()
