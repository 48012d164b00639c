"""Host-side mirror of the reference's Python interface: same names, field order,
validation messages; fails loudly (no CPU fallback) when handed CPU tensors."""
import pytest
import torch

import diff_gaussian_rasterization as dgr
from simple_knn._C import distCUDA2
from monogs_dev_b200 import FusedAdam, synthetic as syn
from helpers import build_case


def test_settings_fields_match_reference_order():
    # DGR/diff_gaussian_rasterization/__init__.py:173-186
    assert dgr.GaussianRasterizationSettings._fields == (
        "image_height", "image_width", "tanfovx", "tanfovy", "bg", "scale_modifier", "viewmatrix",
        "projmatrix", "projmatrix_raw", "sh_degree", "campos", "prefiltered", "debug")


def test_validation_messages():
    sc, rs, inp, _, _ = build_case("tiny")
    r = dgr.GaussianRasterizer(rs)
    m2 = torch.zeros_like(inp["means3D"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(inp["means3D"], m2, inp["opacities"], scales=inp["scales"], rotations=inp["rotations"])
    with pytest.raises(Exception, match="excatly one of either SHs or precomputed colors"):
        r(inp["means3D"], m2, inp["opacities"], shs=inp["shs"], colors_precomp=inp["means3D"],
          scales=inp["scales"], rotations=inp["rotations"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(inp["means3D"], m2, inp["opacities"], shs=inp["shs"], scales=inp["scales"])
    with pytest.raises(Exception, match="exactly one of either scale/rotation pair"):
        r(inp["means3D"], m2, inp["opacities"], shs=inp["shs"], scales=inp["scales"],
          rotations=inp["rotations"], cov3D_precomp=torch.zeros(sc.P, 6))


def test_cpu_tensors_fail_loudly():
    sc, rs, inp, _, _ = build_case("tiny")
    r = dgr.GaussianRasterizer(rs)
    with pytest.raises(RuntimeError, match="no CPU path"):
        r(inp["means3D"], torch.zeros_like(inp["means3D"]), inp["opacities"], shs=inp["shs"],
          scales=inp["scales"], rotations=inp["rotations"])
    with pytest.raises(RuntimeError, match="no CPU path"):
        distCUDA2(torch.rand(10, 3))
    p = torch.nn.Parameter(torch.rand(4))
    p.grad = torch.ones(4)
    with pytest.raises(RuntimeError, match="no CPU path"):
        FusedAdam([p], lr=1e-3).step()


def test_camera_conventions():
    # world_view_transform is W2C^T; campos is the camera centre; tau=(rho,theta) left-multiplies
    tau = torch.tensor([0.1, -0.2, 0.05, 0.02, 0.03, -0.01])
    w2c = syn.se3_exp(tau)
    cam = syn.Camera(64, 48, 50.0, 50.0, 31.5, 23.5, w2c=w2c)
    assert torch.allclose(cam.world_view_transform, w2c.t())
    c = -w2c[:3, :3].t() @ w2c[:3, 3]
    assert torch.allclose(cam.camera_center, c, atol=1e-6)
    cam.cam_trans_delta.data[:] = torch.tensor([0.01, 0.0, 0.0])
    cam.update_pose()
    assert torch.allclose(cam.T, w2c[:3, 3] + torch.tensor([0.01, 0.0, 0.0]), atol=1e-6)
    assert float(cam.cam_trans_delta.abs().sum()) == 0.0


def test_camera_from_reference_viewpoint():
    """Camera.from_reference wraps a MonoGS Camera (utils/camera_utils.py:8-64, mimicked here by a
    namespace with the same attributes): same pose / intrinsics / projection, the SAME parameter
    tensors, the frame dict the loop-level entry points take."""
    import math
    import types
    import numpy as np
    from monogs_dev_b200.camera import Camera, projection_matrix, se3_exp
    W, H = 64, 48
    w2c = se3_exp(torch.tensor([0.1, -0.2, 0.3, 0.02, 0.01, -0.03]))
    P = projection_matrix(0.01, 100.0, 31.5, 23.5, 50.0, 52.0, W, H).t().contiguous()
    par = lambda n: torch.nn.Parameter(torch.zeros(n))
    vp = types.SimpleNamespace(uid=7, R=w2c[:3, :3].clone(), T=w2c[:3, 3].clone(), fx=50.0, fy=52.0, cx=31.5, cy=23.5,
                               FoVx=2 * math.atan(W / 100.0), FoVy=2 * math.atan(H / 104.0), image_width=W,
                               image_height=H, projection_matrix=P, cam_rot_delta=par(3), cam_trans_delta=par(3),
                               exposure_a=par(1), exposure_b=par(1), original_image=torch.rand(3, H, W),
                               depth=np.full((H, W), 2.5, dtype=np.float32), grad_mask=None)
    frame = Camera.from_reference(vp, device="cpu")
    cam = frame["cam"]
    assert cam.uid == 7 and cam.cam_rot_delta is vp.cam_rot_delta and cam.exposure_b is vp.exposure_b
    assert torch.allclose(cam.R, vp.R) and torch.allclose(cam.T, vp.T)
    # utils/camera_utils.py:94-107: world_view_transform = W2C^T, full_proj = view @ projection, centre
    assert torch.allclose(cam.world_view_transform, w2c.t(), atol=1e-6)
    assert torch.allclose(cam.full_proj_transform, w2c.t() @ P, atol=1e-5)
    assert torch.allclose(cam.camera_center, torch.linalg.inv(w2c.t())[3, :3], atol=1e-6)
    assert frame["gt_image"].shape == (3, H, W) and frame["gt_depth"].shape == (1, H, W)
    assert float(frame["gt_depth"].mean()) == 2.5 and cam.original_image is vp.original_image
