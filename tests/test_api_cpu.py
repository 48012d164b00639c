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
