"""oracle/ref_loop.py (the restated reference loop that bench.py --impl reference and the
loop-level GPU parity tests run) against golden vectors generated from the reference's OWN
importable modules (tests/golden/make_loop_golden.py -> loop_golden.pt): SE(3) exponential,
update_pose, the camera properties, and the four mapping / tracking losses with their autograd
gradients.  Also: the product's host-side camera (CPU path of camera.Camera) against the same."""
import os
import sys

import pytest
import torch

from helpers import GOLDEN
import synthdata
from oracle import ref_loop as L

G = torch.load(os.path.join(GOLDEN, "loop_golden.pt"), weights_only=False)


def test_ref_loop_does_not_import_the_product():
    import subprocess
    code = ("import sys; sys.path.insert(0, %r); import oracle.ref_loop, oracle.ref_ext, synthdata; "
            "assert not any(m.startswith('monogs_dev_b200') for m in sys.modules), 'product imported'"
            % os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    subprocess.check_call([sys.executable, "-c", code])


@pytest.mark.parametrize("i", range(len(G["poses"])))
def test_pose_update_and_camera_properties(i):
    p, k = G["poses"][i], G["intrinsics"]
    assert torch.allclose(L.SE3_exp(p["tau"]), p["exp_tau"], atol=1e-7)
    assert torch.allclose(synthdata.se3_exp(p["tau"]), p["exp_tau"], atol=1e-7)
    cam = L.RefCamera(3, k["W"], k["H"], k["fx"], k["fy"], k["cx"], k["cy"], w2c=p["w2c"], device="cpu")
    assert torch.allclose(cam.projection_matrix, p["proj"], atol=0) and cam.FoVx == p["FoVx"]
    for name in ("view", "full", "center"):
        got = dict(view=cam.world_view_transform, full=cam.full_proj_transform, center=cam.camera_center)[name]
        assert torch.allclose(got, p["before"][name], atol=1e-7), name
    cam.cam_trans_delta.data.copy_(p["tau"][:3]); cam.cam_rot_delta.data.copy_(p["tau"][3:])
    conv = L.update_pose(cam)
    assert bool(conv) == p["converged"]
    assert torch.allclose(cam.R, p["R"], atol=1e-7) and torch.allclose(cam.T, p["T"], atol=1e-7)
    assert torch.allclose(cam.world_view_transform, p["view"], atol=1e-7)
    assert torch.allclose(cam.full_proj_transform, p["full"], atol=1e-6)
    assert torch.allclose(cam.camera_center, p["center"], atol=1e-7)
    assert float(cam.cam_rot_delta.abs().sum()) == 0.0


@pytest.mark.parametrize("i", range(len(G["poses"])))
def test_product_camera_host_path(i):
    """camera.Camera on the CPU (host logic; the device path is checked in tests/test_loop_gpu.py)."""
    from monogs_dev_b200.camera import Camera
    p, k = G["poses"][i], G["intrinsics"]
    cam = Camera(k["W"], k["H"], k["fx"], k["fy"], k["cx"], k["cy"], w2c=p["w2c"], device="cpu")
    assert torch.allclose(cam.projection_matrix, p["proj"], atol=0)
    assert torch.allclose(cam.world_view_transform, p["before"]["view"], atol=1e-6)
    assert torch.allclose(cam.full_proj_transform, p["before"]["full"], atol=1e-5)
    assert torch.allclose(cam.camera_center, p["before"]["center"], atol=1e-6)
    cam.cam_trans_delta.data.copy_(p["tau"][:3]); cam.cam_rot_delta.data.copy_(p["tau"][3:])
    assert cam.update_pose() == p["converged"]
    assert torch.allclose(cam.R, p["R"], atol=1e-6) and torch.allclose(cam.T, p["T"], atol=1e-6)
    assert torch.allclose(cam.world_view_transform, p["view"], atol=1e-6)
    assert torch.allclose(cam.camera_center, p["center"], atol=1e-6)


def _run(fn, name, **kw):
    i, c = G["loss_inputs"], G["loss_cases"][name]
    cam = L.RefCamera(1, 53, 37, 100.0, 100.0, 26.5, 18.5, device="cpu")
    cam.exposure_a.data.fill_(i["exposure_a"]); cam.exposure_b.data.fill_(i["exposure_b"])
    im, de, op = (i[k].clone().requires_grad_(True) for k in ("image", "depth", "opacity"))
    loss = fn(im, de, op, cam, i, **kw)
    (c["upstream"] * loss).backward()
    z = lambda t, like: t if t is not None else torch.zeros_like(like)
    assert torch.allclose(loss.detach(), c["loss"], rtol=1e-6), name
    assert torch.allclose(z(im.grad, im), c["d_image"], rtol=1e-6, atol=1e-12), name
    assert torch.allclose(z(de.grad, de), c["d_depth"], rtol=1e-6, atol=1e-12), name
    assert torch.allclose(z(op.grad, op), c["d_opacity"], rtol=1e-6, atol=1e-12), name
    assert torch.allclose(z(cam.exposure_a.grad, cam.exposure_a), c["d_exposure_a"], rtol=1e-5, atol=1e-9), name
    assert torch.allclose(z(cam.exposure_b.grad, cam.exposure_b), c["d_exposure_b"], rtol=1e-5, atol=1e-9), name


def test_ref_loop_losses_match_reference_modules():
    _run(lambda im, de, op, cam, i: L.get_loss_mapping_rgbd(im, de, cam, i["gt_image"], i["gt_depth"]), "mapping_rgbd")
    _run(lambda im, de, op, cam, i: L.get_loss_mapping_rgbd(im, de, cam, i["gt_image"], i["gt_depth"],
                                                            initialization=True), "mapping_rgbd_init")
    _run(lambda im, de, op, cam, i: L.get_loss_tracking_rgbd(im, de, op, cam, i["gt_image"], i["gt_depth"],
                                                             grad_mask=i["grad_mask"]), "tracking_rgbd")


def test_product_torch_losses_match_reference_modules():
    """mapping.get_loss_*_torch (host-side torch restatements kept in the product for CPU logic tests)."""
    from monogs_dev_b200 import mapping
    _run(lambda im, de, op, cam, i: mapping.get_loss_mapping_rgbd_torch(im, de, cam, i["gt_image"], i["gt_depth"]),
         "mapping_rgbd")
