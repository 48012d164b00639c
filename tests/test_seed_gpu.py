"""Keyframe seeding on the GPU (csrc/seed.cu through the C ABI, monogs_dev_b200/seeding.py) against
the CPU oracle (oracle/seed_oracle.py)."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

pytestmark = pytest.mark.gpu

from monogs_dev_b200 import mapping, seeding, synthetic as syn, distCUDA2  # noqa: E402
from oracle import seed_oracle as SO  # noqa: E402

DEV = torch.device("cuda", 0)


def _frame(W, H, seed):
    g = torch.Generator().manual_seed(seed)
    img = torch.rand(3, H, W, generator=g) * 1.2 - 0.1            # exercises the clamp
    depth = torch.rand(H, W, generator=g) * 6.0
    depth[torch.rand(H, W, generator=g) < 0.1] = 0.0              # holes
    depth[0, :5] = 120.0                                          # beyond depth_trunc
    depth[1, 0] = float("nan")
    w2c = syn.se3_exp(torch.tensor([0.3, -0.2, 0.1, 0.05, -0.08, 0.02]))
    return img, depth, w2c


@pytest.mark.parametrize("W,H", [(64, 48), (1200, 680), (7, 2)])
def test_backproject_matches_oracle_bit_exact(W, H):
    img, depth, w2c = _frame(W, H, 7 + W)
    fx, fy, cx, cy = 600.0, 590.0, (W - 1) / 2.0, (H - 1) / 2.0
    xyz, rgb, valid = seeding.backproject_rgbd(img.to(DEV), depth.to(DEV), fx, fy, cx, cy, w2c.double().numpy())
    q = SO.quantise_colour(img.numpy())
    exyz, ergb, evalid, _ = SO.backproject(q, depth.numpy(), fx, fy, cx, cy, w2c.double().numpy())
    assert np.array_equal(valid.cpu().numpy().astype(bool), evalid)
    assert np.array_equal(rgb.cpu().numpy().view(np.int32), ergb.view(np.int32)), "colour bits"
    assert np.array_equal(xyz.cpu().numpy().view(np.int32), exyz.view(np.int32)), "point bits"


def test_exposure_is_applied_before_quantisation():
    img, depth, w2c = _frame(96, 64, 3)
    a, b = torch.tensor([0.2]), torch.tensor([-0.03])
    _, rgb, _ = seeding.backproject_rgbd(img.to(DEV), depth.to(DEV), 100.0, 100.0, 47.5, 31.5, np.eye(4),
                                         exposure_a=a.to(DEV), exposure_b=b.to(DEV))
    q = SO.quantise_colour(img.numpy(), 0.2, -0.03).reshape(-1, 3).astype(np.int32)
    got = torch.round(rgb.cpu() * 255).int().numpy()
    # expf on the device vs numpy may differ in the last bit of exp(a): at most one level, rarely
    assert np.abs(got - q).max() <= 1 and (got != q).mean() < 1e-3


def test_extend_from_pcd_seq_appends_seeded_gaussians():
    sc = syn.make_scene("cfg1_plumbing", device=DEV, n_views=1)
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=0)
    P0 = gm._xyz.shape[0]
    gm.optimizer.state[gm._xyz]["exp_avg"].fill_(0.25)            # old rows keep their moments
    cam = sc.cameras[0]
    img, depth, _ = _frame(sc.W, sc.H, 11)
    cam.original_image, cam.depth = img.to(DEV), depth.numpy()
    gen = torch.Generator(device="cpu").manual_seed(5)
    pcd = gm.create_pcd_from_image(cam, init=True, generator=gen)
    xyz, feats, scales, rots, opac = pcd
    nvalid = int(((depth > 0) & (depth < 100)).sum())
    n = xyz.shape[0]
    assert n == int(nvalid / 32) and feats.shape == (n, 3, 1) and scales.shape == (n, 3)
    # every sampled point is one of the oracle's points, colours follow, attributes follow the oracle
    q = SO.quantise_colour(img.numpy())
    exyz, ergb, evalid, _ = SO.backproject(q, depth.numpy(), cam.fx, cam.fy, cam.cx, cam.cy, seeding._w2c(cam))
    table = {tuple(r.view(np.int32)): i for i, r in enumerate(exyz) if evalid[i]}
    rows = [table[tuple(r.view(np.int32))] for r in xyz.cpu().numpy()]
    assert len(set(rows)) == n, "sampling must be without replacement"
    point_size = min(0.05, 0.05 * float(np.median(depth.numpy()[~np.isnan(depth.numpy())])))
    d2 = distCUDA2(xyz).cpu().numpy()
    ef, es, er, eo = SO.seed_attributes(exyz[rows], ergb[rows], d2, min(0.05, 0.05 * float(torch.median(depth.to(DEV)))))
    assert np.allclose(feats.cpu().numpy(), ef, atol=1e-6) and np.allclose(scales.cpu().numpy(), es, rtol=1e-6, atol=1e-6)
    assert np.array_equal(rots.cpu().numpy(), er) and np.allclose(opac.cpu().numpy(), eo)
    added = gm.extend_from_pcd(*pcd, kf_id=3)
    assert added == n and gm._xyz.shape[0] == P0 + n
    assert torch.equal(gm._xyz[P0:], xyz) and torch.equal(gm._features_dc[P0:, 0], feats[:, :, 0])
    m = gm.optimizer.state[gm._xyz]["exp_avg"]
    assert float(m[:P0].min()) == 0.25 and float(m[P0:].abs().max()) == 0.0
    assert gm.unique_kfIDs.shape[0] == P0 + n and int(gm.unique_kfIDs[P0:].min()) == 3 and int(gm.n_obs.sum()) == 0
    # the extended model renders and steps
    pkg = mapping.render_fused(cam, gm, None, sc.bg)
    assert pkg["render"].shape == (3, sc.H, sc.W) and bool(torch.isfinite(pkg["render"]).all())
