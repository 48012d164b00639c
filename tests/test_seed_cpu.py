"""Oracle of the keyframe seeding path against hand-computed known answers (CPU)."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import seed_oracle as SO  # noqa: E402


def test_backproject_known_answers():
    # 2x3 image, identity pose: x = (u - cx) z / fx, y = (v - cy) z / fy
    depth = np.array([[1.0, 2.0, 0.0], [150.0, np.nan, 4.0]], dtype=np.float32)
    rgb = np.zeros((2, 3, 3), dtype=np.uint8)
    rgb[0, 1] = (255, 128, 0)
    xyz, col, valid, _ = SO.backproject(rgb, depth, fx=2.0, fy=4.0, cx=1.0, cy=0.5, w2c=np.eye(4))
    assert valid.tolist() == [True, True, False, False, False, True]    # 0, >= trunc and NaN are dropped
    assert np.allclose(xyz[0], [(0 - 1.0) * 1 / 2.0, (0 - 0.5) * 1 / 4.0, 1.0])
    assert np.allclose(xyz[1], [0.0, (0 - 0.5) * 2 / 4.0, 2.0])
    assert np.allclose(xyz[5], [(2 - 1.0) * 4 / 2.0, (1 - 0.5) * 4 / 4.0, 4.0])
    assert np.allclose(col[1], [1.0, 128 / 255.0, 0.0]) and col.dtype == np.float32
    # a pure translation of the camera: world = cam - t
    w2c = np.eye(4); w2c[:3, 3] = (1.0, 2.0, 3.0)
    xyz2, _, _, c2w = SO.backproject(rgb, depth, 2.0, 4.0, 1.0, 0.5, w2c)
    assert np.allclose(xyz2[0], xyz[0] - [1.0, 2.0, 3.0]) and np.allclose(c2w @ w2c, np.eye(4))


def test_colour_quantisation_and_attributes():
    img = np.array([[[0.0, 0.5, 1.2]], [[-0.1, 0.999, 0.25]], [[1.0, 0.0, 0.1]]], dtype=np.float32)  # [3,1,3]
    q = SO.quantise_colour(img)
    assert q.shape == (1, 3, 3) and q.dtype == np.uint8
    assert q[0, :, 0].tolist() == [0, 127, 255] and q[0, :, 1].tolist() == [0, 254, 63]   # truncation, clamp
    feats, scales, rots, opac = SO.seed_attributes(np.zeros((2, 3), np.float32), np.array([[0.5, 1, 0]] * 2, np.float32),
                                                   np.array([4.0, 0.0], np.float32), point_size=0.25, max_sh_degree=1)
    assert feats.shape == (2, 3, 4) and np.allclose(feats[0, :, 0], [0.0, 0.5 / SO.C0, -0.5 / SO.C0])
    assert np.all(feats[:, :, 1:] == 0)
    assert np.allclose(scales[0], np.log(np.sqrt(4.0 * 0.25))) and scales.shape == (2, 3)
    assert np.allclose(scales[1], np.log(np.sqrt(1e-7 * 0.25)))                          # clamp_min(1e-7)
    assert rots.tolist() == [[1, 0, 0, 0]] * 2 and np.all(opac == 0.0)
