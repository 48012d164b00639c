"""Keyframe management and the tracking wrapper of monogs_dev_b200/frontend.py (FrontEnd.is_keyframe /
add_to_window / tracking, utils/slam_frontend.py:128-290) on CPU: hand-computed cases for the two
selection rules, an independent numpy evaluation of the window-pruning heuristic, and the tracking
plumbing with a stand-in renderer (the product has no CPU rasterizer)."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from monogs_dev_b200 import frontend, mapping, synthetic as syn  # noqa: E402

H, W, P = 6, 8, 64
CONFIG = {"Training": dict(monocular=False, tracking_itr_num=6, kf_interval=5, window_size=4, kf_translation=0.08,
                           kf_min_translation=0.05, kf_overlap=0.9, kf_cutoff=0.3,
                           lr=dict(cam_rot_delta=0.003, cam_trans_delta=0.001))}


def cam_at(t):
    w2c = torch.eye(4)
    w2c[:3, 3] = torch.tensor(t, dtype=torch.float32)
    return dict(cam=syn.Camera(W, H, 10.0, 10.0, W / 2, H / 2, w2c=w2c, device="cpu"),
                gt_image=torch.rand(3, H, W), gt_depth=torch.rand(1, H, W) + 0.5)


def mask(lo, hi):
    m = torch.zeros(P, dtype=torch.bool)
    m[lo:hi] = True
    return m


def test_get_median_depth():
    depth = torch.tensor([[0.0, 1.0, 2.0, 3.0, 10.0]])
    opacity = torch.tensor([[1.0, 0.96, 0.99, 0.97, 0.5]])
    assert float(frontend.get_median_depth(depth, opacity)) == 2.0          # 0 and the translucent 10 drop out
    assert float(frontend.get_median_depth(depth)) == 2.0                  # lower median of (1, 2, 3, 10)
    med, std, valid = frontend.get_median_depth(depth, opacity, mask=torch.tensor([[True, True, True, False, True]]),
                                                return_std=True)
    assert float(med) == 1.0 and valid.tolist() == [[False, True, True, False, False]]


def test_is_keyframe_rules():
    fe = frontend.FrontEnd(CONFIG, None, torch.zeros(3))
    fe.median_depth = torch.tensor(2.0)
    fe.cameras = {0: cam_at([0, 0, 0]), 1: cam_at([0.12, 0, 0]), 2: cam_at([0.2, 0, 0]), 3: cam_at([0.05, 0, 0])}
    occ = {0: mask(0, 40).long()}
    # |t| = 0.12: between kf_min_translation * median (0.10) and kf_translation * median (0.16)
    assert fe.is_keyframe(1, 0, mask(0, 38), occ) is False                 # IoU 38/40 = 0.95 >= 0.9
    assert fe.is_keyframe(1, 0, mask(10, 44), occ) is True                 # IoU 30/44 < 0.9 and far enough
    # |t| = 0.2 > 0.16: a keyframe whatever the overlap
    assert fe.is_keyframe(2, 0, mask(0, 40), occ) is True
    # |t| = 0.05 < 0.10: never, even without overlap
    assert fe.is_keyframe(3, 0, mask(50, 60), occ) is False
    # the distance is that of the relative pose, not of the camera centres' coordinates
    fe.cameras[4] = dict(cam=syn.Camera(W, H, 10.0, 10.0, W / 2, H / 2,
                                        w2c=syn.se3_exp(torch.tensor([0.3, 0.0, 0.0, 0.0, 0.0, 0.5])), device="cpu"))
    fe.cameras[5] = dict(cam=syn.Camera(W, H, 10.0, 10.0, W / 2, H / 2,
                                        w2c=syn.se3_exp(torch.tensor([0.3, 0.0, 0.0, 0.0, 0.0, 0.5])), device="cpu"))
    assert fe.is_keyframe(5, 4, mask(0, 40), {4: mask(0, 40).long()}) is False


def test_add_to_window_overlap_cutoff_and_distance_heuristic():
    import copy
    cfg5 = copy.deepcopy(CONFIG)
    cfg5["Training"]["window_size"] = 5
    fe = frontend.FrontEnd(cfg5, None, torch.zeros(3))
    fe.initialized = True
    pos = {0: [0, 0, 0], 1: [0.1, 0, 0], 2: [0.25, 0, 0], 3: [0.3, 0.1, 0], 4: [0.6, 0, 0], 5: [0.9, 0, 0.2]}
    fe.cameras = {k: cam_at(v) for k, v in pos.items()}
    cur = mask(0, 40)
    occ = {0: mask(30, 64).long(),          # 10 / min(40, 34) = 0.29 <= 0.3 -> candidate
           1: mask(28, 60).long(),          # 12 / 32 = 0.375
           2: mask(29, 64).long(),          # 11 / 35 = 0.31
           3: mask(0, 40).long(), 4: mask(0, 40).long()}
    # the two newest entries are never touched; of the low-overlap ones the LAST is removed
    window, removed = fe.add_to_window(5, cur, occ, [4, 3, 2, 1, 0])
    assert removed == 0 and window == [5, 4, 3, 2, 1]
    # before initialisation the cut-off is 0.4 whatever the config says: now 1 (0.375) is the last candidate...
    fe.config = CONFIG                                                      # window_size = 4
    fe.initialized = False
    window2, removed2 = fe.add_to_window(5, cur, {**occ, 0: mask(0, 40).long()}, [4, 3, 2, 1, 0])
    # ... and the window is still longer than window_size = 4, so the distance heuristic removes one more
    def heuristic(win):
        t = {k: np.array(v, dtype=np.float64) for k, v in pos.items()}     # R = I: |T_CiCj| = |t_i - t_j|
        score = []
        for i in range(2, len(win)):
            s = sum(1.0 / (np.linalg.norm(t[win[i]] - t[win[j]]) + 1e-6) for j in range(2, len(win)) if j != i)
            score.append(np.sqrt(np.linalg.norm(t[win[i]] - t[win[0]])) * s)
        return win[2 + int(np.argmax(score))]
    after_cut = [5, 4, 3, 2, 0]                                            # 1 removed by the overlap rule
    expect = heuristic(after_cut)
    assert removed2 == expect and window2 == [k for k in after_cut if k != expect]
    assert len(window2) == 4
    # full overlap everywhere and a short window: nothing is removed
    fe.initialized = True
    window3, removed3 = fe.add_to_window(3, cur, {k: mask(0, 40).long() for k in range(3)}, [2, 1, 0])
    assert removed3 is None and window3 == [3, 2, 1, 0]


def standin_render(cam, pc, pipe=None, bg=None):
    n = H * W
    base = pc.get_xyz.detach().sum(1)[:n] * 0.01 + cam.state[9:12].sum()
    img = torch.stack([base, 2 * base, 3 * base]).view(3, H, W) * torch.exp(cam.cam_rot_delta.sum())
    depth = (base.abs() + 1.0 + cam.cam_trans_delta.sum()).view(1, H, W)
    return {"render": img, "depth": depth, "opacity": torch.full((1, H, W), 0.99),
            "radii": torch.ones(P, dtype=torch.int32), "visibility_filter": torch.ones(P, dtype=torch.bool),
            "viewspace_points": None, "n_touched": torch.ones(P, dtype=torch.int32)}


def test_tracking_plumbing():
    g = torch.Generator().manual_seed(2)
    r = lambda *s: torch.randn(*s, generator=g)
    gm = mapping.GaussianModel(r(P, 3), r(P, 1, 3), r(P, 0, 3), r(P, 1), 0.1 * r(P, 3), r(P, 4))
    fe = frontend.FrontEnd(CONFIG, gm, torch.zeros(3), render_fn=standin_render,
                           loss_fn=mapping.get_loss_tracking_rgbd_torch, adam_cls=torch.optim.Adam)
    fe.cameras[0] = cam_at([0.3, -0.1, 0.2])
    new = cam_at([0, 0, 0])
    before = gm.flat.detach().clone()
    pkg = fe.tracking(1, new)
    assert set(pkg) >= {"render", "depth", "opacity", "n_touched"}
    cam = fe.cameras[1]["cam"]
    # started from the previous frame's pose and moved away from it; the deltas are folded into (R, T)
    assert not torch.allclose(cam.T, torch.zeros(3)) and not torch.equal(cam.T, fe.cameras[0]["cam"].T)
    assert float((cam.T - fe.cameras[0]["cam"].T).norm()) < 0.1
    assert float(cam.cam_trans_delta.abs().sum()) == 0.0 and float(cam.cam_rot_delta.abs().sum()) == 0.0
    assert float(cam.exposure_a) != 0.0 and float(cam.exposure_b) != 0.0
    assert torch.equal(gm.flat.detach(), before)                           # the map is frozen while tracking
    assert float(fe.median_depth) == float(frontend.get_median_depth(pkg["depth"], pkg["opacity"]))
    # the next frame starts from this one
    pkg2 = fe.tracking(2, cam_at([0, 0, 0]))
    assert float((fe.cameras[2]["cam"].T - cam.T).norm()) < 0.1
    # sync_backend: poses pushed by the back end replace the front end's
    fe.sync_backend(["sync_backend", gm, {1: torch.ones(P).long()}, [(1, torch.eye(3), torch.tensor([1.0, 2.0, 3.0]))]])
    assert fe.cameras[1]["cam"].T.tolist() == [1.0, 2.0, 3.0] and 1 in fe.occ_aware_visibility


def test_add_new_keyframe_depth_maps():
    import copy
    cfg = copy.deepcopy(CONFIG)
    cfg["Training"]["rgb_boundary_threshold"] = 0.01
    fe = frontend.FrontEnd(cfg, None, torch.zeros(3))
    frame = cam_at([0, 0, 0])
    frame["gt_image"][:, 0, :] = 0.0                          # a black row: below the RGB boundary
    frame["gt_depth"] = torch.full((1, H, W), 3.0)
    fe.cameras[7] = frame
    # RGB-D: the observed depth, 0 where the image is black
    d = fe.add_new_keyframe(7)
    assert fe.kf_indices == [7] and d.shape == (H, W)
    assert float(d[0].abs().sum()) == 0.0 and bool((d[1:] == 3.0).all())
    assert float(frame["gt_depth"].min()) == 3.0              # the frame's own depth is not modified
    # monocular initialisation: 2 + 0.3 N(0,1)
    cfg["Training"]["monocular"] = True
    fm = frontend.FrontEnd(cfg, None, torch.zeros(3))
    fm.cameras[0] = frame
    g = torch.Generator().manual_seed(1)
    d0 = fm.add_new_keyframe(0, init=True, generator=g)
    want = 2 + 0.3 * torch.randn(1, H, W, generator=torch.Generator().manual_seed(1))
    want[:, 0, :] = 0
    assert torch.equal(d0, want[0])
    # monocular, later keyframes: rendered depth, outliers / translucent pixels replaced by the median
    depth = torch.full((1, H, W), 2.0)
    depth[0, 2, 3] = 50.0                                     # outlier
    depth[0, 3, :] = 2.2
    opacity = torch.full((1, H, W), 0.99)
    opacity[0, 4, 4] = 0.2                                    # translucent
    valid = (depth > 0) & (opacity > 0.95) & (frame["gt_image"].sum(0) > 0.01)[None]
    med, std = depth[valid].median(), depth[valid].std()
    g = torch.Generator().manual_seed(2)
    d1 = fm.add_new_keyframe(0, depth=depth, opacity=opacity, generator=g)
    noise = torch.randn(1, H, W, generator=torch.Generator().manual_seed(2))
    inval = (depth > med + std) | (depth < med - std) | ~valid
    base = depth.clone()
    base[inval] = med
    want1 = base + noise * torch.where(inval, std * 0.5, std * 0.2)
    want1[:, 0, :] = 0
    assert torch.allclose(d1, want1[0]) and bool(inval[0, 2, 3]) and bool(inval[0, 4, 4])
    assert abs(float(d1[2, 3]) - float(med)) < 3 * float(std)   # the outlier came back to the median
