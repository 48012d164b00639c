"""The mapping schedule (monogs_dev_b200/backend.py: BackEnd.map / initialize_map) against a literal
restatement of the reference's loop (utils/slam_backend.py:141-318) written here with torch.optim.Adam,
on CPU with a stand-in differentiable renderer (the product has no CPU rasterizer; what is under
test is the order of operations and the reference's caller-level quirks)."""
import copy
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from monogs_dev_b200 import backend, mapping, model_io, synthetic as syn  # noqa: E402

H, W, P = 6, 8, 64
CONFIG = {"Training": dict(monocular=False, init_itr_num=5, init_gaussian_update=100, init_gaussian_reset=3,
                           init_gaussian_th=0.005, init_gaussian_extent=30, mapping_itr_num=3,
                           gaussian_update_every=1000, gaussian_update_offset=50, gaussian_th=0.7,
                           gaussian_extent=1.0, gaussian_reset=4, size_threshold=20, window_size=3,
                           pose_window=2, prune_mode="slam", lr=dict(cam_rot_delta=0.003, cam_trans_delta=0.001))}


def standin_render(cam, pc, pipe=None, bg=None):
    k = float(cam.fx)
    n = H * W
    vsp = torch.zeros(pc.get_xyz.shape[0], 3, requires_grad=True)
    base = (pc.get_xyz * k).sum(1)[:n] + pc.get_opacity[:n, 0] + pc.get_scaling[:n].sum(1) \
        + pc.get_rotation[:n].sum(1) * 0.1 + pc.get_features[:n, 0].sum(1) + (vsp[:n, :2] * k).sum(1)
    img = torch.stack([base, 2 * base, 3 * base]).view(3, H, W) * torch.exp(cam.cam_rot_delta.sum())
    depth = (base * 0.5 + cam.cam_trans_delta.sum()).view(1, H, W)
    Pn = pc.get_xyz.shape[0]
    idx = torch.arange(Pn)
    return {"render": img, "depth": depth, "radii": (idx % 7 + int(k)).int(),
            "visibility_filter": (idx % int(k - 8)) != 0, "viewspace_points": vsp,
            "opacity": torch.ones(1, H, W), "n_touched": (idx % int(k - 7) == 0).int()}


def build(seed=5, n_kf=5):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    gm = mapping.GaussianModel(r(P, 3), r(P, 1, 3), r(P, 0, 3), r(P, 1), 0.1 * r(P, 3), r(P, 4))
    # the fused optimizer is CUDA-only: the same arena under torch's Adam for these CPU tests
    gm.optimizer_cls = torch.optim.Adam
    gm._allocate({k: gm._params[k].detach().clone() for k in mapping._GROUPS})
    gm.unique_kfIDs = (torch.arange(P) % n_kf).int()
    gm.n_obs = torch.zeros(P).int()
    views = {}
    for v in range(n_kf):
        cam = syn.Camera(W, H, 10.0 + v, 10.0 + v, W / 2, H / 2, device="cpu")
        views[v] = dict(cam=cam, gt_image=r(3, H, W).abs(), gt_depth=r(1, H, W).abs())
    return gm, views


def make_backend(gm, views, config=CONFIG, seed=3):
    be = backend.BackEnd(config, gm, torch.zeros(3), cameras_extent=6.0, render_fn=standin_render,
                         loss_fn=mapping.get_loss_mapping_rgbd_torch, adam_cls=torch.optim.Adam, seed=seed)
    be.viewpoints = views
    return be


# ---- the reference loop, restated (utils/slam_backend.py:141-318, :434-471) ---------------------
class RefBackEnd:
    def __init__(self, config, gm, views, seed=3):
        self.config, self.g, self.viewpoints = config, gm, views
        self.iteration_count = 0
        self.gen = torch.Generator().manual_seed(seed)
        self.occ = {}
        t = config["Training"]
        self.every, self.offset, self.reset_every = t["gaussian_update_every"], t["gaussian_update_offset"], t["gaussian_reset"]

    def set_kf_opt(self, window, frames_to_optimize):
        lr, groups = self.config["Training"]["lr"], []
        for cam_idx, kf in enumerate(window):
            if kf == 0:
                continue
            cam = self.viewpoints[kf]["cam"]
            if cam_idx < frames_to_optimize:
                groups.append(dict(params=[cam.cam_rot_delta], lr=lr["cam_rot_delta"] * 0.5))
                groups.append(dict(params=[cam.cam_trans_delta], lr=lr["cam_trans_delta"] * 0.5))
            groups.append(dict(params=[cam.exposure_a], lr=0.01))
            groups.append(dict(params=[cam.exposure_b], lr=0.01))
        self.kf_opt = torch.optim.Adam(groups)

    def map(self, window, prune=False, iters=1):
        g = self.g
        frames_to_optimize = self.config["Training"]["pose_window"]
        rand = [v for v in self.viewpoints if v not in set(window)]
        for _ in range(iters):
            self.iteration_count += 1
            loss, vis_acm, vsp_acm, radii_acm, touched_acm = 0, [], [], [], []
            picks = torch.randperm(len(rand), generator=self.gen)[:2].tolist()
            for n, kf in enumerate(list(window) + sorted(rand[i] for i in picks)):
                vp = self.viewpoints[kf]
                pkg = standin_render(vp["cam"], g)
                loss = loss + mapping.get_loss_mapping_rgbd_torch(pkg["render"], pkg["depth"], vp["cam"],
                                                                  vp["gt_image"], vp["gt_depth"])
                vsp_acm.append(pkg["viewspace_points"]); vis_acm.append(pkg["visibility_filter"])
                radii_acm.append(pkg["radii"])
                if n < len(window):
                    touched_acm.append(pkg["n_touched"])
            scaling = g.get_scaling
            loss = loss + 10 * torch.abs(scaling - scaling.mean(dim=1).view(-1, 1)).mean()
            loss.backward()
            with torch.no_grad():
                self.occ = {kf: (touched_acm[i] > 0).long() for i, kf in enumerate(window)}
                if prune:
                    return False
                acc, den, mr = g._stats()
                for i in range(len(vsp_acm)):
                    f = vis_acm[i]
                    mr[f] = torch.max(mr[f], radii_acm[i][f].float())
                    acc[f] += torch.norm(vsp_acm[i].grad[f, :2], dim=-1, keepdim=True)
                    den[f] += 1
                update_gaussian = self.iteration_count % self.every == self.offset
                skipped = None
                if (self.iteration_count % self.reset_every) == 0 and not update_gaussian:
                    act = g.get_opacity
                    new = model_io.inverse_sigmoid(torch.ones_like(act) * 0.4)
                    for f in vis_acm:
                        new[f] = act[f]
                    p = g._params["opacity"]
                    p.copy_(new)
                    st = g.optimizer.state[p]
                    st["exp_avg"].zero_(); st["exp_avg_sq"].zero_()
                    skipped, p.grad = p.grad, None        # a fresh Parameter has no .grad: Adam skips it
                g.optimizer.step()
                if skipped is not None:
                    g._params["opacity"].grad = skipped
                g.flat_grad.zero_()
                g.update_learning_rate(self.iteration_count)
                self.kf_opt.step()
                self.kf_opt.zero_grad(set_to_none=True)
                for cam_idx in range(min(frames_to_optimize, len(window))):
                    if window[cam_idx] == 0:
                        continue
                    self.viewpoints[window[cam_idx]]["cam"].update_pose()
        return False


def same_state(gm_a, views_a, gm_b, views_b, tol=1e-6):
    assert torch.allclose(gm_a.flat.detach(), gm_b.flat.detach(), rtol=tol, atol=tol)
    assert torch.allclose(gm_a.flat_grad, gm_b.flat_grad, rtol=1e-5, atol=1e-7)
    for k in views_a:
        ca, cb = views_a[k]["cam"], views_b[k]["cam"]
        assert torch.allclose(ca.R, cb.R, atol=1e-6) and torch.allclose(ca.T, cb.T, atol=1e-6), k
        assert torch.allclose(ca.exposure_a, cb.exposure_a, atol=1e-7), k
        assert torch.allclose(ca.exposure_b, cb.exposure_b, atol=1e-7), k


def test_map_matches_reference_loop_incl_random_views_reset_and_prune_leak():
    gm_a, views_a = build()
    gm_b, views_b = build()
    be = make_backend(gm_a, views_a)
    ref = RefBackEnd(CONFIG, gm_b, views_b)
    window = [4, 2, 0]                      # newest first; keyframe 0 is in the window
    be.set_keyframe_optimizers(window)
    ref.set_kf_opt(window, CONFIG["Training"]["pose_window"])
    # 3 plain iterations, the 4th resets the opacity of non-visible Gaussians (gaussian_reset=4)
    out = be.map(window, iters=5)
    ref.map(window, iters=5)
    assert out is False and be.iteration_count == 5
    same_state(gm_a, views_a, gm_b, views_b)
    for kf in window:
        assert torch.equal(be.occ_aware_visibility[kf], ref.occ[kf])
    assert set(be.occ_aware_visibility) == set(window)
    # prune mode: backward only; gradients stay and leak into the next iteration (q7)
    before = gm_a.flat.detach().clone()
    stats_before = gm_a.xyz_gradient_accum.clone()
    assert be.map(window, prune=True) is False
    ref.map(window, prune=True)
    assert torch.equal(gm_a.flat.detach(), before) and float(gm_a.flat_grad.abs().sum()) > 0
    assert torch.equal(gm_a.xyz_gradient_accum, stats_before)
    assert be.initialized
    same_state(gm_a, views_a, gm_b, views_b)
    be.map(window, iters=2)
    ref.map(window, iters=2)
    same_state(gm_a, views_a, gm_b, views_b)
    assert torch.allclose(gm_a.xyz_gradient_accum, gm_b.xyz_gradient_accum, rtol=1e-5)
    assert torch.equal(gm_a.max_radii2D, gm_b.max_radii2D) and torch.equal(gm_a.denom, gm_b.denom)
    # camera bookkeeping: keyframe 0 never moves, the third window slot is beyond pose_window,
    # the random views (1, 3) keep pose and exposure
    for k in (0, 1, 3):
        assert torch.equal(views_a[k]["cam"].R, torch.eye(3)) and float(views_a[k]["cam"].exposure_a) == 0.0
    assert not torch.equal(views_a[4]["cam"].T, torch.zeros(3)) and float(views_a[2]["cam"].exposure_a) != 0.0


def test_reset_iteration_returns_split_and_keeps_opacity():
    gm, views = build()
    cfg = copy.deepcopy(CONFIG)
    cfg["Training"]["gaussian_reset"] = 2
    be = make_backend(gm, views, cfg)
    be.set_keyframe_optimizers([1, 2])
    assert be.map([1, 2], iters=1) is False
    assert be.map([1, 2], iters=1) is True                   # iteration 2: opacity reset
    op = gm._opacity.detach()[:, 0]
    st = gm.optimizer.state[gm._opacity]
    assert float(st["exp_avg"].abs().sum()) == 0.0 and float(st["step"]) == 1.0   # opacity did not step
    # Gaussians outside every view's filter sit at inverse_sigmoid(0.4); the seen ones hold their
    # former ACTIVATED opacity as raw value (in (0, 1)), the reference's own behaviour
    unseen_value = float(model_io.inverse_sigmoid(torch.tensor(0.4)))
    at_reset = (op - unseen_value).abs() < 1e-6
    assert bool(((op > 0) & (op < 1))[~at_reset].all())
    assert float(gm.optimizer.state[gm._xyz]["step"]) == 2.0


def test_densify_iteration_voids_gaussian_step_but_steps_cameras():
    gm, views = build()
    cfg = copy.deepcopy(CONFIG)
    cfg["Training"].update(gaussian_update_every=2, gaussian_update_offset=0, gaussian_reset=1000,
                           gaussian_th=0.0, gaussian_extent=100.0, size_threshold=10 ** 6)
    be = make_backend(gm, views, cfg)
    be.set_keyframe_optimizers([1, 2])
    be.map([1, 2], iters=1)
    gm.xyz_gradient_accum.fill_(1.0)                         # force clones / splits
    T_before = views[1]["cam"].T.clone()
    xyz_step = float(gm.optimizer.state[gm._xyz]["step"])
    n_before = gm._xyz.shape[0]
    assert be.map([1, 2], iters=1) is True
    assert gm._xyz.shape[0] != n_before                      # the map was densified / pruned
    assert float(gm.flat_grad.abs().sum()) == 0.0
    for k in mapping._GROUPS:                                # no Gaussian Adam step on this iteration
        assert float(gm.optimizer.state[gm._params[k]]["step"]) == xyz_step
    assert not torch.equal(views[1]["cam"].T, T_before)      # the cameras stepped and moved
    assert all(m.shape[0] == n_before for m in be.occ_aware_visibility.values())   # masks of the old map
    assert be.map([1, 2], iters=1) is False                  # and the loop carries on with the new arena
    assert float(gm.optimizer.state[gm._xyz]["step"]) == xyz_step + 1


def test_prune_slam_mode_monocular():
    gm, views = build(n_kf=5)
    cfg = copy.deepcopy(CONFIG)
    cfg["Training"]["monocular"] = True
    be = make_backend(gm, views, cfg)
    window = [4, 3, 2]
    be.set_keyframe_optimizers(window)
    assert be.initialized is False
    n_touch = {kf: (torch.arange(P) % int(10 + kf - 7) == 0).long() for kf in window}
    n_obs = sum(n_touch.values())
    want_prune = n_obs <= 3                                   # not initialised: mask = every Gaussian
    assert be.map(window, prune=True) is False
    assert be.initialized is True
    assert gm._xyz.shape[0] == int((~want_prune).sum())
    for kf in window:
        assert torch.equal(be.occ_aware_visibility[kf], n_touch[kf][~want_prune])
    # initialised: only Gaussians seeded by the three newest window keyframes are candidates
    gm2, views2 = build(n_kf=6)
    be2 = make_backend(gm2, views2, cfg)
    be2.initialized = True
    window2 = [5, 4, 3]
    be2.set_keyframe_optimizers(window2)
    n_obs2 = sum((torch.arange(P) % int(10 + kf - 7) == 0).long() for kf in window2)
    want2 = (n_obs2 <= 3) & (gm2.unique_kfIDs >= 3)
    be2.map(window2, prune=True)
    assert gm2._xyz.shape[0] == int((~want2).sum())
    # a window that is not full does not prune
    gm3, views3 = build()
    be3 = make_backend(gm3, views3, cfg)
    be3.set_keyframe_optimizers([4, 3])
    be3.map([4, 3], prune=True)
    assert gm3._xyz.shape[0] == P and be3.initialized is False


def test_initialize_map_schedule():
    gm, views = build()
    be = make_backend(gm, views)
    pkg = be.initialize_map(0, views[0])
    assert be.iteration_count == 5 and pkg is not None
    # iteration 0 densifies (step void), iteration 3 (init_gaussian_reset) resets every opacity and
    # leaves the opacity group out of that step: 4 steps for xyz, 3 for opacity
    assert float(gm.optimizer.state[gm._xyz]["step"]) == 4.0
    assert float(gm.optimizer.state[gm._opacity]["step"]) == 3.0
    assert list(be.occ_aware_visibility) == [0]
    assert be.occ_aware_visibility[0].shape[0] == gm._xyz.shape[0]
    assert torch.equal(views[0]["cam"].R, torch.eye(3)) and float(views[0]["cam"].exposure_a) == 0.0


def test_ssim_and_l1_match_reference_formula():
    """losses.ssim applies the 11-tap Gaussian window as two 1-D passes; the reference convolves with
    the 2-D outer-product window (gaussian_splatting/utils/loss_utils.py:42-101, restated here)."""
    from math import exp
    import torch.nn.functional as F
    from monogs_dev_b200 import losses
    g = torch.Generator().manual_seed(4)
    a, b = torch.rand(3, 37, 53, generator=g), torch.rand(3, 37, 53, generator=g)
    gauss = torch.tensor([exp(-((x - 11 // 2) ** 2) / float(2 * 1.5 ** 2)) for x in range(11)])
    w1 = (gauss / gauss.sum()).unsqueeze(1)
    window = w1.mm(w1.t()).float().unsqueeze(0).unsqueeze(0).expand(3, 1, 11, 11).contiguous()
    conv = lambda t: F.conv2d(t, window, padding=5, groups=3)
    mu1, mu2 = conv(a), conv(b)
    s1, s2, s12 = conv(a * a) - mu1.pow(2), conv(b * b) - mu2.pow(2), conv(a * b) - mu1 * mu2
    ref = (((2 * mu1 * mu2 + 0.01 ** 2) * (2 * s12 + 0.03 ** 2))
           / ((mu1.pow(2) + mu2.pow(2) + 0.01 ** 2) * (s1 + s2 + 0.03 ** 2))).mean()
    assert abs(float(losses.ssim(a, b)) - float(ref)) < 2e-6
    assert abs(float(losses.ssim(a, a)) - 1.0) < 1e-6
    assert losses.ssim(a[None], b[None], size_average=False).shape == (1,)
    assert float(losses.l1_loss(a, b)) == float((a - b).abs().mean())


def test_color_refinement_steps_gaussians_only():
    gm, views = build()
    be = make_backend(gm, views)
    before = gm.flat.detach().clone()
    poses = {k: (v["cam"].R.clone(), v["cam"].T.clone()) for k, v in views.items()}
    first = float(be.color_refinement(iteration_total=1))
    last = float(be.color_refinement(iteration_total=40))
    assert not torch.equal(gm.flat.detach(), before) and float(gm.flat_grad.abs().sum()) == 0.0
    assert float(gm.optimizer.state[gm._xyz]["step"]) == 41.0
    want_lr = model_io.helper(40, lr_init=gm.lr_init, lr_final=gm.lr_final, lr_delay_mult=gm.lr_delay_mult,
                              max_steps=gm.max_steps)
    assert [gr["lr"] for gr in gm.optimizer.param_groups if gr["name"] == "xyz"] == [want_lr]
    for k, v in views.items():                                  # cameras and exposures are not refined
        assert torch.equal(v["cam"].R, poses[k][0]) and torch.equal(v["cam"].T, poses[k][1])
        assert float(v["cam"].exposure_a) == 0.0
    assert float(gm.denom.sum()) == 0.0 and float(gm.max_radii2D.max()) > 0.0
    assert last == last and first == first                      # finite
