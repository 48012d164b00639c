"""GPU parity tests of the loop-level pieces around the rasterizer: fused losses vs the
reference's torch-op losses (utils/slam_utils.py:56-128 restated in mapping.*_torch), the
device pose update vs utils/pose_utils.py:76-93 (restated on the host in synthetic.se3_exp),
capturable Adam vs torch.optim.Adam, and the CUDA-graph mapping step vs the eager one."""
import copy

import pytest
import torch

from monogs_dev_b200 import synthetic as syn, mapping, losses, rasterizer as rz
from monogs_dev_b200.adam import FusedAdam

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _loss_inputs(H=97, W=131, seed=3):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.rand(*s, generator=g).to(DEV)
    image, depth, opacity = r(3, H, W), 3 * r(1, H, W), r(1, H, W)
    opacity[:, : H // 2] = 0.97 + 0.03 * opacity[:, : H // 2]
    gt_image, gt_depth = r(3, H, W), 3 * r(1, H, W)
    gt_image[:, :5] = 0.0                       # below the rgb boundary threshold
    gt_depth[:, :, :7] = 0.0                    # invalid depth
    cam = syn.Camera(W, H, 100.0, 100.0, W / 2, H / 2, device=DEV)
    cam.exposure_a.data.fill_(0.13)
    cam.exposure_b.data.fill_(-0.02)
    return image, depth, opacity, gt_image, gt_depth, cam


@pytest.mark.parametrize("kind", ["mapping", "tracking"])
def test_fused_loss_matches_torch_ops(kind):
    image, depth, opacity, gt_image, gt_depth, cam = _loss_inputs()
    res = []
    for fused in (False, True):
        im, de, op = (t.clone().requires_grad_(True) for t in (image, depth, opacity))
        cam.exposure_a.grad = cam.exposure_b.grad = None
        if kind == "mapping":
            fn = losses.get_loss_mapping_rgbd if fused else mapping.get_loss_mapping_rgbd_torch
            loss = fn(im, de, cam, gt_image, gt_depth)
        else:
            fn = losses.get_loss_tracking_rgbd if fused else mapping.get_loss_tracking_rgbd_torch
            loss = fn(im, de, op, cam, gt_image, gt_depth)
        (2.5 * loss).backward()
        res.append((loss.detach(), im.grad, de.grad, op.grad, cam.exposure_a.grad.clone(),
                    cam.exposure_b.grad.clone()))
    a, b = res
    assert abs(float(a[0]) - float(b[0])) <= 1e-6 * abs(float(a[0]))
    assert torch.allclose(a[1], b[1], rtol=1e-6, atol=1e-12), "dL/dimage"
    assert torch.allclose(a[2], b[2], rtol=1e-6, atol=1e-12), "dL/ddepth"
    if kind == "tracking":
        assert torch.allclose(a[3], b[3], rtol=1e-5, atol=1e-10), "dL/dopacity"
    assert torch.allclose(a[4], b[4], rtol=1e-4, atol=1e-8), "dL/dexposure_a"
    assert torch.allclose(a[5], b[5], rtol=1e-4, atol=1e-8), "dL/dexposure_b"


@pytest.mark.parametrize("name", ["mapping_rgbd", "mapping_rgb", "mapping_rgbd_init", "tracking_rgbd", "tracking_rgb"])
def test_fused_loss_kernels_match_reference_modules_golden(name):
    """csrc/loss.cu against values + autograd gradients produced by the reference's OWN
    utils/slam_utils.py (tests/golden/make_loop_golden.py -> loop_golden.pt): pins §8f row 2."""
    import os
    from helpers import GOLDEN
    G = torch.load(os.path.join(GOLDEN, "loop_golden.pt"), weights_only=False)
    i, c = G["loss_inputs"], G["loss_cases"][name]
    H, W = i["image"].shape[-2:]
    cam = syn.Camera(W, H, 100.0, 100.0, W / 2, H / 2, device=DEV)
    cam.exposure_a.data.fill_(i["exposure_a"]); cam.exposure_b.data.fill_(i["exposure_b"])
    im, de, op = (i[k].to(DEV).clone().requires_grad_(True) for k in ("image", "depth", "opacity"))
    gt_i, gt_d, gmask = i["gt_image"].to(DEV), i["gt_depth"].to(DEV), i["grad_mask"].to(DEV)
    if name == "mapping_rgbd":
        loss = losses.get_loss_mapping_rgbd(im, de, cam, gt_i, gt_d)
    elif name == "mapping_rgbd_init":
        loss = losses.get_loss_mapping_rgbd(im, de, cam, gt_i, gt_d, initialization=True)
    elif name == "mapping_rgb":
        loss = losses.get_loss_mapping_rgb(im, cam, gt_i)
    elif name == "tracking_rgbd":
        loss = losses.get_loss_tracking_rgbd(im, de, op, cam, gt_i, gt_d, grad_mask=gmask)
    else:
        loss = losses.get_loss_tracking_rgb(im, op, cam, gt_i, grad_mask=gmask)
    (c["upstream"] * loss).backward()
    z = lambda t, like: t.cpu() if t is not None else torch.zeros_like(like).cpu()
    assert torch.allclose(loss.detach().cpu(), c["loss"], rtol=2e-6), name
    assert torch.allclose(z(im.grad, im), c["d_image"], rtol=1e-5, atol=1e-10), name
    assert torch.allclose(z(de.grad, de), c["d_depth"], rtol=1e-5, atol=1e-10), name
    # q1 (the rasterizer drops dL/dopacity) is the rasterizer's business: the loss kernel returns it
    assert torch.allclose(z(op.grad, op), c["d_opacity"], rtol=1e-5, atol=1e-9), name
    assert torch.allclose(z(cam.exposure_a.grad, cam.exposure_a), c["d_exposure_a"], rtol=1e-4, atol=1e-8), name
    assert torch.allclose(z(cam.exposure_b.grad, cam.exposure_b), c["d_exposure_b"], rtol=1e-4, atol=1e-8), name


def test_device_pose_update_matches_reference_modules_golden():
    """b2gs_pose_update against utils/pose_utils.update_pose + utils/camera_utils.Camera run from the
    reference's own modules (loop_golden.pt)."""
    import os
    from helpers import GOLDEN
    G = torch.load(os.path.join(GOLDEN, "loop_golden.pt"), weights_only=False)
    k = G["intrinsics"]
    for p in G["poses"]:
        cam = syn.Camera(k["W"], k["H"], k["fx"], k["fy"], k["cx"], k["cy"], w2c=p["w2c"], device=DEV)
        cam.cam_trans_delta.data.copy_(p["tau"][:3]); cam.cam_rot_delta.data.copy_(p["tau"][3:])
        assert cam.update_pose() == p["converged"]
        assert torch.allclose(cam.R, p["R"], atol=2e-6) and torch.allclose(cam.T, p["T"], atol=2e-6)
        assert torch.allclose(cam.world_view_transform.cpu(), p["view"], atol=2e-6)
        assert torch.allclose(cam.full_proj_transform.cpu(), p["full"], atol=2e-5)
        assert torch.allclose(cam.camera_center.cpu(), p["center"], atol=5e-6)


def test_fused_loss_monocular_and_initialization():
    image, depth, opacity, gt_image, gt_depth, cam = _loss_inputs()
    mask = (gt_image.sum(0) > 0.01).view(1, *gt_image.shape[1:])
    ab = torch.exp(cam.exposure_a) * image + cam.exposure_b
    ref = torch.abs(ab * mask - gt_image * mask).mean()
    assert abs(float(losses.get_loss_mapping_rgb(image, cam, gt_image)) - float(ref)) < 1e-6
    ref0 = torch.abs(image * mask - gt_image * mask).mean()
    assert abs(float(losses.get_loss_mapping_rgb(image, cam, gt_image, initialization=True)) - float(ref0)) < 1e-6
    ref_t = (opacity * torch.abs(ab * mask - gt_image * mask)).mean()
    assert abs(float(losses.get_loss_tracking_rgb(image, opacity, cam, gt_image)) - float(ref_t)) < 1e-6
    with pytest.raises(RuntimeError):
        losses.get_loss_mapping_rgb(image.cpu(), cam, gt_image.cpu())


@pytest.mark.parametrize("scale", [0.05, 1e-7])
def test_device_pose_update_matches_host(scale):
    g = torch.Generator().manual_seed(11)
    w2c = syn.se3_exp(0.3 * torch.randn(6, generator=g))
    cam = syn.Camera(640, 480, 517.3, 516.5, 318.6, 255.3, w2c=w2c, device=DEV)
    tau = scale * torch.randn(6, generator=g)
    cam.cam_trans_delta.data.copy_(tau[:3])
    cam.cam_rot_delta.data.copy_(tau[3:])
    want = syn.se3_exp(tau) @ w2c
    conv = cam.update_pose(converged_threshold=1e-4)
    assert conv == bool(tau.norm() < 1e-4)
    assert torch.allclose(cam.R, want[:3, :3], atol=2e-6) and torch.allclose(cam.T, want[:3, 3], atol=2e-6)
    ref = syn.Camera(640, 480, 517.3, 516.5, 318.6, 255.3, w2c=want, device="cpu")
    assert torch.allclose(cam.world_view_transform.cpu(), ref.world_view_transform, atol=2e-6)
    assert torch.allclose(cam.full_proj_transform.cpu(), ref.full_proj_transform, atol=2e-5)
    assert torch.allclose(cam.camera_center.cpu(), ref.camera_center, atol=5e-6)
    assert float(cam.cam_rot_delta.abs().sum()) == 0.0 and float(cam.cam_trans_delta.abs().sum()) == 0.0


def test_capturable_adam_in_graph_matches_torch():
    torch.manual_seed(0)
    ps = [torch.randn(1000, 3, device=DEV), torch.randn(77, device=DEV)]
    a = [p.clone().requires_grad_(True) for p in ps]
    b = [p.clone().requires_grad_(True) for p in ps]
    oa = FusedAdam([dict(params=[a[0]], lr=1e-2), dict(params=[a[1]], lr=3e-3)], lr=0.0, eps=1e-15,
                   capturable=True)
    ob = torch.optim.Adam([dict(params=[b[0]], lr=1e-2), dict(params=[b[1]], lr=3e-3)], lr=0.0, eps=1e-15)
    grads = [[torch.randn_like(p) for p in ps] for _ in range(6)]
    for p in a:
        p.grad = torch.zeros_like(p)

    def load(k, dst):
        for p, g in zip(dst, grads[k]):
            p.grad.copy_(g) if p.grad is not None else setattr(p, "grad", g.clone())

    load(0, a); oa.step()                                   # eager warm-up step (allocates state)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):                           # capture only records: nothing is applied
        oa.step()
    for k in range(1, 6):
        if k == 4:
            oa.param_groups[0]["lr"] = 5e-3                 # schedule change reaches the captured kernel
            oa.sync_lr()
        load(k, a)
        graph.replay()
    for k in range(6):
        if k == 4:
            ob.param_groups[0]["lr"] = 5e-3
        load(k, b)
        ob.step()
    torch.cuda.synchronize()
    for x, y in zip(a, b):
        assert float((x - y).abs().max()) <= 2e-6 * float(y.abs().max()) + 1e-7


def _window(cuda_graph, view_streams=1, n_views=3, lean=None):
    sc = syn.make_scene("cfg1_plumbing", device=DEV, n_views=n_views)
    gt = dict(xyz=sc.xyz + 0.01, scaling=sc.scaling, rotation=sc.rotation, opacity=sc.opacity + 0.1,
              f_dc=sc.f_dc + 0.05, f_rest=sc.f_rest)
    gtm = mapping.GaussianModel(gt["xyz"], gt["f_dc"], gt["f_rest"], gt["opacity"], gt["scaling"],
                                gt["rotation"], sh_degree=sc.sh_degree)
    kfs = []
    with torch.no_grad():
        for cam in sc.cameras:
            pkg = mapping.render(cam, gtm, None, sc.bg)
            kfs.append(dict(cam=cam, gt_image=pkg["render"].clone(), gt_depth=pkg["depth"].clone()))
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation,
                               sh_degree=sc.sh_degree)
    return mapping.MappingWindow(gm, kfs, sc.bg, cuda_graph=cuda_graph, view_streams=view_streams, lean=lean), gm, kfs


def test_lean_window_matches_autograd_route():
    """window.LeanWindow (loss fused into the render epilogue, window_finalize, Adam + zero_grad fused,
    one camera kernel) == the autograd route over the same kernels (render_fused + fused loss kernels +
    FusedAdam + per-keyframe pose update), eager, 3 views, 6 steps."""
    w0, g0, k0 = _window(False, lean=False)
    w1, g1, k1 = _window(False, lean=True)
    assert w1._lean is not None and w0._lean is None
    for s in range(6):
        a, b = float(w0.step()), float(w1.step())
        assert abs(a - b) <= 2e-6 * abs(a) * (1 + 4 * s), (s, a, b)     # same arithmetic up to reduction order
    diff = (g0.flat - g1.flat).abs()
    assert float((diff > 1e-4 * (1 + g0.flat.abs())).float().mean()) < 1e-3
    for a, b in zip(k0[1:], k1[1:]):
        assert torch.allclose(a["cam"].T, b["cam"].T, atol=1e-4)
        assert torch.allclose(a["cam"].exposure_a, b["cam"].exposure_a, atol=1e-4)
        assert torch.allclose(a["cam"].exposure_b, b["cam"].exposure_b, atol=1e-4)
    assert float(g1.flat_grad.abs().max()) == 0.0, "Adam must leave the gradient arena cleared"
    assert torch.equal(g0.denom, g1.denom) and torch.equal(g0.max_radii2D, g1.max_radii2D)
    assert torch.allclose(g0.xyz_gradient_accum, g1.xyz_gradient_accum, rtol=1e-3,
                          atol=1e-5 * float(g0.xyz_gradient_accum.abs().max()))
    pk = w1.last[0]
    assert pk["render"].shape == k1[0]["gt_image"].shape and int((pk["radii"] > 0).sum()) > 0


@pytest.mark.parametrize("view_streams,lean", [(1, True), (2, True), (2, False)])
def test_graph_mapping_step_matches_eager(view_streams, lean):
    try:
        w0, g0, k0 = _window(False, lean=lean)
        w1, g1, k1 = _window(True, view_streams, lean=lean)
        l0 = [float(w0.step()) for _ in range(8)]
        l1 = [float(w1.step()) for _ in range(8)]
        assert w1._g1 is not None and w1.graph_launches > 0, "graph path not taken"
        for a, b in zip(l0, l1):
            assert abs(a - b) <= 2e-4 * abs(a), (l0, l1)
        assert l0[-1] < l0[0], "loss did not go down"
        # parameters after 8 steps: Adam normalises the update, so atomics-order noise can flip a
        # step where the gradient is ~0; compare in the bulk
        diff = (g0.flat - g1.flat).abs()
        assert float((diff > 1e-3 * (1 + g0.flat.abs())).float().mean()) < 2e-3
        for a, b in zip(k0[1:], k1[1:]):
            assert torch.allclose(a["cam"].T, b["cam"].T, atol=2e-3)
            assert torch.allclose(a["cam"].exposure_a, b["cam"].exposure_a, atol=2e-3)
        # overflow path: starve the capacity -> the step discards its own update on the device, the
        # host notices two calls later (or at flush), re-captures and runs the discarded steps again
        w1.flush()
        before = g1.flat.clone()
        if w1._lean is not None:
            for v in w1.local[:1]:              # ONE view starved: the other two fit and are counted
                w1._lean._alloc_binning(v, 64)  # in the discarded step -> the re-run must skip them
        for key in list(rz._SyncFree.cap):
            rz._SyncFree.cap[key] = 64
        w1._g1 = None
        w1.step()
        torch.cuda.synchronize()
        assert torch.equal(g1.flat, before), "an overflowed step must not touch the parameters"
        w1.flush()
        assert not torch.equal(g1.flat, before), "flush() must re-run the discarded step"
        l1b = float(w1._loss)
        l0b = float(w0.step())
        assert abs(l0b - l1b) <= 5e-4 * abs(l0b)
        w1.step(); w1.step(); w1.step(); w1.flush()        # and the deferred path stays clean
        w0.step(); w0.step(); w0.step()
        diff = (g0.flat - g1.flat).abs()
        assert float((diff > 1e-3 * (1 + g0.flat.abs())).float().mean()) < 2e-3
        if lean:
            # densification statistics: every APPLIED step counted exactly once, the discarded
            # launches not at all (B2gsBackwardIO.stats_gate; GS/scene/gaussian_model.py:691-695)
            assert 12 <= int(g0.denom.max()) <= 12 * len(k0)
            mism = float((g0.denom != g1.denom).float().mean())
            assert mism < 2e-3, mism        # a radius can flip 0 <-> 1 on float noise, nothing more
            assert float((g0.denom - g1.denom).abs().max()) <= 2
            assert torch.allclose(g0.xyz_gradient_accum, g1.xyz_gradient_accum, rtol=5e-2,
                                  atol=2e-3 * float(g0.xyz_gradient_accum.abs().max()))
    finally:
        rz.set_sync_free(False)


def test_lean_eager_overflow_retry_counts_statistics_once():
    """eager lean window: one view runs out of binning capacity -> the host repeats the step; the
    views that fitted must not add their densification statistics a second time, the one that
    overflowed (a background-only pass) must not have counted at all."""
    w0, g0, k0 = _window(False, lean=True)
    w1, g1, k1 = _window(False, lean=True)
    w0.step(); w1.step()
    w1._lean._alloc_binning(w1.local[1], 64)
    a, b = float(w0.step()), float(w1.step())
    assert abs(a - b) <= 1e-5 * abs(a)
    assert w1._lean.cap[w1.local[1]] > 64, "the retry did not happen"
    assert torch.equal(g0.denom, g1.denom)
    assert torch.allclose(g0.xyz_gradient_accum, g1.xyz_gradient_accum, rtol=1e-4,
                          atol=1e-6 * float(g0.xyz_gradient_accum.abs().max()))
    assert torch.equal(g0.max_radii2D, g1.max_radii2D)
    assert int(w1._lean.stats_mask.abs().sum()) == 0 and int(w1._lean.guard) == 0
    a, b = float(w0.step()), float(w1.step())
    assert abs(a - b) <= 1e-5 * abs(a) and torch.equal(g0.denom, g1.denom)


@pytest.mark.parametrize("name", ["cfg1_plumbing", "tiny_sh3"])
def test_fused_activation_render_matches_unfused(name):
    """render_fused (raw parameters, activations + grad accumulation + densify stats inside the
    kernels) == render (torch activations, autograd accumulation, torch statistics)."""
    sc = syn.make_scene(name, device=DEV, n_views=2)
    mk = lambda: mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation,
                                       sh_degree=sc.sh_degree)
    ga, gb = mk(), mk()
    gc, gd = syn.upstream_grads(sc, device=DEV)
    outs = []
    for g, fn in ((ga, mapping.render), (gb, mapping.render_fused)):
        g._stats()
        for cam in sc.cameras:                      # two views: exercises the accumulation
            cam.cam_rot_delta.grad = cam.cam_trans_delta.grad = None
            pkg = fn(cam, g, None, sc.bg)
            torch.autograd.backward([pkg["render"], pkg["depth"]], [gc, gd])
            if not pkg.get("stats_fused"):
                vis = pkg["visibility_filter"]
                g.max_radii2D[vis] = torch.max(g.max_radii2D[vis], pkg["radii"][vis])
                g.add_densification_stats(pkg["viewspace_points"], vis)
        outs.append((pkg, cam.cam_rot_delta.grad.clone(), cam.cam_trans_delta.grad.clone()))
    (pa, ra, ta), (pb, rb, tb) = outs
    # activations differ from torch's by <= 1 ulp, so a rare radius / tile flip is possible
    assert float((pa["radii"] != pb["radii"]).float().mean()) < 1e-3
    for k in ("render", "depth", "opacity"):
        assert float(((pa[k] - pb[k]).abs() > 1e-5 + 1e-4 * pa[k].abs()).float().mean()) < 1e-3, k
    scale = float(ga.flat_grad.abs().max())
    bad = (ga.flat_grad - gb.flat_grad).abs() > 2e-4 * ga.flat_grad.abs() + 2e-5 * scale
    assert float(bad.float().mean()) < 1e-3, "parameter gradients"
    assert torch.allclose(ra, rb, rtol=1e-3, atol=1e-4 * float(ra.abs().max()))
    assert torch.allclose(ta, tb, rtol=1e-3, atol=1e-4 * float(ta.abs().max()))
    assert float((ga.denom != gb.denom).float().mean()) < 1e-3
    assert float((ga.max_radii2D != gb.max_radii2D).float().mean()) < 1e-3
    acc_scale = float(ga.xyz_gradient_accum.abs().max())
    bad = (ga.xyz_gradient_accum - gb.xyz_gradient_accum).abs() > 1e-3 * ga.xyz_gradient_accum.abs() + 1e-4 * acc_scale
    assert float(bad.float().mean()) < 1e-3


def test_tracking_loop_recovers_pose_and_graph_matches_eager():
    """BASELINE config 2 shape at reduced size: pose-only tracking converges towards the true
    pose, and the CUDA-graph loop gives the eager loop's result."""
    try:
        sc = syn.make_scene("cfg1_plumbing", device=DEV, n_views=1)
        gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation,
                                   sh_degree=sc.sh_degree)
        true_w2c = syn.se3_exp(torch.tensor([0.02, -0.015, 0.01, 0.004, -0.006, 0.003]))
        cam_true = syn.Camera(sc.W, sc.H, 300.0, 300.0, 159.5, 119.5, w2c=true_w2c, device=DEV)
        with torch.no_grad():
            gt = mapping.render_pose_only(cam_true, gm, None, sc.bg)
        res = []
        for graph in (False, True):
            cam = syn.Camera(sc.W, sc.H, 300.0, 300.0, 159.5, 119.5, device=DEV)
            loop = mapping.TrackingLoop(gm, cam, sc.bg, iters=60, cuda_graph=graph)
            loop.load_frame(torch.eye(3), torch.zeros(3), gt["render"], gt["depth"])
            loop.track()
            res.append((cam.R.clone(), cam.T.clone(), float(loop.loss)))
            if graph:
                assert loop._graph is not None and loop.graph_launches > 0
        e0 = float((torch.zeros(3) - true_w2c[:3, 3]).norm())
        for R, T, _ in res:
            assert float((T - true_w2c[:3, 3]).norm()) < 0.5 * e0, "tracking did not approach the true pose"
        assert torch.allclose(res[0][1], res[1][1], atol=2e-3) and torch.allclose(res[0][0], res[1][0], atol=2e-3)
        # no gradient leaks into the frozen map
        assert float(gm.flat_grad.abs().max()) == 0.0
    finally:
        rz.set_sync_free(False)


def test_lean_tracking_loop_matches_autograd_route():
    """TrackingLoop(lean=True) — tracking loss fused into the render epilogue, pose-only backward with no
    per-Gaussian output, window_finalize, one pose_step kernel — against the autograd route over the
    same kernels (render_pose_only + fused loss kernels + FusedAdam + b2gs_pose_update), 12 iterations."""
    sc = syn.make_scene("cfg1_plumbing", device=DEV, n_views=1)
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=sc.sh_degree)
    true_w2c = syn.se3_exp(torch.tensor([0.02, -0.015, 0.01, 0.004, -0.006, 0.003]))
    cam_true = syn.Camera(sc.W, sc.H, 300.0, 300.0, 159.5, 119.5, w2c=true_w2c, device=DEV)
    with torch.no_grad():
        gt = mapping.render_pose_only(cam_true, gm, None, sc.bg)
    res = []
    for lean, graph in ((False, False), (True, False), (True, True)):
        cam = syn.Camera(sc.W, sc.H, 300.0, 300.0, 159.5, 119.5, device=DEV)
        loop = mapping.TrackingLoop(gm, cam, sc.bg, iters=12, cuda_graph=graph, lean=lean)
        assert (loop._lean is not None) == lean
        loop.load_frame(torch.eye(3), torch.zeros(3), gt["render"], gt["depth"])
        pkg = loop.track()
        torch.cuda.synchronize()
        res.append((cam.R.clone(), cam.T.clone(), float(cam.exposure_a.detach()), float(cam.exposure_b.detach()), float(loop.loss)))
        assert pkg["render"].shape == gt["render"].shape and pkg["visibility_filter"].dtype == torch.bool
    for r in res[1:]:
        assert torch.allclose(r[0], res[0][0], atol=2e-5) and torch.allclose(r[1], res[0][1], atol=2e-5), (r, res[0])
        assert abs(r[2] - res[0][2]) < 2e-5 and abs(r[3] - res[0][3]) < 2e-5
        assert abs(r[4] - res[0][4]) <= 1e-4 * abs(res[0][4])
    assert float(gm.flat_grad.abs().max()) == 0.0
    rz.set_sync_free(False)


def test_densify_prune_config4_size_on_gpu():
    """BASELINE config 4: one densify_and_prune(max_grad=2e-4, min_opacity=0.7, extent=6.0,
    max_screen_size=20) on a 2M-Gaussian model with synthetic statistics, on the device."""
    P = 2_000_000
    g = torch.Generator().manual_seed(1238)
    r = lambda *s: torch.randn(*s, generator=g).to(DEV)
    gm = mapping.GaussianModel(r(P, 3), r(P, 1, 3), torch.zeros(P, 0, 3, device=DEV), 1.5 * r(P, 1),
                               torch.log(0.06 * torch.exp(0.9 * r(P, 3))), r(P, 4))
    acc, den, _ = gm._stats()
    acc.copy_(4e-4 * torch.rand(P, 1, generator=g).to(DEV))
    den.fill_(1.0)
    gm.flat_m.fill_(0.5)
    before = gm._xyz.detach().clone()
    op_before = torch.sigmoid(gm._opacity.detach()).reshape(-1)
    nc, ns, npr = gm.densify_and_prune(2e-4, 0.7, 6.0, 20)
    n = gm._xyz.shape[0]
    assert n == P + nc + ns - npr and nc > 0 and ns > 0 and npr > 0
    assert bool((torch.sigmoid(gm._opacity) >= 0.7).all()), "pruned by opacity"
    assert bool((torch.exp(gm._scaling).max(1).values <= 0.6).all()), "pruned by world size"
    assert gm.flat.is_cuda and gm.flat_grad.numel() == gm.flat.numel()
    # surviving originals come first, in order, with their moments; new rows have zero moments
    m_xyz = gm.optimizer.state[gm._xyz]["exp_avg"]
    n_old = int((m_xyz[:, 0] != 0).sum())
    assert bool((m_xyz[:n_old] == 0.5).all()) and bool((m_xyz[n_old:] == 0).all())
    surv = torch.nonzero(op_before >= 0.7).squeeze(1)
    assert n_old <= surv.numel()
    # an optimizer step works on the rebuilt arena
    gm.flat_grad.fill_(1e-3)
    gm.optimizer.step()
    assert bool(torch.isfinite(gm.flat).all())


@pytest.mark.parametrize("P,sh_rest", [(2_000_000, 0), (60_000, 15)])
def test_densify_kernels_match_reference_sequence_oracle(P, sh_rest):
    """csrc/densify.cu (plan -> scan -> apply, three kernels) against the literal restatement of the
    reference's four-stage sequence (oracle/densify_oracle.py, CPU) at BASELINE config 4 size, given
    a shared random draw: same clone / split / prune counts, same final row order, copied rows and
    Adam moments bit-equal, sampled children within fp32 exp / log rounding (CPU libm vs GPU).
    Inputs sitting within 1e-4 of a decision threshold are nudged off it first: the CPU oracle and
    the kernel evaluate exp / sigmoid with different libraries and may round such a value to
    different sides."""
    from oracle import densify_oracle as DO
    g = torch.Generator().manual_seed(1238)
    r = lambda *s: torch.randn(*s, generator=g)
    extent, max_grad, min_op = 6.0, 2e-4, 0.7
    scaling = torch.log(0.06 * torch.exp(0.9 * r(P, 3)))
    opacity = 1.5 * r(P, 1)
    accum = 4e-4 * torch.rand(P, 1, generator=g)
    smax = torch.exp(scaling).max(1, keepdim=True).values
    for thr in (0.01 * extent, 0.1 * extent, 0.1 * extent * 1.6):       # dense / world-size / child world-size
        scaling = torch.where(((smax / thr) - 1).abs() < 1e-4, scaling + 1e-3, scaling)
    opacity = torch.where((torch.sigmoid(opacity) - min_op).abs() < 1e-4, opacity + 1e-2, opacity)
    accum = torch.where((accum - max_grad).abs() < 1e-8, accum + 1e-6, accum)
    gm = mapping.GaussianModel(r(P, 3).to(DEV), r(P, 1, 3).to(DEV), r(P, sh_rest, 3).to(DEV), opacity.to(DEV),
                               scaling.to(DEV), r(P, 4).to(DEV), sh_degree=3 if sh_rest else 0)
    acc, den, mr = gm._stats()
    acc.copy_(accum.to(DEV)); den.fill_(1.0); den[::17] = 0.0          # never-seen points: 0/0 -> NaN -> 0
    acc[::17] = 0.0
    mr.copy_((30 * torch.rand(P, generator=g)).to(DEV))
    gm.flat_m.copy_(torch.randn(gm.flat_m.shape, generator=g).to(DEV))
    gm.flat_v.copy_(torch.rand(gm.flat_v.shape, generator=g).to(DEV))
    gm.unique_kfIDs = torch.randint(0, 9, (P,), generator=g).int().to(DEV)
    gm.n_obs = torch.randint(0, 5, (P,), generator=g).int().to(DEV)
    st = lambda k, n: gm.optimizer.state[gm._params[k]][n].detach().cpu()
    par = {k: gm._params[k].detach().cpu() for k in mapping._GROUPS}
    om = DO.Model(par, {k: st(k, "exp_avg") for k in mapping._GROUPS}, {k: st(k, "exp_avg_sq") for k in mapping._GROUPS},
                  gm.xyz_gradient_accum.cpu(), gm.denom.cpu(), gm.max_radii2D.cpu())
    kf_old = gm.unique_kfIDs.cpu()
    # the reference's selection, to size the shared draw
    grads = (om.xyz_gradient_accum / om.denom); grads[grads.isnan()] = 0.0
    ns_expect = int(((grads.squeeze() >= max_grad) & (om.scaling.max(1).values > 0.01 * extent)).sum())
    unit = torch.randn(2 * ns_expect, 3, generator=g)
    want = om.densify_and_prune(max_grad, min_op, extent, 20, unit)
    got = gm.densify_and_prune(max_grad, min_op, extent, 20, samples=unit.to(DEV))
    torch.cuda.synchronize()
    assert got == want, (got, want)
    assert gm._xyz.shape[0] == om.p["xyz"].shape[0]
    n_new = want[0] + 2 * want[1]                                      # rows appended before the final prune
    for k in mapping._GROUPS:
        a, b = gm._params[k].detach().cpu(), om.p[k]
        if k in ("xyz", "scaling"):
            assert torch.allclose(a, b, rtol=2e-6, atol=2e-6), k
        else:
            assert torch.equal(a, b), k
        sta = gm.optimizer.state[gm._params[k]]
        assert torch.equal(sta["exp_avg"].cpu(), om.m[k]) and torch.equal(sta["exp_avg_sq"].cpu(), om.v[k]), k
    # kept originals come first with their rows bit-equal (xyz / scaling of originals are copies too)
    n_o = int((om.m["xyz"].abs().sum(1) != 0).sum())
    assert torch.equal(gm._xyz.detach().cpu()[:n_o], om.p["xyz"][:n_o])
    assert gm.unique_kfIDs.shape[0] == gm._xyz.shape[0] and gm.n_obs.shape[0] == gm._xyz.shape[0]
    assert float(gm.xyz_gradient_accum.abs().max()) == 0.0 and gm.flat_grad.numel() == gm.flat.numel()
    gm.flat_grad.fill_(1e-3); gm.optimizer.step()                      # the rebuilt arena is a working optimizer
    assert bool(torch.isfinite(gm.flat).all())


def test_scale_regulariser_matches_torch():
    g = torch.Generator().manual_seed(2)
    raw = (0.7 * torch.randn(5001, 3, generator=g) - 3).to(DEV).requires_grad_(True)
    s = torch.exp(raw)
    ref = 10 * torch.abs(s - s.mean(dim=1).view(-1, 1)).mean()
    ref.backward()
    sink = torch.full_like(raw, 0.25)
    val = losses.scale_regulariser(raw, sink)
    assert abs(float(val) - float(ref)) <= 1e-6 * abs(float(ref))
    assert torch.allclose(sink - 0.25, raw.grad, rtol=1e-4, atol=2e-8)   # sink holds 0.25 + g: fp32 ulp of 0.25


def test_peer_adam_two_ranks_on_one_gpu_matches_allreduce_plus_adam():
    """b2gs_adam_step_peer through the C ABI with two 'ranks' that live on ONE device (two arenas,
    two streams — the flag barriers make the streams wait for each other): equals gradient sum +
    b2gs_adam_step_dev bit for bit, clears every gradient arena, leaves both replicas identical,
    and only touches the moments of the chunks a rank owns."""
    import ctypes as C
    from monogs_dev_b200 import _lib
    lib = _lib.lib
    torch.manual_seed(3)
    sizes, lrs = [5000, 1234, 4099], [1e-2, 3e-3, 5e-4]        # 4099: a ragged, non-float4 tail
    offs, total = [], 0
    for n in sizes:
        offs.append(total)
        total += (n + 3) // 4 * 4
    W = 2
    p0 = torch.randn(total, device=DEV)
    params = [p0.clone() for _ in range(W)]
    grads = [torch.randn(total, device=DEV) for _ in range(W)]
    ms = [torch.zeros(total, device=DEV) for _ in range(W)]
    vs = [torch.zeros(total, device=DEV) for _ in range(W)]
    sig = [torch.zeros(64, dtype=torch.int32, device=DEV) for _ in range(W)]
    fail = [torch.zeros(1, dtype=torch.int32, device=DEV) for _ in range(W)]
    steps = [torch.zeros(1, dtype=torch.int64, device=DEV) for _ in range(W)]
    # expected: sum in rank order, then the single-rank kernel
    pe, me, ve, se = p0.clone(), torch.zeros(total, device=DEV), torch.zeros(total, device=DEV), \
        torch.zeros(1, dtype=torch.int64, device=DEV)
    streams = [torch.cuda.Stream() for _ in range(W)]

    def segs(p, g, m, v):
        a = (_lib.B2gsAdamSegment * len(sizes))()
        for s, o, n, lr in zip(a, offs, sizes, lrs):
            s.param, s.grad, s.exp_avg, s.exp_avg_sq = (p.data_ptr() + 4 * o, g.data_ptr() + 4 * o,
                                                        m.data_ptr() + 4 * o, v.data_ptr() + 4 * o)
            s.n, s.lr = n, lr
        return a

    # world = 1 call first: CUDA loads a kernel lazily at its first launch and that load waits for
    # running kernels — here the other "rank"'s spinning barrier (separate processes do not share
    # this: each loads its own copy while only its own barrier is in flight)
    wp, wg, wm, wv = (torch.zeros(total, device=DEV) for _ in range(4))
    wsig, wst = torch.zeros(64, dtype=torch.int32, device=DEV), torch.zeros(1, dtype=torch.int64, device=DEV)
    one = (C.c_void_p * 1)
    pa1 = _lib.B2gsPeerArena(1, 0, one(wp.data_ptr()), one(wg.data_ptr()), one(wsig.data_ptr()), fail[0].data_ptr())
    _lib.check(lib.b2gs_adam_step_peer(segs(wp, wg, wm, wv), len(sizes), wst.data_ptr(), 0.9, 0.999, 1e-15,
                                       C.byref(pa1), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()

    for it in range(3):
        for g in grads:
            g.normal_()
        gsum = (torch.zeros_like(grads[0]) + grads[0]) + grads[1]
        _lib.check(lib.b2gs_adam_step_dev(segs(pe, gsum, me, ve), len(sizes), se.data_ptr(), 0.9, 0.999, 1e-15,
                                          torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        keep = []
        for r in range(W):
            pb = (C.c_void_p * W)(*[t.data_ptr() for t in params])
            gb = (C.c_void_p * W)(*[t.data_ptr() for t in grads])
            sb = (C.c_void_p * W)(*[t.data_ptr() for t in sig])
            pa = _lib.B2gsPeerArena(W, r, pb, gb, sb, fail[r].data_ptr())
            keep.append((pb, gb, sb, pa))
            _lib.check(lib.b2gs_adam_step_peer(segs(params[r], grads[r], ms[r], vs[r]), len(sizes),
                                               steps[r].data_ptr(), 0.9, 0.999, 1e-15, C.byref(pa),
                                               streams[r].cuda_stream))
        torch.cuda.synchronize()
        assert int(fail[0]) == 0 and int(fail[1]) == 0, "peer barrier timed out"
        assert torch.equal(params[0], params[1]), "replicas differ"
        assert torch.equal(params[0], pe), "peer step != sum + Adam"
        for o, n in zip(offs, sizes):       # every rank's gradients are cleared by the chunk owners
            assert float(grads[0][o:o + n].abs().max()) == 0.0 and float(grads[1][o:o + n].abs().max()) == 0.0
        # padding between segments is never touched
        for o, n in zip(offs, sizes):
            pad = slice(o + n, o + (n + 3) // 4 * 4)
            assert torch.equal(params[0][pad], p0[pad])
    # moments: each element maintained by exactly one rank, the other rank's copy still zero
    own0, own1 = vs[0] != 0, vs[1] != 0
    assert not bool((own0 & own1).any())
    assert bool(own0.any()) and bool(own1.any())
    assert torch.equal(ms[0] + ms[1], me) and torch.equal(vs[0] + vs[1], ve)


def test_peer_barrier_status_ors_the_overflow_flags_of_all_ranks():
    """b2gs_peer_barrier_status through the C ABI, two 'ranks' on one device: each rank's guard word
    ends up as the OR over ALL ranks of the overflow flags in their (instances, overflow) status
    pairs — whatever the guard words held before (b2gs_window_finalize need not have run)."""
    import ctypes as C
    from monogs_dev_b200 import _lib
    lib = _lib.lib
    W = 2
    dummy = [torch.zeros(16, device=DEV) for _ in range(W)]
    sig = [torch.zeros(64, dtype=torch.int32, device=DEV) for _ in range(W)]
    fail = [torch.zeros(1, dtype=torch.int32, device=DEV) for _ in range(W)]
    guard = [torch.full((1,), 7, dtype=torch.int32, device=DEV) for _ in range(W)]     # stale values
    status = [torch.zeros(3, 2, dtype=torch.int32, device=DEV) for _ in range(W)]      # 3 views per rank
    streams = [torch.cuda.Stream() for _ in range(W)]
    one = (C.c_void_p * 1)
    try:
        # world = 1 warm-up call: loads the kernel while no other barrier is spinning (see the test above)
        pa1 = _lib.B2gsPeerArena(1, 0, one(dummy[0].data_ptr()), one(dummy[0].data_ptr()), one(sig[0].data_ptr()),
                                 fail[0].data_ptr())
        lib.b2gs_set_update_guard(guard[0].data_ptr())
        _lib.check(lib.b2gs_peer_barrier_status(C.byref(pa1), status[0].data_ptr(), 3,
                                                torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        assert int(guard[0]) == 0
        for flags, want in (((0, 0), 0), ((0, 1), 1), ((1, 0), 1), ((0, 0), 0)):
            keep = []
            for r in range(W):
                status[r].zero_()
                status[r][:, 0] = 1000 + r
                status[r][2, 1] = flags[r]              # the LAST view of rank r overflowed (or not)
                guard[r].fill_(5)
            torch.cuda.synchronize()
            for r in range(W):
                pb = (C.c_void_p * W)(*[t.data_ptr() for t in dummy])
                sb = (C.c_void_p * W)(*[t.data_ptr() for t in sig])
                pa = _lib.B2gsPeerArena(W, r, pb, pb, sb, fail[r].data_ptr())
                keep.append((pb, sb, pa))
                lib.b2gs_set_update_guard(guard[r].data_ptr())
                _lib.check(lib.b2gs_peer_barrier_status(C.byref(pa), status[r].data_ptr(), 3, streams[r].cuda_stream))
            torch.cuda.synchronize()
            assert int(fail[0]) == 0 and int(fail[1]) == 0, "peer barrier timed out"
            assert int(guard[0]) == want and int(guard[1]) == want, (flags, int(guard[0]), int(guard[1]))
        # without a guard word the call is refused
        lib.b2gs_set_update_guard(None)
        assert lib.b2gs_peer_barrier_status(C.byref(pa1), status[0].data_ptr(), 3, None) != 0
    finally:
        lib.b2gs_set_update_guard(None)
