"""TEST INFRASTRUCTURE — the reference's callers of the rasterizer, restated in torch so that the
reference arm of bench.py and the loop-level parity tests run WITHOUT importing the product
(`monogs_dev_b200` is never imported here; the only native code behind this module is the
unmodified reference extension in oracle/_ref, through oracle/ref_ext.py).

/root/reference does not exist on the GPU box, so the Python side of the reference loop is
restated op for op (same torch calls in the same order, so the same kernels run):

  RefCamera            utils/camera_utils.py:8-112 — R, T tensors; world_view_transform,
                       full_proj_transform and camera_center RECOMPUTED on every property access
                       (getWorld2View2 + transpose, bmm, 4x4 inverse), as the reference does
  getWorld2View2       gaussian_splatting/utils/graphics_utils.py:38-50
  getProjectionMatrix2 gaussian_splatting/utils/graphics_utils.py:71-92
  SE3_exp / update_pose  utils/pose_utils.py:12-93
  get_loss_mapping_rgbd / get_loss_tracking_rgbd   utils/slam_utils.py:56-128
  RefMappingWindow.step  utils/slam_backend.py:156-317 (one iteration of BackEnd.map over a fixed
                       window, no random extra views, no densify / reset iteration)
  RefTrackingLoop.track  utils/slam_frontend.py:128-196

The restated pieces are pinned against the reference's own importable modules by
tests/golden/make_loop_golden.py (run in the build container where /root/reference exists;
fixtures under tests/golden/loop_*.pt, checked by tests/test_ref_loop_cpu.py).
"""
import math

import torch

from . import ref_ext


# -- gaussian_splatting/utils/graphics_utils.py -------------------------------------------------
def getWorld2View2(R, t, translate=torch.tensor([0.0, 0.0, 0.0]), scale=1.0):
    """:38-50"""
    translate = translate.to(R.device)
    Rt = torch.zeros((4, 4), device=R.device)
    Rt[:3, :3] = R
    Rt[:3, 3] = t
    Rt[3, 3] = 1.0
    C2W = torch.linalg.inv(Rt)
    cam_center = C2W[:3, 3]
    cam_center = (cam_center + translate) * scale
    C2W[:3, 3] = cam_center
    Rt = torch.linalg.inv(C2W)
    return Rt


def getProjectionMatrix2(znear, zfar, cx, cy, fx, fy, W, H):
    """:71-92"""
    left = ((2 * cx - W) / W - 1.0) * W / 2.0
    right = ((2 * cx - W) / W + 1.0) * W / 2.0
    top = ((2 * cy - H) / H + 1.0) * H / 2.0
    bottom = ((2 * cy - H) / H - 1.0) * H / 2.0
    left = znear / fx * left
    right = znear / fx * right
    top = znear / fy * top
    bottom = znear / fy * bottom
    P = torch.zeros(4, 4)
    z_sign = 1.0
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = z_sign
    P[2, 2] = z_sign * zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


# -- utils/camera_utils.py ----------------------------------------------------------------------
class RefCamera:
    """utils/camera_utils.py:8-112 without the dataset plumbing."""

    def __init__(self, uid, W, H, fx, fy, cx, cy, w2c=None, device="cuda", znear=0.01, zfar=100.0):
        self.uid = uid
        self.device = device
        T = torch.eye(4, device=device) if w2c is None else w2c.float().to(device)
        self.R = T[:3, :3]
        self.T = T[:3, 3]
        self.fx, self.fy, self.cx, self.cy = fx, fy, cx, cy
        self.FoVx = 2 * math.atan(W / (2 * fx))          # focal2fov, graphics_utils.py:99-100
        self.FoVy = 2 * math.atan(H / (2 * fy))
        self.image_height, self.image_width = H, W
        P = torch.nn.Parameter
        self.cam_rot_delta = P(torch.zeros(3, requires_grad=True, device=device))
        self.cam_trans_delta = P(torch.zeros(3, requires_grad=True, device=device))
        self.exposure_a = P(torch.tensor([0.0], requires_grad=True, device=device))
        self.exposure_b = P(torch.tensor([0.0], requires_grad=True, device=device))
        # utils/slam_frontend.py:318-328 (transposed at construction, camera_utils.py:86-88)
        self.projection_matrix = getProjectionMatrix2(
            znear=znear, zfar=zfar, fx=fx, fy=fy, cx=cx, cy=cy, W=W, H=H).transpose(0, 1).to(device=device)
        self.original_image = None
        self.depth = None
        self.grad_mask = None

    @property
    def world_view_transform(self):          # :94-96
        return getWorld2View2(self.R, self.T).transpose(0, 1)

    @property
    def full_proj_transform(self):           # :98-104
        return (self.world_view_transform.unsqueeze(0).bmm(self.projection_matrix.unsqueeze(0))).squeeze(0)

    @property
    def camera_center(self):                 # :106-107
        return self.world_view_transform.inverse()[3, :3]

    def update_RT(self, R, t):               # :109-111
        self.R = R.to(device=self.device)
        self.T = t.to(device=self.device)


# -- utils/pose_utils.py ------------------------------------------------------------------------
def skew_sym_mat(x):
    ssm = torch.zeros(3, 3, device=x.device, dtype=x.dtype)
    ssm[0, 1] = -x[2]
    ssm[0, 2] = x[1]
    ssm[1, 0] = x[2]
    ssm[1, 2] = -x[0]
    ssm[2, 0] = -x[1]
    ssm[2, 1] = x[0]
    return ssm


def SO3_exp(theta):
    W = skew_sym_mat(theta)
    W2 = W @ W
    angle = torch.norm(theta)
    I = torch.eye(3, device=theta.device, dtype=theta.dtype)
    if angle < 1e-5:
        return I + W + 0.5 * W2
    return I + (torch.sin(angle) / angle) * W + ((1 - torch.cos(angle)) / (angle ** 2)) * W2


def V(theta):
    I = torch.eye(3, device=theta.device, dtype=theta.dtype)
    W = skew_sym_mat(theta)
    W2 = W @ W
    angle = torch.norm(theta)
    if angle < 1e-5:
        return I + 0.5 * W + (1.0 / 6.0) * W2
    return I + W * ((1.0 - torch.cos(angle)) / (angle ** 2)) + W2 * ((angle - torch.sin(angle)) / (angle ** 3))


def SE3_exp(tau):
    rho, theta = tau[:3], tau[3:]
    R = SO3_exp(theta)
    t = V(theta) @ rho
    T = torch.eye(4, device=tau.device, dtype=tau.dtype)
    T[:3, :3] = R
    T[:3, 3] = t
    return T


def update_pose(camera, converged_threshold=1e-4):
    """:76-93"""
    tau = torch.cat([camera.cam_trans_delta, camera.cam_rot_delta], axis=0)
    T_w2c = torch.eye(4, device=tau.device)
    T_w2c[0:3, 0:3] = camera.R
    T_w2c[0:3, 3] = camera.T
    new_w2c = SE3_exp(tau) @ T_w2c
    new_R = new_w2c[0:3, 0:3]
    new_T = new_w2c[0:3, 3]
    converged = tau.norm() < converged_threshold
    camera.update_RT(new_R, new_T)
    camera.cam_rot_delta.data.fill_(0)
    camera.cam_trans_delta.data.fill_(0)
    return converged


# -- utils/slam_utils.py ------------------------------------------------------------------------
def get_loss_mapping_rgbd(image, depth, viewpoint, gt_image, gt_depth, alpha=0.95,
                          rgb_boundary_threshold=0.01, initialization=False):
    """get_loss_mapping (:91-99) -> get_loss_mapping_rgbd (:113-128); gt tensors are passed in
    instead of read from viewpoint.original_image / viewpoint.depth."""
    if initialization:
        image_ab = image
    else:
        image_ab = (torch.exp(viewpoint.exposure_a)) * image + viewpoint.exposure_b
    rgb_pixel_mask = (gt_image.sum(dim=0) > rgb_boundary_threshold).view(*depth.shape)
    depth_pixel_mask = (gt_depth > 0.01).view(*depth.shape)
    l1_rgb = torch.abs(image_ab * rgb_pixel_mask - gt_image * rgb_pixel_mask)
    l1_depth = torch.abs(depth * depth_pixel_mask - gt_depth * depth_pixel_mask)
    return alpha * l1_rgb.mean() + (1 - alpha) * l1_depth.mean()


def get_loss_tracking_rgbd(image, depth, opacity, viewpoint, gt_image, gt_depth, alpha=0.95,
                           rgb_boundary_threshold=0.01, grad_mask=None):
    """get_loss_tracking (:56-61) -> get_loss_tracking_rgbd (:74-88) -> get_loss_tracking_rgb (:64-72);
    grad_mask = None stands for an all-ones mask."""
    image_ab = (torch.exp(viewpoint.exposure_a)) * image + viewpoint.exposure_b
    _, h, w = gt_image.shape
    rgb_pixel_mask = (gt_image.sum(dim=0) > rgb_boundary_threshold).view(1, h, w)
    if grad_mask is not None:
        rgb_pixel_mask = rgb_pixel_mask * grad_mask
    l1 = opacity * torch.abs(image_ab * rgb_pixel_mask - gt_image * rgb_pixel_mask)
    l1_rgb = l1.mean()
    depth_pixel_mask = (gt_depth > 0.01).view(*depth.shape)
    opacity_mask = (opacity > 0.95).view(*depth.shape)
    depth_mask = depth_pixel_mask * opacity_mask
    l1_depth = torch.abs(depth * depth_mask - gt_depth * depth_mask)
    return alpha * l1_rgb + (1 - alpha) * l1_depth.mean()


# -- the loops ----------------------------------------------------------------------------------
def make_cameras(t, device):
    """keyframe cameras of a synthdata.scene_tensors() scene; uid = index (uid 0 is never moved)."""
    return [RefCamera(k, t.W, t.H, t.fx, t.fy, t.cx, t.cy, w2c=m, device=device) for k, m in enumerate(t.w2c)]


def make_model(t, device, params=None):
    p = params if params is not None else dict(xyz=t.xyz, f_dc=t.f_dc, f_rest=t.f_rest, opacity=t.opacity,
                                               scaling=t.scaling, rotation=t.rotation)
    to = lambda x: x.to(device).contiguous()
    return ref_ext.RefGaussianModel(to(p["xyz"]), to(p["f_dc"]), to(p["f_rest"]), to(p["opacity"]),
                                    to(p["scaling"]), to(p["rotation"]), sh_degree=t.sh_degree)


class RefMappingWindow:
    """One iteration of BackEnd.map (utils/slam_backend.py:156-317) over a fixed keyframe window:
    render + mapping loss per keyframe, isotropic regulariser, ONE backward, densification
    statistics, torch.optim.Adam over the six Gaussian groups, the keyframe optimizer
    (utils/slam_backend.py:434-471: pose deltas and exposure for every keyframe but uid 0), update_pose."""

    def __init__(self, gaussians, keyframes, bg, pose_lr=(0.0015, 0.0005, 0.01, 0.01), densify_stats=True):
        """pose_lr: Training.lr.cam_rot_delta * 0.5, cam_trans_delta * 0.5 (configs/rgbd/replica/
        base_config.yaml:59-61, utils/slam_backend.py:443-452), exposure 0.01."""
        self.g, self.kfs, self.bg = gaussians, keyframes, bg
        self.densify_stats = densify_stats
        opt_params = []
        for kf in keyframes:
            vp = kf["cam"]
            if vp.uid == 0:     # :435-436: keyframe 0 has no optimised parameter at all
                continue
            # :438-455 (cam_idx < frames_to_optimize: the whole window here)
            opt_params.append({"params": [vp.cam_rot_delta], "lr": pose_lr[0], "name": f"rot_{vp.uid}"})
            opt_params.append({"params": [vp.cam_trans_delta], "lr": pose_lr[1], "name": f"trans_{vp.uid}"})
            opt_params.append({"params": [vp.exposure_a], "lr": pose_lr[2], "name": f"exposure_a_{vp.uid}"})
            opt_params.append({"params": [vp.exposure_b], "lr": pose_lr[3], "name": f"exposure_b_{vp.uid}"})
        self.keyframe_optimizers = torch.optim.Adam(opt_params)
        self.last = {}

    def step(self):
        g = self.g
        loss_mapping = 0
        pkgs = []
        for kf in self.kfs:
            vp = kf["cam"]
            pkg = ref_ext.ref_render(vp, g, None, self.bg)
            loss_mapping += get_loss_mapping_rgbd(pkg["render"], pkg["depth"], vp, kf["gt_image"], kf["gt_depth"])
            pkgs.append(pkg)
        scaling = g.get_scaling
        isotropic_loss = torch.abs(scaling - scaling.mean(dim=1).view(-1, 1))
        loss_mapping += 10 * isotropic_loss.mean()
        loss_mapping.backward()
        with torch.no_grad():
            if self.densify_stats:
                for pkg in pkgs:       # :277-284
                    vis = pkg["visibility_filter"]
                    g.max_radii2D[vis] = torch.max(g.max_radii2D[vis], pkg["radii"][vis])
                    g.add_densification_stats(pkg["viewspace_points"], vis)
            g.optimizer.step()
            g.optimizer.zero_grad(set_to_none=True)
            self.keyframe_optimizers.step()
            self.keyframe_optimizers.zero_grad(set_to_none=True)
            for kf in self.kfs:        # :312-317
                if kf["cam"].uid == 0:
                    continue
                update_pose(kf["cam"])
        self.last = {i: p for i, p in enumerate(pkgs)}
        return loss_mapping.detach()

    def flush(self):
        pass


class RefTrackingLoop:
    """FrontEnd.tracking (utils/slam_frontend.py:128-196) on one frame: `iters` iterations of
    render -> tracking loss -> backward -> Adam(rot 0.003, trans 0.001, exposure 0.01) -> update_pose;
    the convergence break is optional (BASELINE config 2 disables it)."""

    def __init__(self, gaussians, camera, bg, lr_rot=0.003, lr_trans=0.001, lr_exposure=0.01, iters=100):
        self.g, self.cam, self.bg, self.iters = gaussians, camera, bg, iters
        self.lrs = (lr_rot, lr_trans, lr_exposure)
        self.opt = None
        self.gt_image = self.gt_depth = None

    def load_frame(self, R, T, gt_image, gt_depth, exposure=(0.0, 0.0)):
        vp = self.cam
        vp.update_RT(R, T)                  # :129-130
        self.gt_image, self.gt_depth = gt_image, gt_depth
        with torch.no_grad():
            vp.cam_rot_delta.zero_(); vp.cam_trans_delta.zero_()
            vp.exposure_a.fill_(exposure[0]); vp.exposure_b.fill_(exposure[1])
        self.opt = torch.optim.Adam([       # :132-161
            {"params": [vp.cam_rot_delta], "lr": self.lrs[0], "name": "rot"},
            {"params": [vp.cam_trans_delta], "lr": self.lrs[1], "name": "trans"},
            {"params": [vp.exposure_a], "lr": self.lrs[2], "name": "exposure_a"},
            {"params": [vp.exposure_b], "lr": self.lrs[2], "name": "exposure_b"}])

    def track(self, iters=None, check_converged=False):
        vp = self.cam
        for _ in range(self.iters if iters is None else iters):
            pkg = ref_ext.ref_render(vp, self.g, None, self.bg)
            self.opt.zero_grad()            # :171 (set_to_none by default)
            loss = get_loss_tracking_rgbd(pkg["render"], pkg["depth"], pkg["opacity"], vp, self.gt_image,
                                          self.gt_depth)
            loss.backward()
            with torch.no_grad():
                self.opt.step()
                converged = update_pose(vp)
            self.loss, self.pkg = loss.detach(), pkg
            if check_converged and bool(converged):
                break
        return self.pkg


def build_window_gt(t, cams, bg, device, seed=77):
    """keyframes with ground truth rendered (by the reference rasterizer) from the 5 %-perturbed copy
    of the scene (SURVEY.md §8d 'Loop-level')."""
    import synthdata
    dev_t = lambda x: x.to(device)
    from types import SimpleNamespace
    on = SimpleNamespace(xyz=dev_t(t.xyz), scaling=dev_t(t.scaling), rotation=dev_t(t.rotation),
                         opacity=dev_t(t.opacity), f_dc=dev_t(t.f_dc), f_rest=dev_t(t.f_rest))
    gt = synthdata.perturbed_params(on, seed=seed)
    gm = make_model(t, device, gt)
    kfs = []
    with torch.no_grad():
        for cam in cams:
            pkg = ref_ext.ref_render(cam, gm, None, bg)
            kfs.append(dict(cam=cam, gt_image=pkg["render"].detach().clone(), gt_depth=pkg["depth"].detach().clone()))
    del gm
    return kfs
