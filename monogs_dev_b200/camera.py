"""`Camera`: the part of the reference's utils/camera_utils.py:Camera (:8-112) that the rasterizer
path reads, with the pose-derived matrices in ONE device block so that a captured CUDA graph stays
valid when the pose changes.

Conventions restated from the reference:
  world_view_transform = getWorld2View2(R, T)^T      utils/camera_utils.py:94-96
  projection_matrix    = getProjectionMatrix2(...)^T utils/slam_frontend.py:318-328,
                                                     gaussian_splatting/utils/graphics_utils.py:71-92
  full_proj_transform  = view @ proj                 utils/camera_utils.py:98-104
  camera_center        = inverse(view)[3, :3]        utils/camera_utils.py:106-107
  pose update          = Exp(tau) @ T_w2c, tau=(rho, theta)   utils/pose_utils.py:66-93
`Camera.from_reference(vp)` wraps a MonoGS viewpoint for MappingWindow / TrackingLoop.
"""
import math

import torch

from .rasterizer import GaussianRasterizationSettings


def projection_matrix(znear, zfar, cx, cy, fx, fy, W, H):
    left = ((2 * cx - W) / W - 1.0) * W / 2.0
    right = ((2 * cx - W) / W + 1.0) * W / 2.0
    top = ((2 * cy - H) / H + 1.0) * H / 2.0
    bottom = ((2 * cy - H) / H - 1.0) * H / 2.0
    left, right = znear / fx * left, znear / fx * right
    top, bottom = znear / fy * top, znear / fy * bottom
    P = torch.zeros(4, 4)
    P[0, 0] = 2.0 * znear / (right - left)
    P[1, 1] = 2.0 * znear / (top - bottom)
    P[0, 2] = (right + left) / (right - left)
    P[1, 2] = (top + bottom) / (top - bottom)
    P[3, 2] = 1.0
    P[2, 2] = zfar / (zfar - znear)
    P[2, 3] = -(zfar * znear) / (zfar - znear)
    return P


def skew(x):
    m = torch.zeros(3, 3, dtype=x.dtype)
    m[0, 1], m[0, 2], m[1, 0], m[1, 2], m[2, 0], m[2, 1] = -x[2], x[1], x[2], -x[0], -x[1], x[0]
    return m


def se3_exp(tau):
    """tau = (rho, theta) -> 4x4 (utils/pose_utils.py:24-73)."""
    rho, theta = tau[:3], tau[3:]
    W = skew(theta)
    W2 = W @ W
    a = torch.norm(theta)
    I = torch.eye(3, dtype=tau.dtype)
    if a < 1e-5:
        R = I + W + 0.5 * W2
        V = I + 0.5 * W + (1.0 / 6.0) * W2
    else:
        R = I + (torch.sin(a) / a) * W + ((1 - torch.cos(a)) / a ** 2) * W2
        V = I + W * ((1.0 - torch.cos(a)) / a ** 2) + W2 * ((a - torch.sin(a)) / a ** 3)
    T = torch.eye(4, dtype=tau.dtype)
    T[:3, :3] = R
    T[:3, 3] = V @ rho
    return T


class Camera:
    """The part of utils/camera_utils.py:Camera that the rasterizer path reads.

    Pose-derived tensors live in ONE 64-float device block (layout of b2gs_pose_update,
    include/b2gs.h): R | T | viewmatrix | full projection | camera centre.  The properties
    below are stable views into it, so a CUDA graph that captured them stays valid when
    the pose changes; on CUDA `update_pose()` runs on the device (no .cpu() round trip)."""

    def __init__(self, W, H, fx, fy, cx, cy, w2c=None, device="cpu", znear=0.01, zfar=100.0):
        self.image_width, self.image_height = W, H
        self.fx, self.fy, self.cx, self.cy = fx, fy, cx, cy
        self.FoVx = 2 * math.atan(W / (2 * fx))
        self.FoVy = 2 * math.atan(H / (2 * fy))
        self.device = torch.device(device)
        w2c = torch.eye(4) if w2c is None else w2c.float()
        self.projection_matrix = projection_matrix(znear, zfar, cx, cy, fx, fy, W, H).t().contiguous().to(device)
        self.state = torch.zeros(64, device=device)
        self._host_RT = None
        self.update_RT(w2c[:3, :3], w2c[:3, 3])
        self.cam_rot_delta = torch.zeros(3, device=device, requires_grad=True)
        self.cam_trans_delta = torch.zeros(3, device=device, requires_grad=True)
        self.exposure_a = torch.zeros(1, device=device, requires_grad=True)
        self.exposure_b = torch.zeros(1, device=device, requires_grad=True)

    @classmethod
    def from_reference(cls, viewpoint, device=None):
        """wrap a MonoGS `Camera` (utils/camera_utils.py:8-64) for the loop-level entry points:
        same intrinsics and current pose, the SAME cam_rot_delta / cam_trans_delta / exposure tensors
        (an optimizer built on the reference object keeps working), uid / original_image / depth /
        grad_mask carried along.  Returns the frame dict {cam, gt_image, gt_depth} that
        MappingWindow, TrackingLoop, BackEnd and FrontEnd take."""
        vp = viewpoint
        dev = torch.device(device if device is not None else vp.cam_rot_delta.device)
        w2c = torch.eye(4)
        w2c[:3, :3], w2c[:3, 3] = vp.R.detach().float().cpu(), vp.T.detach().float().cpu()
        cam = cls(int(vp.image_width), int(vp.image_height), float(vp.fx), float(vp.fy), float(vp.cx),
                  float(vp.cy), w2c=w2c, device=dev)
        cam.FoVx, cam.FoVy = float(vp.FoVx), float(vp.FoVy)
        if getattr(vp, "projection_matrix", None) is not None:        # already transposed, as stored there
            cam.projection_matrix = vp.projection_matrix.detach().float().to(dev).contiguous()
            cam.update_RT(w2c[:3, :3], w2c[:3, 3])
        for name in ("cam_rot_delta", "cam_trans_delta", "exposure_a", "exposure_b"):
            setattr(cam, name, getattr(vp, name))
        cam.uid = getattr(vp, "uid", None)
        image, depth = getattr(vp, "original_image", None), getattr(vp, "depth", None)
        cam.original_image, cam.depth, cam.grad_mask = image, depth, getattr(vp, "grad_mask", None)
        frame = dict(cam=cam, gt_image=None, gt_depth=None)
        if image is not None:
            frame["gt_image"] = image.detach().float().to(dev)
        if depth is not None:                                         # a numpy [H,W] array in the reference
            frame["gt_depth"] = torch.as_tensor(depth, dtype=torch.float32).to(dev).reshape(
                1, int(vp.image_height), int(vp.image_width))
        return frame

    def _host(self):
        if self._host_RT is None:       # pose was last changed on the device: read it back
            h = self.state[:12].detach().cpu()
            self._host_RT = (h[:9].view(3, 3).clone(), h[9:12].clone())
        return self._host_RT

    R = property(lambda self: self._host()[0])
    T = property(lambda self: self._host()[1])

    @property
    def w2c(self):
        m = torch.eye(4)
        m[:3, :3], m[:3, 3] = self.R, self.T
        return m

    def update_RT(self, R, T):
        """view / full-projection / centre for (R, T), computed on the host and uploaded in one
        copy: the reference recomputes them (two 4x4 inversions, a bmm) on every property
        access (utils/camera_utils.py:94-107)."""
        R, T = R.detach().float().cpu().clone(), T.detach().float().cpu().clone()
        self._host_RT = (R, T)
        w2c = torch.eye(4)
        w2c[:3, :3], w2c[:3, 3] = R, T
        view = w2c.t().contiguous()
        full = view @ self.projection_matrix.cpu()
        center = torch.linalg.inv(view)[3, :3]
        h = torch.zeros(64)
        h[:9], h[9:12] = R.reshape(-1), T
        h[16:32], h[32:48], h[48:51] = view.reshape(-1), full.reshape(-1), center
        self.state.copy_(h)

    world_view_transform = property(lambda self: self.state[16:32].view(4, 4))
    full_proj_transform = property(lambda self: self.state[32:48].view(4, 4))
    camera_center = property(lambda self: self.state[48:51])

    def update_pose(self, converged_threshold=1e-4, check_converged=True):
        """utils/pose_utils.py:76-93.  On CUDA the update is one kernel on the current stream;
        `check_converged=False` skips the only host sync (the convergence read-back)."""
        if self.state.is_cuda:
            from ._lib import lib, check
            norm = torch.empty(1, device=self.state.device) if check_converged else None
            with torch.cuda.device(self.state.device):
                check(lib.b2gs_pose_update(self.state.data_ptr(), self.projection_matrix.data_ptr(),
                                           self.cam_rot_delta.data_ptr(), self.cam_trans_delta.data_ptr(),
                                           norm.data_ptr() if norm is not None else None,
                                           torch.cuda.current_stream().cuda_stream))
            self._host_RT = None
            return bool(norm.item() < converged_threshold) if check_converged else False
        tau = torch.cat([self.cam_trans_delta, self.cam_rot_delta]).detach().cpu()
        new = se3_exp(tau) @ self.w2c
        self.update_RT(new[:3, :3], new[:3, 3])
        self.cam_rot_delta.data.fill_(0)
        self.cam_trans_delta.data.fill_(0)
        return bool(tau.norm() < converged_threshold)

    def settings(self, bg, sh_degree=0, scale_modifier=1.0, debug=False):
        return GaussianRasterizationSettings(
            image_height=int(self.image_height), image_width=int(self.image_width),
            tanfovx=math.tan(self.FoVx * 0.5), tanfovy=math.tan(self.FoVy * 0.5), bg=bg,
            scale_modifier=scale_modifier, viewmatrix=self.world_view_transform,
            projmatrix=self.full_proj_transform, projmatrix_raw=self.projection_matrix,
            sh_degree=sh_degree, campos=self.camera_center, prefiltered=False, debug=debug)
