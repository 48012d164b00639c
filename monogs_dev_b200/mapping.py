"""Keyframe mapping window: the one path sharded across GPUs.

Host-side mirror of the reference's callers of the rasterizer:
  render()                      gaussian_splatting/gaussian_renderer/__init__.py:24-153
  GaussianModel (params/Adam)   gaussian_splatting/scene/gaussian_model.py:76-101,245-310
  get_loss_mapping(_rgb/_rgbd)  utils/slam_utils.py:91-128
  BackEnd.map() inner loop      utils/slam_backend.py:156-317
  FrontEnd.tracking() loop      utils/slam_frontend.py:163-193, utils/slam_utils.py:56-88

B200-first changes: all Gaussian parameters live in ONE packed fp32 arena (and their
gradients / Adam moments in three more), so that
  * the optimizer step is one fused kernel launch (libb2gs `b2gs_adam_step`),
  * with N GPUs (one process per GPU, keyframes sharded rank-sticky) the per-Gaussian
    gradients of all groups travel in ONE NCCL all-reduce over NVLink instead of six.
The loss is a plain sum over views (utils/slam_backend.py:160-231), so keyframe sharding
with a SUM all-reduce is exact up to fp32 summation order.
"""
import contextlib
import math

import torch
import torch.distributed as dist

from .adam import FusedAdam
from .losses import get_loss_mapping_rgbd, get_loss_tracking_rgbd, scale_regulariser  # csrc/loss.cu
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, GradSinks,
                         rasterize_gaussians_raw)

_GROUPS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


class GaussianModel:
    """Parameter container with the attribute surface render() and BackEnd.map() use.
    Learning rates follow configs/rgbd/replica/base_config.yaml:64-80 with
    spatial_lr_scale = 6.0 (slam.py:54)."""

    def __init__(self, xyz, f_dc, f_rest, opacity, scaling, rotation, sh_degree=0,
                 spatial_lr_scale=6.0, lrs=None):
        self.max_sh_degree = self.active_sh_degree = sh_degree
        self.spatial_lr_scale = spatial_lr_scale
        self.percent_dense = 0.01          # opt_params.percent_dense (GS/arguments/__init__.py)
        lr = dict(xyz=0.00016 * spatial_lr_scale, f_dc=0.0025, f_rest=0.0025 / 20.0, opacity=0.05,
                  scaling=0.001 * spatial_lr_scale, rotation=0.001)
        lr.update(lrs or {})
        self._lr = lr
        # xyz schedule of update_learning_rate (configs/rgbd/replica/base_config.yaml:66-69);
        # training_setup() overrides it from an opt_params object
        self.lr_init, self.lr_final = lr["xyz"], 0.0000016 * spatial_lr_scale
        self.lr_delay_mult, self.max_steps = 0.01, 30000
        self.optimizer = None
        self._allocate(dict(xyz=xyz, f_dc=f_dc, f_rest=f_rest, opacity=opacity, scaling=scaling,
                            rotation=rotation))

    def _allocate(self, src, moments=None, steps=None, shapes=None, device=None):
        """(re)build the four packed arenas (parameters, gradients, Adam m / v) from per-group
        tensors; used at construction and after densify / prune (the reference re-wraps six
        nn.Parameters and twelve moment tensors, GS/scene/gaussian_model.py:468-555).
        src=None with `shapes` (group -> shape) and `device`: arenas are allocated and wired up but
        left for the caller to fill (the densify kernels write the new rows in place)."""
        if src is None:
            src = {k: torch.empty(tuple(shapes[k]), device="meta") for k in _GROUPS}
        fill = src["xyz"].device.type != "meta"
        dev = src["xyz"].device if fill else torch.device(device)
        capturable = bool(getattr(self.optimizer, "capturable", False))
        step_dev = getattr(self.optimizer, "_step_dev", None)     # device step count survives re-allocation
        if self.optimizer is not None:
            self._lr = {g["name"]: g["lr"] for g in self.optimizer.param_groups}
        sizes = [src[k].numel() for k in _GROUPS]
        offs, total = [], 0
        for n in sizes:
            offs.append(total)
            total += (n + 3) // 4 * 4          # keep every segment 16 B aligned for float4 access
        peer = getattr(self, "peer", None)
        if peer is not None:                   # arenas every rank of the node can address (peer.py)
            self.flat, self.flat_grad = peer.alloc(total)
        else:
            self.flat = torch.zeros(total, device=dev)
            self.flat_grad = torch.zeros(total, device=dev)
        self.flat_m = torch.zeros(total, device=dev)
        self.flat_v = torch.zeros(total, device=dev)
        self._params = {}
        for k, o, n in zip(_GROUPS, offs, sizes):
            view = self.flat[o:o + n].view(src[k].shape)
            if fill:
                view.copy_(src[k])
            p = torch.nn.Parameter(view)
            p.grad = self.flat_grad[o:o + n].view(src[k].shape)
            self._params[k] = p
        # optimizer_cls: the CPU host-logic tests run the same arena under torch.optim.Adam (the fused
        # optimizer has no CPU path); torch's Adam keeps its step count as a tensor
        cls = getattr(self, "optimizer_cls", FusedAdam)
        fused = cls is FusedAdam
        self.optimizer = cls([dict(params=[self._params[k]], lr=self._lr[k], name=k) for k in _GROUPS],
                             lr=0.0, eps=1e-15, **(dict(capturable=capturable) if fused else {}))
        for k, o, n in zip(_GROUPS, offs, sizes):   # Adam moments are views of the arenas too
            st = self.optimizer.state[self._params[k]]
            step = (steps or {}).get(k, 0)
            st["step"] = step if fused else torch.tensor(float(step))
            st["exp_avg"] = self.flat_m[o:o + n].view(src[k].shape)
            st["exp_avg_sq"] = self.flat_v[o:o + n].view(src[k].shape)
            if moments is not None:
                st["exp_avg"].copy_(moments[k][0])
                st["exp_avg_sq"].copy_(moments[k][1])
        self.optimizer.peer = peer
        self.optimizer._step_dev = step_dev
        for a in ("xyz_gradient_accum", "denom", "max_radii2D"):
            if hasattr(self, a):
                delattr(self, a)

    def enable_peer_arena(self, peer):
        """move the arenas into peer-mapped memory and make optimizer.step() the sharded
        all-reduce + Adam + zero_grad kernel (csrc/adam.cu: adam_peer_kernel).  Collective: every
        rank calls it with identical parameters.  The Adam moments of a chunk then live on the
        rank that owns it; `gather_moments()` rebuilds full copies (needed before densify)."""
        src = {k: self._params[k].detach().clone() for k in _GROUPS}
        mom = {k: (self.optimizer.state[self._params[k]]["exp_avg"].clone(),
                   self.optimizer.state[self._params[k]]["exp_avg_sq"].clone()) for k in _GROUPS}
        steps = {k: self.optimizer.state[self._params[k]]["step"] for k in _GROUPS}
        self.peer = peer
        self.optimizer.capturable = True
        self._allocate(src, mom, steps)

    def gather_moments(self):
        """peer mode: moments are only maintained by each chunk's owner (the other ranks' copies
        keep their value from before enable_peer_arena); sum the owners' copies back together."""
        peer = getattr(self, "peer", None)
        if peer is None:
            return
        chunk = 1024                            # csrc/adam.cu: PEER_CHUNK
        idx = torch.arange(self.flat_m.numel(), device=self.flat_m.device)
        # chunk ownership is per segment (chunks restart at every segment start)
        own = torch.zeros_like(idx, dtype=torch.bool)
        c0 = 0
        for k in _GROUPS:
            p = self._params[k]
            o = (p.data_ptr() - self.flat.data_ptr()) // 4
            n = p.numel()
            ch = c0 + torch.arange(n, device=idx.device) // chunk
            own[o:o + n] = (ch % peer.world) == peer.rank
            c0 += (n + chunk - 1) // chunk
        for t in (self.flat_m, self.flat_v):
            t.mul_(own)
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=peer.group)

    def sync_densification_stats(self):
        """data-parallel window: every rank has accumulated the statistics of ITS views only; the
        densification decision must be the same on all replicas -> sum / max them over ranks."""
        grp = getattr(self, "_dist_group", None)
        if grp is None or not hasattr(self, "xyz_gradient_accum"):
            return
        dist.all_reduce(self.xyz_gradient_accum, op=dist.ReduceOp.SUM, group=grp)
        dist.all_reduce(self.denom, op=dist.ReduceOp.SUM, group=grp)
        dist.all_reduce(self.max_radii2D, op=dist.ReduceOp.MAX, group=grp)

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, **kw):
        from .densify import densify_and_prune
        self.sync_densification_stats()
        self.gather_moments()              # peer mode: full Adam moments before they are re-packed
        return densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, **kw)

    # keyframe seeding (gaussian_model.py:107-227): back-projection, down-sample, distCUDA2 scales
    def create_pcd_from_image(self, cam_info, init=False, scale=2.0, depthmap=None, generator=None):
        from .seeding import create_pcd_from_image
        return create_pcd_from_image(self, cam_info, init, scale, depthmap, generator)

    def extend_from_pcd(self, fused_point_cloud, features, scales, rots, opacities, kf_id):
        from .seeding import extend_from_pcd
        return extend_from_pcd(self, fused_point_cloud, features, scales, rots, opacities, kf_id)

    def extend_from_pcd_seq(self, cam_info, kf_id=-1, init=False, scale=2.0, depthmap=None, generator=None):
        from .seeding import extend_from_pcd_seq
        return extend_from_pcd_seq(self, cam_info, kf_id, init, scale, depthmap, generator)

    def prune_points(self, mask):
        from .densify import prune_points
        self.gather_moments()
        return prune_points(self, mask)

    # schedule, opacity resets, SH degree, PLY (gaussian_model.py:103-105,205-206,245-465): model_io.py
    def init_lr(self, spatial_lr_scale):
        self.spatial_lr_scale = spatial_lr_scale

    def training_setup(self, training_args):
        from .model_io import training_setup
        return training_setup(self, training_args)

    def update_learning_rate(self, iteration):
        from .model_io import update_learning_rate
        return update_learning_rate(self, iteration)

    def oneupSHdegree(self):
        from .model_io import oneupSHdegree
        return oneupSHdegree(self)

    def reset_opacity(self):
        from .model_io import reset_opacity
        return reset_opacity(self)

    def reset_opacity_nonvisible(self, visibility_filters):
        from .model_io import reset_opacity_nonvisible
        return reset_opacity_nonvisible(self, visibility_filters)

    def construct_list_of_attributes(self):
        from .model_io import construct_list_of_attributes
        return construct_list_of_attributes(self)

    def save_ply(self, path):
        from .model_io import save_ply
        return save_ply(self, path)

    def load_ply(self, path):
        """replaces the whole map by the file's Gaussians: fresh arenas, zero Adam moments,
        active_sh_degree = max_sh_degree, statistics and keyframe ids reset (:377-466)."""
        from .model_io import load_ply_tensors
        src = load_ply_tensors(path, self.max_sh_degree, self.flat.device)
        self._allocate(src)
        self.optimizer._step_dev = None         # a new map: the device step count restarts too
        self.active_sh_degree = self.max_sh_degree
        P = src["xyz"].shape[0]
        self.unique_kfIDs = torch.zeros(P, dtype=torch.int32, device=self.flat.device)
        self.n_obs = torch.zeros(P, dtype=torch.int32, device=self.flat.device)
        self._stats()

    _xyz = property(lambda s: s._params["xyz"])
    _features_dc = property(lambda s: s._params["f_dc"])
    _features_rest = property(lambda s: s._params["f_rest"])
    _opacity = property(lambda s: s._params["opacity"])
    _scaling = property(lambda s: s._params["scaling"])
    _rotation = property(lambda s: s._params["rotation"])

    # activations (gaussian_model.py:76-101)
    get_xyz = property(lambda s: s._xyz)
    get_scaling = property(lambda s: torch.exp(s._scaling))
    get_rotation = property(lambda s: torch.nn.functional.normalize(s._rotation))
    get_opacity = property(lambda s: torch.sigmoid(s._opacity))
    get_features = property(lambda s: torch.cat((s._features_dc, s._features_rest), dim=1))

    def get_covariance(self, scaling_modifier=1):
        """[P,6] upper triangle of R S S^T R^T with the normalised quaternion
        (build_covariance_from_scaling_rotation, gaussian_model.py:42-49,98-101); only used by
        render() under pipe.compute_cov3D_python."""
        from .densify import build_rotation
        L = build_rotation(self._rotation) * (scaling_modifier * self.get_scaling).unsqueeze(1)   # R @ diag(s)
        cov = L @ L.transpose(1, 2)
        return torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], dim=-1)

    def zero_grad(self):
        self.flat_grad.zero_()

    # densification statistics (GS/scene/gaussian_model.py:233-235,691-695)
    def _stats(self):
        if not hasattr(self, "xyz_gradient_accum"):
            P, dev = self._xyz.shape[0], self.flat.device
            self.xyz_gradient_accum = torch.zeros(P, 1, device=dev)
            self.denom = torch.zeros(P, 1, device=dev)
            self.max_radii2D = torch.zeros(P, device=dev)
        return self.xyz_gradient_accum, self.denom, self.max_radii2D

    def add_densification_stats(self, viewspace_point_tensor, update_filter):
        acc, den, _ = self._stats()
        acc[update_filter] += torch.norm(viewspace_point_tensor.grad[update_filter, :2], dim=-1, keepdim=True)
        den[update_filter] += 1

    def grad_sinks(self, with_stats=True):
        """gradient arena views + statistics buffers for the fused-activation rasterizer."""
        st = self._stats() if with_stats else (None, None, None)
        g = {k: self._params[k].grad for k in _GROUPS}
        return GradSinks(g["xyz"], g["f_dc"], g["f_rest"], g["opacity"], g["scaling"], g["rotation"], *st)


def render(viewpoint_camera, pc, pipe=None, bg_color=None, scaling_modifier=1.0, override_color=None):
    """Same contract as the reference's render(): returns the dict with keys render,
    viewspace_points, visibility_filter, radii, depth, opacity, n_touched."""
    screenspace_points = torch.zeros_like(pc.get_xyz, dtype=pc.get_xyz.dtype, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()
    except Exception:
        pass
    cam = viewpoint_camera
    rs = GaussianRasterizationSettings(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, projmatrix_raw=cam.projection_matrix,
        sh_degree=pc.active_sh_degree, campos=cam.camera_center, prefiltered=False, debug=False)
    rasterizer = GaussianRasterizer(raster_settings=rs)
    # the two PipelineParams switches (gaussian_renderer/__init__.py:82-111); both default to False
    scales = rotations = cov3D = None
    if getattr(pipe, "compute_cov3D_python", False):
        cov3D = pc.get_covariance(scaling_modifier)
    else:
        scales = pc.get_scaling
        if scales.shape[-1] == 1:              # isotropic model: one scale per Gaussian
            scales = scales.repeat(1, 3)
        rotations = pc.get_rotation
    shs, colors = (pc.get_features, None) if override_color is None else (None, override_color)
    if colors is None and getattr(pipe, "convert_SHs_python", False):
        from .sh_utils import eval_sh
        feats = pc.get_features
        dirs = pc.get_xyz - cam.camera_center.reshape(1, 3)
        dirs = dirs / dirs.norm(dim=1, keepdim=True)
        shs, colors = None, torch.clamp_min(eval_sh(pc.active_sh_degree, feats.transpose(1, 2), dirs) + 0.5, 0.0)
    rendered_image, radii, depth, opacity, n_touched = rasterizer(
        means3D=pc.get_xyz, means2D=screenspace_points, shs=shs, colors_precomp=colors,
        opacities=pc.get_opacity, scales=scales, rotations=rotations,
        cov3D_precomp=cov3D, theta=cam.cam_rot_delta, rho=cam.cam_trans_delta)
    return {"render": rendered_image, "viewspace_points": screenspace_points,
            "visibility_filter": radii > 0, "radii": radii, "depth": depth, "opacity": opacity,
            "n_touched": n_touched}


def render_fused(viewpoint_camera, pc, pipe=None, bg_color=None, scaling_modifier=1.0, with_stats=True):
    """render() with the parameter activations, the gradient accumulation into the optimiser's
    arena and the per-view densification statistics fused into the rasterizer kernels
    (SURVEY.md §8f row 1).  Same return dict; `viewspace_points` is None (its only consumer,
    add_densification_stats, is fused: pc.xyz_gradient_accum / denom / max_radii2D are updated
    by the backward)."""
    cam = viewpoint_camera
    rs = GaussianRasterizationSettings(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, projmatrix_raw=cam.projection_matrix,
        sh_degree=pc.active_sh_degree, campos=cam.camera_center, prefiltered=False, debug=False)
    rendered_image, radii, depth, opacity, n_touched = rasterize_gaussians_raw(
        pc._xyz, pc._features_dc, pc._features_rest, pc._opacity, pc._scaling, pc._rotation, rs,
        theta=cam.cam_rot_delta, rho=cam.cam_trans_delta, sinks=pc.grad_sinks(with_stats))
    return {"render": rendered_image, "viewspace_points": None, "visibility_filter": radii > 0,
            "radii": radii, "depth": depth, "opacity": opacity, "n_touched": n_touched,
            "stats_fused": with_stats}


def get_loss_mapping_rgbd_torch(image, depth, viewpoint, gt_image, gt_depth, alpha=0.95,
                                rgb_boundary_threshold=0.01, initialization=False):
    """utils/slam_utils.py:91-96,113-128 restated op for op in torch (what the reference arm of
    bench.py runs; the product path uses the fused kernels of losses.py)."""
    image_ab = image if initialization else torch.exp(viewpoint.exposure_a) * image + viewpoint.exposure_b
    rgb_pixel_mask = (gt_image.sum(dim=0) > rgb_boundary_threshold).view(*depth.shape)
    depth_pixel_mask = (gt_depth > 0.01).view(*depth.shape)
    l1_rgb = torch.abs(image_ab * rgb_pixel_mask - gt_image * rgb_pixel_mask)
    l1_depth = torch.abs(depth * depth_pixel_mask - gt_depth * depth_pixel_mask)
    return alpha * l1_rgb.mean() + (1 - alpha) * l1_depth.mean()


def get_loss_tracking_rgbd_torch(image, depth, opacity, viewpoint, gt_image, gt_depth, alpha=0.95,
                                 rgb_boundary_threshold=0.01):
    """utils/slam_utils.py:56-88 in torch ops (grad_mask taken as all-ones)."""
    image_ab = torch.exp(viewpoint.exposure_a) * image + viewpoint.exposure_b
    rgb_pixel_mask = (gt_image.sum(dim=0) > rgb_boundary_threshold).view(1, *gt_image.shape[1:])
    l1_rgb = (opacity * torch.abs(image_ab * rgb_pixel_mask - gt_image * rgb_pixel_mask)).mean()
    depth_mask = (gt_depth > 0.01).view(*depth.shape) * (opacity > 0.95).view(*depth.shape)
    l1_depth = torch.abs(depth * depth_mask - gt_depth * depth_mask)
    return alpha * l1_rgb + (1 - alpha) * l1_depth.mean()


def shard_views(num_views, rank, world):
    """rank-sticky round-robin assignment of window keyframes."""
    return [v for v in range(num_views) if v % world == rank]


def auto_view_streams(gaussians, n_views, most=8):
    """How many forked streams the views of a window are issued on (view_streams=None).
    Concurrent view chains pay while one view's kernels leave SMs idle (cfg3, 500k Gaussians: 8
    streams 3.83 ms per step, 4: 3.88, 2: 3.97) and cost once a single view's per-Gaussian
    working set (geometry records + accumulator rows, ~96 B per Gaussian) no longer fits L2:
    every kernel then fills the GPU for many waves on its own and the concurrent views only evict
    each other's rows (cfg5, 4M Gaussians: 1 stream 60.8 ms, 2: 69.4, 8: 67.9, 16: 65.9)."""
    flat = getattr(gaussians, "flat", None)
    if flat is None or not flat.is_cuda:
        return 1
    P = int(gaussians.get_xyz.shape[0]) if hasattr(gaussians, "get_xyz") else 0
    l2 = int(getattr(torch.cuda.get_device_properties(flat.device), "L2_cache_size", 0)) or (96 << 20)
    return 1 if 96 * P > l2 else max(1, min(int(n_views), int(most)))


class MappingWindow:
    """One BackEnd.map() iteration over a keyframe window (utils/slam_backend.py:156-317),
    optionally data-parallel over keyframes and optionally replayed as CUDA graphs.

    keyframes: list of dicts {cam, gt_image[3,H,W], gt_depth[1,H,W]} — every rank holds
    the full list (only its shard's images are touched).

    cuda_graph=True: the step is captured once as two graphs —
        G1: every local view's render -> fused loss, the regulariser, one backward, and the OR of
            the views' binning-overflow flags into one device word;
        G2: [gradient exchange] -> fused Adam -> pose/exposure Adam -> pose updates -> zero grads,
            every state-changing kernel guarded by that word (b2gs_set_update_guard)
    — and replayed back to back with NO host sync inside or between steps.  The sync-free forward
    renders background-only when its instance capacity is exceeded; such a step discards its own
    update on the device.  The host reads the flags asynchronously and looks at them two step()
    calls later (or at flush()): G1 is then re-captured with a larger capacity and the discarded
    steps are run again, so after flush() every step() has been applied exactly once."""

    def __init__(self, gaussians, keyframes, bg, pose_lr=(0.003, 0.001, 0.01, 0.01),
                 optimize_pose=True, group=None, render_fn=None, adam_cls=FusedAdam,
                 loss_fn=None, cuda_graph=False, densify_stats=True, view_streams=1, peer_adam=False,
                 kf_optimizer=None, regulariser=True, lean=None):
        """keyframe dict flags (optional): `optimize_pose` (default: every view but the first, as
        viewpoint.uid == 0 is never moved, utils/slam_backend.py:315-316,434-436) and
        `optimize_exposure` (default True; the reference gives keyframe 0 and the two random
        views of an iteration no optimised parameter at all, :196-227,434-471).
        kf_optimizer: an existing optimizer over the camera parameters (BackEnd.keyframe_optimizers
        outlives a map() call); by default one is built for the local views.
        lean: run the step as straight libb2gs calls (window.LeanWindow: loss fused into the render
        epilogue, one finalize kernel, Adam + zero_grad fused, one camera kernel — no autograd graph,
        no torch kernel inside a step).  Default: whenever the window is the stock one (packed-arena
        GaussianModel on the GPU, FusedAdam, default renderer / RGB-D mapping loss, no external
        keyframe optimizer); lean=False keeps the autograd route over the same kernels."""
        self.g, self.kfs, self.bg = gaussians, keyframes, bg
        stock = (isinstance(gaussians, GaussianModel) and gaussians.flat.is_cuda and render_fn is None
                 and loss_fn is None and adam_cls is FusedAdam and kf_optimizer is None
                 and isinstance(gaussians.optimizer, FusedAdam)
                 and all(hasattr(kf["cam"], "state") for kf in keyframes))
        if lean and not stock:
            raise RuntimeError("lean=True needs the stock window (GaussianModel on CUDA, FusedAdam, default "
                               "render / loss functions, no external keyframe optimizer)")
        self._lean = None
        use_lean = stock if lean is None else bool(lean)
        self.regulariser = bool(regulariser)     # BackEnd.initialize_map has no isotropic term (:86-139)
        self.render_fn = render_fn or (render_fused if isinstance(gaussians, GaussianModel) else render)
        self.loss_fn = loss_fn or get_loss_mapping_rgbd
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.local = shard_views(len(keyframes), self.rank, self.world)
        self.optimize_pose = optimize_pose
        self.cuda_graph = bool(cuda_graph)
        self.densify_stats = bool(densify_stats)
        self.fused_reg = isinstance(gaussians, GaussianModel) and gaussians.flat.is_cuda
        if self.densify_stats and hasattr(gaussians, "_stats"):
            gaussians._stats()
        groups = []
        for v in ([] if use_lean else self.local):       # utils/slam_backend.py:434-471 (one Adam over the kf params)
            cam = keyframes[v]["cam"]
            if self._moves(v):
                groups += [dict(params=[cam.cam_rot_delta], lr=pose_lr[0]),
                           dict(params=[cam.cam_trans_delta], lr=pose_lr[1])]
            if keyframes[v].get("optimize_exposure", True):
                groups += [dict(params=[cam.exposure_a], lr=pose_lr[2]),
                           dict(params=[cam.exposure_b], lr=pose_lr[3])]
        kw = dict(capturable=True) if (self.cuda_graph and adam_cls is FusedAdam) else {}
        if kf_optimizer is not None:
            self.kf_optimizer, groups = kf_optimizer, []
        else:
            self.kf_optimizer = adam_cls(groups, lr=0.0, eps=1e-8, **kw) if groups else None
        # the window's camera / exposure gradients live in ONE buffer: zeroing them is one fill
        # instead of four per keyframe
        self._kf_grad = None
        kf_params = [p for gr in groups for p in gr["params"]]
        if kf_params and all(p.is_cuda for p in kf_params):
            self._kf_grad = torch.zeros(sum(p.numel() for p in kf_params), device=kf_params[0].device)
            o = 0
            for p in kf_params:
                p.grad = self._kf_grad[o:o + p.numel()].view_as(p)
                o += p.numel()
        if self.cuda_graph:
            if not isinstance(gaussians.optimizer, FusedAdam):
                raise RuntimeError("cuda_graph=True needs the FusedAdam optimizer")
            gaussians.optimizer.capturable = True
        # view_streams > 1: the local views are issued round-robin on that many side streams
        # (forked from / joined into the calling stream, also inside the captured graph), so one
        # view's single-CTA tile scan, kernel tails and loss reductions overlap another view's
        # render kernels.  Gradients meet in the arena through atomics either way.
        if view_streams is None:
            view_streams = auto_view_streams(gaussians, len(keyframes))
        self.view_streams = max(1, int(view_streams))
        self._side = self._reg_stream = None
        # peer_adam: gradient exchange + optimiser step as one kernel over NVLink peer memory
        # (GaussianModel.enable_peer_arena) instead of NCCL all-reduce + replicated Adam
        if self.world > 1 and hasattr(gaussians, "sync_densification_stats"):
            gaussians._dist_group = group if group is not None else dist.group.WORLD
        self.peer = None
        if peer_adam and self.world > 1 and self.fused_reg:
            have = getattr(gaussians, "peer", None)
            if have is not None and have.world == self.world:
                self.peer = have                 # an earlier window of this map already moved the arenas
                gaussians.optimizer.capturable = True
            else:
                from .peer import PeerArena
                # peer_adam="nomc": plain peer loads / stores even where NVLS multicast is available
                self.peer = PeerArena(gaussians.flat.device, group, multicast=peer_adam != "nomc")
                gaussians.enable_peer_arena(self.peer)
        if use_lean:
            from .window import LeanWindow
            self._lean = LeanWindow(gaussians, keyframes, self.local, bg, pose_lr, self._moves,
                                    lambda v: bool(self.kfs[v].get("optimize_exposure", True)),
                                    densify_stats=self.densify_stats, regulariser=self.regulariser,
                                    rank=self.rank, world=self.world)
            if self.peer is not None:
                self.peer.set_fail_word(self._lean.fail)
                self._lean.defer_finalize = bool(self.local)
        self.last = {}
        self._g1 = self._g2 = None
        self._ovf = self._status_all = None
        self._pending, self._launched = [], 0
        self.graph_launches = 0
        self._loss = None
        self._steps = 0

    def _moves(self, v):
        return bool(self.optimize_pose and self.kfs[v].get("optimize_pose", v > 0))

    def _zero_kf_grads(self, set_to_none):
        if self._kf_grad is not None:
            self._kf_grad.zero_()
        elif self.kf_optimizer is not None:
            self.kf_optimizer.zero_grad(set_to_none=set_to_none)

    # -- the two halves of a step ------------------------------------------------------
    def _forward_backward(self):
        if self._lean is not None:
            side = []
            if self.view_streams > 1 and len(self.local) > 1:
                if self._side is None:
                    self._side = [torch.cuda.Stream(self.bg.device) for _ in range(self.view_streams)]
                side = self._side
            loss = self._lean.forward_backward(side)
            return loss, {v: self._lean.package(v) for v in self.local}
        g = self.g
        loss = 0
        stats = {}
        side = []
        cuda = self.bg.is_cuda
        main = torch.cuda.current_stream() if cuda else None
        if self.view_streams > 1 and len(self.local) > 1 and cuda:
            if self._side is None:
                self._side = [torch.cuda.Stream(self.bg.device) for _ in range(self.view_streams)]
            side = self._side
            for s in side:
                s.wait_stream(main)
        reg = None
        reg_stream = None
        if self.rank == 0 and self.fused_reg and self.regulariser:
            # view-independent regulariser, counted once (slam_backend.py:229-231): one kernel,
            # value + gradient added (atomically) to the scaling grad arena — on its own stream, next
            # to the views instead of after them
            if self._reg_stream is None:
                self._reg_stream = torch.cuda.Stream(self.bg.device)
            reg_stream = self._reg_stream
            reg_stream.wait_stream(main)
            with torch.cuda.stream(reg_stream):
                reg = scale_regulariser(g._scaling, g._scaling.grad)
        terms = []
        # prune-mode iterations leave the densification statistics alone (utils/slam_backend.py:241-275)
        rkw = dict(with_stats=False) if (self.render_fn is render_fused and not self.densify_stats) else {}
        for i, v in enumerate(self.local):
            kf = self.kfs[v]
            with (torch.cuda.stream(side[i % len(side)]) if side else contextlib.nullcontext()):
                pkg = self.render_fn(kf["cam"], g, None, self.bg, **rkw)
                terms.append(self.loss_fn(pkg["render"], pkg["depth"], kf["cam"], kf["gt_image"],
                                          kf["gt_depth"]))
            stats[v] = pkg
        for s in side:
            main.wait_stream(s)
        for t in terms:
            loss = loss + t
        if self.rank == 0 and not self.fused_reg and self.regulariser:
            scaling = g.get_scaling
            loss = loss + 10 * torch.abs(scaling - scaling.mean(dim=1).view(-1, 1)).mean()
        if torch.is_tensor(loss):
            loss.backward()      # every node runs on the stream of its forward
            for s in side:
                main.wait_stream(s)
            loss = loss.detach()
        else:
            loss = torch.zeros((), device=self.bg.device)
        if reg is not None:
            main.wait_stream(reg_stream)
            loss = loss + reg
        # densification statistics of every view (utils/slam_backend.py:277-284); the fused
        # renderer has already accumulated them inside its backward kernel
        if self.densify_stats and hasattr(g, "add_densification_stats"):
            with torch.no_grad():
                for pkg in stats.values():
                    if pkg.get("stats_fused"):
                        continue
                    vis = pkg["visibility_filter"]
                    g.max_radii2D[vis] = torch.max(g.max_radii2D[vis], pkg["radii"][vis])
                    g.add_densification_stats(pkg["viewspace_points"], vis)
        return loss, stats

    def _update(self, set_to_none, gaussians=True):
        g = self.g
        if self._lean is not None:
            return self._lean.update(peer=self.peer, world=self.world, group=self.group, gaussians=gaussians)
        if not gaussians:
            # the reference's optimizer.step() right after densify_and_prune finds only fresh
            # Parameters without .grad and changes nothing (utils/slam_backend.py:290-309)
            pass
        elif self.peer is not None:
            g.optimizer.step()            # all-reduce + Adam + zero_grad: one kernel over peer memory
        elif self.world > 1:
            if hasattr(g, "flat_grad"):   # ONE message for all six parameter groups
                dist.all_reduce(g.flat_grad, op=dist.ReduceOp.SUM, group=self.group)
            else:                         # stock layout: one message per parameter tensor
                for p in g.parameters():
                    if p.grad is None:
                        p.grad = torch.zeros_like(p)
                    dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=self.group)
        if self.peer is None and gaussians:
            g.optimizer.step()
            g.zero_grad()
        if self.kf_optimizer is not None:
            self.kf_optimizer.step()
            self._zero_kf_grads(set_to_none)
            if self.optimize_pose:
                for v in self.local:
                    if self._moves(v):
                        cam = self.kfs[v]["cam"]
                        try:
                            cam.update_pose(check_converged=False)
                        except TypeError:
                            cam.update_pose()

    def _states(self, stats):
        out = []
        for pkg in stats.values():
            fn = pkg["render"].grad_fn
            out.append(getattr(fn, "state", None) if fn is not None else None)
        return out

    # -- eager -------------------------------------------------------------------------
    def _eager_forward_backward(self, set_to_none=True):
        from . import rasterizer as rz
        if self._lean is not None:
            for attempt in range(3):
                loss, stats = self._forward_backward()
                vals = self._lean.status_all.tolist()      # one host sync per step
                if not self._lean.note_overflow(vals):
                    break
                self.g.zero_grad()              # discard the background-only pass, redo it with more room
                self._lean.begin_rerun(vals)    # the views that fitted are not counted a second time
                if attempt == 2:
                    raise RuntimeError("binning capacity overflow persisted after two retries")
            self._lean.end_rerun()
            self.last = stats
            return loss
        loss, stats = self._forward_backward()
        if rz._SyncFree.enabled and rz.check_overflow(self._states(stats)):
            self.g.zero_grad()              # discard the background-only pass and redo it
            self._zero_kf_grads(set_to_none)
            loss, stats = self._forward_backward()
            if rz.check_overflow(self._states(stats)):
                raise RuntimeError("binning capacity overflow persisted after a retry")
        self.last = stats
        return loss

    def _eager_step(self, set_to_none=True):
        loss = self._eager_forward_backward(set_to_none)
        self._update(set_to_none)
        return loss

    # the two halves on their own, for a caller that decides between them (backend.py: prune
    # mode returns after the backward; densify / opacity resets sit between backward and step)
    def forward_backward(self):
        self.flush()
        return self._eager_forward_backward(set_to_none=not self.cuda_graph)

    def update(self, gaussians=True):
        self._update(not self.cuda_graph, gaussians=gaussians)

    # -- CUDA graphs -------------------------------------------------------------------
    def _capture(self):
        from . import rasterizer as rz
        rz.set_sync_free(True, rz._SyncFree.headroom)
        from ._lib import lib
        torch.cuda.synchronize()
        dev = self.bg.device
        if self._ovf is None:
            self._ovf = torch.zeros(1, dtype=torch.int32, device=dev)
        if self._lean is not None:
            # the whole step as ONE graph: views -> finalize -> [exchange] -> Adam -> camera step
            self._lean.ensure()
            self._g1, self._g2 = torch.cuda.CUDAGraph(), None
            with torch.cuda.graph(self._g1):
                self._loss, self._stats = self._forward_backward()
                self._update(set_to_none=False)
            self._status_all, self._ovf, self._gstates = self._lean.status_all, self._lean.guard, None
            self.graph_launches = self._lean.launches_fb + self._lean.launches_up
            self._flag_host = [torch.zeros(self._lean.flags.shape, dtype=torch.int32).pin_memory() for _ in range(3)]
            return
        l0 = lib.b2gs_kernel_launches()
        self._g1 = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._g1):
            self._loss, self._stats = self._forward_backward()
            # did any local view run out of binning capacity?  -> the device word that guards G2
            sts = [s for s in self._states(self._stats) if s is not None and s.status is not None]
            if sts:
                self._status_all = torch.stack([s.status for s in sts])
                self._ovf.copy_(self._status_all[:, 1].max().reshape(1))
            else:
                self._status_all = None
        self._gstates = sts
        l1 = lib.b2gs_kernel_launches()
        if self._g2 is None:
            self._g2 = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._g2):
                if self.world > 1 and self.peer is None:     # the ranks must agree (peer mode: the
                    dist.all_reduce(self._ovf, op=dist.ReduceOp.MAX, group=self.group)   # barrier ORs)
                lib.b2gs_set_update_guard(self._ovf.data_ptr())
                try:
                    self._update(set_to_none=False)
                finally:
                    lib.b2gs_set_update_guard(None)
            self._l2 = lib.b2gs_kernel_launches() - l1
        self.graph_launches = (l1 - l0) + self._l2      # libb2gs kernels per replayed step
        if self._status_all is not None:
            self._flag_host = [(torch.empty(self._status_all.shape, dtype=torch.int32).pin_memory(),
                                torch.empty(1, dtype=torch.int32).pin_memory(),
                                torch.zeros(1, dtype=torch.int32).pin_memory()) for _ in range(3)]

    def _launch(self):
        """one replay of the captured step + an asynchronous read-back of its overflow flags"""
        for opt in (self.g.optimizer, self.kf_optimizer):
            if opt is not None:
                opt.sync_lr()
        self._g1.replay()
        if self._g2 is not None:
            self._g2.replay()
        if self._lean is not None:
            # ONE read-back per step: per-view (instances, overflow), the guard (after the exchange: the OR
            # over all ranks), the peer barrier's time-out word
            h = self._flag_host[self._launched % 3]
            h.copy_(self._lean.flags, non_blocking=True)
            n = self._status_all.shape[0]
            ev = torch.cuda.Event()
            ev.record()
            self._pending.append((ev, h[:2 * n].view(n, 2), h[2 * n:2 * n + 1], h[2 * n + 1:2 * n + 2]))
        elif self._status_all is not None:
            st, ov, fl = self._flag_host[self._launched % 3]
            st.copy_(self._status_all, non_blocking=True)      # this rank's views: sizes the retry
            ov.copy_(self._ovf, non_blocking=True)             # after G2: the OR over all ranks
            if self.peer is not None:
                fl.copy_(self.peer.fail, non_blocking=True)    # did a peer barrier time out?
            ev = torch.cuda.Event()
            ev.record()
            self._pending.append((ev, st, ov, fl))
        self._launched += 1

    def _settle(self, keep):
        """look at the flags of all but the `keep` newest launches; on overflow: drain, enlarge the
        capacities, re-capture G1 and run the discarded steps again (their updates were no-ops: the
        guard word made every Adam / pose kernel return, on every rank — the word is the OR over
        all ranks, so all ranks take this path together)."""
        from . import rasterizer as rz
        over, first = False, None
        while len(self._pending) > keep:
            ev, st, ov, fl = self._pending.pop(0)
            ev.synchronize()
            if int(fl[0]):
                raise RuntimeError("libb2gs peer barrier timed out: a rank died or fell out of step; the "
                                   "replicas were left untouched by that step but the window cannot continue")
            self._note(st.tolist())
            if int(ov[0]) and first is None:
                first = st.tolist()             # which local views fitted in the first discarded step
            over = over or bool(int(ov[0]))
        if not over:
            return
        torch.cuda.synchronize()
        lost = 1 + len(self._pending)           # every launch since the overflowing one was discarded
        for ev, st, ov, fl in self._pending:
            self._note(st.tolist())
        self._pending.clear()
        for k in range(lost):
            self.g.zero_grad()
            self._zero_kf_grads(False)
            self._g1 = None
            if self._lean is not None:
                # densification statistics: the first discarded step counted the views that fitted
                # (the replays after it counted nothing: B2gsBackwardIO.stats_gate), so the first
                # repeat, still at the same parameters, counts only the views that did not
                if k == 0:
                    self._lean.begin_rerun(first)
                else:
                    self._lean.end_rerun()
            self._capture()                     # larger capacities are now in rz._SyncFree.cap
            self._launch()
            torch.cuda.synchronize()
            ev, st, ov, fl = self._pending.pop(0)
            self._note(st.tolist())
            if int(ov[0]):
                raise RuntimeError("binning capacity overflow persisted after re-capture")
        if self._lean is not None:
            self._lean.end_rerun()

    def _note(self, vals):
        from . import rasterizer as rz
        if self._lean is not None:
            return self._lean.note_overflow(vals)
        return rz.note_overflow(self._gstates, vals)

    def flush(self):
        """settle every launched step (one host sync): afterwards all of them have been applied."""
        if self.cuda_graph and self._g1 is not None:
            self._settle(0)

    def _graph_step(self):
        from . import rasterizer as rz
        if self._steps < 3:                 # warm-up: exact sizing, optimizer state, .grad buffers
            rz.set_sync_free(True, rz._SyncFree.headroom)
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                loss = self._eager_step(set_to_none=False)
            torch.cuda.current_stream().wait_stream(s)
            return loss
        if self._g1 is None or (self._lean is not None and self._lean.ensure()):
            self._capture()
        # no host sync inside the step: the flags of the step launched two calls ago are looked at
        # now (its event completed long ago, the GPU is busy with the previous step meanwhile)
        self._settle(1)
        self._launch()
        self.last = self._stats
        return self._loss

    def step(self):
        loss = self._graph_step() if self.cuda_graph else self._eager_step()
        self._steps += 1
        return loss

    def release_graphs(self):
        """drop the captured graphs (they pin communicator resources in multi-rank runs)."""
        self._g1 = self._g2 = None

    # -- measurement helpers (bench.py) ------------------------------------------------------
    def profile_steps(self, n):
        """`n` extra EAGER steps with the views serialised on one stream, for per-kernel CUDA-event
        timing (b2gs_profile_enable): a kernel timed while another view's kernels share the SMs
        would not be that kernel's duration."""
        self.flush()
        vs, self.view_streams = self.view_streams, 1
        try:
            for _ in range(n):
                self._eager_step(set_to_none=False)
        finally:
            self.view_streams = vs

    def work_counters(self, v):
        """the work one fwd+bwd of keyframe `v` executes, read from the run itself: visible
        Gaussians, the true instance count, sum of n_contrib, and the render kernels' debug
        counters (b2gs_debug_set_counters).  Leaves parameters and gradients untouched."""
        import ctypes as C
        from . import _lib
        self.flush()
        kf, g, dev = self.kfs[v], self.g, self.bg.device
        cnt = torch.zeros(8, dtype=torch.int64, device=dev)
        _lib.lib.b2gs_debug_set_counters(cnt.data_ptr())
        try:
            pkg = self.render_fn(kf["cam"], g, None, self.bg, **(
                dict(with_stats=False) if self.render_fn is render_fused else {}))
            self.loss_fn(pkg["render"], pkg["depth"], kf["cam"], kf["gt_image"], kf["gt_depth"]).backward()
            torch.cuda.synchronize()
        finally:
            _lib.lib.b2gs_debug_set_counters(None)
        g.zero_grad()
        self._zero_kf_grads(False)
        st = pkg["render"].grad_fn.state
        inst = int(st.status[0].item()) if st.status is not None else int(st.num_rendered)
        sc = _lib.B2gsScene()
        sc.P, sc.W, sc.H = 1, int(kf["cam"].image_width), int(kf["cam"].image_height)
        ws = _lib.B2gsWorkspace(None, 0, _lib.ptr(st.image), st.image.numel(), None, 0)
        nc = torch.zeros(sc.W * sc.H, dtype=torch.int32, device=dev)
        _lib.check(_lib.lib.b2gs_debug_export_image(C.byref(sc), C.byref(ws), None, nc.data_ptr(),
                                                    torch.cuda.current_stream().cuda_stream))
        c = cnt.tolist()
        return dict(visible=int((pkg["radii"] > 0).sum().item()), instances=inst,
                    effective_pairs=int(nc.long().sum().item()),
                    fwd_visit_px=64 * c[0], fwd_eval_px=64 * c[1], fwd_contributing=c[2], fwd_block_px=32 * c[3],
                    bwd_visit_px=64 * c[4], bwd_blend_px=64 * c[5], bwd_contributing=c[6])


def render_pose_only(viewpoint_camera, pc, pipe=None, bg_color=None, scaling_modifier=1.0):
    """render() for tracking (BASELINE config 2): the map is frozen, only the camera deltas
    (theta / rho -> dL_dtau) receive gradients, so the backward skips every per-Gaussian
    gradient output it can (SURVEY.md §7.2 'pose-only mode')."""
    cam = viewpoint_camera
    rs = GaussianRasterizationSettings(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, projmatrix_raw=cam.projection_matrix,
        sh_degree=pc.active_sh_degree, campos=cam.camera_center, prefiltered=False, debug=False)
    d = lambda t: t.detach()
    image, radii, depth, opacity, n_touched = rasterize_gaussians_raw(
        d(pc._xyz), d(pc._features_dc), d(pc._features_rest), d(pc._opacity), d(pc._scaling),
        d(pc._rotation), rs, theta=cam.cam_rot_delta, rho=cam.cam_trans_delta, sinks=None)
    return {"render": image, "viewspace_points": None, "visibility_filter": radii > 0, "radii": radii,
            "depth": depth, "opacity": opacity, "n_touched": n_touched}


class TrackingLoop:
    """FrontEnd.tracking() (utils/slam_frontend.py:128-196): `iters` iterations of
    render -> tracking loss -> backward -> Adam over (cam_rot_delta, cam_trans_delta, exposure_a,
    exposure_b) -> update_pose, on one GPU (tracking does not shard).

    The loop owns ONE persistent camera and ground-truth buffers; `track()` loads a frame into
    them.  With cuda_graph=True one iteration is captured once and replayed `iters` times per
    frame with no host synchronisation inside the loop (the reference syncs every iteration for
    its convergence test); binning overflow is checked once per frame and the frame re-run."""

    def __init__(self, gaussians, camera, bg, lr_rot=0.003, lr_trans=0.001, lr_exposure=0.01, iters=100,
                 cuda_graph=False, render_fn=None, loss_fn=None, adam_cls=FusedAdam, lean=None, grad_mask=None):
        """lean (default: whenever the loop is the stock one — packed-arena GaussianModel on the GPU,
        camera.Camera, default renderer / RGB-D tracking loss, FusedAdam): an iteration is five libb2gs
        launches sequences with no autograd graph and no torch kernel — b2gs_forward with the tracking
        loss fused into the render epilogue, a pose-only b2gs_backward (no per-Gaussian output),
        b2gs_window_finalize, b2gs_pose_step (window.LeanWindow(pose_only=True)).
        grad_mask: optional [1,H,W] viewpoint.grad_mask buffer (utils/slam_utils.py:70)."""
        self.g, self.cam, self.bg, self.iters = gaussians, camera, bg, iters
        stock = (isinstance(gaussians, GaussianModel) and gaussians.flat.is_cuda and render_fn is None
                 and loss_fn is None and adam_cls is FusedAdam and hasattr(camera, "state"))
        if lean and not stock:
            raise RuntimeError("lean=True needs the stock tracking loop")
        self._lean = None
        use_lean = stock if lean is None else bool(lean)
        self.render_fn = render_fn or render_pose_only
        self.loss_fn = loss_fn or get_loss_tracking_rgbd
        self.cuda_graph = bool(cuda_graph)
        dev, H, W = bg.device, camera.image_height, camera.image_width
        self.gt_image = torch.zeros(3, H, W, device=dev)
        self.gt_depth = torch.zeros(1, H, W, device=dev)
        groups = [dict(params=[camera.cam_rot_delta], lr=lr_rot), dict(params=[camera.cam_trans_delta], lr=lr_trans),
                  dict(params=[camera.exposure_a], lr=lr_exposure), dict(params=[camera.exposure_b], lr=lr_exposure)]
        kw = dict(capturable=True) if (self.cuda_graph and adam_cls is FusedAdam) else {}
        self.opt = None if use_lean else adam_cls(groups, lr=0.0, eps=1e-8, **kw)
        if use_lean:
            from .window import LeanWindow
            from .losses import TRACKING_RGBD
            self._kf = dict(cam=camera, gt_image=self.gt_image, gt_depth=self.gt_depth, grad_mask=grad_mask)
            self._lean = LeanWindow(gaussians, [self._kf], [0], bg, (lr_rot, lr_trans, lr_exposure, lr_exposure),
                                    lambda v: True, lambda v: True, mode=TRACKING_RGBD, pose_only=True)
        self._graph = None
        self.graph_launches = 0
        self.pkg = None
        self._ever, self._key = None, None

    def _iteration(self, check_converged=False):
        cam = self.cam
        pkg = self.render_fn(cam, self.g, None, self.bg)
        st = getattr(pkg["render"].grad_fn, "state", None)
        if st is not None and st.status is not None:   # sticky (max instances, overflow flag) of the frame
            if self._ever is None:
                self._ever = torch.zeros_like(st.status)
            self._ever.copy_(torch.maximum(self._ever, st.status))
            self._key = st.key
        loss = self.loss_fn(pkg["render"], pkg["depth"], pkg["opacity"], cam, self.gt_image, self.gt_depth)
        loss.backward()
        self.opt.step()
        self.opt.zero_grad(set_to_none=False)
        try:
            conv = cam.update_pose(check_converged=check_converged)
        except TypeError:
            conv = cam.update_pose()
        return pkg, loss.detach(), conv

    def _reset_optimizer(self):
        if self._lean is not None:
            return self._lean.reset_camera_optimizer()
        for st in self.opt.state.values():
            if "exp_avg" in st:
                st["exp_avg"].zero_()
                st["exp_avg_sq"].zero_()
                st["step"] = st["step"].zero_() if torch.is_tensor(st["step"]) else 0
        if getattr(self.opt, "_step_dev", None) is not None:
            self.opt._step_dev.zero_()

    def load_frame(self, R, T, gt_image, gt_depth, exposure=(0.0, 0.0)):
        """viewpoint.update_RT(prev.R, prev.T) + a fresh pose optimizer (slam_frontend.py:129-161)."""
        self.cam.update_RT(R, T)
        self.gt_image.copy_(gt_image, non_blocking=True)
        self.gt_depth.copy_(gt_depth, non_blocking=True)
        with torch.no_grad():
            self.cam.cam_rot_delta.zero_(); self.cam.cam_trans_delta.zero_()
            self.cam.exposure_a.fill_(exposure[0]); self.cam.exposure_b.fill_(exposure[1])
        self._reset_optimizer()

    def _track_lean(self, iters, check_converged):
        L = self._lean
        L.ensure()
        done = 0
        if not self.cuda_graph:
            while done < iters:
                self.loss = L.forward_backward()
                L.update(gaussians=False)
                done += 1
                if check_converged and float(L.tau_norm[0]) < 1e-4:      # the reference's per-iteration sync
                    break
        else:
            if self._graph is None or L.ensure():
                torch.cuda.synchronize()
                self._graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(self._graph):
                    self.loss = L.forward_backward()
                    L.update(gaussians=False)
                self.graph_launches = L.launches_fb + L.launches_up
            for _ in range(iters):
                self._graph.replay()
        # one host sync per frame: did any iteration run out of binning capacity?  (an overflowed
        # iteration rendered background only and skipped its own pose step on the device)
        if int(L.tickets[2].item()):
            L.tickets[2:3].zero_()
            L.note_overflow([[int(L.status_all[0, 0].item()) * 2, 1]])
            self._graph = None
            self.cam.state.copy_(self._snap[0]); self.cam._host_RT = None
            with torch.no_grad():
                for p, q in zip((self.cam.cam_rot_delta, self.cam.cam_trans_delta, self.cam.exposure_a,
                                 self.cam.exposure_b), self._snap[1]):
                    p.copy_(q)
            self._reset_optimizer()
            return self._track_lean(iters, check_converged)
        self.pkg = L.package(0)
        self.pkg["visibility_filter"] = self.pkg["radii"] > 0      # once per frame, outside the iteration loop
        return self.pkg

    def track(self, iters=None, check_converged=False):
        """runs the loop on the loaded frame; returns the last render package."""
        from . import rasterizer as rz
        from ._lib import lib
        iters = self.iters if iters is None else iters
        if self._lean is not None:
            self._snap = (self.cam.state.clone(), [p.detach().clone() for p in (
                self.cam.cam_rot_delta, self.cam.cam_trans_delta, self.cam.exposure_a, self.cam.exposure_b)])
            return self._track_lean(iters, check_converged)
        if not self.cuda_graph:
            for _ in range(iters):
                self.pkg, self.loss, conv = self._iteration(check_converged)
                if check_converged and conv:
                    break
            return self.pkg
        rz.set_sync_free(True, rz._SyncFree.headroom)
        # the frame as loaded: an overflow anywhere in the loop restores THIS state (pose, exposure,
        # fresh optimizer) and runs all iterations again — not the state after the warm-up iterations
        # with zeroed moments (advisor finding)
        snap = self.cam.state.clone(), [p.detach().clone() for g in self.opt.param_groups for p in g["params"]]
        done = 0
        if self._graph is None:
            s = torch.cuda.Stream()
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                for _ in range(min(3, iters)):      # exact sizing, optimizer state, .grad buffers
                    self.pkg, self.loss, _ = self._iteration()
                    done += 1
            torch.cuda.current_stream().wait_stream(s)
            torch.cuda.synchronize()
            l0 = lib.b2gs_kernel_launches()
            self._graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self._graph):
                self.pkg, self.loss, _ = self._iteration()
            self.graph_launches = lib.b2gs_kernel_launches() - l0
        for _ in range(iters - done):
            self._graph.replay()
        from types import SimpleNamespace
        worst = SimpleNamespace(status=self._ever, key=self._key) if self._ever is not None else None
        over = rz.check_overflow([worst])           # one host sync per frame
        if self._ever is not None:
            self._ever.zero_()
        if over:
            # capacity was exceeded somewhere in the loop: restore the frame, re-capture, re-run
            self.cam.state.copy_(snap[0]); self.cam._host_RT = None
            with torch.no_grad():
                for p, q in zip([p for g in self.opt.param_groups for p in g["params"]], snap[1]):
                    p.copy_(q)
            self._reset_optimizer()
            self._graph = None
            return self.track(iters, check_converged)
        return self.pkg
