"""Keyframe seeding on the device: RGB-D back-projection -> random down-sample -> distCUDA2 scales
-> new rows of the packed arena (the caller of distCUDA2, §8a a16).

Mirrors, with the reference's names and argument meaning,
  GaussianModel.create_pcd_from_image            gaussian_splatting/scene/gaussian_model.py:107-131
  GaussianModel.create_pcd_from_image_and_depth  :133-193
  GaussianModel.extend_from_pcd / _seq           :198-227
The reference quantises the colour on the GPU, copies image and depth to the HOST, back-projects
and down-samples with open3d, and copies the cloud back; here nothing leaves the device
(csrc/seed.cu: b2gs_backproject_rgbd) and the only host sync is the size of the sampled cloud.
"""
import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import check, lib

C0 = 0.28209479177387814                       # gaussian_splatting/utils/sh_utils.py:24
DATASET_DEFAULTS = dict(pcd_downsample=64, pcd_downsample_init=32, point_size=0.05,
                        adaptive_pointsize=True, sensor_type="depth")   # configs/rgbd/replica/base_config.yaml:12-15


def _w2c(cam):
    """getWorld2View2(cam.R, cam.T) (gaussian_splatting/utils/graphics_utils.py:38-49, no recentring)."""
    m = np.eye(4, dtype=np.float64)
    m[:3, :3] = cam.R.detach().cpu().double().numpy()
    m[:3, 3] = cam.T.detach().cpu().double().numpy()
    return m


def backproject_rgbd(image, depth, fx, fy, cx, cy, w2c, exposure_a=None, exposure_b=None, depth_trunc=100.0):
    """dense (xyz [N,3], rgb [N,3], valid [N] uint8) of one frame, row-major pixel order."""
    if not image.is_cuda:
        raise RuntimeError("backproject_rgbd needs CUDA tensors: libb2gs has no CPU path")
    image = image.detach().float().contiguous()
    depth = depth.detach().float().contiguous().view(image.shape[-2], image.shape[-1])
    H, W = depth.shape
    dev = image.device
    xyz = torch.empty(H * W, 3, device=dev)
    rgb = torch.empty(H * W, 3, device=dev)
    valid = torch.empty(H * W, dtype=torch.uint8, device=dev)
    c2w = np.linalg.inv(np.asarray(w2c, dtype=np.float64))
    m = (C.c_double * 12)(*[float(x) for x in c2w[:3].reshape(-1)])
    ea = exposure_a.detach().float().contiguous() if exposure_a is not None else None
    eb = exposure_b.detach().float().contiguous() if exposure_b is not None else None
    with torch.cuda.device(dev):
        check(lib.b2gs_backproject_rgbd(W, H, image.data_ptr(), _lib.ptr(ea), _lib.ptr(eb), depth.data_ptr(),
                                        float(fx), float(fy), float(cx), float(cy), m, float(depth_trunc),
                                        xyz.data_ptr(), rgb.data_ptr(), valid.data_ptr(),
                                        torch.cuda.current_stream().cuda_stream))
    return xyz, rgb, valid


def create_pcd_from_image(model, cam_info, init=False, scale=2.0, depthmap=None, generator=None):
    """-> (fused_point_cloud [n,3], features [n,3,(deg+1)^2], scales [n,3|1], rots [n,4], opacities [n,1])"""
    cfg = dict(DATASET_DEFAULTS)
    cfg.update((getattr(model, "config", None) or {}).get("Dataset", {}))
    cam = cam_info
    image = cam.original_image
    dev = image.device
    H, W = image.shape[-2:]
    if depthmap is not None:
        depth = torch.as_tensor(depthmap, dtype=torch.float32).to(dev)
    else:
        depth_raw = getattr(cam, "depth", None)
        if cfg["sensor_type"] == "monocular" or depth_raw is None:
            # gaussian_model.py:121-126: (1 + (randn - 0.5) * 0.05) * scale
            noise = torch.randn(H, W, generator=generator, device=dev if generator is None or
                                generator.device.type == "cuda" else "cpu").to(dev)
            depth = (torch.ones(H, W, device=dev) + (noise - 0.5) * 0.05) * scale
        else:
            depth = torch.as_tensor(depth_raw, dtype=torch.float32).to(dev)
    return create_pcd_from_image_and_depth(model, cam, image, depth, init=init, cfg=cfg, generator=generator)


def create_pcd_from_image_and_depth(model, cam, image, depth, init=False, cfg=None, generator=None):
    cfg = cfg or dict(DATASET_DEFAULTS)
    downsample_factor = cfg["pcd_downsample_init"] if init else cfg["pcd_downsample"]
    point_size = cfg["point_size"]
    if cfg.get("adaptive_pointsize"):
        point_size = min(0.05, point_size * float(torch.median(depth)))      # np.median(depth), :140-142
    xyz, rgb, valid = backproject_rgbd(image, depth, cam.fx, cam.fy, cam.cx, cam.cy, _w2c(cam),
                                       getattr(cam, "exposure_a", None), getattr(cam, "exposure_b", None))
    idx = torch.nonzero(valid, as_tuple=False).squeeze(1)           # open3d's cloud: valid pixels, row-major
    # PointCloud.random_down_sample(1 / factor): a uniformly random subset of int(ratio * n) points
    k = int((1.0 / downsample_factor) * idx.numel())
    perm = torch.randperm(idx.numel(), generator=generator,
                          device=idx.device if generator is None or generator.device.type == "cuda" else "cpu")
    idx = idx[perm[:k].to(idx.device)]
    new_xyz, new_rgb = xyz[idx].contiguous(), rgb[idx]
    return (new_xyz,) + seed_attributes(model, new_xyz, new_rgb, point_size)


def seed_attributes(model, new_xyz, new_rgb, point_size):
    """gaussian_model.py:161-193: SH DC from the colour, log-scale from the mean squared distance to
    the three nearest neighbours, identity rotation, opacity 0.5."""
    from .knn import distCUDA2
    n, dev = new_xyz.shape[0], new_xyz.device
    features = torch.zeros(n, 3, (model.max_sh_degree + 1) ** 2, device=dev)
    features[:, :3, 0] = (new_rgb - 0.5) / C0                       # RGB2SH
    dist2 = torch.clamp_min(distCUDA2(new_xyz), 0.0000001) * point_size
    scales = torch.log(torch.sqrt(dist2))[..., None]
    if not getattr(model, "isotropic", False):
        scales = scales.repeat(1, 3)
    rots = torch.zeros(n, 4, device=dev)
    rots[:, 0] = 1
    opacities = torch.log(0.5 * torch.ones(n, 1, device=dev) / (1 - 0.5 * torch.ones(n, 1, device=dev)))
    return features, scales, rots, opacities


@torch.no_grad()
def extend_from_pcd(model, fused_point_cloud, features, scales, rots, opacities, kf_id):
    """append the seeded Gaussians (new rows get zero Adam moments: cat_tensors_to_optimizer,
    gaussian_model.py:534-555); unique_kfIDs / n_obs ride along on the device."""
    from .densify import _state
    from .mapping import _GROUPS
    if hasattr(model, "gather_moments"):
        model.gather_moments()
    par, mom, steps = _state(model)
    old_p = par["xyz"].shape[0]
    new = dict(xyz=fused_point_cloud, f_dc=features[:, :, 0:1].transpose(1, 2).contiguous(),
               f_rest=features[:, :, 1:].transpose(1, 2).contiguous(), opacity=opacities,
               scaling=scales, rotation=rots)
    if new["scaling"].shape[1] != par["scaling"].shape[1]:
        new["scaling"] = new["scaling"].expand(-1, par["scaling"].shape[1]).contiguous()
    stats = [getattr(model, a) for a in ("xyz_gradient_accum", "denom", "max_radii2D")] \
        if hasattr(model, "xyz_gradient_accum") else None
    kf_ids = getattr(model, "unique_kfIDs", None)
    n_obs = getattr(model, "n_obs", None)
    dev = par["xyz"].device
    n = fused_point_cloud.shape[0]
    model._allocate({k: torch.cat([par[k], new[k].to(dev).reshape((n,) + tuple(par[k].shape[1:]))]) for k in _GROUPS},
                    {k: tuple(torch.cat([m, torch.zeros((n,) + tuple(m.shape[1:]), device=dev)])
                              for m in mom[k]) for k in _GROUPS}, steps)
    # densification_postfix (gaussian_model.py:583-585) resets the statistics of the whole model
    if stats is not None:
        model._stats()
    z = lambda t, fill: torch.cat([t.to(dev) if t is not None else torch.zeros(old_p, dtype=torch.int32, device=dev),
                                   torch.full((n,), fill, dtype=torch.int32, device=dev)])
    model.unique_kfIDs = z(kf_ids, int(kf_id))
    model.n_obs = z(n_obs, 0)
    return n


def extend_from_pcd_seq(model, cam_info, kf_id=-1, init=False, scale=2.0, depthmap=None, generator=None):
    """gaussian_model.py:222-227"""
    pcd = create_pcd_from_image(model, cam_info, init, scale=scale, depthmap=depthmap, generator=generator)
    return extend_from_pcd(model, *pcd, kf_id)
