"""TEST INFRASTRUCTURE — CPU restatement of the keyframe seeding path.
Only tests/ may import this; the product path is csrc/seed.cu + monogs_dev_b200/seeding.py.

Follows GaussianModel.create_pcd_from_image / create_pcd_from_image_and_depth
(gaussian_splatting/scene/gaussian_model.py:107-207 of the reference).  The back-projection
itself lives in open3d (pinned 0.17.0 in the reference's environment.yml), a third-party
dependency that is neither vendored under /root/reference nor installed here, so its published
algorithm is restated (cpp/open3d/geometry/RGBDImageFactory.cpp: CreateFromColorAndDepth,
ImageFactory / Image.cpp: ConvertDepthToFloatImage, PointCloudFactory.cpp:
CreatePointCloudFromRGBDImageT).  PARITY UNPINNED for this row: the reference holds no test or
fixture for it and open3d cannot be run here; the restatement is anchored on the reference's call
site (depth_scale=1.0, depth_trunc=100.0, convert_rgb_to_intensity=False,
project_valid_depth_only=True) and the documented semantics.
"""
import numpy as np

C0 = 0.28209479177387814           # gaussian_splatting/utils/sh_utils.py:24


def quantise_colour(original_image, exposure_a=0.0, exposure_b=0.0):
    """gaussian_model.py:109-111: clamp(exp(a) * img + b, 0, 1) in float32, * 255, .byte()
    -> uint8 [H,W,3]."""
    img = np.asarray(original_image, dtype=np.float32)
    ea = np.exp(np.float32(exposure_a)).astype(np.float32)
    ab = (ea * img).astype(np.float32) + np.float32(exposure_b)
    ab = np.clip(ab.astype(np.float32), np.float32(0), np.float32(1))
    return (ab * np.float32(255)).astype(np.float32).astype(np.uint8).transpose(1, 2, 0)


def backproject(rgb_u8, depth, fx, fy, cx, cy, w2c, depth_trunc=100.0):
    """open3d 0.17: RGBDImage.create_from_color_and_depth(depth_scale=1, depth_trunc) +
    PointCloud.create_from_rgbd_image(intrinsic, extrinsic=w2c, project_valid_depth_only=True).
    Returns dense (xyz float32 [N,3], rgb float32 [N,3], valid bool [N]) in row-major pixel order;
    the valid rows, in order, are open3d's point cloud (converted to float32 as the reference does,
    gaussian_model.py:158-159)."""
    d = np.asarray(depth, dtype=np.float32).copy()
    H, W = d.shape
    d[d >= np.float32(depth_trunc)] = 0.0          # ConvertDepthToFloatImage: truncated to 0
    valid = d > 0                                   # NaN -> False
    c2w = np.linalg.inv(np.asarray(w2c, dtype=np.float64))
    v, u = np.meshgrid(np.arange(H, dtype=np.float64), np.arange(W, dtype=np.float64), indexing="ij")
    z = d.astype(np.float64)
    x = (u - cx) * z / fx
    y = (v - cy) * z / fy
    pts = np.empty((H, W, 3), dtype=np.float64)
    for r in range(3):                              # left to right, every operation rounded
        pts[..., r] = ((c2w[r, 0] * x + c2w[r, 1] * y) + c2w[r, 2] * z) + c2w[r, 3]
    xyz = pts.reshape(-1, 3).astype(np.float32)
    xyz[~valid.reshape(-1)] = 0.0
    rgb = (rgb_u8.reshape(-1, 3).astype(np.float64) / 255.0).astype(np.float32)
    return xyz, rgb, valid.reshape(-1), c2w


def seed_attributes(xyz, rgb, dist2, point_size, max_sh_degree=0, isotropic=False):
    """gaussian_model.py:158-193 after the down-sample: SH DC from colour, log-scale from the mean
    3-NN distance, identity rotation, opacity logit 0."""
    n = xyz.shape[0]
    feats = np.zeros((n, 3, (max_sh_degree + 1) ** 2), dtype=np.float32)
    feats[:, :, 0] = (rgb.astype(np.float32) - np.float32(0.5)) / np.float32(C0)
    d2 = np.maximum(np.asarray(dist2, dtype=np.float32), np.float32(1e-7)) * np.float32(point_size)
    scales = np.log(np.sqrt(d2)).astype(np.float32)[:, None]
    if not isotropic:
        scales = np.repeat(scales, 3, axis=1)
    rots = np.zeros((n, 4), dtype=np.float32)
    rots[:, 0] = 1
    opac = np.log(np.float32(0.5) / (np.float32(1) - np.float32(0.5))) * np.ones((n, 1), dtype=np.float32)
    return feats, scales, rots, opac
