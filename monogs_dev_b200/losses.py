"""Fused per-pixel losses of the mapping / tracking loops, backed by libb2gs (csrc/loss.cu).

Same math, names and argument meaning as the reference's utils/slam_utils.py:
  get_loss_mapping(config, image, depth, viewpoint, opacity, initialization)   :91-99
  get_loss_mapping_rgb / get_loss_mapping_rgbd                                  :101-128
  get_loss_tracking(config, image, depth, opacity, viewpoint)                   :56-61
  get_loss_tracking_rgb / get_loss_tracking_rgbd                                :64-88
but ONE forward kernel (+ a one-block finalize) instead of ~15 torch kernels, and ONE
backward kernel producing dL/dimage, dL/ddepth, dL/dopacity and the exposure gradients.
No CPU path: CPU tensors raise.
"""
import ctypes as C

import torch

from . import _lib
from ._lib import lib, check, ptr

MAPPING_RGBD, TRACKING_RGBD, MAPPING_RGB, TRACKING_RGB = 0, 1, 2, 3


def _args(mode, alpha, thr, image, depth, opacity, gt_image, gt_depth, exp_a, exp_b, grad_mask):
    a = _lib.B2gsLossArgs()
    a.N = image.shape[-1] * image.shape[-2]
    a.mode, a.alpha, a.rgb_boundary_threshold = mode, float(alpha), float(thr)
    a.image, a.depth, a.opacity = ptr(image), ptr(depth), ptr(opacity)
    a.gt_image, a.gt_depth = ptr(gt_image), ptr(gt_depth)
    a.exposure_a, a.exposure_b, a.grad_mask = ptr(exp_a), ptr(exp_b), ptr(grad_mask)
    return a


def _c(t, name):
    if t is None:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: libb2gs has no CPU path")
    if t.dtype != torch.float32:
        t = t.float()
    return t.contiguous()


class _FusedLoss(torch.autograd.Function):
    @staticmethod
    def forward(ctx, image, depth, opacity, exp_a, exp_b, gt_image, gt_depth, grad_mask, mode, alpha, thr):
        image, depth, opacity = _c(image, "image"), _c(depth, "depth"), _c(opacity, "opacity")
        gt_image, gt_depth, grad_mask = _c(gt_image, "gt_image"), _c(gt_depth, "gt_depth"), _c(grad_mask, "grad_mask")
        exp_a, exp_b = _c(exp_a, "exposure_a"), _c(exp_b, "exposure_b")
        dev = image.device
        with torch.cuda.device(dev):
            loss = torch.empty((), dtype=torch.float32, device=dev)
            n = image.shape[-1] * image.shape[-2]
            scratch = torch.empty(lib.b2gs_loss_scratch_bytes(n), dtype=torch.uint8, device=dev)
            a = _args(mode, alpha, thr, image, depth, opacity, gt_image, gt_depth, exp_a, exp_b, grad_mask)
            check(lib.b2gs_loss_forward(C.byref(a), loss.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                        torch.cuda.current_stream().cuda_stream))
        e = lambda t: t if t is not None else torch.empty(0, device=dev)
        ctx.save_for_backward(image, e(depth), e(opacity), e(exp_a), e(exp_b), gt_image, e(gt_depth),
                              e(grad_mask))
        ctx.cfg = (mode, alpha, thr)
        return loss

    @staticmethod
    def backward(ctx, g):
        image, depth, opacity, exp_a, exp_b, gt_image, gt_depth, grad_mask = \
            [t if t.numel() else None for t in ctx.saved_tensors]
        mode, alpha, thr = ctx.cfg
        dev = image.device
        with torch.cuda.device(dev):
            g = g.contiguous().float()
            d_image = torch.empty_like(image)
            n = image.shape[-1] * image.shape[-2]
            d_depth = torch.empty(depth.shape if depth is not None else (1,) + tuple(image.shape[-2:]),
                                  dtype=torch.float32, device=dev)
            want_op = opacity is not None and ctx.needs_input_grad[2]
            d_op = torch.empty_like(opacity) if want_op else None
            d_exp = torch.empty(2, dtype=torch.float32, device=dev)
            scratch = torch.empty(lib.b2gs_loss_scratch_bytes(n), dtype=torch.uint8, device=dev)
            a = _args(mode, alpha, thr, image, depth, opacity, gt_image, gt_depth, exp_a, exp_b, grad_mask)
            check(lib.b2gs_loss_backward(C.byref(a), g.data_ptr(), d_image.data_ptr(), d_depth.data_ptr(),
                                         ptr(d_op), d_exp.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                         torch.cuda.current_stream().cuda_stream))
        return (d_image, d_depth if depth is not None else None, d_op,
                d_exp[0:1] if exp_a is not None else None, d_exp[1:2] if exp_b is not None else None,
                None, None, None, None, None, None)


def fused_loss(mode, image, depth, opacity, exposure_a, exposure_b, gt_image, gt_depth, grad_mask=None,
               alpha=0.95, rgb_boundary_threshold=0.01):
    return _FusedLoss.apply(image, depth, opacity, exposure_a, exposure_b, gt_image, gt_depth, grad_mask,
                            int(mode), float(alpha), float(rgb_boundary_threshold))


def get_loss_mapping_rgbd(image, depth, viewpoint, gt_image, gt_depth, alpha=0.95,
                          rgb_boundary_threshold=0.01, initialization=False):
    """utils/slam_utils.py:91-99,113-128; `image` is the raw render (the exposure affine is fused)."""
    ea, eb = (None, None) if initialization else (viewpoint.exposure_a, viewpoint.exposure_b)
    return fused_loss(MAPPING_RGBD, image, depth, None, ea, eb, gt_image, gt_depth, None, alpha,
                      rgb_boundary_threshold)


def get_loss_mapping_rgb(image, viewpoint, gt_image, rgb_boundary_threshold=0.01, initialization=False):
    """utils/slam_utils.py:91-110 (monocular)."""
    ea, eb = (None, None) if initialization else (viewpoint.exposure_a, viewpoint.exposure_b)
    return fused_loss(MAPPING_RGB, image, None, None, ea, eb, gt_image, None, None, 0.95,
                      rgb_boundary_threshold)


def get_loss_tracking_rgbd(image, depth, opacity, viewpoint, gt_image, gt_depth, alpha=0.95,
                           rgb_boundary_threshold=0.01, grad_mask=None):
    """utils/slam_utils.py:56-61,74-88."""
    return fused_loss(TRACKING_RGBD, image, depth, opacity, viewpoint.exposure_a, viewpoint.exposure_b,
                      gt_image, gt_depth, grad_mask, alpha, rgb_boundary_threshold)


def get_loss_tracking_rgb(image, opacity, viewpoint, gt_image, rgb_boundary_threshold=0.01, grad_mask=None):
    """utils/slam_utils.py:56-72 (monocular)."""
    return fused_loss(TRACKING_RGB, image, None, opacity, viewpoint.exposure_a, viewpoint.exposure_b,
                      gt_image, None, grad_mask, 0.95, rgb_boundary_threshold)


def scale_regulariser(raw_scaling, grad_sink=None, weight=10.0):
    """10 * |exp(s) - mean_row(exp(s))|.mean() (utils/slam_backend.py:229-231) as one kernel; the
    gradient wrt the raw log-scales is ADDED to `grad_sink` (the optimiser's gradient buffer) — not
    an autograd op: the caller adds the returned value to the reported loss."""
    if not raw_scaling.is_cuda:
        raise RuntimeError("scale_regulariser needs CUDA tensors: libb2gs has no CPU path")
    s = raw_scaling.detach()
    s = s if s.is_contiguous() else s.contiguous()
    dev = s.device
    with torch.cuda.device(dev):
        out = torch.empty((), dtype=torch.float32, device=dev)
        scratch = torch.empty(lib.b2gs_loss_scratch_bytes(s.shape[0]), dtype=torch.uint8, device=dev)
        check(lib.b2gs_scale_regulariser(s.shape[0], s.data_ptr(), float(weight), ptr(grad_sink),
                                         out.data_ptr(), scratch.data_ptr(), scratch.numel(),
                                         torch.cuda.current_stream().cuda_stream))
    return out


# -- colour refinement loss (BackEnd.color_refinement, utils/slam_backend.py:320-355) ---------------
# Plain torch ops on whatever device the image lives on: this loss runs once per refinement
# iteration on one view and is outside the §8 hot path (the rasterizer and Adam under it are not).

def l1_loss(network_output, gt):
    """gaussian_splatting/utils/loss_utils.py:21-22"""
    return torch.abs(network_output - gt).mean()


def _gaussian_window(window_size, sigma, like):
    x = torch.arange(window_size, dtype=torch.float32) - window_size // 2
    g = torch.exp(-(x * x) / (2.0 * sigma * sigma))
    return (g / g.sum()).to(device=like.device, dtype=like.dtype)


def ssim(img1, img2, window_size=11, size_average=True):
    """loss_utils.py:59-101: Gaussian-windowed (11 taps, sigma 1.5, zero padding) SSIM per channel.
    The 2-D window is the outer product of a 1-D one, so it is applied as a row pass and a column
    pass (22 taps per pixel instead of 121)."""
    c = img1.shape[-3]
    a, b = img1.reshape(-1, c, *img1.shape[-2:]), img2.reshape(-1, c, *img2.shape[-2:])
    w = _gaussian_window(window_size, 1.5, a)
    wr = w.view(1, 1, 1, -1).expand(c, 1, 1, window_size).contiguous()
    wc = w.view(1, 1, -1, 1).expand(c, 1, window_size, 1).contiguous()
    pad = window_size // 2
    F = torch.nn.functional

    def blur(t):
        return F.conv2d(F.conv2d(t, wr, padding=(0, pad), groups=c), wc, padding=(pad, 0), groups=c)

    mu1, mu2 = blur(a), blur(b)
    mu1_sq, mu2_sq, mu1_mu2 = mu1 * mu1, mu2 * mu2, mu1 * mu2
    sigma1_sq = blur(a * a) - mu1_sq
    sigma2_sq = blur(b * b) - mu2_sq
    sigma12 = blur(a * b) - mu1_mu2
    C1, C2 = 0.01 ** 2, 0.03 ** 2
    ssim_map = ((2 * mu1_mu2 + C1) * (2 * sigma12 + C2)) / ((mu1_sq + mu2_sq + C1) * (sigma1_sq + sigma2_sq + C2))
    if size_average:
        return ssim_map.mean()
    return ssim_map.mean(1).mean(1).mean(1)
