"""Drop-in Python surface of the rasterizer: same names, argument meaning, return
tuple and error behaviour as the reference's
`submodules/diff-gaussian-rasterization/diff_gaussian_rasterization/__init__.py`
(GaussianRasterizationSettings :173-186, GaussianRasterizer :188-244,
rasterize_gaussians :21-46, _RasterizeGaussians :50-171), backed by libb2gs.so.

Host responsibilities here (the C library never allocates): output / gradient /
workspace tensors, stream + device selection, and the autograd bridge.
"""
from typing import NamedTuple
import ctypes as C

import torch
import torch.nn as nn

from . import _lib
from ._lib import lib, check, ptr


class GaussianRasterizationSettings(NamedTuple):
    image_height: int
    image_width: int
    tanfovx: float
    tanfovy: float
    bg: torch.Tensor
    scale_modifier: float
    viewmatrix: torch.Tensor
    projmatrix: torch.Tensor
    projmatrix_raw: torch.Tensor
    sh_degree: int
    campos: torch.Tensor
    prefiltered: bool
    debug: bool


def _snapshot(args):
    return tuple(a.detach().cpu().clone() if isinstance(a, torch.Tensor) else a for a in args)


def _f32c(t, name):
    """contiguous float32 CUDA tensor or None for absent (empty) inputs."""
    if t is None or t.numel() == 0:
        return None
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor: libb2gs has no CPU path")
    if t.dtype != torch.float32:
        raise RuntimeError(f"{name} must be float32, got {t.dtype}")
    return t.contiguous()


def _scene_struct(rs, P, M, means3D, sh, colors, opac, scales, rots, cov3D, keep, sh_rest=None,
                  raw=False):
    dev = means3D.device
    def mat(t, name):
        t = _f32c(t, name)
        if t is not None and t.device != dev:
            t = t.to(dev)
        keep.append(t)
        return ptr(t)
    s = _lib.B2gsScene()
    s.P, s.D, s.M = P, int(rs.sh_degree), M
    s.W, s.H = int(rs.image_width), int(rs.image_height)
    s.tan_fovx, s.tan_fovy = float(rs.tanfovx), float(rs.tanfovy)
    s.scale_modifier = float(rs.scale_modifier)
    s.prefiltered, s.debug = int(bool(rs.prefiltered)), int(bool(rs.debug))
    s.bg = mat(rs.bg, "bg")
    s.viewmatrix = mat(rs.viewmatrix, "viewmatrix")
    s.projmatrix = mat(rs.projmatrix, "projmatrix")
    s.projmatrix_raw = mat(rs.projmatrix_raw, "projmatrix_raw")
    s.campos = mat(rs.campos, "campos")
    s.means3D = ptr(means3D)
    s.shs, s.colors_precomp, s.opacities = ptr(sh), ptr(colors), ptr(opac)
    s.scales, s.rotations, s.cov3D_precomp = ptr(scales), ptr(rots), ptr(cov3D)
    s.raw_params, s.shs_rest = int(bool(raw)), ptr(sh_rest)
    return s


class RasterState:
    """What the forward hands to the backward (the reference's three opaque byte
    buffers + num_rendered, DGR/diff_gaussian_rasterization/__init__.py:102-104).
    `num_rendered` is the binning capacity the backward must carve with: the exact
    instance count in the default two-call mode, the workspace capacity in sync-free mode
    (then `status` is the device [2] = (instances, overflow flag) the caller must check)."""
    __slots__ = ("geom", "image", "binning", "num_rendered", "status", "key", "scene", "keep")

    def __init__(self):
        self.status = None
        self.key = None
        self.scene = None      # the forward's B2gsScene (same pointers in backward: not rebuilt)
        self.keep = None


class _SyncFree:
    """Sync-free forward (b2gs_forward): no D2H read of the instance count, so the whole
    render -> loss -> backward chain can be enqueued ahead of the GPU or captured in a CUDA
    graph.  The binning workspace is sized from the last exactly-measured instance count of
    the same (device, W, H, P) times `headroom`; the caller checks `check_overflow(states)`
    once per step and re-runs the step if a view overflowed."""
    enabled = False
    headroom = 1.5
    cap = {}


def set_sync_free(enabled=True, headroom=1.5):
    _SyncFree.enabled = bool(enabled)
    _SyncFree.headroom = float(headroom)
    if not enabled:
        _SyncFree.cap.clear()


def check_overflow(states):
    """One host sync for a whole batch of forwards: returns True (and raises the stored
    capacities) if any sync-free forward ran out of binning capacity — its outputs are
    background-only and the step must be re-run."""
    sts = [s for s in states if s is not None and s.status is not None]
    if not sts:
        return False
    return note_overflow(sts, torch.stack([s.status for s in sts]).cpu().tolist())


def note_overflow(sts, vals):
    """host half of check_overflow: `vals` = the (instances, overflow flag) pairs of `sts` already
    on the host (MappingWindow reads them back asynchronously, two steps late)."""
    over = False
    for s, v in zip(sts, vals):
        need = int(v[0] * _SyncFree.headroom) + 4096
        if v[1]:
            over = True
            _SyncFree.cap[s.key] = max(_SyncFree.cap.get(s.key, 0), need)
    return over


def _forward_native(rs, means3D, sh, colors, opac, scales, rots, cov3D, sh_rest=None, raw=False):
    if means3D.dim() != 2 or means3D.size(1) != 3:
        # message of DGR/rasterize_points.cu:58-60
        raise RuntimeError("means3D must have dimensions (num_points, 3)")
    P = means3D.size(0)
    H, W = int(rs.image_height), int(rs.image_width)
    dev = means3D.device
    M = sh.size(1) if sh is not None else 0
    if raw:
        M = 1 + (sh_rest.size(1) if sh_rest is not None else 0)
    st = RasterState()
    if P == 0:
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=dev)
        st.geom = st.image = st.binning = torch.empty(0, dtype=torch.uint8, device=dev)
        st.num_rendered = 0
        zi = torch.zeros(0, dtype=torch.int32, device=dev)
        return st, z(3, H, W), zi, z(1, H, W), z(1, H, W), zi.clone()
    with torch.cuda.device(dev):
        stream = torch.cuda.current_stream().cuda_stream
        f32 = dict(dtype=torch.float32, device=dev)
        color = torch.empty(3, H, W, **f32)
        depth = torch.empty(1, H, W, **f32)
        opacity = torch.empty(1, H, W, **f32)
        radii = torch.empty(P, dtype=torch.int32, device=dev)
        n_touched = torch.empty(P, dtype=torch.int32, device=dev)
        st.geom = torch.empty(lib.b2gs_geom_bytes(P), dtype=torch.uint8, device=dev)
        st.image = torch.empty(lib.b2gs_image_bytes(W, H), dtype=torch.uint8, device=dev)
        keep = []
        scene = _scene_struct(rs, P, M, means3D, sh, colors, opac, scales, rots, cov3D, keep,
                              sh_rest=sh_rest, raw=raw)
        st.scene, st.keep = scene, (keep, means3D, sh, colors, opac, scales, rots, cov3D, sh_rest)
        ws = _lib.B2gsWorkspace(st.geom.data_ptr(), st.geom.numel(), st.image.data_ptr(),
                                st.image.numel(), None, 0)
        out = _lib.B2gsForwardOut(color.data_ptr(), depth.data_ptr(), opacity.data_ptr(),
                                  radii.data_ptr(), n_touched.data_ptr(), None)
        key = (dev.index, W, H, P)
        cap = _SyncFree.cap.get(key) if _SyncFree.enabled else None
        if cap is not None:
            # sync-free: fixed-capacity binning workspace, status stays on the device
            st.binning = torch.empty(lib.b2gs_binning_bytes(cap, W, H), dtype=torch.uint8, device=dev)
            st.status = torch.empty(2, dtype=torch.int32, device=dev)
            st.key = key
            ws.binning, ws.binning_bytes = st.binning.data_ptr(), st.binning.numel()
            out.status = st.status.data_ptr()
            check(lib.b2gs_forward(C.byref(scene), C.byref(ws), C.byref(out), stream))
            st.num_rendered = lib.b2gs_binning_capacity(st.binning.numel())
            return st, color, radii, depth, opacity, n_touched
        R = C.c_int64(0)
        check(lib.b2gs_forward_plan(C.byref(scene), C.byref(ws), C.byref(out), stream, C.byref(R)))
        R = R.value
        st.binning = torch.empty(lib.b2gs_binning_bytes(R, W, H) if R > 0 else 0,
                                 dtype=torch.uint8, device=dev)
        ws.binning, ws.binning_bytes = ptr(st.binning), st.binning.numel()
        check(lib.b2gs_forward_render(C.byref(scene), C.byref(ws), C.byref(out), R, stream))
        st.num_rendered = R
        if _SyncFree.enabled:      # first call for this shape: measured exactly, remembered
            _SyncFree.cap[key] = int(R * _SyncFree.headroom) + 4096
    return st, color, radii, depth, opacity, n_touched


class _RasterizeGaussians(torch.autograd.Function):
    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, theta, rho, raster_settings):
        rs = raster_settings
        m3 = _f32c(means3D, "means3D")
        if m3 is None:
            if not means3D.is_cuda:
                raise RuntimeError("means3D must be a CUDA tensor: libb2gs has no CPU path")
            m3 = means3D.contiguous()
        ins = (m3, _f32c(sh, "shs"), _f32c(colors_precomp, "colors_precomp"),
               _f32c(opacities, "opacities"), _f32c(scales, "scales"),
               _f32c(rotations, "rotations"), _f32c(cov3Ds_precomp, "cov3D_precomp"))
        if rs.debug:
            snap = _snapshot((rs.bg, means3D, colors_precomp, opacities, scales, rotations,
                              rs.scale_modifier, cov3Ds_precomp, rs.viewmatrix, rs.projmatrix,
                              rs.projmatrix_raw, rs.tanfovx, rs.tanfovy, rs.image_height,
                              rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug))
            try:
                st, color, radii, depth, opacity, n_touched = _forward_native(rs, *ins)
            except Exception:
                torch.save(snap, "snapshot_fw.dump")
                print("\nAn error occured in forward. Please forward snapshot_fw.dump for debugging.")
                raise
        else:
            st, color, radii, depth, opacity, n_touched = _forward_native(rs, *ins)
        ctx.raster_settings = rs
        ctx.state = st
        ctx.save_for_backward(*[t if t is not None else torch.empty(0) for t in ins], radii)
        ctx.mark_non_differentiable(radii, n_touched)
        return color, radii, depth, opacity, n_touched

    @staticmethod
    def backward(ctx, grad_out_color, grad_out_radii, grad_out_depth, grad_out_opacity,
                 grad_n_touched):
        # grad_out_opacity / radii / n_touched are dropped, as in the reference
        # (DGR/diff_gaussian_rasterization/__init__.py:108,116-138) — quirk q1.
        rs = ctx.raster_settings
        st = ctx.state
        means3D, sh, colors, opac, scales, rots, cov3D, radii = ctx.saved_tensors
        none_if_empty = lambda t: t if t.numel() else None
        sh, colors, scales, rots, cov3D = map(none_if_empty, (sh, colors, scales, rots, cov3D))
        need = ctx.needs_input_grad
        P = means3D.size(0)
        dev = means3D.device
        H, W = int(rs.image_height), int(rs.image_width)
        M = sh.size(1) if sh is not None else 0
        f32 = dict(dtype=torch.float32, device=dev)
        if grad_out_color is None:
            grad_out_color = torch.zeros(3, H, W, **f32)
        if grad_out_depth is None:
            grad_out_depth = torch.zeros(1, H, W, **f32)
        g_color = grad_out_color.contiguous()
        g_depth = grad_out_depth.contiguous()

        def run():
            with torch.cuda.device(dev):
                stream = torch.cuda.current_stream().cuda_stream
                # every gradient output is a slice of ONE allocation (rows 16 B aligned): the drop-in
                # path at small P is bound by host time, and nine torch.empty calls are a third of it
                want_sr = scales is not None and (need[5] or need[6])
                widths = [3, 3, 1, 3 if (colors is not None and need[3]) else 0,
                          3 * M if (sh is not None and need[2]) else 0, 3 if want_sr else 0, 4 if want_sr else 0,
                          6 if (cov3D is not None and need[7]) else 0]
                offs, total = [], 0
                for w in widths:
                    offs.append(total)
                    total += (P * w + 3) // 4 * 4
                flat = torch.empty(total + 8, **f32)
                cut = lambda k, *shape: flat[offs[k]:offs[k] + P * widths[k]].view(*shape) if widths[k] else None
                g_means3D, g_means2D, g_opac = cut(0, P, 3), cut(1, P, 3), cut(2, P, 1)
                g_colors, g_sh = cut(3, P, 3), cut(4, P, M, 3)
                g_scales, g_rots, g_cov3D = cut(5, P, 3), cut(6, P, 4), cut(7, P, 6)
                g_tau = flat[total:total + 6]
                if P == 0:
                    g_tau.zero_()
                    return g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rots, g_cov3D, g_tau
                scratch = torch.empty(lib.b2gs_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
                scene = st.scene        # built by the forward from the very tensors saved for backward
                if scene is None:
                    keep = []
                    scene = _scene_struct(rs, P, M, means3D, sh, colors, opac, scales, rots, cov3D, keep)
                ws = _lib.B2gsWorkspace(ptr(st.geom), st.geom.numel(), ptr(st.image),
                                        st.image.numel(), ptr(st.binning), st.binning.numel())
                io = _lib.B2gsBackwardIO(
                    g_color.data_ptr(), g_depth.data_ptr(), radii.data_ptr(), g_means3D.data_ptr(),
                    g_means2D.data_ptr(), g_opac.data_ptr(), ptr(g_colors), ptr(g_sh), ptr(g_scales),
                    ptr(g_rots), ptr(g_cov3D), g_tau.data_ptr(), scratch.data_ptr(), scratch.numel())
                check(lib.b2gs_backward(C.byref(scene), C.byref(ws), st.num_rendered, C.byref(io),
                                        stream))
                return g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rots, g_cov3D, g_tau

        if rs.debug:
            snap = _snapshot((rs.bg, means3D, radii, colors, scales, rots, rs.scale_modifier, cov3D,
                              rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.tanfovx,
                              rs.tanfovy, grad_out_color, grad_out_depth, sh, rs.sh_degree,
                              rs.campos, st.num_rendered, rs.debug))
            try:
                res = run()
            except Exception:
                torch.save(snap, "snapshot_bw.dump")
                print("\nAn error occured in backward. Writing snapshot_bw.dump for debugging.\n")
                raise
        else:
            res = run()
        g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rots, g_cov3D, g_tau = res
        # tau = (rho, theta); the reference returns both as [1,3]
        # (DGR/diff_gaussian_rasterization/__init__.py:152-154)
        grad_rho = g_tau[:3].view(1, -1)
        grad_theta = g_tau[3:].view(1, -1)
        return (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rots, g_cov3D,
                grad_theta if need[8] else None, grad_rho if need[9] else None, None)


class GradSinks:
    """Where the fused-activation rasterizer puts its per-Gaussian results in backward: the
    optimiser's gradient buffers (accumulated in place, one `+=` per live Gaussian per view —
    replaces autograd's AccumulateGrad adds) and, optionally, the densification statistics of
    GaussianModel.add_densification_stats / max_radii2D (GS/scene/gaussian_model.py:691-695,
    utils/slam_backend.py:277-284)."""

    def __init__(self, xyz, f_dc, f_rest, opacity, scaling, rotation, xyz_gradient_accum=None,
                 denom=None, max_radii2D=None):
        self.xyz, self.f_dc, self.f_rest = xyz, f_dc, f_rest
        self.opacity, self.scaling, self.rotation = opacity, scaling, rotation
        self.xyz_gradient_accum, self.denom, self.max_radii2D = xyz_gradient_accum, denom, max_radii2D


class _RasterizeRaw(torch.autograd.Function):
    """Rasterizer over the RAW optimiser parameters with the activations fused into the
    preprocess kernels: exp(_scaling), normalize(_rotation),
    sigmoid(_opacity), cat(_features_dc, _features_rest) and their backward never exist as
    separate kernels.  With `sinks`, gradients are accumulated straight into the given
    buffers and autograd receives None for the six parameter inputs."""

    @staticmethod
    def forward(ctx, xyz, f_dc, f_rest, opacity, scaling, rotation, means2D, theta, rho, rs, sinks):
        ins = (_f32c(xyz, "xyz"), _f32c(f_dc, "features_dc"), _f32c(f_rest, "features_rest"),
               _f32c(opacity, "opacity"), _f32c(scaling, "scaling"), _f32c(rotation, "rotation"))
        m3, dc, rest, op, sc, rot = ins
        if m3 is None:
            raise RuntimeError("xyz must be a non-empty CUDA tensor")
        st, color, radii, depth, opac_img, n_touched = _forward_native(
            rs, m3, dc, None, op, sc, rot, None, sh_rest=rest, raw=True)
        ctx.raster_settings, ctx.state, ctx.sinks = rs, st, sinks
        ctx.save_for_backward(*[t if t is not None else torch.empty(0) for t in ins], radii)
        ctx.mark_non_differentiable(radii, n_touched)
        return color, radii, depth, opac_img, n_touched

    @staticmethod
    def backward(ctx, g_color, g_radii, g_depth, g_opacity, g_nt):
        rs, st, sinks = ctx.raster_settings, ctx.state, ctx.sinks
        m3, dc, rest, op, sc, rot, radii = ctx.saved_tensors
        rest = rest if rest.numel() else None
        P, dev = m3.size(0), m3.device
        H, W = int(rs.image_height), int(rs.image_width)
        M = 1 + (rest.size(1) if rest is not None else 0)
        f32 = dict(dtype=torch.float32, device=dev)
        g_color = (g_color if g_color is not None else torch.zeros(3, H, W, **f32)).contiguous()
        g_depth = (g_depth if g_depth is not None else torch.zeros(1, H, W, **f32)).contiguous()
        need = ctx.needs_input_grad
        with torch.cuda.device(dev):
            stream = torch.cuda.current_stream().cuda_stream
            g_tau = torch.empty(6, **f32)
            scratch = torch.empty(lib.b2gs_backward_scratch_bytes(P), dtype=torch.uint8, device=dev)
            g_m2d = torch.empty(P, 3, **f32) if need[6] else None
            if sinks is not None:
                outs = (sinks.xyz, sinks.f_dc, sinks.f_rest, sinks.opacity, sinks.scaling, sinks.rotation)
                acc = 1
            else:
                # pose-only (tracking shape): parameter inputs that need no gradient get no output
                w_sh, w_sr = need[1] or need[2], need[4] or need[5]
                outs = (torch.empty(P, 3, **f32), torch.empty(P, 1, 3, **f32) if w_sh else None,
                        torch.empty(P, M - 1, 3, **f32) if (M > 1 and w_sh) else None,
                        torch.empty(P, 1, **f32), torch.empty(P, 3, **f32) if w_sr else None,
                        torch.empty(P, 4, **f32) if w_sr else None)
                acc = 0
            o_xyz, o_dc, o_rest, o_op, o_sc, o_rot = outs
            keep = []
            scene = _scene_struct(rs, P, M, m3, dc, None, op, sc, rot, None, keep, sh_rest=rest, raw=True)
            ws = _lib.B2gsWorkspace(ptr(st.geom), st.geom.numel(), ptr(st.image), st.image.numel(),
                                    ptr(st.binning), st.binning.numel())
            io = _lib.B2gsBackwardIO(
                g_color.data_ptr(), g_depth.data_ptr(), radii.data_ptr(), o_xyz.data_ptr(), ptr(g_m2d),
                o_op.data_ptr(), None, ptr(o_dc), ptr(o_sc), ptr(o_rot), None,
                g_tau.data_ptr(), scratch.data_ptr(), scratch.numel(), acc, ptr(o_rest),
                ptr(sinks.xyz_gradient_accum) if sinks is not None else None,
                ptr(sinks.denom) if sinks is not None else None,
                ptr(sinks.max_radii2D) if sinks is not None else None)
            check(lib.b2gs_backward(C.byref(scene), C.byref(ws), st.num_rendered, C.byref(io), stream))
        grads = (None,) * 6 if sinks is not None else tuple(
            o if n else None for o, n in zip((o_xyz, o_dc, o_rest, o_op, o_sc, o_rot), need[:6]))
        return grads + (g_m2d, g_tau[3:].view(1, -1) if need[7] else None,
                        g_tau[:3].view(1, -1) if need[8] else None, None, None)


def rasterize_gaussians_raw(xyz, features_dc, features_rest, opacity, scaling, rotation, raster_settings,
                            means2D=None, theta=None, rho=None, sinks=None):
    """5-tuple (color, radii, depth, opacity, n_touched) like GaussianRasterizer.forward, from the raw
    parameters of GaussianModel (_xyz, _features_dc, _features_rest, _opacity, _scaling, _rotation)."""
    e = lambda t: torch.Tensor([]) if t is None else t
    return _RasterizeRaw.apply(xyz, features_dc, e(features_rest), opacity, scaling, rotation,
                               e(means2D), e(theta), e(rho), raster_settings, sinks)


def rasterize_gaussians(means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                        cov3Ds_precomp, theta, rho, raster_settings):
    return _RasterizeGaussians.apply(means3D, means2D, sh, colors_precomp, opacities, scales,
                                     rotations, cov3Ds_precomp, theta, rho, raster_settings)


class GaussianRasterizer(nn.Module):
    def __init__(self, raster_settings):
        super().__init__()
        self.raster_settings = raster_settings

    def markVisible(self, positions):
        """bool[P]: view-space z > 0.2 (DGR/diff_gaussian_rasterization/__init__.py:193-202)."""
        with torch.no_grad():
            rs = self.raster_settings
            pos = _f32c(positions, "positions")
            P = positions.size(0)
            present = torch.zeros(P, dtype=torch.bool, device=positions.device)
            if P:
                if pos is None:
                    raise RuntimeError("positions must be a CUDA tensor: libb2gs has no CPU path")
                with torch.cuda.device(pos.device):
                    view = _f32c(rs.viewmatrix, "viewmatrix")
                    proj = _f32c(rs.projmatrix, "projmatrix")
                    check(lib.b2gs_mark_visible(P, pos.data_ptr(), view.data_ptr(), ptr(proj),
                                                present.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream))
        return present

    def forward(self, means3D, means2D, opacities, shs=None, colors_precomp=None, scales=None,
                rotations=None, cov3D_precomp=None, theta=None, rho=None):
        rs = self.raster_settings
        if (shs is None) == (colors_precomp is None):
            raise Exception('Please provide excatly one of either SHs or precomputed colors!')
        if ((scales is None or rotations is None) and cov3D_precomp is None) or \
                ((scales is not None or rotations is not None) and cov3D_precomp is not None):
            raise Exception('Please provide exactly one of either scale/rotation pair or precomputed 3D covariance!')
        e = lambda t: torch.Tensor([]) if t is None else t
        return rasterize_gaussians(means3D, means2D, e(shs), e(colors_precomp), opacities, e(scales),
                                   e(rotations), e(cov3D_precomp), e(theta), e(rho), rs)
