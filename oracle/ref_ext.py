"""TEST INFRASTRUCTURE — thin access to the UNMODIFIED reference CUDA extensions built by
oracle/build_ref.py into oracle/_ref/ (GPU box only; nothing in the product imports this).

`forward` / `backward` call the reference's pybind entry points directly
(`rasterize_gaussians`, `rasterize_gaussians_backward`, DGR/ext.cpp:15-19) with the
argument order of DGR/diff_gaussian_rasterization/__init__.py:66-87,119-142, and decode
the three opaque byte buffers with the layout of fromChunk (CR/rasterizer_impl.cu:155-194,
128-byte aligned segments, CR/rasterizer_impl.h:21-27) so tests can compare keys, sorted
lists and tile ranges bit for bit.
"""
import glob
import importlib.util
import os

import torch

_REF = os.path.join(os.path.dirname(os.path.abspath(__file__)), "_ref")


def _load(name):
    hits = glob.glob(os.path.join(_REF, name + "*.so"))
    if not hits:
        return None
    spec = importlib.util.spec_from_file_location(name, hits[0])
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


_dgr = None
_knn = None


def dgr():
    global _dgr
    if _dgr is None:
        _dgr = _load("dgr_ref_C")
    return _dgr


def knn():
    global _knn
    if _knn is None:
        _knn = _load("knn_ref_C")
    return _knn


def available():
    return torch.cuda.is_available() and dgr() is not None


def _e(t):
    return torch.Tensor([]) if t is None else t


def forward(rs, means3D, opacities, shs=None, colors_precomp=None, scales=None, rotations=None,
            cov3D_precomp=None):
    args = (rs.bg, means3D, _e(colors_precomp), opacities, _e(scales), _e(rotations),
            rs.scale_modifier, _e(cov3D_precomp), rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw,
            rs.tanfovx, rs.tanfovy, rs.image_height, rs.image_width, _e(shs), rs.sh_degree,
            rs.campos, rs.prefiltered, rs.debug)
    (R, color, radii, geom, binning, img, depth, opacity, n_touched) = dgr().rasterize_gaussians(*args)
    return dict(num_rendered=R, color=color, radii=radii, geom=geom, binning=binning, img=img,
                depth=depth, opacity=opacity, n_touched=n_touched)


def backward(rs, fwd, g_color, g_depth, means3D, shs=None, colors_precomp=None, scales=None,
             rotations=None, cov3D_precomp=None):
    args = (rs.bg, means3D, fwd["radii"], _e(colors_precomp), _e(scales), _e(rotations),
            rs.scale_modifier, _e(cov3D_precomp), rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw,
            rs.tanfovx, rs.tanfovy, g_color, g_depth, _e(shs), rs.sh_degree, rs.campos,
            fwd["geom"], fwd["num_rendered"], fwd["binning"], fwd["img"], rs.debug)
    (g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales, g_rots,
     g_tau) = dgr().rasterize_gaussians_backward(*args)
    tau = g_tau.view(-1, 6).sum(0)      # DGR/diff_gaussian_rasterization/__init__.py:152
    return dict(dL_dmeans2D=g_means2D, dL_dcolors=g_colors, dL_dopacity=g_opac,
                dL_dmeans3D=g_means3D, dL_dcov3D=g_cov3D, dL_dsh=g_sh, dL_dscales=g_scales,
                dL_drotations=g_rots, dL_dtau_per=g_tau, dL_dtau=tau)


class _Chunk:
    def __init__(self, buf):
        self.buf, self.off = buf, 0
        assert buf.data_ptr() % 128 == 0

    def take(self, dtype, count, shape=None):
        self.off = (self.off + 127) // 128 * 128
        nbytes = count * torch.empty(0, dtype=dtype).element_size()
        t = self.buf[self.off:self.off + nbytes].view(dtype)
        self.off += nbytes
        return t.reshape(shape) if shape else t


def decode_geom(fwd, P):
    c = _Chunk(fwd["geom"])
    return dict(depths=c.take(torch.float32, P), clamped=c.take(torch.bool, 3 * P, (P, 3)),
                internal_radii=c.take(torch.int32, P), means2D=c.take(torch.float32, 2 * P, (P, 2)),
                cov3D=c.take(torch.float32, 6 * P, (P, 6)),
                conic_opacity=c.take(torch.float32, 4 * P, (P, 4)),
                rgb=c.take(torch.float32, 3 * P, (P, 3)), tiles_touched=c.take(torch.int32, P))


def decode_binning(fwd):
    R = fwd["num_rendered"]
    if R == 0:
        z = torch.zeros(0, dtype=torch.int64, device=fwd["binning"].device)
        return dict(point_list=z, keys=z, point_list_unsorted=z, keys_unsorted=z)
    c = _Chunk(fwd["binning"])
    return dict(point_list=c.take(torch.int32, R), point_list_unsorted=c.take(torch.int32, R),
                keys=c.take(torch.int64, R), keys_unsorted=c.take(torch.int64, R))


def decode_image(fwd, W, H):
    N = W * H
    c = _Chunk(fwd["img"])
    out = dict(final_T=c.take(torch.float32, N), n_contrib=c.take(torch.int32, N))
    tiles = ((W + 15) // 16) * ((H + 15) // 16)
    out["ranges"] = c.take(torch.int32, 2 * N, (N, 2))[:tiles]
    return out


class RefRasterizeFunction(torch.autograd.Function):
    """Autograd bridge over the reference extension with the semantics of the reference's
    _RasterizeGaussians (DGR/diff_gaussian_rasterization/__init__.py:50-171), restated here
    because /root/reference does not exist on the GPU box.  Used by bench.py --impl reference."""

    @staticmethod
    def forward(ctx, means3D, means2D, sh, colors_precomp, opacities, scales, rotations,
                cov3Ds_precomp, theta, rho, rs):
        args = (rs.bg, means3D, colors_precomp, opacities, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.tanfovx, rs.tanfovy,
                rs.image_height, rs.image_width, sh, rs.sh_degree, rs.campos, rs.prefiltered, rs.debug)
        (R, color, radii, geom, binning, img, depth, opacity, n_touched) = dgr().rasterize_gaussians(*args)
        ctx.rs, ctx.R = rs, R
        ctx.save_for_backward(colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh,
                              geom, binning, img)
        return color, radii, depth, opacity, n_touched

    @staticmethod
    def backward(ctx, g_color, g_radii, g_depth, g_opacity, g_touched):
        rs = ctx.rs
        colors_precomp, means3D, scales, rotations, cov3Ds_precomp, radii, sh, geom, binning, img = ctx.saved_tensors
        args = (rs.bg, means3D, radii, colors_precomp, scales, rotations, rs.scale_modifier,
                cov3Ds_precomp, rs.viewmatrix, rs.projmatrix, rs.projmatrix_raw, rs.tanfovx, rs.tanfovy,
                g_color, g_depth, sh, rs.sh_degree, rs.campos, geom, ctx.R, binning, img, rs.debug)
        (g_means2D, g_colors, g_opac, g_means3D, g_cov3D, g_sh, g_scales, g_rots,
         g_tau) = dgr().rasterize_gaussians_backward(*args)
        g_tau = torch.sum(g_tau.view(-1, 6), dim=0)
        return (g_means3D, g_means2D, g_sh, g_colors, g_opac, g_scales, g_rots, g_cov3D,
                g_tau[3:].view(1, -1), g_tau[:3].view(1, -1), None)


def ref_render(viewpoint_camera, pc, pipe=None, bg_color=None, scaling_modifier=1.0):
    """render() of the reference (gaussian_splatting/gaussian_renderer/__init__.py:24-153)
    over RefRasterizeFunction; same return dict."""
    import math
    from types import SimpleNamespace
    cam = viewpoint_camera
    screenspace_points = torch.zeros_like(pc.get_xyz, requires_grad=True) + 0
    try:
        screenspace_points.retain_grad()  # gaussian_renderer/__init__.py:50-52
    except Exception:
        pass
    rs = SimpleNamespace(
        image_height=int(cam.image_height), image_width=int(cam.image_width),
        tanfovx=math.tan(cam.FoVx * 0.5), tanfovy=math.tan(cam.FoVy * 0.5), bg=bg_color,
        scale_modifier=scaling_modifier, viewmatrix=cam.world_view_transform,
        projmatrix=cam.full_proj_transform, projmatrix_raw=cam.projection_matrix,
        sh_degree=pc.active_sh_degree, campos=cam.camera_center, prefiltered=False, debug=False)
    e = torch.Tensor([])
    color, radii, depth, opacity, n_touched = RefRasterizeFunction.apply(
        pc.get_xyz, screenspace_points, pc.get_features, e, pc.get_opacity, pc.get_scaling,
        pc.get_rotation, e, cam.cam_rot_delta, cam.cam_trans_delta, rs)
    return {"render": color, "viewspace_points": screenspace_points, "visibility_filter": radii > 0,
            "radii": radii, "depth": depth, "opacity": opacity, "n_touched": n_touched}


class RefGaussianModel:
    """Stock layout of the reference's GaussianModel: six separate nn.Parameters and
    torch.optim.Adam(l, lr=0.0, eps=1e-15) (gaussian_splatting/scene/gaussian_model.py:245-283)."""

    def __init__(self, xyz, f_dc, f_rest, opacity, scaling, rotation, sh_degree=0, spatial_lr_scale=6.0):
        P = torch.nn.Parameter
        self.max_sh_degree = self.active_sh_degree = sh_degree
        self._xyz, self._features_dc, self._features_rest = P(xyz.clone()), P(f_dc.clone()), P(f_rest.clone())
        self._opacity, self._scaling, self._rotation = P(opacity.clone()), P(scaling.clone()), P(rotation.clone())
        l = [dict(params=[self._xyz], lr=0.00016 * spatial_lr_scale, name="xyz"),
             dict(params=[self._features_dc], lr=0.0025, name="f_dc"),
             dict(params=[self._features_rest], lr=0.0025 / 20.0, name="f_rest"),
             dict(params=[self._opacity], lr=0.05, name="opacity"),
             dict(params=[self._scaling], lr=0.001 * spatial_lr_scale, name="scaling"),
             dict(params=[self._rotation], lr=0.001, name="rotation")]
        self.optimizer = torch.optim.Adam(l, lr=0.0, eps=1e-15)
        # gaussian_splatting/scene/gaussian_model.py:233-235
        n = xyz.shape[0]
        self.xyz_gradient_accum = torch.zeros((n, 1), device=xyz.device)
        self.denom = torch.zeros((n, 1), device=xyz.device)
        self.max_radii2D = torch.zeros((n), device=xyz.device)

    def _stats(self):
        return self.xyz_gradient_accum, self.denom, self.max_radii2D

    def add_densification_stats(self, viewspace_point_tensor, update_filter):
        """gaussian_splatting/scene/gaussian_model.py:691-695, verbatim semantics."""
        self.xyz_gradient_accum[update_filter] += torch.norm(
            viewspace_point_tensor.grad[update_filter, :2], dim=-1, keepdim=True)
        self.denom[update_filter] += 1

    get_xyz = property(lambda s: s._xyz)
    get_scaling = property(lambda s: torch.exp(s._scaling))
    get_rotation = property(lambda s: torch.nn.functional.normalize(s._rotation))
    get_opacity = property(lambda s: torch.sigmoid(s._opacity))
    get_features = property(lambda s: torch.cat((s._features_dc, s._features_rest), dim=1))

    def parameters(self):
        return [self._xyz, self._features_dc, self._features_rest, self._opacity, self._scaling, self._rotation]

    def zero_grad(self):
        self.optimizer.zero_grad(set_to_none=True)
