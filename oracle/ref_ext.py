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
