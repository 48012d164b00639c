"""Real spherical harmonics up to degree 3 in torch: the host-side colour evaluation render() uses
when `pipe.convert_SHs_python` is set (gaussian_splatting/gaussian_renderer/__init__.py:100-109,
gaussian_splatting/utils/sh_utils.py: eval_sh / RGB2SH / SH2RGB).  The default path evaluates SH
inside preprocess_fwd_kernel; this is the same basis (CR/forward.cu:33-61) as a [P,M] matrix."""
import torch

C0 = 0.28209479177387814
C1 = 0.4886025119029199
C2 = (1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792, 0.5462742152960396)
C3 = (-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154, -0.4570457994644658,
      1.445305721320277, -0.5900435899266435)


def sh_basis(deg, dirs):
    """[..., (deg+1)^2] basis values for unit directions dirs [..., 3]"""
    if not 0 <= deg <= 3:
        raise ValueError("SH degree must be 0..3")
    x, y, z = dirs.unbind(-1)
    cols = [torch.full_like(x, C0)]
    if deg >= 1:
        cols += [-C1 * y, C1 * z, -C1 * x]
    if deg >= 2:
        xx, yy, zz = x * x, y * y, z * z
        cols += [C2[0] * x * y, C2[1] * y * z, C2[2] * (2.0 * zz - xx - yy), C2[3] * x * z, C2[4] * (xx - yy)]
    if deg >= 3:
        cols += [C3[0] * y * (3 * xx - yy), C3[1] * x * y * z, C3[2] * y * (4 * zz - xx - yy),
                 C3[3] * z * (2 * zz - 3 * xx - 3 * yy), C3[4] * x * (4 * zz - xx - yy),
                 C3[5] * z * (xx - yy), C3[6] * x * (xx - 3 * yy)]
    return torch.stack(cols, dim=-1)


def eval_sh(deg, sh, dirs):
    """sh [..., C, M'] with M' >= (deg+1)^2 coefficients per channel, dirs [..., 3] unit -> [..., C]"""
    n = (deg + 1) ** 2
    if sh.shape[-1] < n:
        raise ValueError("eval_sh: %d coefficients given, degree %d needs %d" % (sh.shape[-1], deg, n))
    return (sh[..., :n] * sh_basis(deg, dirs).unsqueeze(-2)).sum(-1)


def RGB2SH(rgb):
    return (rgb - 0.5) / C0


def SH2RGB(sh):
    return sh * C0 + 0.5
