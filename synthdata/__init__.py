"""Deterministic synthetic scenes for tests and bench — NEUTRAL plumbing.

Pure torch on the CPU generator; imports neither the product (`monogs_dev_b200`) nor the
checker (`oracle`), so that both arms of bench.py and both sides of a parity test are fed
from the same tensors: raw (pre-activation) Gaussian parameters, the world-to-camera pose of
every keyframe, the intrinsics.  Each side wraps the poses in its own camera class
(`monogs_dev_b200.camera.Camera` / `oracle.ref_loop.RefCamera`).
"""
from types import SimpleNamespace

import torch

# name -> (P, W, H, fx, fy, cx, cy, sh_degree, seed)   (BASELINE.json configs / SURVEY.md §8d)
CONFIGS = {
    "cfg1_plumbing": (10_000, 320, 240, 300.0, 300.0, 159.5, 119.5, 0, 1235),
    "cfg2_tracking": (100_000, 640, 480, 517.306, 516.469, 318.643, 255.314, 0, 1236),
    "cfg3_replica": (500_000, 1200, 680, 600.0, 600.0, 599.5, 339.5, 0, 1237),
    "cfg5_large": (4_000_000, 1920, 1080, 1000.0, 1000.0, 959.5, 539.5, 3, 1239),
    "tiny": (1500, 160, 112, 150.0, 150.0, 79.5, 55.5, 0, 7),
    "tiny_sh3": (1200, 150, 100, 140.0, 140.0, 74.5, 49.5, 3, 8),
}


def skew(x):
    m = torch.zeros(3, 3, dtype=x.dtype)
    m[0, 1], m[0, 2], m[1, 0], m[1, 2], m[2, 0], m[2, 1] = -x[2], x[1], x[2], -x[0], -x[1], x[0]
    return m


def se3_exp(tau):
    """tau = (rho, theta) -> 4x4 (the closed form of utils/pose_utils.py:24-73)."""
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


def scene_tensors(name=None, *, P=None, W=None, H=None, fx=None, fy=None, cx=None, cy=None,
                  sh_degree=None, seed=None, n_views=1, max_sh_degree=None):
    """Raw Gaussian parameters (CPU tensors) + keyframe poses, per SURVEY.md §8d.  The draw order
    of the generator is part of the contract: goldens under tests/golden/ depend on it."""
    if name is not None:
        cP, cW, cH, cfx, cfy, ccx, ccy, cdeg, cseed = CONFIGS[name]
        P = P or cP; W = W or cW; H = H or cH
        fx = fx or cfx; fy = fy or cfy
        cx = ccx if cx is None else cx; cy = ccy if cy is None else cy
        sh_degree = cdeg if sh_degree is None else sh_degree
        seed = cseed if seed is None else seed
    g = torch.Generator().manual_seed(int(seed))
    U = lambda *s: torch.rand(*s, generator=g)
    N = lambda *s: torch.randn(*s, generator=g)
    near = U(P) < 0.05
    z = torch.where(near, -1.0 + 1.2 * U(P), 0.5 + 5.5 * U(P))
    u = (-0.1 + 1.2 * U(P)) * W
    v = (-0.1 + 1.2 * U(P)) * H
    xyz = torch.stack([(u - cx) * z / fx, (v - cy) * z / fy, z], -1)
    s0 = 1.2 / fx
    scaling = torch.log(s0 * z.abs().clamp_min(0.05)[:, None] * torch.exp(0.7 * N(P, 3)))
    rotation = N(P, 4)
    opacity = 1.5 * N(P, 1)
    Mmax = ((max_sh_degree if max_sh_degree is not None else sh_degree) + 1) ** 2
    f_dc = N(P, 1, 3)
    f_rest = 0.1 * N(P, Mmax - 1, 3)
    w2c = [torch.eye(4) if k == 0 else se3_exp(0.05 * N(6)) for k in range(n_views)]
    c = lambda t: t.contiguous()
    return SimpleNamespace(P=P, W=W, H=H, fx=fx, fy=fy, cx=cx, cy=cy, sh_degree=sh_degree,
                           xyz=c(xyz), scaling=c(scaling), rotation=c(rotation), opacity=c(opacity),
                           f_dc=c(f_dc), f_rest=c(f_rest), w2c=w2c, generator=g)


def perturbed_params(sc, seed=77, rel=0.05):
    """the 'ground-truth' scene of the loop-level workloads: parameters perturbed by 5 % noise
    (SURVEY.md §8d); noise drawn on the CPU, moved to wherever the parameters live."""
    g = torch.Generator().manual_seed(seed)
    n = lambda t: torch.randn(t.shape, generator=g).to(t.device)
    return dict(xyz=sc.xyz + rel * torch.exp(sc.scaling).mean(1, keepdim=True) * n(sc.xyz),
                scaling=sc.scaling + rel * n(sc.scaling), rotation=sc.rotation + rel * n(sc.rotation),
                opacity=sc.opacity + rel * n(sc.opacity), f_dc=sc.f_dc + rel * n(sc.f_dc),
                f_rest=sc.f_rest)


def upstream_grads(W, H, device="cpu"):
    """kernel-level upstream gradients: dL/dcolor ~ N(0,1)/(3N), dL/ddepth ~ N(0,1)/N."""
    N = W * H
    g = torch.Generator().manual_seed(4242)
    gc = torch.randn(3, H, W, generator=g) / (3 * N)
    gd = torch.randn(1, H, W, generator=g) / N
    return gc.to(device), gd.to(device)


def param_checksum(tensors):
    """fp64 (sum, sum of |.|) over a list of parameter tensors: what bench.py prints so that the two
    arms and N = 1/2/4/8 can be compared by the driver."""
    s = a = 0.0
    for t in tensors:
        d = t.detach().double()
        s += float(d.sum())
        a += float(d.abs().sum())
    return s, a
