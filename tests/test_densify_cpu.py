"""densify / prune on the packed arena (monogs_dev_b200/densify.py, one gather plan) against the
literal restatement of the reference's four-stage sequence (oracle/densify_oracle.py).  The plan
is torch index arithmetic, so this parity runs on the CPU; the GPU run of the same code at
BASELINE config 4 size is in test_loop_gpu.py."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from monogs_dev_b200 import mapping, densify  # noqa: E402
from oracle import densify_oracle as DO  # noqa: E402


def make(P=4000, seed=3, sh_rest=0, device="cpu"):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    scaling = torch.log(0.06 * torch.exp(0.9 * r(P, 3)))              # around percent_dense * extent = 0.06
    gm = mapping.GaussianModel(r(P, 3).to(device), r(P, 1, 3).to(device), r(P, sh_rest, 3).to(device),
                               (1.5 * r(P, 1)).to(device), scaling.to(device), r(P, 4).to(device))
    acc, den, mr = gm._stats()
    acc.copy_(4e-4 * torch.rand(P, 1, generator=g))
    den.fill_(1.0)
    den[::17] = 0.0                                                    # never-seen points -> NaN -> 0
    mr.copy_(30 * torch.rand(P, generator=g))
    gm.flat_m.copy_(torch.randn(gm.flat_m.shape, generator=g))
    gm.flat_v.copy_(torch.rand(gm.flat_v.shape, generator=g))
    for k in mapping._GROUPS:
        gm.optimizer.state[gm._params[k]]["step"] = 7
    gm.unique_kfIDs = torch.randint(0, 9, (P,), generator=g).int()
    gm.n_obs = torch.randint(0, 5, (P,), generator=g).int()
    return gm, g


def oracle_of(gm):
    st = lambda k, n: gm.optimizer.state[gm._params[k]][n].detach().cpu()
    par = {k: gm._params[k].detach().cpu() for k in mapping._GROUPS}
    m = {k: st(k, "exp_avg") for k in mapping._GROUPS}
    v = {k: st(k, "exp_avg_sq") for k in mapping._GROUPS}
    return DO.Model(par, m, v, gm.xyz_gradient_accum.cpu(), gm.denom.cpu(), gm.max_radii2D.cpu())


def check_equal(gm, om):
    assert gm._xyz.shape[0] == om.p["xyz"].shape[0]
    for k in mapping._GROUPS:
        st = gm.optimizer.state[gm._params[k]]
        assert torch.allclose(gm._params[k].detach().cpu(), om.p[k], rtol=1e-6, atol=1e-6), k
        assert torch.equal(st["exp_avg"].cpu(), om.m[k]) and torch.equal(st["exp_avg_sq"].cpu(), om.v[k]), k
        assert st["step"] == 7
        assert gm._params[k].grad is not None and gm._params[k].grad.shape == gm._params[k].shape


def test_densify_and_prune_matches_reference_sequence():
    for sh_rest in (0, 15):
        gm, g = make(sh_rest=sh_rest)
        om = oracle_of(gm)
        P = gm._xyz.shape[0]
        unit = torch.randn(2 * P, 3, generator=g)                      # more than needed; sliced below
        grads = (om.xyz_gradient_accum / om.denom); grads[grads.isnan()] = 0
        ns = int(((grads.squeeze() >= 2e-4) & (om.scaling.max(1).values > 0.01 * 6.0)).sum())
        unit = unit[: 2 * ns]
        kf = gm.unique_kfIDs.clone()
        want = om.densify_and_prune(2e-4, 0.7, 6.0, 20, unit)
        got = gm.densify_and_prune(2e-4, 0.7, 6.0, 20, samples=unit)
        assert want == got and got[0] > 0 and got[1] > 0 and got[2] > 0
        check_equal(gm, om)
        n = gm._xyz.shape[0]
        assert gm.xyz_gradient_accum.shape == (n, 1) and float(gm.xyz_gradient_accum.abs().sum()) == 0.0
        assert gm.max_radii2D.shape == (n,) and gm.unique_kfIDs.shape == (n,) and gm.n_obs.shape == (n,)
        assert set(gm.unique_kfIDs.tolist()) <= set(kf.tolist())
        assert gm.flat.numel() % 4 == 0 and float(gm.flat_grad.abs().sum()) == 0.0


def test_prune_points_and_quirk_q9():
    gm, g = make(P=500)
    om = oracle_of(gm)
    mask = torch.rand(500, generator=g) < 0.3
    om.prune_points(mask)
    kept = gm.prune_points(mask)
    assert kept == int((~mask).sum())
    check_equal(gm, om)
    assert torch.equal(gm.max_radii2D, om.max_radii2D)
    # q9: max_radii2D is zeroed by densification_postfix before the final prune, so a huge
    # screen-space radius alone never prunes a point
    gm2, _ = make(P=300, seed=9)
    gm2.xyz_gradient_accum.zero_()                                     # nothing to clone / split
    gm2._params["opacity"].data.fill_(5.0)                             # opaque: survives min_opacity
    gm2._params["scaling"].data.fill_(-5.0)                            # small in world space
    gm2.max_radii2D.fill_(1000.0)
    assert gm2.densify_and_prune(2e-4, 0.7, 6.0, 20) == (0, 0, 0) and gm2._xyz.shape[0] == 300
