"""distCUDA2 and the fused Adam against their oracles (brute force / torch.optim.Adam) and,
when built, the reference simple-knn extension (bit-exact)."""
import pytest
import torch

from helpers import assert_grad_close
from monogs_dev_b200 import distCUDA2, FusedAdam
from oracle import gs_oracle as O, ref_ext

pytestmark = pytest.mark.gpu


def _cloud(P, seed):
    g = torch.Generator().manual_seed(seed)
    n_surf = int(0.7 * P)
    c = torch.rand(200, 3, generator=g) * 10
    a = torch.randn(200, 3, generator=g)
    b = torch.randn(200, 3, generator=g)
    k = torch.randint(200, (n_surf,), generator=g)
    uv = torch.rand(n_surf, 2, generator=g) - 0.5
    surf = c[k] + uv[:, :1] * a[k] + uv[:, 1:] * b[k]
    return torch.cat([surf, torch.rand(P - n_surf, 3, generator=g) * 10]).contiguous()


@pytest.mark.parametrize("P", [1, 3, 4, 31, 33, 1000, 20_000])
def test_knn_vs_bruteforce(P):
    pts = _cloud(P, 5) if P >= 100 else torch.rand(P, 3)
    if P >= 1000:
        pts[7] = pts[3]                       # coincident points: self excluded by index -> 0
    out = distCUDA2(pts.cuda()).cpu()
    if P < 4:       # fewer than 3 neighbours: FLT_MAX slots overflow to inf, as in the reference
        assert bool(torch.isinf(out).all()) or bool((out > 1e30).all())
        return
    ref = O.knn3_mean_dist2(pts)
    assert torch.equal(out, ref) or float(((out - ref).abs() / ref.clamp_min(1e-20)).max()) < 2e-6


@pytest.mark.skipif(ref_ext.knn() is None or not torch.cuda.is_available(), reason="oracle/_ref knn not built")
@pytest.mark.parametrize("P", [5000, 300_000])
def test_knn_vs_reference_ext_bit_exact(P):
    pts = _cloud(P, 11).cuda()
    out = distCUDA2(pts)
    ref = ref_ext.knn().distCUDA2(pts)
    torch.cuda.synchronize()
    assert torch.equal(out, ref)


def test_knn_full_size_property():
    pts = _cloud(2_000_000, 1238).cuda()
    out = distCUDA2(pts)
    assert bool(torch.isfinite(out).all()) and float(out.min()) >= 0
    idx = torch.randint(2_000_000, (512,))
    d = ((pts[idx][:, None, :] - pts[None, :, :]) ** 2).sum(-1)
    d[torch.arange(512), idx] = float("inf")
    ref = d.topk(3, dim=1, largest=False).values.sum(1) / 3
    assert float(((out[idx] - ref).abs() / ref.clamp_min(1e-12)).max()) < 1e-4


def test_fused_adam_vs_torch():
    torch.manual_seed(0)
    shapes = [(5000, 3), (5000, 1, 3), (5000, 0, 3), (5000, 1), (5000, 3), (5000, 4), (3,), (1,)]
    lrs = [1.6e-4 * 6, 2.5e-3, 2.5e-3 / 20, 0.05, 0.001 * 6, 0.001, 0.003, 0.01]
    A = [torch.nn.Parameter(torch.randn(s, device="cuda")) for s in shapes]
    B = [torch.nn.Parameter(a.detach().clone()) for a in A]
    oa = FusedAdam([dict(params=[p], lr=lr) for p, lr in zip(A, lrs)], lr=0.0, eps=1e-15)
    ob = torch.optim.Adam([dict(params=[p], lr=lr) for p, lr in zip(B, lrs)], lr=0.0, eps=1e-15)
    for it in range(25):
        for a, b in zip(A, B):
            g = torch.randn_like(a) * (10.0 ** torch.randint(-6, 1, (1,)).item())
            a.grad, b.grad = g.clone(), g.clone()
        if it == 10:
            oa.param_groups[0]["lr"] = ob.param_groups[0]["lr"] = 3e-4   # xyz lr schedule
        oa.step(); ob.step()
    for a, b in zip(A, B):
        if a.numel():
            assert float((a - b).abs().max()) <= 2e-6 * float(b.abs().max()) + 1e-7
            sa, sb = oa.state[a], ob.state[b]
            assert_grad_close(sa["exp_avg"], sb["exp_avg"], tol=1e-5, name="exp_avg")
            assert_grad_close(sa["exp_avg_sq"], sb["exp_avg_sq"], tol=1e-5, name="exp_avg_sq")
            assert int(sa["step"]) == int(sb["step"]) == 25
