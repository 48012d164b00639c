"""b2gs_knn3_dist2_static — the no-allocation variant of distCUDA2 (scratch in a static device array
inside libb2gs) — against b2gs_knn3_dist2 with a caller workspace, through the C ABI, on two calls
in a row that share the array (the second one a different, smaller cloud: it must not see the
first one's leftovers).  Integer / comparison work end to end plus one fp32 mean in a fixed order:
the bar is bit-exact, as for the reference extension in tests/test_knn_adam_gpu.py.

Collected last on purpose."""
import pytest
import torch

from monogs_dev_b200._lib import lib, check

pytestmark = pytest.mark.gpu


def _cloud(P, seed):
    g = torch.Generator().manual_seed(seed)
    n_surf = int(0.7 * P)
    c = torch.rand(64, 3, generator=g) * 10
    a = torch.randn(64, 3, generator=g)
    b = torch.randn(64, 3, generator=g)
    k = torch.randint(64, (n_surf,), generator=g)
    uv = torch.rand(n_surf, 2, generator=g) - 0.5
    surf = c[k] + uv[:, :1] * a[k] + uv[:, 1:] * b[k]
    return torch.cat([surf, torch.rand(P - n_surf, 3, generator=g) * 10]).contiguous()


def _with_workspace(pts):
    P = pts.shape[0]
    out = torch.empty(P, device=pts.device)
    ws = torch.empty(lib.b2gs_knn_bytes(P), dtype=torch.uint8, device=pts.device)
    check(lib.b2gs_knn3_dist2(pts.data_ptr(), P, out.data_ptr(), ws.data_ptr(), ws.numel(),
                              torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    return out


def test_static_scratch_knn_matches_workspace_knn_twice():
    cap = lib.b2gs_knn3_dist2_static_max_points()
    assert cap == 262144
    st = torch.cuda.current_stream().cuda_stream
    for P, seed in ((cap, 21), (70_001, 22)):             # full capacity, then a ragged smaller cloud
        pts = _cloud(P, seed).cuda()
        ref = _with_workspace(pts)
        out = torch.full((P,), -1.0, device="cuda")
        before = torch.cuda.memory_allocated()
        check(lib.b2gs_knn3_dist2_static(pts.data_ptr(), P, out.data_ptr(), st))
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated() == before
        assert bool(torch.isfinite(out).all()) and float(out.min()) >= 0
        assert torch.equal(out, ref), float((out - ref).abs().max())


def test_static_scratch_knn_rejects_oversize():
    cap = lib.b2gs_knn3_dist2_static_max_points()
    pts = torch.zeros(8, 3, device="cuda")
    out = torch.zeros(8, device="cuda")
    rc = lib.b2gs_knn3_dist2_static(pts.data_ptr(), cap + 1, out.data_ptr(), None)
    assert rc != 0 and b"exceed" in lib.b2gs_last_error()
    assert lib.b2gs_knn3_dist2_static(None, 0, None, None) == 0
