"""GaussianModel bookkeeping around the mapping loop (monogs_dev_b200/model_io.py): learning-rate
schedule, opacity resets on the packed arena, SH degree, PLY export / import.  Host logic only —
runs on CPU tensors; each expectation restates the reference line it pins."""
import math
import os
import sys
import types

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from monogs_dev_b200 import mapping, model_io  # noqa: E402


def make(P=257, sh_degree=0, seed=5):
    g = torch.Generator().manual_seed(seed)
    r = lambda *s: torch.randn(*s, generator=g)
    M = (sh_degree + 1) ** 2
    gm = mapping.GaussianModel(r(P, 3), r(P, 1, 3), r(P, M - 1, 3), 1.5 * r(P, 1), 0.3 * r(P, 3), r(P, 4),
                               sh_degree=sh_degree)
    gm.flat_m.copy_(torch.randn(gm.flat_m.shape, generator=g))
    gm.flat_v.copy_(torch.rand(gm.flat_v.shape, generator=g))
    for k in mapping._GROUPS:
        gm.optimizer.state[gm._params[k]]["step"] = 11
    return gm, g


def test_lr_schedule_matches_reference_formula():
    # gaussian_splatting/utils/general_utils.py:79-94 with numpy, restated literally
    def ref(step, lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
        if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
            return 0.0
        if lr_delay_steps > 0:
            delay = lr_delay_mult + (1 - lr_delay_mult) * np.sin(0.5 * np.pi * np.clip(step / lr_delay_steps, 0, 1))
        else:
            delay = 1.0
        t = np.clip(step / max_steps, 0, 1)
        return delay * np.exp(np.log(lr_init) * (1 - t) + np.log(lr_final) * t)
    for step in (-1, 0, 1, 150, 29999, 30000, 10 ** 6):
        for kw in (dict(), dict(lr_delay_steps=100, lr_delay_mult=0.01)):
            a = model_io.helper(step, 9.6e-4, 9.6e-6, max_steps=30000, **kw)
            b = ref(step, 9.6e-4, 9.6e-6, max_steps=30000, **kw)
            assert math.isclose(a, float(b), rel_tol=1e-12, abs_tol=0.0), (step, kw)
    assert model_io.helper(5, 0.0, 0.0) == 0.0
    assert model_io.get_expon_lr_func(1.0, 0.1) is model_io.helper


def test_training_setup_and_update_learning_rate():
    gm, _ = make()
    gm.init_lr(6.0)
    opt = types.SimpleNamespace(position_lr_init=0.00016, position_lr_final=0.0000016, position_lr_delay_mult=0.01,
                                position_lr_max_steps=30000, feature_lr=0.0025, opacity_lr=0.05,
                                scaling_lr=0.001, rotation_lr=0.001, percent_dense=0.02)
    gm.training_setup(opt)
    lr = {g["name"]: g["lr"] for g in gm.optimizer.param_groups}
    assert lr == dict(xyz=0.00016 * 6.0, f_dc=0.0025, f_rest=0.0025 / 20.0, opacity=0.05,
                      scaling=0.001 * 6.0, rotation=0.001)
    assert gm.percent_dense == 0.02 and gm.xyz_gradient_accum.shape == (257, 1)
    got = gm.update_learning_rate(1500)
    want = math.exp(math.log(0.00096) * 0.95 + math.log(0.0000096) * 0.05)
    assert math.isclose(got, want, rel_tol=1e-12)
    lr2 = {g["name"]: g["lr"] for g in gm.optimizer.param_groups}
    assert lr2["xyz"] == got and all(lr2[k] == lr[k] for k in lr if k != "xyz")     # xyz only (:299-310)
    # the rates survive a re-allocation of the arena (densify / prune)
    gm.prune_points(torch.zeros(257, dtype=torch.bool))
    assert {g["name"]: g["lr"] for g in gm.optimizer.param_groups} == lr2


def test_reset_opacity_in_place():
    gm, _ = make()
    before = {k: gm._params[k].detach().clone() for k in mapping._GROUPS}
    m_before = gm.flat_m.clone()
    ptr = gm._opacity.data_ptr()
    gm.reset_opacity()
    want = math.log(0.01 / 0.99)                                        # inverse_sigmoid(0.01), :363
    assert torch.allclose(gm._opacity.detach(), torch.full((257, 1), want), rtol=1e-6, atol=0)
    assert torch.allclose(gm.get_opacity, torch.full((257, 1), 0.01), rtol=1e-5)
    assert gm._opacity.data_ptr() == ptr                                # arena slice, not a new tensor
    st = gm.optimizer.state[gm._opacity]
    assert float(st["exp_avg"].abs().sum()) == 0.0 and float(st["exp_avg_sq"].abs().sum()) == 0.0
    assert st["step"] == 11                                             # stored_state keeps "step" (:472-478)
    for k in mapping._GROUPS:
        if k != "opacity":
            assert torch.equal(gm._params[k].detach(), before[k])
            o = (gm._params[k].data_ptr() - gm.flat.data_ptr()) // 4
            assert torch.equal(gm.flat_m[o:o + before[k].numel()], m_before[o:o + before[k].numel()])


def test_reset_opacity_nonvisible_keeps_reference_quirk():
    gm, g = make()
    act = gm.get_opacity.clone()
    f1 = torch.rand(257, generator=g) < 0.3
    f2 = torch.rand(257, generator=g) < 0.3
    gm.reset_opacity_nonvisible([f1, f2])
    seen = f1 | f2
    new = gm._opacity.detach()
    # unseen -> inverse_sigmoid(0.4); seen -> the ACTIVATED opacity written as the raw value (:370-373)
    assert torch.allclose(new[~seen], torch.full_like(new[~seen], math.log(0.4 / 0.6)), rtol=1e-6)
    assert torch.equal(new[seen], act[seen])
    assert float(gm.optimizer.state[gm._opacity]["exp_avg_sq"].abs().sum()) == 0.0


def test_oneup_sh_degree():
    gm, _ = make(P=9, sh_degree=2)
    gm.active_sh_degree = 0
    for want in (1, 2, 2):
        gm.oneupSHdegree()
        assert gm.active_sh_degree == want


def test_ply_round_trip_and_layout(tmp_path):
    for deg in (0, 2):
        gm, _ = make(P=101, sh_degree=deg)
        path = str(tmp_path / "sub" / ("pc%d.ply" % deg))
        gm.save_ply(path)
        M = (deg + 1) ** 2
        names = gm.construct_list_of_attributes()
        assert names[:9] == ["x", "y", "z", "nx", "ny", "nz", "f_dc_0", "f_dc_1", "f_dc_2"]
        assert len(names) == 6 + 3 + 3 * (M - 1) + 1 + 3 + 4 and names[-4:] == ["rot_0", "rot_1", "rot_2", "rot_3"]
        raw = open(path, "rb").read()
        head, body = raw.split(b"end_header\n", 1)
        lines = head.decode().split("\n")
        assert lines[0] == "ply" and lines[1] == "format binary_little_endian 1.0"
        assert lines[2] == "element vertex 101" and lines[3] == "property float x"
        assert len(body) == 101 * len(names) * 4
        v = model_io.read_ply_vertices(path)
        assert v.dtype.names == tuple(names) and v.shape == (101,)
        assert float(np.abs(v["nx"]).sum() + np.abs(v["ny"]).sum() + np.abs(v["nz"]).sum()) == 0.0
        # SH storage is channel-major: f_rest_j = _features_rest[:, j % (M-1), j // (M-1)]  (:331-346)
        if M > 1:
            j = M                                                        # channel 1, coefficient 1
            assert np.array_equal(v["f_rest_%d" % j], gm._features_rest[:, j % (M - 1), j // (M - 1)].detach().numpy())
        assert np.array_equal(v["f_dc_2"], gm._features_dc[:, 0, 2].detach().numpy())
        want = {k: gm._params[k].detach().clone() for k in mapping._GROUPS}
        other, _ = make(P=7, sh_degree=deg, seed=99)
        other.active_sh_degree = 0
        other.load_ply(path)
        for k in mapping._GROUPS:
            assert torch.equal(other._params[k].detach(), want[k]), k
            st = other.optimizer.state[other._params[k]]
            assert st["step"] == 0 and float(st["exp_avg"].abs().sum()) == 0.0
            assert other._params[k].grad.shape == want[k].shape
        assert other.active_sh_degree == deg and other.max_radii2D.shape == (101,)
        assert other.unique_kfIDs.shape == (101,) and other.n_obs.shape == (101,)


def test_ply_reader_ascii_and_errors(tmp_path):
    p = tmp_path / "a.ply"
    p.write_text("ply\nformat ascii 1.0\ncomment made by hand\nelement vertex 2\nproperty float x\n"
                 "property double y\nproperty uchar z\nelement face 0\nproperty list uchar int vertex_indices\n"
                 "end_header\n1.5 2.5 3\n-1 0.25 7\n")
    v = model_io.read_ply_vertices(str(p))
    assert v["x"].tolist() == [1.5, -1.0] and v["y"].tolist() == [2.5, 0.25] and v["z"].tolist() == [3, 7]
    bad = tmp_path / "b.ply"
    bad.write_bytes(b"plx\n")
    for path, exc in ((bad, ValueError),):
        try:
            model_io.read_ply_vertices(str(path))
            raise AssertionError("no error")
        except exc:
            pass
    gm, _ = make(P=5, sh_degree=0)
    q = str(tmp_path / "c.ply")
    gm.save_ply(q)
    wrong, _ = make(P=5, sh_degree=1)
    try:
        wrong.load_ply(q)
        raise AssertionError("no error")
    except ValueError:
        pass
    open(q, "r+b").truncate(os.path.getsize(q) - 3)
    try:
        model_io.read_ply_vertices(q)
        raise AssertionError("no error")
    except ValueError:
        pass


def test_eval_sh_and_covariance_match_oracle():
    """host-side SH colour (render() under pipe.convert_SHs_python) and get_covariance (under
    pipe.compute_cov3D_python) against the oracle's restatement of the kernels (CR/forward.cu:22-73,
    120-154): same basis, same covariance once the quaternion is normalised."""
    from monogs_dev_b200 import sh_utils
    from oracle import gs_oracle as O
    g = torch.Generator().manual_seed(8)
    d = torch.randn(50, 3, generator=g)
    d = d / d.norm(dim=1, keepdim=True)
    for deg in range(4):
        M = (deg + 1) ** 2
        assert torch.allclose(sh_utils.sh_basis(deg, d), O._sh_basis(deg, d)[:, :M], atol=1e-7)
        sh = torch.randn(50, 16, 3, generator=g)                      # [P, M, 3] as the model stores it
        want = (O._sh_basis(deg, d)[:, :M, None] * sh[:, :M]).sum(1)
        assert torch.allclose(sh_utils.eval_sh(deg, sh.transpose(1, 2), d), want, atol=1e-6)
    rgb = torch.rand(5, 3)
    assert torch.allclose(sh_utils.SH2RGB(sh_utils.RGB2SH(rgb)), rgb, atol=1e-7)
    try:
        sh_utils.eval_sh(2, torch.zeros(4, 3, 4), d[:4])
        raise AssertionError("no error")
    except ValueError:
        pass
    gm, _ = make(P=40)
    cov = gm.get_covariance(1.5)
    q = torch.nn.functional.normalize(gm._rotation.detach())
    want = O.compute_cov3d(torch.exp(gm._scaling.detach()), 1.5, q)
    assert torch.allclose(cov.detach(), want, rtol=1e-5, atol=1e-7)


def test_render_pipeline_switches(monkeypatch):
    """render() (gaussian_renderer/__init__.py:82-111): by default the rasterizer gets scales /
    rotations / SH; pipe.compute_cov3D_python hands it the [P,6] covariance instead,
    pipe.convert_SHs_python the clamped host-side colours; an isotropic model's single scale is
    repeated.  The rasterizer itself is replaced by a recorder (no CPU rasterizer exists)."""
    from monogs_dev_b200 import synthetic as syn
    seen = {}

    class Recorder:
        def __init__(self, raster_settings):
            seen["sh_degree"] = raster_settings.sh_degree

        def __call__(self, **kw):
            seen.update(kw)
            P = kw["means3D"].shape[0]
            return torch.zeros(3, 2, 2), torch.zeros(P), torch.zeros(1, 2, 2), torch.zeros(1, 2, 2), torch.zeros(P)

    monkeypatch.setattr(mapping, "GaussianRasterizer", Recorder)
    gm, _ = make(P=9, sh_degree=1)
    cam = syn.Camera(8, 6, 10.0, 10.0, 4.0, 3.0)
    pkg = mapping.render(cam, gm, None, torch.zeros(3))
    assert set(pkg) == {"render", "viewspace_points", "visibility_filter", "radii", "depth", "opacity", "n_touched"}
    assert seen["cov3D_precomp"] is None and seen["colors_precomp"] is None and seen["sh_degree"] == 1
    assert torch.equal(seen["scales"], gm.get_scaling) and torch.equal(seen["shs"], gm.get_features)
    assert seen["theta"] is cam.cam_rot_delta and seen["rho"] is cam.cam_trans_delta
    pipe = types.SimpleNamespace(compute_cov3D_python=True, convert_SHs_python=True)
    mapping.render(cam, gm, pipe, torch.zeros(3), scaling_modifier=2.0)
    assert seen["scales"] is None and seen["rotations"] is None
    assert torch.allclose(seen["cov3D_precomp"], gm.get_covariance(2.0))
    assert seen["shs"] is None and seen["colors_precomp"].shape == (9, 3)
    assert float(seen["colors_precomp"].min()) >= 0.0
    iso = mapping.GaussianModel(torch.randn(4, 3), torch.randn(4, 1, 3), torch.randn(4, 0, 3), torch.randn(4, 1),
                                torch.randn(4, 1), torch.randn(4, 4))
    mapping.render(cam, iso, None, torch.zeros(3))
    assert seen["scales"].shape == (4, 3) and torch.equal(seen["scales"][:, 0], seen["scales"][:, 2])


def test_auto_view_streams_is_serial_without_a_gpu_and_bounded():
    """mapping.auto_view_streams: one stream for anything that is not a CUDA arena; the GPU rule
    (8 streams unless one view's working set exceeds L2) is exercised by bench.py's default run."""
    from types import SimpleNamespace
    from monogs_dev_b200 import mapping
    g = SimpleNamespace(flat=torch.zeros(10), get_xyz=torch.zeros(5, 3))
    assert mapping.auto_view_streams(g, 8) == 1
    assert mapping.auto_view_streams(SimpleNamespace(), 8) == 1
