"""Timings of BASELINE configs 1, 2, 4, 5 — native library vs the reference
extensions built under oracle/_ref, on one GPU.  The headline config (3) is bench.py's job; these
are the parity-test configurations, timed here for the record (profiles/*_secondary_configs.json).

usage: python tools/secondary_configs.py [out.json]
"""
import json, math, os, sys, time
import torch
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from helpers import build_case
from monogs_dev_b200 import rasterizer as rz, mapping, synthetic as syn, distCUDA2
from oracle import ref_ext

DEV = torch.device("cuda", 0)
OUT = {}


def timeit(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / n


def single_view(name, n=20, **kw):
    """fwd and fwd+bwd of one view through the drop-in operator API, both implementations."""
    sc, rs, inp, gc, gd = build_case(name, device=DEV, **kw)
    r = rz.GaussianRasterizer(rs)
    m2 = torch.zeros_like(inp["means3D"])
    leaves = {k: v.clone().requires_grad_(True) for k, v in inp.items()}
    call = lambda src: r(means3D=src["means3D"], means2D=m2, opacities=src["opacities"], shs=src["shs"],
                         scales=src["scales"], rotations=src["rotations"])

    def fwd():
        with torch.no_grad():
            call(inp)

    def fwdbwd():
        c, rad, d, o, nt = call(leaves)
        torch.autograd.backward([c, d], [gc, gd])
        for v in leaves.values():
            v.grad = None
    c = call(leaves)[0]
    row = dict(P=sc.P, W=sc.W, H=sc.H, sh_degree=sc.sh_degree, instances=c.grad_fn.state.num_rendered,
               native_fwd_ms=timeit(fwd, n), native_fwd_bwd_ms=timeit(fwdbwd, n))
    if ref_ext.available():
        kws = dict(shs=inp["shs"], scales=inp["scales"], rotations=inp["rotations"])

        def rfb():
            f = ref_ext.forward(rs, **inp)
            ref_ext.backward(rs, f, gc, gd, inp["means3D"], **kws)
        row["reference_fwd_ms"] = timeit(lambda: ref_ext.forward(rs, **inp), n)
        row["reference_fwd_bwd_ms"] = timeit(rfb, n)
        row["speedup_fwd_bwd"] = row["reference_fwd_bwd_ms"] / row["native_fwd_bwd_ms"]
    OUT[name] = row
    print(name, json.dumps(row), flush=True)
    del leaves, inp
    torch.cuda.empty_cache()


def tracking():
    """config 2: 100 pose-only iterations on a 100k-Gaussian 640x480 frame."""
    sc = syn.make_scene("cfg2_tracking", device=DEV, n_views=1)
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation,
                               sh_degree=sc.sh_degree)
    cam0 = sc.cameras[0]
    true_w2c = syn.se3_exp(torch.tensor([0.02, -0.015, 0.01, 0.004, -0.006, 0.003]))
    cam_true = syn.Camera(sc.W, sc.H, cam0.fx, cam0.fy, cam0.cx, cam0.cy, w2c=true_w2c, device=DEV)
    with torch.no_grad():
        gt = mapping.render_pose_only(cam_true, gm, None, sc.bg)
    row = dict(P=sc.P, W=sc.W, H=sc.H, iters=100)
    for graph in (False, True):
        cam = syn.Camera(sc.W, sc.H, cam0.fx, cam0.fy, cam0.cx, cam0.cy, device=DEV)
        loop = mapping.TrackingLoop(gm, cam, sc.bg, iters=100, cuda_graph=graph)
        best = 1e9
        for rep in range(3):
            loop.load_frame(torch.eye(3), torch.zeros(3), gt["render"], gt["depth"])
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            loop.track()
            torch.cuda.synchronize()
            best = min(best, (time.perf_counter() - t0) * 1e3 / 100)
        row["native_graph_ms_per_iter" if graph else "native_eager_ms_per_iter"] = best
    rz.set_sync_free(False)
    if ref_ext.available():
        # the reference's loop (utils/slam_frontend.py:163-193) as restated in oracle/ref_loop.py: reference
        # rasterizer, torch loss, torch Adam on the 4 camera tensors, torch update_pose, reference camera
        from oracle import ref_loop
        import synthdata
        t = synthdata.scene_tensors("cfg2_tracking", n_views=1)
        rgm = ref_loop.make_model(t, DEV)
        rcam = ref_loop.RefCamera(1, t.W, t.H, t.fx, t.fy, t.cx, t.cy, device=DEV)
        rloop = ref_loop.RefTrackingLoop(rgm, rcam, sc.bg, iters=100)
        try:
            best = 1e9
            for rep in range(2):
                rloop.load_frame(torch.eye(3, device=DEV), torch.zeros(3, device=DEV), gt["render"], gt["depth"])
                torch.cuda.synchronize()
                t0 = time.perf_counter()
                rloop.track()
                torch.cuda.synchronize()
                best = min(best, (time.perf_counter() - t0) * 1e3 / 100)
            row["reference_ms_per_iter"] = best
            row["speedup"] = row["reference_ms_per_iter"] / row["native_graph_ms_per_iter"]
        except Exception as e:       # the reference-side loop is a convenience, never a blocker
            row["reference_error"] = repr(e)[:200]
    OUT["cfg2_tracking_loop"] = row
    print("cfg2_tracking_loop", json.dumps(row), flush=True)


def init_densify():
    """config 4: distCUDA2 on a 2M-point cloud, one densify_and_prune on a 2M-Gaussian model."""
    P = 2_000_000
    g = torch.Generator().manual_seed(1238)
    n_pl = int(0.7 * P)
    c = torch.rand(200, 3, generator=g) * 10
    u = torch.nn.functional.normalize(torch.randn(200, 3, generator=g), dim=1)
    v = torch.nn.functional.normalize(torch.cross(u, torch.randn(200, 3, generator=g), dim=1), dim=1)
    k = torch.randint(0, 200, (n_pl,), generator=g)
    ab = torch.rand(n_pl, 2, generator=g) * 2 - 1
    pts = torch.cat([c[k] + ab[:, :1] * u[k] + ab[:, 1:] * v[k], torch.rand(P - n_pl, 3, generator=g) * 10]).to(DEV)
    row = dict(P=P, native_distCUDA2_ms=timeit(lambda: distCUDA2(pts), 5, 2))
    if ref_ext.knn() is not None:
        row["reference_distCUDA2_ms"] = timeit(lambda: ref_ext.knn().distCUDA2(pts), 3, 1)
        row["speedup_distCUDA2"] = row["reference_distCUDA2_ms"] / row["native_distCUDA2_ms"]
    r = lambda *s: torch.randn(*s, generator=g).to(DEV)
    times = []
    for rep in range(3):
        gm = mapping.GaussianModel(r(P, 3), r(P, 1, 3), torch.zeros(P, 0, 3, device=DEV), 1.5 * r(P, 1),
                                   torch.log(0.06 * torch.exp(0.9 * r(P, 3))), r(P, 4))
        acc, den, _ = gm._stats()
        acc.copy_(4e-4 * torch.rand(P, 1, generator=g).to(DEV))
        den.fill_(1.0)
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        gm.densify_and_prune(2e-4, 0.7, 6.0, 20)
        torch.cuda.synchronize()
        times.append((time.perf_counter() - t0) * 1e3)
        row["P_after"] = int(gm._xyz.shape[0])
        del gm
    row["native_densify_and_prune_ms"] = min(times)
    OUT["cfg4_init_densify"] = row
    print("cfg4_init_densify", json.dumps(row), flush=True)


def seeding_step():
    """keyframe insertion at the Replica shape: back-projection + down-sample + distCUDA2 scales +
    append to a 500k-Gaussian arena (the reference goes through the host and open3d: not timed here)."""
    sc = syn.make_scene("cfg3_replica", device=DEV, n_views=1)
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=0)
    cam = sc.cameras[0]
    g = torch.Generator().manual_seed(3)
    cam.original_image = torch.rand(3, sc.H, sc.W, generator=g).to(DEV)
    cam.depth = (torch.rand(sc.H, sc.W, generator=g) * 5 + 0.5).numpy()
    from monogs_dev_b200 import seeding
    depth_dev = torch.as_tensor(cam.depth).to(DEV)
    row = dict(W=sc.W, H=sc.H,
               backproject_kernel_ms=timeit(lambda: seeding.backproject_rgbd(cam.original_image, depth_dev, cam.fx, cam.fy,
                                                                             cam.cx, cam.cy, seeding._w2c(cam)), 20, 3))
    ts = []
    for rep in range(4):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        pcd = gm.create_pcd_from_image(cam, init=False)
        torch.cuda.synchronize(); t1 = time.perf_counter()
        n = gm.extend_from_pcd(*pcd, kf_id=rep)
        torch.cuda.synchronize(); t2 = time.perf_counter()
        ts.append(((t1 - t0) * 1e3, (t2 - t1) * 1e3, n))
    row["create_pcd_from_image_ms"], row["extend_from_pcd_ms"], row["seeded"] = min(t[0] for t in ts[1:]), min(t[1] for t in ts[1:]), ts[-1][2]
    OUT["cfg3_keyframe_seeding"] = row
    print("cfg3_keyframe_seeding", json.dumps(row), flush=True)


if __name__ == "__main__":
    out = sys.argv[1] if len(sys.argv) > 1 and not sys.argv[1].startswith("--") else os.path.join(ROOT, "gpurun_out", "secondary_configs.json")
    steps = (seeding_step,) if "--seeding-only" in sys.argv else None
    for step in steps or (lambda: single_view("cfg1_plumbing", 50), lambda: single_view("cfg2_tracking", 50),
                 lambda: single_view("cfg3_replica", 20), tracking, init_densify,
                 lambda: single_view("cfg5_large", 5), seeding_step):
        try:
            step()
        except Exception as e:
            print("step failed:", repr(e)[:300], flush=True)
        with open(out, "w") as f:
            json.dump(OUT, f, indent=1)
