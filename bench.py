#!/usr/bin/env python
"""bench.py — the headline metric of BASELINE.json on B200.

    python bench.py --gpus N --steps K --warmup W [--impl reference]

A STEP is one iteration of the keyframe mapping window (utils/slam_backend.py:156-317 of
the reference) at BASELINE config 3 (Replica shape): 1200x680, 500k Gaussians, SH degree 0,
8 keyframes: 8 x (render -> RGB-D mapping loss) -> one backward -> [N>1: the packed per-Gaussian gradients
are summed over ranks — by the sharded peer-memory optimiser kernel, or with --no-peer-adam by ONE
NCCL all-reduce] -> fused Adam over the packed parameter arena ->
per-keyframe pose/exposure Adam + pose update.  Keyframes are sharded over ranks, so the
total work is fixed ("strong" scaling).

  value        = keyframe renders (fwd+bwd) per second, whole job  [iters/s]
  ms_per_step  = mapping-window step time                           [ms]
  e2e          = same, with this step's ground-truth RGB-D frames copied from pinned host
                 memory and the loss read back inside the timed region
  roofline     = dominant kernel of the step against its bound
  cpu_baseline = the pure-PyTorch CPU oracle (oracle/gs_oracle.py) on a bounded 1/10-scale
                 sample, rank 0, N=1 only

`--impl reference` times the UNMODIFIED reference CUDA rasterizer (oracle/_ref, built for
sm_100 by oracle/build_ref.py) + torch.optim.Adam through the reference's own loop restated in
oracle/ref_loop.py (camera properties recomputed per access, torch-op loss, torch update_pose).
That arm never imports `monogs_dev_b200` (no libb2gs.so in its process) and is single-GPU, as the
reference is: under torchrun rank 0 alone runs it on the full window, the other ranks exit 0.
If the extension cannot be loaded it falls back to the CPU oracle port on a bounded sample.

Both arms start from the same seeded scene (synthdata), run the same number of steps and print
`state_check` = {steps_applied, loss_last, param_sum, param_abs_sum (fp64), pose_sum}: the two
arms and N = 1/2/4/8 must agree ("the 1-GPU step on the same window").
"""
import argparse
import json
import os
import subprocess
import sys
import time

import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import synthdata

WORKLOADS = {   # name -> (keyframes, description)
    "cfg3_replica": (8, "cfg3_replica: 1200x680, 500k Gaussians, SH0, 8-keyframe mapping window "
                        "(fwd+bwd per keyframe, fused Adam, pose Adam)"),
    "cfg5_large": (16, "cfg5_large: 1920x1080, 4M Gaussians, SH degree 3, 16-keyframe mapping window "
                       "(fwd+bwd per keyframe, fused Adam, pose Adam)"),
}
METRIC = "fwd+bwd render iters/s @1200x680, 500k Gaussians (8-keyframe map-window step)"
# keyframe optimizer of BackEnd (utils/slam_backend.py:434-471): cam_rot_delta * 0.5, cam_trans_delta * 0.5
# (configs/rgbd/replica/base_config.yaml:59-61), exposure 0.01 — the same in both arms
POSE_LR = (0.0015, 0.0005, 0.01, 0.01)
# executed flop per pixel of the render kernels' inner loops, counted on the source
FWD_POWER, FWD_BLEND = 9.5, 24.0     # power tree | expf (11) + alpha/test_T (4) + 4 FMA blend + weight (9)
BWD_HEAD, BWD_BLEND = 22.5, 39.0     # power + expf + alpha | rcp, colour . dL_dpixel, dL_dalpha, 6 moments + 4 colour sums, 12-shuffle fold
#                                      (67.0 until the kernel folded moments instead of finished conic / mean gradients, 52.0
#                                      until it carried accum . dL_dpixel as one scalar instead of four blended channels)
GRAPH_WARMUP = 4   # native graph mode: 3 eager sizing steps + the capture; the reference arm runs them too


def env_int(k, d):
    return int(os.environ.get(k, d))


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.p = None
        try:
            self.p = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                       "-lms", "100", "-i", str(index)], stdout=subprocess.PIPE,
                                      stderr=subprocess.DEVNULL, text=True)
        except Exception:
            self.p = None

    def stop(self):
        out = {"sm_mhz": None, "sm_max_mhz": None, "reasons": []}
        if self.p is None:
            return out
        time.sleep(0.15)
        self.p.terminate()
        try:
            txt, _ = self.p.communicate(timeout=5)
        except Exception:
            self.p.kill()
            return out
        sm, reasons, mx = [], set(), None
        for line in txt.strip().splitlines():
            f = [x.strip() for x in line.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0]))
                mx = float(f[1])
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[3:7]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        if sm:
            sm.sort()
            out.update(sm_mhz=sm[len(sm) // 2], sm_max_mhz=mx, reasons=sorted(reasons), samples=len(sm))
        return out


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm_gbs=float(d["hbm_gbs"]), sm_max_mhz=float(d.get("sm_max_mhz", 1965.0)), source="measured")
    return dict(hbm_gbs=6650.0, sm_max_mhz=1965.0, source="fallback")


def dist_setup(args):
    world = env_int("WORLD_SIZE", 1)
    rank = env_int("RANK", 0)
    local = env_int("LOCAL_RANK", 0)
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        torch.cuda.set_device(local)
        dist.init_process_group("nccl", rank=rank, world_size=world,
                                device_id=torch.device("cuda", local))
    else:
        torch.cuda.set_device(local)
    return rank, world, local


def barrier_sync(world):
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x, world, device):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(x, world, device):
    if world == 1:
        return x
    import torch.distributed as dist
    t = torch.tensor([x], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())


def timed_steps(step_fn, K, world, settle=None):
    """`settle`: called inside the timed region after the K steps — the mapping window's flush(),
    which confirms that no step was discarded by a binning overflow (and re-runs any that were)."""
    barrier_sync(world)
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(K):
        step_fn()
    if settle is not None:
        settle()
    b.record()
    barrier_sync(world)
    return a.elapsed_time(b)


def build_window(device, sc):
    """native arm: keyframes with ground truth rendered (untimed) from a 5 %-perturbed copy of the scene."""
    from monogs_dev_b200 import mapping
    gt = synthdata.perturbed_params(sc)
    gt_model = mapping.GaussianModel(gt["xyz"], gt["f_dc"], gt["f_rest"], gt["opacity"], gt["scaling"],
                                     gt["rotation"], sh_degree=sc.sh_degree)
    kfs = []
    with torch.no_grad():
        for v, cam in enumerate(sc.cameras):
            pkg = mapping.render_fused(cam, gt_model, None, sc.bg)
            # keyframe 0 has no optimised parameter at all (utils/slam_backend.py:435-436)
            kfs.append(dict(cam=cam, gt_image=pkg["render"].detach().clone(), gt_depth=pkg["depth"].detach().clone(),
                            optimize_pose=v > 0, optimize_exposure=v > 0))
    del gt_model
    torch.cuda.synchronize()
    return kfs


def pinned_host_copies(kfs):
    return [(kf["gt_image"].cpu().pin_memory(), kf["gt_depth"].cpu().pin_memory()) for kf in kfs]


def sum_n_contrib(state, W, H, device):
    """effective (pixel, Gaussian) pairs of one view = sum over pixels of n_contrib."""
    import ctypes as C
    from monogs_dev_b200 import _lib
    sc = _lib.B2gsScene()
    sc.P, sc.W, sc.H = 1, W, H
    ws = _lib.B2gsWorkspace(None, 0, _lib.ptr(state.image), state.image.numel(), None, 0)
    nc = torch.zeros(W * H, dtype=torch.int32, device=device)
    _lib.check(_lib.lib.b2gs_debug_export_image(C.byref(sc), C.byref(ws), None, nc.data_ptr(),
                                                torch.cuda.current_stream().cuda_stream))
    return int(nc.long().sum().item())


def rooflines(prof, units, pk):
    """per-kernel achieved vs bound.  units: dict of per-launch work:
    bytes for the HBM-bound kernels, EXECUTED flops for the FP32-pipe-bound ones."""
    fp32_peak = 148 * 128 * 2 * pk["sm_max_mhz"] * 1e6 / 1e12        # TFLOP/s at max SM clock
    table = {}
    for name, (cnt, ms) in prof.items():
        if cnt == 0 or name not in units:
            continue
        kind, work = units[name]
        avg_s = ms / cnt / 1e3
        if kind == "hbm":
            ach = work / avg_s / 1e9
            table[name] = dict(bound="hbm", achieved=round(ach, 1), peak=pk["hbm_gbs"], unit="GB/s",
                               frac=round(ach / pk["hbm_gbs"], 4), traffic=None, launches=cnt,
                               avg_ms=round(ms / cnt, 5), peak_source=pk["source"])
        else:
            ach = work / avg_s / 1e12
            table[name] = dict(bound="fp32", achieved=round(ach, 3), peak=round(fp32_peak, 1), unit="TFLOP/s",
                               frac=round(ach / fp32_peak, 4), traffic=None, launches=cnt,
                               avg_ms=round(ms / cnt, 5),
                               peak_source="148 SM x 128 lanes x 2 x sm_max_mhz (%s)" % pk["source"])
    return table


def _ncu_file():
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_render_full.txt")))
    return files[-1] if files else None


def ncu_metrics(kernel):
    """issue-slot and FMA-pipe utilisation of `kernel` from the newest committed `ncu --set full`
    summary under profiles/ (context next to the measured `frac`; never a bench value)."""
    import re
    f = _ncu_file()
    if not f:
        return None
    want = {"smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slot_pct",
            "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
            "gpu__time_duration.sum": "ncu_us"}
    out, on = {}, False
    for line in open(f):
        if line.startswith("----"):
            on = kernel + "_kernel" in line
        elif on:
            m = re.match(r"\s+(\S+)\s+([0-9.]+)", line)
            if m and m.group(1) in want:
                out[want[m.group(1)]] = round(float(m.group(2)), 1)
    if out:
        out["source"] = os.path.basename(f)
    return out or None


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the newest
    `ncu --set full` summary committed under profiles/ (tools/ncu_summary.py); None if absent."""
    import re
    f = _ncu_file()
    if not f:
        return None
    vals, cur = [], None
    for line in open(f):
        if line.startswith("----"):
            cur = {} if kernel + "_kernel" in line else None
            if cur is not None:
                vals.append(cur)
        elif cur is not None:
            m = re.match(r"\s+dram__bytes_(read|write)\.sum\s+([0-9.]+)\s+(\w+)", line)
            if m:
                mult = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(m.group(3), 1.0)
                cur[m.group(1)] = float(m.group(2)) * mult
    tot = [v.get("read", 0.0) + v.get("write", 0.0) for v in vals if v]
    return round(sum(tot) / len(tot)) if tot else None


def cpu_oracle_sample():
    """bounded CPU sample: two fwd+bwd of a 1/10-scale config 3 (50k Gaussians, 380x216,
    focal / sqrt(10)) through the PyTorch oracle; scaled to full-size iters/s by 1/10."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from oracle import gs_oracle as O
    s = 10 ** 0.5
    t = synthdata.scene_tensors("cfg3_replica", P=50_000, W=380, H=216, fx=600 / s, fy=600 / s, cx=189.5, cy=107.5)
    from oracle import ref_loop
    cam = ref_loop.RefCamera(0, t.W, t.H, t.fx, t.fy, t.cx, t.cy, device="cpu")
    from types import SimpleNamespace
    import math
    rs = SimpleNamespace(image_height=t.H, image_width=t.W, tanfovx=math.tan(cam.FoVx * 0.5),
                         tanfovy=math.tan(cam.FoVy * 0.5), bg=torch.zeros(3), scale_modifier=1.0,
                         viewmatrix=cam.world_view_transform, projmatrix=cam.full_proj_transform,
                         projmatrix_raw=cam.projection_matrix, sh_degree=t.sh_degree, campos=cam.camera_center,
                         prefiltered=False, debug=False)
    gc, gd = synthdata.upstream_grads(t.W, t.H)
    torch.set_num_threads(os.cpu_count() or 1)
    reps = 2            # about 15 s of CPU work on the box's 16 cores
    t0 = time.time()
    for _ in range(reps):
        sc = O.scene_from_settings(rs, t.xyz, torch.sigmoid(t.opacity), shs=torch.cat([t.f_dc, t.f_rest], 1).contiguous(),
                                   scales=torch.exp(t.scaling), rotations=torch.nn.functional.normalize(t.rotation))
        st = O.rasterize_forward(sc)
        O.rasterize_backward(sc, st, gc, gd)
    dt = (time.time() - t0) / reps
    return dict(value=round(1.0 / dt / 10.0, 5), unit="iters/s", cores=torch.get_num_threads(), kind="port",
                sample="%d fwd+bwd of a 1/10-scale cfg3 view (50k Gaussians, 380x216), %.1f s each, on the "
                       "pure-PyTorch oracle; value = measured rate / 10" % (reps, dt))


class E2E:
    """e2e harness shared by both arms (imports neither): every step's ground-truth RGB-D frames
    come from pinned host memory.  The H2D copy of step k+1 runs on a copy stream while step k
    computes (double buffering through a staging set; one D2D hand-over per step), and the loss is
    read back to the host every step, one step late."""

    def __init__(self, win, kfs, host, local_views):
        self.win, self.kfs, self.host, self.local = win, kfs, host, local_views
        self.copy_stream = torch.cuda.Stream()
        self.stage = {v: (torch.empty_like(kfs[v]["gt_image"]), torch.empty_like(kfs[v]["gt_depth"]))
                      for v in local_views}
        self.ev_ready, self.ev_free = torch.cuda.Event(), torch.cuda.Event()
        self.pending = False
        self.loss_host = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
        self.loss_ev = [torch.cuda.Event(), torch.cuda.Event()]
        self.n, self.losses = 0, []

    def prefetch(self):
        self.copy_stream.wait_event(self.ev_free)
        with torch.cuda.stream(self.copy_stream):
            for v in self.local:
                self.stage[v][0].copy_(self.host[v][0], non_blocking=True)
                self.stage[v][1].copy_(self.host[v][1], non_blocking=True)
            self.ev_ready.record(self.copy_stream)
        self.pending = True

    def step(self):
        cur = torch.cuda.current_stream()
        if not self.pending:
            self.ev_free.record(cur)
            self.prefetch()
        cur.wait_event(self.ev_ready)
        for v in self.local:                      # staging -> the buffers the (captured) step reads
            self.kfs[v]["gt_image"].copy_(self.stage[v][0], non_blocking=True)
            self.kfs[v]["gt_depth"].copy_(self.stage[v][1], non_blocking=True)
        self.ev_free.record(cur)
        self.prefetch()                           # next step's inputs, overlapped with this step
        loss = self.win.step()
        # result read back, device -> host, every step: an async copy into pinned memory that the
        # host reads after it has enqueued the NEXT step (settle() reads the last one)
        i = self.n % 2
        self.loss_host[i].copy_(loss.detach().reshape(1), non_blocking=True)
        self.loss_ev[i].record(cur)
        if self.n > 0:
            self.loss_ev[1 - i].synchronize()
            self.losses.append(float(self.loss_host[1 - i][0]))
        self.n += 1

    def settle(self):
        self.win.flush()
        if self.n > 0:
            i = (self.n - 1) % 2
            self.loss_ev[i].synchronize()
            self.losses.append(float(self.loss_host[i][0]))

    def reset(self):
        torch.cuda.synchronize()
        self.pending = False                      # the timed region issues its own first copy
        self.n = 0
        del self.losses[:]


def state_check(steps, loss, params, cams_local, world, device):
    """what the driver compares between the arms and across N: the loss of the last timed step and
    fp64 checksums of the Gaussian parameters / keyframe poses+exposures after `steps` steps."""
    ps, pa = synthdata.param_checksum(params)
    pose = 0.0
    for cam in cams_local:
        pose += float(cam.T.detach().double().abs().sum()) + float(cam.exposure_a.detach().double().sum()) \
            + float(cam.exposure_b.detach().double().sum())
    return dict(steps_applied=steps, loss_last=round(sum_over_ranks(float(loss), world, device), 8),
                param_sum=round(ps, 6), param_abs_sum=round(pa, 6),
                pose_sum=round(sum_over_ranks(pose, world, device), 8))


def config_common(name, n_views):
    return {"workload": WORKLOADS[name][1], "keyframes": n_views, "pose_lr": list(POSE_LR),
            "seed": synthdata.CONFIGS[name][8], "gt": "render of the 5 %-perturbed scene (seed 77)",
            "l2_policy": "inputs larger than L2: >400 MB touched per step (GT frames, parameter/moment "
                         "arenas, per-view workspaces+gradients) vs 126 MB L2"}


def run_native(args):
    rank, world, local = dist_setup(args)
    device = torch.device("cuda", local)
    from monogs_dev_b200 import synthetic as syn, mapping, _lib
    name = args.workload
    n_views = WORKLOADS[name][0]
    K, W_ = args.steps, max(args.warmup, 3)
    graph = not args.no_graph

    sc = syn.make_scene(name, device=device, n_views=n_views)
    kfs = build_window(device, sc)
    gm = mapping.GaussianModel(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation,
                               sh_degree=sc.sh_degree)
    win = mapping.MappingWindow(gm, kfs, sc.bg, pose_lr=POSE_LR, cuda_graph=graph, view_streams=(args.view_streams or None),
                                peer_adam=(args.peer_mode if args.peer_mode == "nomc" else True)
                                if (world > 1 and not args.no_peer_adam) else False)
    local_views = win.local
    host = pinned_host_copies(kfs)
    e2e = E2E(win, kfs, host, local_views)

    loss = None
    for _ in range(W_ + GRAPH_WARMUP):            # graph mode: 3 eager sizing steps + the capture
        loss = win.step()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = _lib.lib.b2gs_kernel_launches()
    last = [loss]

    def step():
        last[0] = win.step()
    ms = timed_steps(step, K, world, settle=win.flush)
    launches = _lib.lib.b2gs_kernel_launches() - l0
    if graph:      # replays do not pass through the host-side counter: launches captured per step x K
        launches = win.graph_launches * K
    clocks = sampler.stop() if sampler else {}
    ms = max_over_ranks(ms, world, device)
    launches = int(sum_over_ranks(launches, world, device))
    check = state_check(W_ + GRAPH_WARMUP + K, last[0], [gm.flat], [kfs[v]["cam"] for v in local_views],
                        world, device)
    if getattr(win, "peer", None) is not None:
        win.peer.check()                          # a barrier time-out would have corrupted the replicas

    e2e.step()
    e2e.settle()
    e2e.reset()
    ms_e2e = max_over_ranks(timed_steps(e2e.step, K, world, settle=e2e.settle), world, device)
    assert len(e2e.losses) == K, "every step's loss must have been read back inside the timed region"

    # headline companion: single-view fwd+bwd only (no optimizer), rank-local, N=1 only
    single = None
    if world == 1:
        def one_view():
            pkg = mapping.render_fused(kfs[0]["cam"], gm, None, sc.bg)
            mapping.get_loss_mapping_rgbd(pkg["render"], pkg["depth"], kfs[0]["cam"], kfs[0]["gt_image"],
                                          kfs[0]["gt_depth"]).backward()
            gm.zero_grad()
            for t in (kfs[0]["cam"].cam_rot_delta, kfs[0]["cam"].cam_trans_delta,
                      kfs[0]["cam"].exposure_a, kfs[0]["cam"].exposure_b):
                t.grad = None
        for _ in range(3):
            one_view()
        n1 = max(20, K * n_views // 2)
        single = n1 / (timed_steps(one_view, n1, 1) / 1e3)

    roof, rooftab, work = native_rooflines(win, gm, sc, kfs, local_views, world, device)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_oracle_sample()

    if rank != 0:
        return finish(world, win)
    h2d = sum(h[0].numel() * 4 + h[1].numel() * 4 for h in host)
    line = {
        "metric": METRIC, "value": round(n_views * K / (ms / 1e3), 3), "unit": "iters/s", "n_gpus": world,
        "steps": K, "warmup": W_, "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {**config_common(name, n_views),
                   "parallelism": "keyframes sharded round-robin over %d rank(s)" % world + (
                       "; gradient exchange + Adam + zero_grad as one kernel over NVLink peer memory" + (
                           " (NVLS multicast: multimem.ld_reduce / multimem.st)" if win.peer.multicast else " (plain P2P)")
                       if getattr(win, "peer", None) is not None else ""),
                   "launch": ("cuda graphs (no host sync inside a step: overflow flags are read back "
                              "asynchronously, flush() inside the timed region)" if graph else "eager") +
                             ("" if win.view_streams < 2 else
                              ", views issued on %d forked streams" % win.view_streams),
                   **work},
        "clocks": clocks,
        "state_check": check,
        "e2e": {"value": round(n_views * K / (ms_e2e / 1e3), 3), "unit": "iters/s",
                "ms_per_step": round(ms_e2e / K, 4), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4 * world,
                "readback": "every step's loss copied to pinned host memory and read by the host one step "
                            "later (after the next step is enqueued); the last one before the region ends"},
        "gpu_launches": launches,
        "single_view_fwd_bwd_iters_s": round(single, 2) if single else None,
        "roofline": roof, "rooflines_all": rooftab, "cpu_baseline": cpu,
    }
    print(json.dumps(line), flush=True)
    finish(world, win)


def native_rooflines(win, gm, sc, kfs, local_views, world, device):
    """per-kernel durations: CUDA events on the launching stream, 2 extra (untimed) steps with the
    views serialised on one stream (a kernel timed while another view's kernels share the SMs would
    not be that kernel's duration).  Work units are read from THIS run: visible Gaussians, the true
    instance count (status[0], not the workspace capacity), evaluated / contributing pixel-splat
    pairs counted by the render kernels' debug counters."""
    from monogs_dev_b200 import _lib, knn
    _lib.lib.b2gs_profile_enable(1)
    win.profile_steps(2)
    knn_evals = 0
    if world == 1:       # distCUDA2 is not part of a mapping step: one call on the scene's points, for its line
        cnt = torch.zeros(8, dtype=torch.int64, device=device)
        _lib.lib.b2gs_debug_set_counters(cnt.data_ptr())
        knn.distCUDA2(gm._xyz.detach())
        torch.cuda.synchronize()
        _lib.lib.b2gs_debug_set_counters(None)
        knn_evals = int(cnt[7])
    torch.cuda.synchronize()
    prof = _lib.profile_read()
    _lib.lib.b2gs_profile_enable(0)
    info = win.work_counters(local_views[0])
    P, N = sc.P, sc.W * sc.H
    pvis, R = info["visible"], info["instances"]
    M = (sc.sh_degree + 1) ** 2
    n_par = gm.flat.numel()
    tiles = ((sc.W + 15) // 16) * ((sc.H + 15) // 16)
    # EXECUTED FP32 work of the render kernels (FMA = 2 flop, a packed fp32x2 op = its two
    # lanes, MUFU = 1).  A warp visit makes all 64 pixels of the quadrant evaluate `power` (FWD_POWER /
    # BWD_HEAD flop per pixel); a visit that survives the warp vote runs exp + alpha + blend (forward) or
    # the gradient arithmetic + 12-shuffle fold (backward) on every pixel, contributing or not.
    fwd_flop = FWD_POWER * info["fwd_visit_px"] + FWD_BLEND * info["fwd_eval_px"]
    bwd_flop = BWD_HEAD * info["bwd_visit_px"] + BWD_BLEND * info["bwd_blend_px"]
    units = {   # algorithmic work per launch
        "preprocess_fwd": ("hbm", (119.0 + 12.0 * M) * pvis + 20.0 * (P - pvis)),
        "scatter_instances": ("hbm", 12.0 * P + 40.0 * pvis + 8.0 * R),
        "tile_sort": ("hbm", 24.0 * R),               # the sort's lower bound: one read + one write of a 12 B pair
        "render_fwd": ("fp32", fwd_flop),
        "render_bwd": ("fp32", bwd_flop),
        "preprocess_bwd": ("hbm", (283.0 + 36.0 * M) * pvis + 60.0 * (P - pvis)),
        "adam": ("hbm", 32.0 * n_par),                # read p,g,m,v; write p,m,v and the cleared g
        "adam_peer": ("hbm", 28.0 * n_par / max(world, 1)),
        "knn_search": ("fp32", 8.0 * knn_evals),      # 3 sub, 3 mul/fma, compare-exchange chain ~ 8 flop / evaluation
    }
    rooftab = rooflines(prof, units, peaks())
    for k in ("render_fwd", "render_bwd"):
        if k in rooftab:
            rooftab[k]["ncu"] = ncu_metrics(k)
    roof = None
    own = {k: v for k, v in prof.items() if k in rooftab}
    if own:
        top = max(own, key=lambda k: own[k][1])
        roof = dict(kernel=top, **rooftab[top])
        roof["traffic"] = ncu_traffic(top)
        tot = sum(v[1] for v in prof.values())
        roof["share_of_gpu_kernel_time"] = round(prof[top][1] / tot, 3) if tot else None
        roof["unit_definition"] = (
            "EXECUTED flops from the kernels' debug counters in this run: forward %.1f/visited pixel + %.1f/"
            "evaluated pixel; backward %.1f/visited pixel + %.1f/blended pixel (a visit = 64 pixels x 1 splat). "
            "useful_frac = the same with only contributing pairs charged" % (FWD_POWER, FWD_BLEND, BWD_HEAD, BWD_BLEND))
        if top in ("render_fwd", "render_bwd"):
            useful = (FWD_POWER + FWD_BLEND) * info["fwd_contributing"] if top == "render_fwd" else \
                (BWD_HEAD + BWD_BLEND) * info["bwd_contributing"]
            roof["useful_frac"] = round(useful / (roof["avg_ms"] * 1e-3) / 1e12 / roof["peak"], 4)
    work = dict(P=P, P_visible=pvis, instances=R, effective_pairs=info["effective_pairs"],
                **{k: info[k] for k in ("fwd_visit_px", "fwd_eval_px", "fwd_contributing", "fwd_block_px",
                                        "bwd_visit_px", "bwd_blend_px", "bwd_contributing")})
    return roof, rooftab, work


def run_reference(args):
    """the clean reference arm: oracle/_ref (unmodified reference CUDA extension) + oracle/ref_loop.py;
    `monogs_dev_b200` is never imported in this process."""
    rank, local = env_int("RANK", 0), env_int("LOCAL_RANK", 0)
    world = env_int("WORLD_SIZE", 1)
    if rank != 0:
        return                                    # the reference is single-GPU: rank 0 runs the whole window
    from oracle import ref_ext, ref_loop
    if not ref_ext.available():
        return run_reference_cpu(args, rank, world)
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    name = args.workload
    n_views = WORKLOADS[name][0]
    K, W_ = args.steps, max(args.warmup, 3)
    t = synthdata.scene_tensors(name, n_views=n_views)
    bg = torch.zeros(3, device=device)
    cams = ref_loop.make_cameras(t, device)
    kfs = ref_loop.build_window_gt(t, cams, bg, device)
    gm = ref_loop.make_model(t, device)
    win = ref_loop.RefMappingWindow(gm, kfs, bg, pose_lr=POSE_LR)
    host = pinned_host_copies(kfs)
    e2e = E2E(win, kfs, host, list(range(n_views)))
    loss = None
    for _ in range(W_ + GRAPH_WARMUP):
        loss = win.step()
    sampler = ClockSampler(local)
    last = [loss]

    def step():
        last[0] = win.step()
    ms = timed_steps(step, K, 1)
    clocks = sampler.stop()
    check = state_check(W_ + GRAPH_WARMUP + K, last[0], gm.parameters(), cams, 1, device)
    e2e.step()
    e2e.settle()
    e2e.reset()
    ms_e2e = timed_steps(e2e.step, K, 1, settle=e2e.settle)

    def one_view():
        pkg = ref_ext.ref_render(cams[0], gm, None, bg)
        ref_loop.get_loss_mapping_rgbd(pkg["render"], pkg["depth"], cams[0], kfs[0]["gt_image"],
                                       kfs[0]["gt_depth"]).backward()
        gm.zero_grad()
        for p in (cams[0].cam_rot_delta, cams[0].cam_trans_delta, cams[0].exposure_a, cams[0].exposure_b):
            p.grad = None
    for _ in range(3):
        one_view()
    n1 = max(20, K * n_views // 2)
    single = n1 / (timed_steps(one_view, n1, 1) / 1e3)

    value = round(n_views * K / (ms / 1e3), 3)
    h2d = sum(h[0].numel() * 4 + h[1].numel() * 4 for h in host)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": "iters/s", "n_gpus": world, "steps": K,
        "warmup": W_, "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {**config_common(name, n_views),
                   "parallelism": "single GPU (the reference has no multi-GPU path): rank 0 runs the whole window",
                   "launch": "eager, the reference's own loop (oracle/ref_loop.py restates utils/slam_backend.py:156-317)"},
        "clocks": clocks,
        "state_check": check,
        "e2e": {"value": round(n_views * K / (ms_e2e / 1e3), 3), "unit": "iters/s",
                "ms_per_step": round(ms_e2e / K, 4), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4},
        "gpu_launches": None,
        "single_view_fwd_bwd_iters_s": round(single, 2),
        "cpu_baseline": {"value": value, "unit": "iters/s", "cores": 0, "kind": "reference",
                         "sample": "UNMODIFIED reference CUDA rasterizer (oracle/_ref, sm_100 build) + "
                                   "torch.optim.Adam on ONE GPU: full workload, not a CPU run "
                                   "(the reference has no CPU rasterizer)"},
    }
    print(json.dumps(line), flush=True)


def finish(world, win):
    """multi-rank teardown.  Captured CUDA graphs that contain NCCL kernels keep communicator
    resources alive and destroy_process_group() then blocks, so the graphs are released first and
    the process leaves through os._exit once every rank has reached the final barrier."""
    if world == 1:
        return
    import torch.distributed as dist
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    win.release_graphs()
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


def run_reference_cpu(args, rank, world):
    """fallback of the reference arm: the oracle port on the host cores, bounded sample."""
    if rank != 0:
        return
    cpu = cpu_oracle_sample()
    n_views = WORKLOADS["cfg3_replica"][0]
    line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": "iters/s", "n_gpus": world,
            "steps": 1, "warmup": 0, "ms_per_step": round(1e3 * n_views / cpu["value"], 1),
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic", "config": {"workload": "cfg3_replica (1/10-scale sample)"},
            "cpu_baseline": cpu,
            "e2e": {"value": cpu["value"], "unit": "iters/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


def main():
    if os.environ.get("B2GS_BENCH_WATCHDOG"):   # dev aid: dump every thread's stack and exit if stuck
        import faulthandler
        faulthandler.dump_traceback_later(int(os.environ["B2GS_BENCH_WATCHDOG"]), exit=True)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--workload", default="cfg3_replica", choices=sorted(WORKLOADS))
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="native arm: eager launches instead of CUDA graphs")
    ap.add_argument("--no-peer-adam", action="store_true",
                    help="native arm, N>1: NCCL all-reduce + replicated Adam instead of the peer-memory step")
    ap.add_argument("--peer-mode", default="mc", choices=["mc", "nomc"],
                    help="native arm, N>1: peer-memory Adam over NVLS multicast (multimem.ld_reduce / multimem.st) "
                         "where available, or plain peer loads / stores")
    ap.add_argument("--view-streams", type=int, default=0,
                    help="native arm: side streams the window's views are issued on (1 = serial; 0 = the "
                         "library's choice, mapping.auto_view_streams: 8 unless one view's working set exceeds L2)")
    args = ap.parse_args()
    if not torch.cuda.is_available():
        if args.impl == "reference":
            return run_reference_cpu(args, env_int("RANK", 0), env_int("WORLD_SIZE", 1))
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    if args.impl == "reference":
        return run_reference(args)
    run_native(args)


if __name__ == "__main__":
    main()
