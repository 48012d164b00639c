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
sm_100 by oracle/build_ref.py) + torch.optim.Adam through the same loop; if that extension
cannot be loaded it falls back to the CPU oracle port on a bounded sample.
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

WORKLOAD = "cfg3_replica"
N_VIEWS = 8
METRIC = "fwd+bwd render iters/s @1200x680, 500k Gaussians (8-keyframe map-window step)"


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


def perturbed_copy(sc, seed=77):
    """the 'ground-truth' scene: parameters perturbed by 5 % noise."""
    g = torch.Generator().manual_seed(seed)
    n = lambda t: torch.randn(t.shape, generator=g).to(t.device)
    return dict(xyz=sc.xyz + 0.05 * torch.exp(sc.scaling).mean(1, keepdim=True) * n(sc.xyz),
                scaling=sc.scaling + 0.05 * n(sc.scaling), rotation=sc.rotation + 0.05 * n(sc.rotation),
                opacity=sc.opacity + 0.05 * n(sc.opacity), f_dc=sc.f_dc + 0.05 * n(sc.f_dc),
                f_rest=sc.f_rest)


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


def build_window(device, render_fn, model_cls, window_cls, sc):
    """keyframes with ground truth rendered (untimed) from a 5 %-perturbed copy of the scene."""
    from monogs_dev_b200 import mapping
    gt = perturbed_copy(sc)
    gt_model = mapping.GaussianModel(gt["xyz"], gt["f_dc"], gt["f_rest"], gt["opacity"], gt["scaling"],
                                     gt["rotation"], sh_degree=sc.sh_degree) if model_cls is None else \
        model_cls(gt["xyz"], gt["f_dc"], gt["f_rest"], gt["opacity"], gt["scaling"], gt["rotation"],
                  sh_degree=sc.sh_degree)
    kfs, host = [], []
    with torch.no_grad():
        for cam in sc.cameras:
            pkg = render_fn(cam, gt_model, None, sc.bg)
            img, dep = pkg["render"].detach().clone(), pkg["depth"].detach().clone()
            kfs.append(dict(cam=cam, gt_image=img, gt_depth=dep))
            host.append((img.cpu().pin_memory(), dep.cpu().pin_memory()))
    del gt_model
    torch.cuda.synchronize()
    return kfs, host


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
    bytes for the HBM-bound kernels, flops for the FP32-pipe-bound ones."""
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


def ncu_metrics(kernel):
    """issue-slot and FMA-pipe utilisation of `kernel` from the newest committed `ncu --set full`
    summary under profiles/ (what the FP32 'frac' does not say: see DESIGN.md §7)."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_render_full.txt")))
    if not files:
        return None
    want = {"smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_slot_pct",
            "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
            "gpu__time_duration.sum": "ncu_us"}
    out, on = {}, False
    for line in open(files[-1]):
        if line.startswith("----"):
            on = kernel + "_kernel" in line
        elif on:
            m = re.match(r"\s+(\S+)\s+([0-9.]+)", line)
            if m and m.group(1) in want:
                out[want[m.group(1)]] = round(float(m.group(2)), 1)
    if out:
        out["source"] = os.path.basename(files[-1])
    return out or None


def ncu_traffic(kernel):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of `kernel`, from the newest
    `ncu --set full` summary committed under profiles/ (tools/ncu_summary.py); None if absent."""
    import glob
    import re
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_render_full.txt")))
    if not files:
        return None
    vals, cur = [], None
    for line in open(files[-1]):
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
    """bounded CPU sample: ONE fwd+bwd of a 1/10-scale config 3 (50k Gaussians, 380x216,
    focal / sqrt(10)) through the PyTorch oracle; scaled to full-size iters/s by 1/10."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from helpers import build_case, run_oracle
    s = 10 ** 0.5
    sc, rs, inp, gc, gd = build_case(WORKLOAD, P=50_000, W=380, H=216, fx=600 / s, fy=600 / s,
                                     cx=189.5, cy=107.5)
    torch.set_num_threads(os.cpu_count() or 1)
    t = time.time()
    run_oracle(rs, inp, gc, gd)
    dt = time.time() - t
    return dict(value=round(1.0 / dt / 10.0, 5), unit="iters/s", cores=torch.get_num_threads(), kind="port",
                sample="1 fwd+bwd of a 1/10-scale cfg3 view (50k Gaussians, 380x216) in %.1f s on the "
                       "pure-PyTorch oracle; value = measured rate / 10" % dt)


def run(args):
    impl_ref = args.impl == "reference"
    rank, world, local = dist_setup(args)
    device = torch.device("cuda", local)
    from monogs_dev_b200 import synthetic as syn, mapping, _lib
    K, W_ = args.steps, max(args.warmup, 3)

    if impl_ref:
        from oracle import ref_ext
        if not ref_ext.available():
            return run_reference_cpu(args, rank, world)
        render_fn, model_cls, adam_cls = ref_ext.ref_render, ref_ext.RefGaussianModel, torch.optim.Adam
        loss_fn, graph = mapping.get_loss_mapping_rgbd_torch, False     # the reference's own torch-op loss
    else:
        render_fn, model_cls, adam_cls = mapping.render_fused, mapping.GaussianModel, mapping.FusedAdam
        loss_fn, graph = mapping.get_loss_mapping_rgbd, not args.no_graph

    sc = syn.make_scene(WORKLOAD, device=device, n_views=N_VIEWS)
    kfs, host = build_window(device, render_fn, model_cls, None, sc)
    gm = model_cls(sc.xyz, sc.f_dc, sc.f_rest, sc.opacity, sc.scaling, sc.rotation, sh_degree=sc.sh_degree)
    win = mapping.MappingWindow(gm, kfs, sc.bg, render_fn=render_fn, adam_cls=adam_cls, loss_fn=loss_fn,
                                cuda_graph=graph, view_streams=1 if impl_ref else args.view_streams,
                                peer_adam=(not impl_ref) and world > 1 and not args.no_peer_adam)
    local_views = win.local

    def step():
        return win.step()

    # e2e: every step's ground-truth RGB-D frames come from pinned host memory.  The H2D copy of
    # step k+1 runs on a copy stream while step k computes (double buffering through a staging
    # set; one D2D hand-over per step), and the loss is read back to the host every step.
    copy_stream = torch.cuda.Stream()
    stage = {v: (torch.empty_like(kfs[v]["gt_image"]), torch.empty_like(kfs[v]["gt_depth"])) for v in local_views}
    ev_ready, ev_free = torch.cuda.Event(), torch.cuda.Event()
    pending = [False]
    loss_host = [torch.empty(1, dtype=torch.float32).pin_memory() for _ in range(2)]
    loss_ev = [torch.cuda.Event(), torch.cuda.Event()]
    n_e2e, losses_read = [0], []

    def prefetch():
        copy_stream.wait_event(ev_free)
        with torch.cuda.stream(copy_stream):
            for v in local_views:
                stage[v][0].copy_(host[v][0], non_blocking=True)
                stage[v][1].copy_(host[v][1], non_blocking=True)
            ev_ready.record(copy_stream)
        pending[0] = True

    def step_e2e():
        cur = torch.cuda.current_stream()
        if not pending[0]:
            ev_free.record(cur)
            prefetch()
        cur.wait_event(ev_ready)
        for v in local_views:                      # staging -> the buffers the (captured) step reads
            kfs[v]["gt_image"].copy_(stage[v][0], non_blocking=True)
            kfs[v]["gt_depth"].copy_(stage[v][1], non_blocking=True)
        ev_free.record(cur)
        prefetch()                                 # next step's inputs, overlapped with this step
        loss = win.step()
        # result read back, device -> host, every step: an async copy into pinned memory that the
        # host reads after it has enqueued the NEXT step (settle_e2e reads the last one)
        i = n_e2e[0] % 2
        loss_host[i].copy_(loss.detach().reshape(1), non_blocking=True)
        loss_ev[i].record(cur)
        if n_e2e[0] > 0:
            loss_ev[1 - i].synchronize()
            losses_read.append(float(loss_host[1 - i][0]))
        n_e2e[0] += 1

    def settle_e2e():
        win.flush()
        if n_e2e[0] > 0:
            i = (n_e2e[0] - 1) % 2
            loss_ev[i].synchronize()
            losses_read.append(float(loss_host[i][0]))

    for _ in range(W_ + (4 if graph else 0)):      # graph mode: 3 eager sizing steps + the capture
        step()
    sampler = ClockSampler(local) if rank == 0 else None
    l0 = _lib.lib.b2gs_kernel_launches()
    ms = timed_steps(step, K, world, settle=win.flush)
    launches = _lib.lib.b2gs_kernel_launches() - l0
    if graph:      # replays do not pass through the host-side counter: launches captured per step x K
        launches = win.graph_launches * K
    clocks = sampler.stop() if sampler else {}
    ms = max_over_ranks(ms, world, device)
    launches = int(sum_over_ranks(launches, world, device))
    step_e2e()
    settle_e2e()
    torch.cuda.synchronize()
    pending[0] = False                             # the timed region issues its own first copy
    n_e2e[0] = 0
    del losses_read[:]
    ms_e2e = max_over_ranks(timed_steps(step_e2e, K, world, settle=settle_e2e), world, device)
    assert len(losses_read) == K, "every step's loss must have been read back inside the timed region"

    # headline companion: single-view fwd+bwd only (no optimizer), rank-local, N=1 only
    single = None
    if world == 1:
        def one_view():
            pkg = render_fn(kfs[0]["cam"], gm, None, sc.bg)
            loss_fn(pkg["render"], pkg["depth"], kfs[0]["cam"], kfs[0]["gt_image"], kfs[0]["gt_depth"]).backward()
            gm.zero_grad()
            for t in (kfs[0]["cam"].cam_rot_delta, kfs[0]["cam"].cam_trans_delta,
                      kfs[0]["cam"].exposure_a, kfs[0]["cam"].exposure_b):
                t.grad = None
        for _ in range(3):
            one_view()
        n1 = max(20, K * N_VIEWS // 2)
        single = n1 / (timed_steps(one_view, n1, 1) / 1e3)

    roof, rooftab = None, {}
    if not impl_ref:
        # per-kernel durations: CUDA events on the launching stream, 2 extra (untimed) steps
        # (views serialised on one stream for this pass: a kernel timed while another view's
        # kernels share the SMs would not be that kernel's duration)
        _lib.lib.b2gs_profile_enable(1)
        vs, win.view_streams = win.view_streams, 1
        win._eager_step(set_to_none=False); win._eager_step(set_to_none=False)   # eager: events per launch
        win.view_streams = vs
        torch.cuda.synchronize()
        prof = _lib.profile_read()
        _lib.lib.b2gs_profile_enable(0)
        pkg = win.last[local_views[0]]
        st = pkg["render"].grad_fn.state if pkg["render"].grad_fn is not None else None
        P = sc.P
        pvis = int((pkg["radii"] > 0).sum().item())
        pairs = sum_n_contrib(st, sc.W, sc.H, device) if st is not None else 0
        R = st.num_rendered if st is not None else 0
        n_par = gm.flat.numel()
        N = sc.W * sc.H
        tiles = ((sc.W + 15) // 16) * ((sc.H + 15) // 16)
        units = {   # algorithmic work per launch
            "preprocess_fwd": ("hbm", 131.0 * pvis + 20.0 * (P - pvis)),
            "tile_scan": ("hbm", 16.0 * tiles),
            "scatter_instances": ("hbm", 12.0 * P + 40.0 * pvis + 8.0 * R),
            "tile_sort": ("hbm", 24.0 * R),                       # the sort's lower bound: one read + one write of a 12 B pair
            "render_fwd": ("fp32", 26.0 * pairs),
            "render_bwd": ("fp32", 80.0 * pairs),
            "preprocess_bwd": ("hbm", 319.0 * pvis + 60.0 * (P - pvis)),
            "loss_fwd": ("hbm", 32.0 * N),
            "loss_bwd": ("hbm", 48.0 * N),
            "iso_reg": ("hbm", 36.0 * P),
            "adam": ("hbm", 28.0 * n_par),
        }
        rooftab = rooflines(prof, units, peaks())
        own = {k: v for k, v in prof.items() if k in rooftab}
        if own:
            top = max(own, key=lambda k: own[k][1])
            roof = dict(kernel=top, **rooftab[top])
            roof["traffic"] = ncu_traffic(top)
            roof["ncu"] = ncu_metrics(top)
            tot = sum(v[1] for v in prof.values())
            roof["share_of_gpu_kernel_time"] = round(prof[top][1] / tot, 3) if tot else None
        work = dict(P=P, P_visible=pvis, instances=R, effective_pairs=pairs)
    else:
        work = {}

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline and not impl_ref:
        cpu = cpu_oracle_sample()

    if rank != 0:
        return finish(world, win)
    views_per_s = N_VIEWS * K / (ms / 1e3)
    h2d = sum(host[v][0].numel() * 4 + host[v][1].numel() * 4 for v in range(N_VIEWS))
    line = {
        "metric": METRIC, "value": round(views_per_s, 3), "unit": "iters/s", "n_gpus": world, "steps": K,
        "warmup": W_, "ms_per_step": round(ms / K, 4), "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": "cfg3_replica: 1200x680, 500k Gaussians, SH0, 8-keyframe mapping window "
                               "(fwd+bwd per keyframe, fused Adam, pose Adam)",
                   "keyframes": N_VIEWS, "parallelism": "keyframes sharded round-robin over %d rank(s)" % world + (
                       "; gradient exchange + Adam + zero_grad as one kernel over NVLink peer memory"
                       if getattr(win, "peer", None) is not None else ""),
                   "launch": ("cuda graphs (2 per step, no host sync inside a step: overflow flags are read back "
                              "asynchronously, flush() inside the timed region)" if graph else "eager") +
                             ("" if impl_ref or args.view_streams < 2 else
                              ", views issued on %d forked streams" % args.view_streams),
                   "l2_policy": "inputs larger than L2: >400 MB touched per step (8 GT frames 104 MB, "
                                "4 parameter/moment arenas 112 MB, per-view workspaces+gradients) vs 126 MB L2",
                   **work},
        "clocks": clocks,
        "e2e": {"value": round(N_VIEWS * K / (ms_e2e / 1e3), 3), "unit": "iters/s",
                "ms_per_step": round(ms_e2e / K, 4), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4 * world,
                "readback": "every step's loss copied to pinned host memory and read by the host one step "
                            "later (after the next step is enqueued); the last one before the region ends"},
        "gpu_launches": launches,
        "single_view_fwd_bwd_iters_s": round(single, 2) if single else None,
    }
    if impl_ref:
        line["impl"] = "reference"
        line["cpu_baseline"] = {"value": line["value"], "unit": "iters/s", "cores": 0, "kind": "reference",
                                "sample": "UNMODIFIED reference CUDA rasterizer (oracle/_ref, sm_100 build) + "
                                          "torch.optim.Adam on the GPU: full workload, not a CPU run "
                                          "(the reference has no CPU rasterizer)"}
        line["gpu_launches"] = None
    else:
        line["roofline"] = roof
        line["rooflines_all"] = rooftab
        line["cpu_baseline"] = cpu
    print(json.dumps(line), flush=True)
    finish(world, win)


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
    win._g1 = win._g2 = None
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


def run_reference_cpu(args, rank, world):
    """fallback of the reference arm: the oracle port on the host cores, bounded sample."""
    if rank != 0:
        return
    cpu = cpu_oracle_sample()
    line = {"impl": "reference", "metric": METRIC, "value": cpu["value"], "unit": "iters/s", "n_gpus": world,
            "steps": 1, "warmup": 0, "ms_per_step": round(1e3 * N_VIEWS / cpu["value"], 1),
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-graph", action="store_true", help="native arm: eager launches instead of CUDA graphs")
    ap.add_argument("--no-peer-adam", action="store_true",
                    help="native arm, N>1: NCCL all-reduce + replicated Adam instead of the peer-memory step")
    ap.add_argument("--view-streams", type=int, default=8,
                    help="native arm: side streams the window's views are issued on (1 = serial)")
    args = ap.parse_args()
    if not torch.cuda.is_available():
        if args.impl == "reference":
            return run_reference_cpu(args, env_int("RANK", 0), env_int("WORLD_SIZE", 1))
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU fallback")
    run(args)


if __name__ == "__main__":
    main()
