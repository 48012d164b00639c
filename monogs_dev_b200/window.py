"""The mapping-window step as straight libb2gs calls (no autograd graph, no torch kernel inside a step).

One iteration of BackEnd.map (utils/slam_backend.py:156-317) over the local keyframes is, per view,

    b2gs_forward   preprocess -> tile scan -> instance scatter -> per-tile sort -> render, with the
                   mapping loss and its gradient formed in the render kernel's epilogue
                   (B2gsFusedLoss; utils/slam_utils.py:91-128) and the backward's accumulator rows
                   cleared by the preprocess kernel
    b2gs_backward  render backward -> fused preprocess backward, gradients accumulated straight into
                   the optimiser's arena, densification statistics and (for one view of the step) the
                   isotropic-scale regulariser (utils/slam_backend.py:229-231) in the same kernel

then ONE `b2gs_window_finalize` (losses, exposure / pose gradients, total loss, overflow guard) and
the update: `b2gs_adam_step_fused` (Adam + zero_grad) — or the peer-memory / NCCL variants — and ONE
`b2gs_pose_step` (camera Adam + update_pose + step ticks).  8 launches per view + 3 per step, all
from persistent buffers: what `MappingWindow` captures into its CUDA graph.

The generic autograd route (`mapping.render_fused` + `losses.get_loss_mapping_rgbd` +
`FusedAdam`) computes the same step from the same kernels and stays available
(`MappingWindow(lean=False)`); tests/test_loop_gpu.py compares the two.
"""
import ctypes as C
import math

import torch
import torch.distributed as dist

from . import _lib
from ._lib import lib, check, ptr
from .losses import MAPPING_RGBD

ADAM_ZERO_GRAD, ADAM_NO_TICK, ADAM_TICK_LATE, ADAM_NO_ZERO, ADAM_NO_FIRST_BARRIER = 1, 2, 4, 8, 16
HEADROOM = 1.5


class _View:
    """persistent buffers + prebuilt argument structs of one local keyframe"""
    pass


class LeanWindow:
    def __init__(self, gaussians, keyframes, local, bg, pose_lr, moves, exposures, densify_stats=True,
                 regulariser=True, rank=0, alpha=0.95, rgb_boundary_threshold=0.01, mode=MAPPING_RGBD,
                 pose_only=False, world=1):
        """pose_only: the map is frozen (tracking, utils/slam_frontend.py:128-196): the backward writes
        no per-Gaussian gradient at all, only the pose / exposure gradients of the views."""
        self.g, self.kfs, self.local, self.bg = gaussians, keyframes, list(local), bg
        self.pose_only = bool(pose_only)
        import os
        # software pipeline over the views (B2gsForwardOut.event_after_binning): measured SLOWER than
        # letting all views' kernels compete (cfg3, N=1: 4.77 vs 4.60 ms per step) -> opt-in only
        self.stagger = os.environ.get("B2GS_STAGGER", "0") == "1"
        if self.pose_only:
            densify_stats = regulariser = False
        self.pose_lr, self.moves, self.exposures = pose_lr, moves, exposures
        self.densify_stats, self.regulariser, self.rank, self.world = densify_stats, regulariser, rank, max(1, world)
        self.alpha, self.thr, self.mode = float(alpha), float(rgb_boundary_threshold), int(mode)
        self.dev = bg.device
        n = max(len(self.local), 1)
        i32 = dict(dtype=torch.int32, device=self.dev)
        # one block of flag words, read back by the host with ONE async copy per step:
        # (instances, overflow) per local view | the guard (OR of the overflow flags, window_finalize)
        # | the peer barrier's time-out word (MappingWindow points PeerArena.fail at it)
        self.flags = torch.zeros(2 * n + 2, **i32)
        self.status_all = self.flags[:2 * n].view(n, 2)
        self.guard = self.flags[2 * n:2 * n + 1]
        self.fail = self.flags[2 * n + 1:2 * n + 2]
        self.tickets = torch.zeros(4, **i32)                 # [0] finalize, [1] pose step
        # per view: "already counted in the densification statistics of the step being re-run"
        # (B2gsBackwardIO.stats_gate[2]; see begin_rerun)
        self.stats_mask = torch.zeros(n, **i32)
        self.losses = torch.zeros(n, device=self.dev)
        self.loss_total = torch.zeros((), device=self.dev)
        self.kf_grads = torch.zeros(n, 8, device=self.dev)   # per view: dL_dtau[6], dL_dexposure[2]
        self.kf_moments = torch.zeros(n, 16, device=self.dev)
        self.kf_step = torch.zeros(1, dtype=torch.int64, device=self.dev)
        self.tau_norm = torch.zeros(n, device=self.dev)      # |tau| of each view's last pose update
        self.cap = {}                                        # view -> binning capacity (instances)
        self.views = {}
        self._arena_key = None
        self.launches_fb = self.launches_up = 0
        # peer mode (MappingWindow sets it): forward_backward leaves b2gs_window_finalize to update()
        self.defer_finalize = False
        self._fin_pending = False
        # peer mode: the gradient arena of a step is cleared at the START of the next one, on a side
        # branch under the forward kernels (it is first needed by the backward), and the camera step
        # runs beside the exchange: neither sits on the critical path of a one-view rank
        self._aux_stream = None
        self._zero_pending = False
        self._zero_event = None

    # -- (re)build: the argument structs hold raw pointers into the arenas ------------------------
    def reset_camera_optimizer(self):
        """a fresh keyframe / pose optimizer (zero moments, step 0), as torch.optim.Adam(opt_params)
        is rebuilt per tracked frame (utils/slam_frontend.py:132-161)"""
        self.kf_moments.zero_()
        self.kf_step.zero_()

    def _key(self):
        g = self.g
        kf = tuple((self.kfs[v]["gt_image"].data_ptr(), self.kfs[v]["cam"].state.data_ptr(),
                    bool(self.kfs[v].get("initialization", False))) for v in self.local)
        return (g.flat.data_ptr(), g.flat_grad.data_ptr(), g._xyz.shape[0], g.active_sh_degree,
                getattr(g, "xyz_gradient_accum", g.flat).data_ptr(), self.densify_stats, self.regulariser, kf)

    def _build(self):
        g, dev = self.g, self.dev
        P = g._xyz.shape[0]
        M = 1 + g._features_rest.shape[1]
        u8 = dict(dtype=torch.uint8, device=dev)
        f32 = dict(dtype=torch.float32, device=dev)
        sinks = None if self.pose_only else g.grad_sinks(self.densify_stats)
        # the view-independent regulariser is counted once per step: every rank's first local view
        # carries the Gaussians i with i % world == rank (its cost is spread over the ranks)
        # (fewer keyframes than ranks: some rank has no view to carry its slice -> rank 0 takes all)
        spread = len(self.kfs) >= self.world
        reg_view = self.local[0] if (self.regulariser and self.local and (spread or self.rank == 0)) else None
        self._keep = []
        for i, v in enumerate(self.local):
            kf = self.kfs[v]
            cam = kf["cam"]
            if not hasattr(cam, "state") or not cam.state.is_cuda:
                raise RuntimeError("the lean window needs monogs_dev_b200.camera.Camera objects on the GPU")
            W, H = int(cam.image_width), int(cam.image_height)
            tiles = ((W + 15) // 16) * ((H + 15) // 16)
            old = self.views.get(v)
            V = _View()
            V.W, V.H, V.tiles, V.P = W, H, tiles, P
            if old is not None and (old.W, old.H) == (W, H):     # image-sized buffers survive a re-build
                for a in ("color", "depth", "opacity", "image_ws", "dL_dimage", "dL_ddepth", "partials"):
                    setattr(V, a, getattr(old, a))
            else:
                V.color, V.depth, V.opacity = torch.empty(3, H, W, **f32), torch.empty(1, H, W, **f32), torch.empty(1, H, W, **f32)
                V.image_ws = torch.zeros(lib.b2gs_image_bytes(W, H), **u8)
                V.dL_dimage, V.dL_ddepth = torch.empty(3, H, W, **f32), torch.empty(1, H, W, **f32)
                V.partials = torch.empty(tiles, 4, **f32)
            V.radii = torch.empty(P, dtype=torch.int32, device=dev)
            V.n_touched = torch.empty(P, dtype=torch.int32, device=dev)
            V.geom = torch.empty(lib.b2gs_geom_bytes(P), **u8)
            V.scratch = torch.empty(lib.b2gs_backward_scratch_bytes(P), **u8)
            V.binning = None
            # scene: raw parameters, camera block views (stable addresses: a pose update rewrites them in place)
            s = _lib.B2gsScene()
            s.P, s.D, s.M, s.W, s.H = P, int(g.active_sh_degree), M, W, H
            s.tan_fovx, s.tan_fovy = math.tan(cam.FoVx * 0.5), math.tan(cam.FoVy * 0.5)
            s.scale_modifier, s.prefiltered, s.debug = 1.0, 0, 0
            s.bg = self.bg.data_ptr()
            s.viewmatrix = cam.state.data_ptr() + 4 * 16
            s.projmatrix = cam.state.data_ptr() + 4 * 32
            s.campos = cam.state.data_ptr() + 4 * 48
            s.projmatrix_raw = cam.projection_matrix.data_ptr()
            s.means3D, s.shs, s.shs_rest = g._xyz.data_ptr(), g._features_dc.data_ptr(), ptr(g._features_rest)
            s.opacities, s.scales, s.rotations = g._opacity.data_ptr(), g._scaling.data_ptr(), g._rotation.data_ptr()
            s.raw_params = 1
            V.scene = s
            init = bool(kf.get("initialization", False))
            fl = _lib.B2gsFusedLoss()
            fl.args.N, fl.args.mode, fl.args.alpha, fl.args.rgb_boundary_threshold = W * H, self.mode, self.alpha, self.thr
            fl.args.gt_image, fl.args.gt_depth = kf["gt_image"].data_ptr(), ptr(kf.get("gt_depth"))
            fl.args.exposure_a = None if init else cam.exposure_a.data_ptr()
            fl.args.exposure_b = None if init else cam.exposure_b.data_ptr()
            fl.args.grad_mask = ptr(kf.get("grad_mask"))      # tracking modes (viewpoint.grad_mask)
            fl.dL_dimage, fl.dL_ddepth, fl.partials = V.dL_dimage.data_ptr(), V.dL_ddepth.data_ptr(), V.partials.data_ptr()
            V.fused = fl
            V.gt = (kf["gt_image"], kf.get("gt_depth"))      # the captured pointers must stay alive
            V.out = _lib.B2gsForwardOut(V.color.data_ptr(), V.depth.data_ptr(), V.opacity.data_ptr(),
                                        V.radii.data_ptr(), V.n_touched.data_ptr(),
                                        self.status_all[i].data_ptr(), C.addressof(fl), V.scratch.data_ptr(),
                                        V.scratch.numel(), None, 1)     # tile counters: zero-filled, kept clean
            io = _lib.B2gsBackwardIO()
            io.dL_dcolor, io.dL_ddepth, io.radii = V.dL_dimage.data_ptr(), V.dL_ddepth.data_ptr(), V.radii.data_ptr()
            if sinks is not None:
                io.dL_dmeans3D, io.dL_dopacity = sinks.xyz.data_ptr(), sinks.opacity.data_ptr()
                io.dL_dsh, io.dL_dsh_rest = sinks.f_dc.data_ptr(), ptr(sinks.f_rest)
                io.dL_dscales, io.dL_drotations = sinks.scaling.data_ptr(), sinks.rotation.data_ptr()
                io.densify_grad_accum, io.densify_denom = ptr(sinks.xyz_gradient_accum), ptr(sinks.denom)
                io.max_radii2D = ptr(sinks.max_radii2D)
                if self.densify_stats:
                    # a discarded step must not be counted twice (GS/scene/gaussian_model.py:691-695
                    # counts every applied iteration once): own overflow flag | the guard, which
                    # still holds the previous step's verdict here | the host's re-run mask
                    io.stats_gate[0] = self.status_all[i].data_ptr() + 4
                    io.stats_gate[1] = self.guard.data_ptr()
                    io.stats_gate[2] = self.stats_mask[i:i + 1].data_ptr()
            io.scratch, io.scratch_bytes = V.scratch.data_ptr(), V.scratch.numel()
            io.accumulate, io.scratch_zeroed, io.defer_tau = int(sinks is not None), 1, 1
            io.iso_reg_weight = 10.0 if v == reg_view else 0.0
            io.iso_reg_stride, io.iso_reg_offset = (self.world, self.rank) if spread else (0, 0)
            V.io = io
            nb = C.c_int32(0)
            tau_ptr = lib.b2gs_backward_tau_partials(V.scratch.data_ptr(), P, C.byref(nb))
            fin = _lib.B2gsViewFinalize()
            fin.loss_partials, fin.loss_rows, fin.N, fin.mode, fin.alpha = V.partials.data_ptr(), tiles, W * H, self.mode, self.alpha
            fin.exposure_a = fl.args.exposure_a
            fin.tau_partials, fin.tau_rows = tau_ptr, nb.value
            fin.reg_scale = 10.0 / (3.0 * P) if v == reg_view else 0.0
            fin.status, fin.loss = self.status_all[i].data_ptr(), self.losses[i:i + 1].data_ptr()
            fin.dL_dtau, fin.dL_dexposure = self.kf_grads[i].data_ptr(), self.kf_grads[i].data_ptr() + 4 * 6
            V.fin = fin
            pc = _lib.B2gsPoseCam()
            pc.cam_state, pc.projmatrix_raw = cam.state.data_ptr(), cam.projection_matrix.data_ptr()
            pc.rot_delta, pc.trans_delta = cam.cam_rot_delta.data_ptr(), cam.cam_trans_delta.data_ptr()
            pc.exposure_a, pc.exposure_b = cam.exposure_a.data_ptr(), cam.exposure_b.data_ptr()
            pc.dL_dtau, pc.dL_dexposure = fin.dL_dtau, fin.dL_dexposure
            pc.moments = self.kf_moments[i].data_ptr()
            pc.lr_rot, pc.lr_trans, pc.lr_exposure_a, pc.lr_exposure_b = [float(x) for x in self.pose_lr]
            pc.optimize_pose, pc.optimize_exposure = int(self.moves(v)), int(self.exposures(v) and not init)
            pc.tau_norm = self.tau_norm[i:i + 1].data_ptr()
            V.pose = pc
            V.cam = cam
            if old is not None and hasattr(old, "event"):
                V.event = old.event
            else:
                V.event = torch.cuda.Event()
                V.event.record()      # torch creates the CUDA event lazily: the library needs the handle
            self.views[v] = V
        n = len(self.local)
        self._fin_arr = (_lib.B2gsViewFinalize * max(n, 1))(*[self.views[v].fin for v in self.local])
        self._pose_arr = (_lib.B2gsPoseCam * max(n, 1))(*[self.views[v].pose for v in self.local])
        self._segs = None if self.pose_only else self._segments()
        self._arena_key = self._key()
        for v in self.local:                                  # binning workspaces at the known capacities
            self._alloc_binning(v, self.cap.get(v))

    def _segments(self):
        g = self.g
        opt = g.optimizer
        opt.capturable = True
        opt.sync_lr()
        items = []
        for gi, group in enumerate(opt.param_groups):
            for p in group["params"]:
                if p.numel() == 0:
                    continue
                st = opt.state[p]
                items.append((p, p.grad, st["exp_avg"], st["exp_avg_sq"], float(group["lr"]), gi, group))
        segs = (_lib.B2gsAdamSegment * len(items))()
        for s, (p, gr, m, v, lr, gi, group) in zip(segs, items):
            s.param, s.grad, s.exp_avg, s.exp_avg_sq = p.data_ptr(), gr.data_ptr(), m.data_ptr(), v.data_ptr()
            s.n, s.lr = p.numel(), lr
            if opt._lr_dev is not None:
                s.lr_dev = opt._lr_dev.data_ptr() + 4 * gi
        if opt._step_dev is None:
            step0 = max([int(opt.state[p]["step"]) for gr in opt.param_groups for p in gr["params"]] or [0])
            opt._step_dev = torch.full((1,), step0, dtype=torch.int64, device=self.dev)
        self._betas = opt.param_groups[0]["betas"]
        self._eps = float(opt.param_groups[0]["eps"])
        return segs

    def ensure(self):
        """(re)build the argument structs if an arena, a camera block or a ground-truth tensor moved
        (densify / prune re-allocate the arenas); returns True if it did — a captured graph is then stale."""
        if self._arena_key != self._key():
            self._build()
            return True
        return False

    # -- binning capacity ----------------------------------------------------------------------
    def _alloc_binning(self, v, cap):
        V = self.views[v]
        if cap is None:
            # first time: exact instance count through the two-call form (one sync), then head-room
            with torch.cuda.device(self.dev):
                st = torch.cuda.current_stream().cuda_stream
                ws = _lib.B2gsWorkspace(V.geom.data_ptr(), V.geom.numel(), V.image_ws.data_ptr(),
                                        V.image_ws.numel(), None, 0)
                out = _lib.B2gsForwardOut(None, None, None, V.radii.data_ptr(), V.n_touched.data_ptr(), None)
                R = C.c_int64(0)
                check(lib.b2gs_forward_plan(C.byref(V.scene), C.byref(ws), C.byref(out), st, C.byref(R)))
                V.image_ws.zero_()         # the plan leaves scan offsets in the tile counters
            cap = int(R.value * HEADROOM) + 4096
        self.cap[v] = int(cap)
        V.binning = torch.empty(lib.b2gs_binning_bytes(self.cap[v], V.W, V.H), dtype=torch.uint8, device=self.dev)
        V.cap_real = lib.b2gs_binning_capacity(V.binning.numel())
        V.ws = _lib.B2gsWorkspace(V.geom.data_ptr(), V.geom.numel(), V.image_ws.data_ptr(), V.image_ws.numel(),
                                  V.binning.data_ptr(), V.binning.numel())

    def note_overflow(self, vals):
        """vals: [[instances, overflow], ...] of the local views (host).  Grows the capacities of the
        views that overflowed; returns True if any did (the step was discarded on the device)."""
        over = False
        for v, (inst, flag) in zip(self.local, vals):
            if flag:
                over = True
                self._alloc_binning(v, max(self.cap[v], int(inst * HEADROOM) + 4096))
        return over

    def begin_rerun(self, vals):
        """before the FIRST repeat of a discarded step (parameters unchanged since): the views
        that did fit have added their densification statistics already, the ones that overflowed
        have not (stats_gate[0]); the replays after it added nothing (stats_gate[1])."""
        self.guard.zero_()
        self.stats_mask.copy_(torch.tensor([0 if flag else 1 for _, flag in vals] or [0],
                                           dtype=torch.int32))

    def end_rerun(self):
        self.stats_mask.zero_()

    # -- the step --------------------------------------------------------------------------------
    def forward_backward(self, side=()):
        """enqueue every local view's forward + backward (round-robin on the `side` streams, forked
        from / joined into the current stream) and the window finalize.  Returns the device scalar
        that will hold the step's loss."""
        self.ensure()
        n0 = lib.b2gs_kernel_launches()
        with torch.cuda.device(self.dev):
            main = torch.cuda.current_stream()
            side = list(side) if len(self.local) > 1 else []
            for s in side:
                s.wait_stream(main)
            if self._zero_pending and not self.pose_only:
                aux = self._aux()
                aux.wait_stream(main)
                g = self.g
                check(lib.b2gs_memset_async(g.flat_grad.data_ptr(), 0, 4 * g.flat_grad.numel(), aux.cuda_stream))
                self._zero_event.record(aux)
            prev_ev = None
            for i, v in enumerate(self.local):
                V = self.views[v]
                stream = side[i % len(side)] if side else main
                st = stream.cuda_stream
                # software pipeline over the views: view i's front end (preprocess .. sort: HBM- and
                # latency-bound) starts when view i-1 has finished ITS front end, i.e. runs under view
                # i-1's issue-bound render kernels, instead of all front ends competing first
                if side and self.stagger and prev_ev is not None:
                    stream.wait_event(prev_ev)
                if side and self.stagger:
                    V.out.event_after_binning = V.event.cuda_event
                    prev_ev = V.event
                else:
                    V.out.event_after_binning = None
                check(lib.b2gs_forward(C.byref(V.scene), C.byref(V.ws), C.byref(V.out), st))
                if self._zero_pending and not self.pose_only:
                    stream.wait_event(self._zero_event)      # the arena is clear before anything is added
                check(lib.b2gs_backward(C.byref(V.scene), C.byref(V.ws), V.cap_real, C.byref(V.io), st))
            for s in side:
                main.wait_stream(s)
            if self._zero_pending and not self.pose_only and not self.local:
                main.wait_event(self._zero_event)
            if self.local:
                if self.defer_finalize:
                    self._fin_pending = True       # update() runs it beside the first peer barrier
                else:
                    check(lib.b2gs_window_finalize(self._fin_arr, len(self.local), self.loss_total.data_ptr(),
                                                   self.guard.data_ptr(), self.tickets.data_ptr(), main.cuda_stream))
        self.launches_fb = lib.b2gs_kernel_launches() - n0
        return self.loss_total

    def _aux(self):
        if self._aux_stream is None:
            self._aux_stream = torch.cuda.Stream(self.dev)
            self._zero_event = torch.cuda.Event()
        return self._aux_stream

    def update(self, peer=None, world=1, group=None, gaussians=True):
        """[gradient exchange ->] Adam + zero_grad -> camera Adam + pose updates + step ticks, every
        state-changing kernel guarded by the overflow word of this step."""
        self.ensure()
        g = self.g
        gaussians = gaussians and not self.pose_only
        opt = g.optimizer if gaussians else None
        n0 = lib.b2gs_kernel_launches()
        b1, b2 = self._betas if gaussians else (0.9, 0.999)
        with torch.cuda.device(self.dev):
            st = torch.cuda.current_stream().cuda_stream
            if world > 1 and peer is None and gaussians:
                dist.all_reduce(g.flat_grad, op=dist.ReduceOp.SUM, group=group)
                dist.all_reduce(self.guard, op=dist.ReduceOp.MAX, group=group)
            lib.b2gs_set_update_guard(self.guard.data_ptr())
            try:
                tick = None
                if gaussians and peer is not None:
                    # barrier (ORs the guard over the ranks) -> camera step on a side branch || gradient
                    # sum + Adam + parameter broadcast -> second barrier (advances the step counter);
                    # this rank's gradient arena is cleared at the start of the next step
                    main = torch.cuda.current_stream()
                    aux = self._aux()
                    if self._fin_pending:
                        # the barrier forms this rank's guard from the views' overflow flags itself, so
                        # the finalize kernel (losses, camera gradients) leaves the critical path too
                        self._fin_pending = False
                        aux.wait_stream(main)
                        check(lib.b2gs_window_finalize(self._fin_arr, len(self.local), self.loss_total.data_ptr(),
                                                       None, self.tickets.data_ptr(), aux.cuda_stream))
                        check(lib.b2gs_peer_barrier_status(peer.struct(), self.status_all.data_ptr(),
                                                           len(self.local), st))
                    else:
                        check(lib.b2gs_peer_barrier(peer.struct(), st))
                    aux.wait_stream(main)          # the camera step needs the OR'd guard
                    check(lib.b2gs_pose_step(self._pose_arr, len(self.local), self.kf_step.data_ptr(), None, 0.9, 0.999,
                                             1e-8, self.tickets.data_ptr() + 4, aux.cuda_stream))
                    check(lib.b2gs_adam_step_peer_fused(
                        self._segs, len(self._segs), opt._step_dev.data_ptr(), b1, b2, self._eps, peer.struct(),
                        ADAM_NO_TICK | ADAM_TICK_LATE | ADAM_NO_ZERO | ADAM_NO_FIRST_BARRIER, st))
                    main.wait_stream(aux)
                    self._zero_pending = True
                    for v in self.local:
                        self.views[v].cam._host_RT = None
                    self.launches_up = lib.b2gs_kernel_launches() - n0
                    return
                if self._fin_pending:              # deferred, but this update has no exchange to hide it under
                    self._fin_pending = False
                    check(lib.b2gs_window_finalize(self._fin_arr, len(self.local), self.loss_total.data_ptr(),
                                                   self.guard.data_ptr(), self.tickets.data_ptr(), st))
                if gaussians:
                    if peer is not None:
                        pass
                    else:
                        check(lib.b2gs_adam_step_fused(self._segs, len(self._segs), opt._step_dev.data_ptr(), b1, b2,
                                                       self._eps, ADAM_ZERO_GRAD | ADAM_NO_TICK, st))
                    tick = opt._step_dev.data_ptr()
                check(lib.b2gs_pose_step(self._pose_arr, len(self.local), self.kf_step.data_ptr(), tick, 0.9, 0.999,
                                         1e-8, self.tickets.data_ptr() + 4, st))
            finally:
                lib.b2gs_set_update_guard(None)
        # (the step count lives on the device: opt._step_dev; GaussianModel._allocate carries it over)
        for v in self.local:
            self.views[v].cam._host_RT = None      # pose changed on the device
        self.launches_up = lib.b2gs_kernel_launches() - n0

    def package(self, v):
        """what render() returns for keyframe v (gaussian_renderer/__init__.py:145-153), from the
        persistent buffers of the last step; `visibility_filter` is computed on request."""
        V = self.views[v]
        return {"render": V.color, "viewspace_points": None, "radii": V.radii, "depth": V.depth,
                "opacity": V.opacity, "n_touched": V.n_touched, "stats_fused": self.densify_stats}
