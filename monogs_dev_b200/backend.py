"""The mapping schedule around the keyframe window: host-side mirror of the reference's BackEnd
(utils/slam_backend.py), same method names, argument meaning and return values, on top of
MappingWindow (mapping.py) and the packed-arena GaussianModel.

  BackEnd.set_hyperparams      utils/slam_backend.py:40-60
  BackEnd.add_next_kf          :67-70
  BackEnd.reset                :72-84      (queues are the control plane: out of scope)
  BackEnd.initialize_map       :86-139
  BackEnd.map                  :141-318    (window + two random older keyframes, prune mode,
                                            densify / opacity-reset schedule, pose updates)
  keyframe optimizer set-up    :434-471    (`set_keyframe_optimizers`, inline in BackEnd.run)

  BackEnd.color_refinement     :320-355    (single random view per iteration, L1 + D-SSIM)

Not here: the process / queue plumbing of BackEnd.run (control plane, out of scope).

Data-parallel windows (one process per GPU): the views of an iteration are sharded by
MappingWindow; every rank runs this schedule on its replica with identical decisions — the random
views come from a generator seeded identically on all ranks, `n_touched` masks and updated
poses / exposures are exchanged with one collective each.

Reference behaviours kept on purpose (each pinned in tests/test_backend_cpu.py):
  * prune mode returns after the backward: no optimizer step, gradients stay in the arena and
    are added to by the next iteration (q7), densification statistics are not touched;
  * the Gaussian optimizer step that follows densify_and_prune changes nothing (the reference
    steps over fresh Parameters whose .grad is None); after reset_opacity_nonvisible the same
    holds for the opacity group only;
    (with per-parameter step counts that is exact; a capturable / peer-memory FusedAdam keeps ONE
    device step count, so the opacity group's bias correction then runs one step ahead per reset —
    a factor of 1 - 4e-4 at the reference's gaussian_reset = 2001);
  * keyframe 0 and the random views have no optimised camera / exposure parameters; only the first
    `pose_window` views of the window have their pose updated;
  * `occ_aware_visibility` is rebuilt from the window views of the LAST iteration only.
"""
import functools
import random

import torch
import torch.distributed as dist

from .adam import FusedAdam
from .mapping import MappingWindow


class BackEnd:
    def __init__(self, config, gaussians=None, background=None, cameras_extent=6.0, opt_params=None,
                 group=None, render_fn=None, loss_fn=None, adam_cls=FusedAdam, cuda_graph=False,
                 view_streams=1, peer_adam=False, seed=0):
        self.config = config
        self.gaussians = gaussians
        self.background = background
        self.cameras_extent = cameras_extent
        self.opt_params = opt_params
        self.monocular = config["Training"]["monocular"]
        self.iteration_count = 0
        self.last_sent = 0
        self.occ_aware_visibility = {}
        self.viewpoints = {}            # kf id -> dict(cam=, gt_image=, gt_depth=)  (Camera + its frame)
        self.current_window = []
        self.initialized = not self.monocular
        self.keyframe_optimizers = None
        self._group, self._render_fn, self._loss_fn, self._adam = group, render_fn, loss_fn, adam_cls
        self._graph, self._streams, self._peer = bool(cuda_graph), view_streams, peer_adam
        # randperm source for the two extra views: identical on every rank
        self._gen = torch.Generator().manual_seed(seed)
        self._seed, self._dev_gen = seed, None       # densify_and_split samples: same on every rank
        self._seed_gen = torch.Generator().manual_seed(seed + 2)   # keyframe seeding: same on every rank
        self._py_rng = random.Random(seed)           # color_refinement's view choice: same on every rank
        self._win_key, self._win = None, None
        self._kf_flags = {}
        self._vis_iter, self._last_ids = -1, None
        if gaussians is not None:
            self.set_hyperparams()

    # -- configuration --------------------------------------------------------------------------
    def set_hyperparams(self):
        t = self.config["Training"]
        self.init_itr_num = t["init_itr_num"]
        self.init_gaussian_update = t["init_gaussian_update"]
        self.init_gaussian_reset = t["init_gaussian_reset"]
        self.init_gaussian_th = t["init_gaussian_th"]
        self.init_gaussian_extent = self.cameras_extent * t["init_gaussian_extent"]
        self.mapping_itr_num = t["mapping_itr_num"]
        self.gaussian_update_every = t["gaussian_update_every"]
        self.gaussian_update_offset = t["gaussian_update_offset"]
        self.gaussian_th = t["gaussian_th"]
        self.gaussian_extent = self.cameras_extent * t["gaussian_extent"]
        self.gaussian_reset = t["gaussian_reset"]
        self.size_threshold = t["size_threshold"]
        self.window_size = t["window_size"]

    def _densify_grad_threshold(self):
        return getattr(self.opt_params, "densify_grad_threshold", 0.0002)

    def add_next_kf(self, frame_idx, viewpoint, init=False, scale=2.0, depth_map=None):
        """viewpoint: a frame dict {cam, gt_image, gt_depth} (Camera.from_reference(vp) makes one from a
        MonoGS Camera) or a camera object carrying original_image / depth itself."""
        cam = viewpoint
        if isinstance(viewpoint, dict):
            cam = viewpoint["cam"]
            if getattr(cam, "original_image", None) is None:
                cam.original_image = viewpoint["gt_image"]
                cam.depth = viewpoint["gt_depth"].reshape(viewpoint["gt_image"].shape[-2:])
            self.viewpoints.setdefault(frame_idx, viewpoint)
        # the random down-sample subset (and the monocular depth noise) must be the SAME on every rank
        # of a data-parallel run: replicas that seed different Gaussians keep equal sizes and are then
        # silently mixed by the gradient exchange.  A host generator seeded alike on all ranks.
        self.gaussians.extend_from_pcd_seq(cam, kf_id=frame_idx, init=init, scale=scale, depthmap=depth_map,
                                           generator=self._seed_gen)
        self._win_key = None

    def reset(self):
        self.iteration_count = 0
        self.occ_aware_visibility = {}
        self.viewpoints = {}
        self.current_window = []
        self.initialized = not self.monocular
        self.keyframe_optimizers = None
        if hasattr(self.gaussians, "unique_kfIDs"):
            self.gaussians.prune_points(self.gaussians.unique_kfIDs >= 0)     # remove all gaussians
        self._win_key = None

    def set_keyframe_optimizers(self, current_window, frames_to_optimize=None):
        """what BackEnd.run does for every new keyframe (:434-471): rot / trans deltas of the first
        `frames_to_optimize` window views at half the tracking rates, exposure of every window
        view, nothing for keyframe 0."""
        lr = self.config["Training"]["lr"]
        if frames_to_optimize is None:
            frames_to_optimize = self.config["Training"]["pose_window"]
        groups, flags = [], {}
        world, rank = self._world_rank()
        for cam_idx, kf_idx in enumerate(current_window):
            if kf_idx == 0:
                continue
            cam = self.viewpoints[kf_idx]["cam"]
            flags[kf_idx] = cam_idx < frames_to_optimize
            if cam_idx % world != rank:        # the view's owner optimises its camera (mapping.shard_views)
                continue
            if cam_idx < frames_to_optimize:
                groups += [dict(params=[cam.cam_rot_delta], lr=lr["cam_rot_delta"] * 0.5, name="rot_%d" % kf_idx),
                           dict(params=[cam.cam_trans_delta], lr=lr["cam_trans_delta"] * 0.5, name="trans_%d" % kf_idx)]
            groups += [dict(params=[cam.exposure_a], lr=0.01, name="exposure_a_%d" % kf_idx),
                       dict(params=[cam.exposure_b], lr=0.01, name="exposure_b_%d" % kf_idx)]
        kw = dict(capturable=True) if (self._graph and self._adam is FusedAdam) else {}
        self.keyframe_optimizers = self._adam(groups, **kw) if groups else None
        self._kf_flags = flags
        self._win_key = None
        return self.keyframe_optimizers

    # -- helpers --------------------------------------------------------------------------------
    def _densify(self, min_opacity, extent, max_screen_size):
        g = self.gaussians
        dev = g._xyz.device
        if self._dev_gen is None or self._dev_gen.device != dev:
            self._dev_gen = torch.Generator(device=dev).manual_seed(self._seed + 1)
        return g.densify_and_prune(self._densify_grad_threshold(), min_opacity, extent, max_screen_size,
                                   generator=self._dev_gen)

    def _world_rank(self):
        if dist.is_available() and dist.is_initialized():
            return dist.get_world_size(self._group), dist.get_rank(self._group)
        return 1, 0

    def _window(self, ids, n_window, frames_to_optimize, stats):
        """MappingWindow over the views `ids` (the first n_window are the window, the rest the random
        extras); kept across iterations while views, flags and the parameter arena are unchanged, so
        that a captured step is replayed instead of re-captured."""
        arena = lambda: getattr(self.gaussians, "flat", torch.empty(0)).data_ptr()
        key = (tuple(ids), n_window, frames_to_optimize, stats, id(self.keyframe_optimizers), arena())
        if key == self._win_key:
            return self._win
        if self._win is not None:
            self._win.flush()
        kfs = []
        for i, kf_idx in enumerate(ids):
            kf = dict(self.viewpoints[kf_idx])
            in_window = i < n_window and kf_idx != 0
            kf["optimize_pose"] = bool(in_window and i < frames_to_optimize and self._kf_flags.get(kf_idx, True))
            kf["optimize_exposure"] = in_window
            kfs.append(kf)
        kw = {}
        if self.keyframe_optimizers is not None:
            kw["kf_optimizer"] = self.keyframe_optimizers
        lr = self.config["Training"]["lr"]
        win = MappingWindow(self.gaussians, kfs, self.background, lean=False,
                            pose_lr=(lr["cam_rot_delta"] * 0.5, lr["cam_trans_delta"] * 0.5, 0.01, 0.01),
                            group=self._group, render_fn=self._render_fn, adam_cls=self._adam, loss_fn=self._loss_fn,
                            cuda_graph=self._graph, densify_stats=stats, view_streams=self._streams,
                            peer_adam=self._peer, **kw)
        if self.keyframe_optimizers is None:
            self.keyframe_optimizers = win.kf_optimizer
        # (the first peer-memory window moves the arenas: the key holds what the NEXT call will see)
        self._win_key, self._win = key[:4] + (id(self.keyframe_optimizers), arena()), win
        return win

    def _gather_touched(self, win, ids, n_window):
        """(n_touched > 0) of every WINDOW view, on every rank: the owner's mask, summed over ranks
        as one int32 [n_window, P] message (§8e item 3)."""
        g = self.gaussians
        P = g._xyz.shape[0]
        dev = g._xyz.device
        world, _ = self._world_rank()
        masks = torch.zeros(n_window, P, dtype=torch.int32, device=dev)
        for v, pkg in win.last.items():
            if v < n_window:
                masks[v] = (pkg["n_touched"] > 0).to(torch.int32)
        if world > 1:
            dist.all_reduce(masks, op=dist.ReduceOp.SUM, group=self._group)
        return {ids[v]: masks[v].long() for v in range(n_window)}

    def _gather_visibility(self, win):
        """visibility filters of all views of the iteration (window + random), OR-ed over ranks"""
        g = self.gaussians
        seen = torch.zeros(g._xyz.shape[0], dtype=torch.int32, device=g._xyz.device)
        for pkg in win.last.values():
            seen |= pkg["visibility_filter"].to(torch.int32)
        world, _ = self._world_rank()
        if world > 1:
            dist.all_reduce(seen, op=dist.ReduceOp.MAX, group=self._group)
        return seen.bool()

    def sync_keyframes(self, ids):
        """every rank gets the pose and exposure its peers optimised (§8e item 4): one SUM
        all-reduce of a [V,14] block in which only a view's owner fills its row."""
        world, rank = self._world_rank()
        if world == 1:
            return
        dev = self.gaussians._xyz.device
        blk = torch.zeros(len(ids), 14, device=dev)
        for v, kf_idx in enumerate(ids):
            if v % world == rank:
                cam = self.viewpoints[kf_idx]["cam"]
                blk[v, :9], blk[v, 9:12] = cam.R.reshape(-1).to(dev), cam.T.to(dev)
                blk[v, 12], blk[v, 13] = cam.exposure_a.detach()[0], cam.exposure_b.detach()[0]
        dist.all_reduce(blk, op=dist.ReduceOp.SUM, group=self._group)
        host = blk.cpu()
        for v, kf_idx in enumerate(ids):
            if v % world != rank:
                cam = self.viewpoints[kf_idx]["cam"]
                cam.update_RT(host[v, :9].view(3, 3), host[v, 9:12])
                with torch.no_grad():
                    cam.exposure_a.fill_(float(host[v, 12]))
                    cam.exposure_b.fill_(float(host[v, 13]))

    # -- the schedule ---------------------------------------------------------------------------
    def initialize_map(self, cur_frame_idx, viewpoint=None):
        """:86-139 — init_itr_num single-view iterations on the first keyframe with the
        initialisation loss (no exposure, no isotropic term) and thresholds; returns the render
        package of the last iteration (None on a rank that does not own the view)."""
        g = self.gaussians
        if viewpoint is not None:
            self.viewpoints[cur_frame_idx] = viewpoint
        win = self._window_init(cur_frame_idx)
        for mapping_iteration in range(self.init_itr_num):
            self.iteration_count += 1
            win.forward_backward()
            self._last_ids = ([cur_frame_idx], 1, win)
            last = win.last
            densified = mapping_iteration % self.init_gaussian_update == 0
            if densified or mapping_iteration == self.init_itr_num - 1:
                self._set_visibility()          # masks are as long as the map the view was rendered from
            if densified:
                self._densify(self.init_gaussian_th, self.init_gaussian_extent, None)
            reset = (self.iteration_count == self.init_gaussian_reset
                     or self.iteration_count == getattr(self.opt_params, "densify_from_iter", 500))
            if reset:
                g.reset_opacity()
            if densified:
                g.zero_grad()           # fresh Parameters in the reference: its step is void (:133-134)
            else:
                self._step_gaussians(win, skip_opacity=reset)
        self._win_key = None
        return next(iter(last.values()), None)

    def _window_init(self, kf_idx):
        kf = dict(self.viewpoints[kf_idx])
        kf["optimize_pose"] = kf["optimize_exposure"] = False
        from . import mapping
        loss = self._loss_fn or mapping.get_loss_mapping_rgbd
        return MappingWindow(self.gaussians, [kf], self.background, optimize_pose=False, group=self._group, lean=False,
                             render_fn=self._render_fn, adam_cls=self._adam,
                             loss_fn=functools.partial(loss, initialization=True),
                             cuda_graph=False, densify_stats=True, regulariser=False)

    def _step_gaussians(self, win, skip_opacity=False):
        """optimizer.step() + zero_grad; with skip_opacity the opacity group keeps its value and its
        moments (its Parameter was just replaced in the reference, so it has no .grad there).
        Per-parameter step counts: the group is left out of the launch.  One device step count for
        the whole optimizer (capturable / peer-memory Adam, whose chunk ownership follows the
        segment list): the group's three arena slices are put back after the step."""
        g = self.gaussians
        p = g._params["opacity"] if skip_opacity else None
        keep = grad = None
        if p is not None:
            st = g.optimizer.state.get(p, {})
            if getattr(g.optimizer, "capturable", False):
                keep = (p.detach().clone(), st["exp_avg"].clone(), st["exp_avg_sq"].clone())
            else:
                grad, p.grad = p.grad, None
        try:
            win.update(gaussians=True)     # [gradient exchange ->] Adam -> zero_grad -> camera steps
        finally:
            if grad is not None:
                p.grad = grad
                grad.zero_()
            if keep is not None:
                with torch.no_grad():
                    p.copy_(keep[0])
                    st["exp_avg"].copy_(keep[1])
                    st["exp_avg_sq"].copy_(keep[2])

    def map(self, current_window, prune=False, iters=1):
        if len(current_window) == 0:
            return
        g = self.gaussians
        frames_to_optimize = self.config["Training"]["pose_window"]
        window_set = set(current_window)
        random_ids = [idx for idx in self.viewpoints if idx not in window_set]
        n_window = len(current_window)
        gaussian_split = False
        for _ in range(iters):
            self.iteration_count += 1
            self.last_sent += 1
            picks = torch.randperm(len(random_ids), generator=self._gen)[:2].tolist()
            # (the two extras in ascending id order: the loss is a sum, and a canonical order lets an
            # unchanged view set reuse its window / captured graphs)
            ids = list(current_window) + sorted(random_ids[i] for i in picks)
            win = self._window(ids, n_window, frames_to_optimize, stats=not prune)
            update_gaussian = (self.iteration_count % self.gaussian_update_every == self.gaussian_update_offset)
            reset = (self.iteration_count % self.gaussian_reset) == 0 and not update_gaussian
            gaussian_split = False
            if not (prune or update_gaussian or reset):
                # the plain iteration: render -> loss -> backward -> Adam -> pose updates, as one
                # (replayable) step
                win.step()
                g.update_learning_rate(self.iteration_count)
                self._last_ids = (ids, n_window, win)
                continue
            win.forward_backward()
            self._last_ids = (ids, n_window, win)
            self._set_visibility()              # before the map changes size under the masks
            if prune:
                if n_window == self.config["Training"]["window_size"]:
                    self._prune(current_window)
                return False
            if update_gaussian:
                self._densify(self.gaussian_th, self.gaussian_extent, self.size_threshold)
                gaussian_split = True
                # fresh Parameters: the Gaussian step is void; cameras still step (:309-316)
                g.zero_grad()
                win.update(gaussians=False)
                self._win_key = None
            else:
                seen = self._gather_visibility(win)
                g.reset_opacity_nonvisible([seen])
                gaussian_split = True
                self._step_gaussians(win, skip_opacity=True)
            g.update_learning_rate(self.iteration_count)
        if self._vis_iter != self.iteration_count:
            self._set_visibility()
        self.sync_keyframes(list(current_window))
        return gaussian_split

    def color_refinement(self, iteration_total=26000):
        """:320-355 — one random keyframe per iteration, loss (1 - lambda) L1 + lambda (1 - SSIM) on the
        raw render (no exposure), Gaussian Adam step, xyz learning-rate schedule on the refinement's own
        iteration count; densification statistics are left alone except max_radii2D.  With several ranks
        the iterations are dealt round-robin (one view per iteration does not shard); the gradient
        exchange of the optimizer step keeps the replicas identical."""
        from . import mapping
        from .losses import l1_loss, ssim
        g = self.gaussians
        lam = getattr(self.opt_params, "lambda_dssim", 0.2)
        render = self._render_fn or functools.partial(mapping.render_fused, with_stats=False)
        ids = list(self.viewpoints.keys())
        world, rank = self._world_rank()
        peer = getattr(g, "peer", None)
        loss = None
        for iteration in range(1, iteration_total + 1):
            kf = self.viewpoints[ids[self._py_rng.randint(0, len(ids) - 1)]]
            if iteration % world == rank:
                pkg = render(kf["cam"], g, None, self.background)
                image = pkg["render"]
                loss = (1.0 - lam) * l1_loss(image, kf["gt_image"]) + lam * (1.0 - ssim(image, kf["gt_image"]))
                loss.backward()
                with torch.no_grad():
                    vis = pkg["visibility_filter"]
                    mr = g._stats()[2]
                    mr[vis] = torch.max(mr[vis], pkg["radii"][vis].to(mr.dtype))
            with torch.no_grad():
                if world > 1 and peer is None:
                    dist.all_reduce(g.flat_grad, op=dist.ReduceOp.SUM, group=self._group)
                g.optimizer.step()
                if peer is None:
                    g.zero_grad()
                g.update_learning_rate(iteration)
        self._win_key = None
        return None if loss is None else loss.detach()

    def _set_visibility(self):
        """occ_aware_visibility = {kf: n_touched > 0} of the last iteration's window views (:232-236)"""
        ids, n_window, win = self._last_ids
        self._vis_iter = self.iteration_count
        win.flush()
        self.occ_aware_visibility = self._gather_touched(win, ids, n_window)

    def _prune(self, current_window):
        """:241-273 — n_obs = in how many window views a Gaussian was touched; "odometry" prunes
        n_obs < 3, "slam" prunes n_obs <= 3 among the Gaussians seeded by the three newest
        keyframes (all of them before initialisation); applied for monocular input only."""
        g = self.gaussians
        prune_mode = self.config["Training"]["prune_mode"]
        prune_coviz = 3
        n_obs = torch.zeros_like(next(iter(self.occ_aware_visibility.values())))
        for visibility in self.occ_aware_visibility.values():
            n_obs += visibility
        g.n_obs = n_obs.to(torch.int32)
        to_prune = None
        if prune_mode == "odometry":
            to_prune = n_obs < 3
        if prune_mode == "slam":
            sorted_window = sorted(current_window, reverse=True)
            mask = g.unique_kfIDs.to(n_obs.device) >= sorted_window[2]
            if not self.initialized:
                mask = g.unique_kfIDs.to(n_obs.device) >= 0
            to_prune = torch.logical_and(n_obs <= prune_coviz, mask)
        if to_prune is not None and self.monocular:
            # fresh arenas, as the reference's fresh Parameters: no gradient leaks past a real prune
            g.prune_points(to_prune)
            self._win_key = None
            for kf_idx in current_window:
                self.occ_aware_visibility[kf_idx] = self.occ_aware_visibility[kf_idx][~to_prune]
        if not self.initialized:
            self.initialized = True
