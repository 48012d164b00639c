"""Tracking and keyframe management around the rasterizer: host-side mirror of the reference's
FrontEnd (utils/slam_frontend.py), same method names, argument meaning and return values
(SURVEY.md §8a a20).  These are the consumers of the rasterizer's `depth`, `opacity` and
`n_touched` outputs on the tracking side.

  FrontEnd.tracking        utils/slam_frontend.py:128-196   (pose-only loop on ONE GPU: TrackingLoop)
  FrontEnd.add_new_keyframe :57-109   (the depth map a new keyframe is seeded from)
  FrontEnd.is_keyframe     :198-226   (translation against the median depth, covisibility IoU)
  FrontEnd.add_to_window   :228-290   (overlap-coefficient cut-off, then the distance heuristic)
  FrontEnd.sync_backend    :307-315
  get_median_depth         utils/slam_utils.py:131-142

Not here: the queues / GUI packets / dataset loop of FrontEnd.run (control plane, out of scope).

B200-first: the reference builds a new Adam and syncs the host every iteration for its
convergence test; here ONE persistent camera block + ground-truth buffers carry every frame, so a
captured tracking iteration is replayed `tracking_itr_num` times per frame without a host sync
(`cuda_graph=True`; the eager loop keeps the reference's early break on convergence), and the
result is written back into the frame's own viewpoint.
"""
import math

import torch

from .adam import FusedAdam
from .mapping import TrackingLoop


def get_median_depth(depth, opacity=None, mask=None, return_std=False):
    depth = depth.detach().clone()
    valid = depth > 0
    if opacity is not None:
        valid = torch.logical_and(valid, opacity.detach() > 0.95)
    if mask is not None:
        valid = torch.logical_and(valid, mask)
    valid_depth = depth[valid]
    if return_std:
        return valid_depth.median(), valid_depth.std(), valid
    return valid_depth.median()


def _w2c(cam):
    """getWorld2View2(R, T) (gaussian_splatting/utils/graphics_utils.py:38-49, no recentring)."""
    m = torch.eye(4)
    m[:3, :3], m[:3, 3] = cam.R, cam.T
    return m


class FrontEnd:
    def __init__(self, config, gaussians=None, background=None, cuda_graph=False, render_fn=None,
                 loss_fn=None, adam_cls=FusedAdam):
        self.config = config
        self.gaussians = gaussians
        self.background = background
        self.monocular = config["Training"]["monocular"]
        self.initialized = False
        self.kf_indices = []
        self.occ_aware_visibility = {}
        self.current_window = []
        self.cameras = {}                  # frame idx -> dict(cam=, gt_image=, gt_depth=)
        self.use_every_n_frames = 1
        self.median_depth = 1.0
        self._graph, self._render_fn, self._loss_fn, self._adam = bool(cuda_graph), render_fn, loss_fn, adam_cls
        self._loop = None
        self.set_hyperparams()

    def set_hyperparams(self):
        t = self.config["Training"]
        self.tracking_itr_num = t["tracking_itr_num"]
        self.kf_interval = t.get("kf_interval", 5)
        self.window_size = t["window_size"]

    # -- tracking ---------------------------------------------------------------------------------
    def _tracking_loop(self, viewpoint):
        if self._loop is None or self._loop.g is not self.gaussians:
            from .camera import Camera
            c = viewpoint
            cam = Camera(c.image_width, c.image_height, c.fx, c.fy, c.cx, c.cy, device=self.background.device)
            lr = self.config["Training"]["lr"]
            kw = {}
            if self._render_fn is not None:
                kw["render_fn"] = self._render_fn
            if self._loss_fn is not None:
                kw["loss_fn"] = self._loss_fn
            self._loop = TrackingLoop(self.gaussians, cam, self.background, lr_rot=lr["cam_rot_delta"],
                                      lr_trans=lr["cam_trans_delta"], lr_exposure=0.01,
                                      iters=self.tracking_itr_num, cuda_graph=self._graph, adam_cls=self._adam, **kw)
        return self._loop

    def tracking(self, cur_frame_idx, viewpoint=None):
        """starts from the previous frame's pose, optimises (rot, trans, exposure a / b) for
        tracking_itr_num iterations (eager: until update_pose reports convergence), stores the median
        rendered depth and returns the last render package."""
        frame = self.cameras[cur_frame_idx] if viewpoint is None else viewpoint
        self.cameras[cur_frame_idx] = frame
        cam = frame["cam"]
        prev = self.cameras[cur_frame_idx - self.use_every_n_frames]["cam"]
        loop = self._tracking_loop(cam)
        loop.load_frame(prev.R, prev.T, frame["gt_image"], frame["gt_depth"],
                        exposure=(float(cam.exposure_a.detach()), float(cam.exposure_b.detach())))
        render_pkg = loop.track(self.tracking_itr_num, check_converged=not self._graph)
        cam.update_RT(loop.cam.R, loop.cam.T)
        with torch.no_grad():
            cam.exposure_a.copy_(loop.cam.exposure_a)
            cam.exposure_b.copy_(loop.cam.exposure_b)
        self.median_depth = get_median_depth(render_pkg["depth"], render_pkg["opacity"])
        return render_pkg

    # -- keyframe management ----------------------------------------------------------------------
    def add_new_keyframe(self, cur_frame_idx, depth=None, opacity=None, init=False, generator=None):
        """registers the keyframe and returns the [H,W] depth map its Gaussians are seeded from
        (GaussianModel.extend_from_pcd_seq(depthmap=...)): RGB-D -> the observed depth; monocular ->
        2 + 0.3 N(0,1) at initialisation, afterwards the rendered depth with outliers (beyond one
        standard deviation of the opaque pixels' median, or not opaque) set to the median, plus
        0.2 / 0.5 std of noise; pixels under the RGB boundary threshold are 0 (= not seeded).
        Stays on the device (the reference returns a numpy array that is uploaded again)."""
        thr = self.config["Training"]["rgb_boundary_threshold"]
        self.kf_indices.append(cur_frame_idx)
        frame = self.cameras[cur_frame_idx]
        gt_img = frame["gt_image"]
        valid_rgb = (gt_img.sum(dim=0) > thr)[None]
        randn = lambda like: torch.randn(like.shape, generator=generator, dtype=like.dtype,
                                         device=like.device if generator is None else generator.device).to(like.device)
        if self.monocular:
            if depth is None:
                initial_depth = 2 * torch.ones(1, gt_img.shape[1], gt_img.shape[2], device=gt_img.device)
                initial_depth = initial_depth + randn(initial_depth) * 0.3
            else:
                depth = depth.detach().clone()
                median_depth, std, valid_mask = get_median_depth(depth, opacity, mask=valid_rgb, return_std=True)
                invalid = torch.logical_or(depth > median_depth + std, depth < median_depth - std)
                invalid = torch.logical_or(invalid, ~valid_mask)
                depth[invalid] = median_depth
                initial_depth = depth + randn(depth) * torch.where(invalid, std * 0.5, std * 0.2)
            initial_depth[~valid_rgb] = 0
            return initial_depth[0]
        initial_depth = frame["gt_depth"].detach().clone().reshape(1, *gt_img.shape[1:])
        initial_depth[~valid_rgb] = 0
        return initial_depth[0]

    def is_keyframe(self, cur_frame_idx, last_keyframe_idx, cur_frame_visibility_filter, occ_aware_visibility):
        t = self.config["Training"]
        kf_translation, kf_min_translation, kf_overlap = t["kf_translation"], t["kf_min_translation"], t["kf_overlap"]
        pose_CW = _w2c(self.cameras[cur_frame_idx]["cam"])
        last_kf_WC = torch.linalg.inv(_w2c(self.cameras[last_keyframe_idx]["cam"]))
        dist = torch.norm((pose_CW @ last_kf_WC)[0:3, 3])
        median = float(self.median_depth)
        dist_check = bool(dist > kf_translation * median)
        dist_check2 = bool(dist > kf_min_translation * median)
        last = occ_aware_visibility[last_keyframe_idx].to(cur_frame_visibility_filter.device)
        union = torch.logical_or(cur_frame_visibility_filter, last).count_nonzero()
        intersection = torch.logical_and(cur_frame_visibility_filter, last).count_nonzero()
        point_ratio_2 = float(intersection / union)
        return (point_ratio_2 < kf_overlap and dist_check2) or dist_check

    def add_to_window(self, cur_frame_idx, cur_frame_visibility_filter, occ_aware_visibility, window):
        N_dont_touch = 2
        window = [cur_frame_idx] + window
        curr = self.cameras[cur_frame_idx]["cam"]
        # frames with little overlap with the current one (Szymkiewicz-Simpson coefficient): the LAST
        # such frame is dropped
        to_remove, removed_frame = [], None
        cut_off = self.config["Training"].get("kf_cutoff", 0.4) if self.initialized else 0.4
        n_cur = cur_frame_visibility_filter.count_nonzero()
        for i in range(N_dont_touch, len(window)):
            kf_idx = window[i]
            vis = occ_aware_visibility[kf_idx].to(cur_frame_visibility_filter.device)
            intersection = torch.logical_and(cur_frame_visibility_filter, vis).count_nonzero()
            denom = min(n_cur, vis.count_nonzero())
            if float(intersection / denom) <= cut_off:
                to_remove.append(kf_idx)
        if to_remove:
            window.remove(to_remove[-1])
            removed_frame = to_remove[-1]
        if len(window) > self.config["Training"]["window_size"]:
            # still too long: drop the frame that is closest to the others and far from the current one
            kf_0_WC = torch.linalg.inv(_w2c(curr))
            inv_dist = []
            for i in range(N_dont_touch, len(window)):
                kf_i_CW = _w2c(self.cameras[window[i]]["cam"])
                inv_dists = []
                for j in range(N_dont_touch, len(window)):
                    if i == j:
                        continue
                    kf_j_WC = torch.linalg.inv(_w2c(self.cameras[window[j]]["cam"]))
                    inv_dists.append(1.0 / (torch.norm((kf_i_CW @ kf_j_WC)[0:3, 3]) + 1e-6).item())
                k = math.sqrt(torch.norm((kf_i_CW @ kf_0_WC)[0:3, 3]).item())
                inv_dist.append(k * sum(inv_dists))
            idx = max(range(len(inv_dist)), key=lambda n: (inv_dist[n], -n))     # np.argmax: first maximum
            removed_frame = window[N_dont_touch + idx]
            window.remove(removed_frame)
        return window, removed_frame

    def sync_backend(self, data):
        """data = [tag, gaussians, occ_aware_visibility, [(kf_id, R, T), ...]] as pushed by the back end"""
        self.gaussians = data[1]
        self.occ_aware_visibility = data[2]
        for kf_id, kf_R, kf_T in data[3]:
            self.cameras[kf_id]["cam"].update_RT(kf_R.clone(), kf_T.clone())
