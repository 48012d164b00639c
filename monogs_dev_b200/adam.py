"""Fused Adam: one kernel launch for all parameter groups.

Stands in for `torch.optim.Adam(l, lr=0.0, eps=1e-15)` as GaussianModel builds it
(gaussian_splatting/scene/gaussian_model.py:245-283) and as BackEnd steps it
(utils/slam_backend.py:307-310).  It IS a torch.optim.Optimizer with Adam's
param_groups / state layout ("step", "exp_avg", "exp_avg_sq"), so the reference's
optimizer surgery (replace_tensor_to_optimizer / _prune_optimizer /
cat_tensors_to_optimizer, gaussian_model.py:468-555) and its per-group
`param_group["lr"] = ...` schedule (:296-310) keep working unchanged.
No weight decay, no amsgrad (the reference uses neither).
"""
import ctypes as C

import torch

from . import _lib
from ._lib import lib, check


class FusedAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, capturable=False):
        """capturable=True keeps ONE step counter per optimizer on the device
        (b2gs_adam_step_dev) so that step() can be captured in a CUDA graph and replayed;
        learning rates are then baked into the captured launch."""
        if lr < 0.0 or eps < 0.0:
            raise ValueError("invalid lr / eps")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        self.capturable = bool(capturable)
        self.peer = None        # peer.PeerArena: step() = all-reduce + Adam + zero_grad over NVLink peers
        self._step_dev = None
        self._lr_dev = None
        self._lr_host = None

    def sync_lr(self):
        """capturable mode: push the current param_group learning rates to the device array the
        captured kernel reads (call after changing param_group["lr"]; no-op if unchanged)."""
        lrs = tuple(float(g["lr"]) for g in self.param_groups)
        if not self.capturable or lrs == self._lr_host:
            return
        dev = self.param_groups[0]["params"][0].device
        if self._lr_dev is None or self._lr_dev.numel() != len(lrs):
            self._lr_dev = torch.zeros(len(lrs), dtype=torch.float32, device=dev)
        self._lr_dev.copy_(torch.tensor(lrs, dtype=torch.float32), non_blocking=False)
        self._lr_host = lrs

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # group by (device, step, betas, eps): one launch per distinct combination
        buckets = {}
        if self.capturable and not torch.cuda.is_current_stream_capturing():
            self.sync_lr()
        for gi, group in enumerate(self.param_groups):
            beta1, beta2 = group["betas"]
            for p in group["params"]:
                if p.grad is None:
                    continue
                if not p.is_cuda:
                    raise RuntimeError("FusedAdam needs CUDA parameters: libb2gs has no CPU path")
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise RuntimeError("FusedAdam needs contiguous float32 parameters")
                state = self.state[p]
                if len(state) == 0:
                    state["step"] = 0
                    state["exp_avg"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                    state["exp_avg_sq"] = torch.zeros_like(p, memory_format=torch.preserve_format)
                step = state["step"]
                step = int(step.item()) if isinstance(step, torch.Tensor) else int(step)
                step += 1
                state["step"] = step
                if self.capturable:
                    self._host_step = step
                    step = -1          # one device-side counter for the whole optimizer
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                key = (p.device, step, float(beta1), float(beta2), float(group["eps"]))
                buckets.setdefault(key, []).append(
                    (p, g, state["exp_avg"], state["exp_avg_sq"], float(group["lr"]), gi))
        for (dev, step, beta1, beta2, eps), items in buckets.items():
            items = [it for it in items if it[0].numel() > 0]
            if not items:
                continue
            segs = (_lib.B2gsAdamSegment * len(items))()
            for s, (p, g, m, v, lr, gi) in zip(segs, items):
                s.param, s.grad, s.exp_avg, s.exp_avg_sq = (p.data_ptr(), g.data_ptr(),
                                                            m.data_ptr(), v.data_ptr())
                s.n, s.lr = p.numel(), lr
                if self.capturable and self._lr_dev is not None:
                    s.lr_dev = self._lr_dev.data_ptr() + 4 * gi
            with torch.cuda.device(dev):
                stream = torch.cuda.current_stream().cuda_stream
                if self.capturable:
                    if len(buckets) != 1:
                        raise RuntimeError("capturable FusedAdam needs one (betas, eps) for all groups")
                    if self._step_dev is None:
                        # continues from the host-side count (state restored after densify / resume)
                        self._step_dev = torch.full((1,), int(getattr(self, "_host_step", 1)) - 1,
                                                    dtype=torch.int64, device=dev)
                    if self.peer is not None:
                        check(lib.b2gs_adam_step_peer(segs, len(items), self._step_dev.data_ptr(), beta1,
                                                      beta2, eps, self.peer.struct(), stream))
                    else:
                        check(lib.b2gs_adam_step_dev(segs, len(items), self._step_dev.data_ptr(), beta1,
                                                     beta2, eps, stream))
                elif self.peer is not None:
                    raise RuntimeError("the peer-memory step needs FusedAdam(capturable=True)")
                else:
                    check(lib.b2gs_adam_step(segs, len(items), step, beta1, beta2, eps, stream))
        return loss
