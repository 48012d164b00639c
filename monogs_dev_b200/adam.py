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
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        if lr < 0.0 or eps < 0.0:
            raise ValueError("invalid lr / eps")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        # group by (device, step, betas, eps): one launch per distinct combination
        buckets = {}
        for group in self.param_groups:
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
                g = p.grad if p.grad.is_contiguous() else p.grad.contiguous()
                key = (p.device, step, float(beta1), float(beta2), float(group["eps"]))
                buckets.setdefault(key, []).append(
                    (p, g, state["exp_avg"], state["exp_avg_sq"], float(group["lr"])))
        for (dev, step, beta1, beta2, eps), items in buckets.items():
            items = [it for it in items if it[0].numel() > 0]
            if not items:
                continue
            segs = (_lib.B2gsAdamSegment * len(items))()
            for s, (p, g, m, v, lr) in zip(segs, items):
                s.param, s.grad, s.exp_avg, s.exp_avg_sq = (p.data_ptr(), g.data_ptr(),
                                                            m.data_ptr(), v.data_ptr())
                s.n, s.lr = p.numel(), lr
            with torch.cuda.device(dev):
                check(lib.b2gs_adam_step(segs, len(items), step, beta1, beta2, eps,
                                         torch.cuda.current_stream().cuda_stream))
        return loss
