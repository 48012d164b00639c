"""TEST INFRASTRUCTURE — literal restatement of the reference's densify / prune sequence on plain
tensors (gaussian_splatting/scene/gaussian_model.py:468-689): every intermediate model the
reference materialises is materialised here too (clone cat -> split cat -> split prune -> final
prune), with the Adam moments carried the way cat_tensors_to_optimizer / _prune_optimizer do.
Only tests/ import this; the product path is monogs_dev_b200/densify.py."""
import torch

KEYS = ("xyz", "f_dc", "f_rest", "opacity", "scaling", "rotation")


def build_rotation(r):  # gaussian_splatting/utils/general_utils.py:113-135
    norm = torch.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    R = torch.zeros((q.size(0), 3, 3))
    r_, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R[:, 0, 0] = 1 - 2 * (y * y + z * z); R[:, 0, 1] = 2 * (x * y - r_ * z); R[:, 0, 2] = 2 * (x * z + r_ * y)
    R[:, 1, 0] = 2 * (x * y + r_ * z); R[:, 1, 1] = 1 - 2 * (x * x + z * z); R[:, 1, 2] = 2 * (y * z - r_ * x)
    R[:, 2, 0] = 2 * (x * z - r_ * y); R[:, 2, 1] = 2 * (y * z + r_ * x); R[:, 2, 2] = 1 - 2 * (x * x + y * y)
    return R


class Model:
    def __init__(self, params, m, v, accum, denom, max_radii, percent_dense=0.01):
        self.p = {k: params[k].clone() for k in KEYS}
        self.m = {k: m[k].clone() for k in KEYS}
        self.v = {k: v[k].clone() for k in KEYS}
        self.xyz_gradient_accum, self.denom, self.max_radii2D = accum.clone(), denom.clone(), max_radii.clone()
        self.percent_dense = percent_dense

    scaling = property(lambda s: torch.exp(s.p["scaling"]))
    opacity = property(lambda s: torch.sigmoid(s.p["opacity"]))

    def prune_points(self, mask):                                   # :500-517, _prune_optimizer :483-498
        valid = ~mask
        for d in (self.p, self.m, self.v):
            for k in KEYS:
                d[k] = d[k][valid]
        self.xyz_gradient_accum = self.xyz_gradient_accum[valid]
        self.denom = self.denom[valid]
        self.max_radii2D = self.max_radii2D[valid]

    def postfix(self, new):                                         # :556-589, cat_tensors_to_optimizer :519-554
        for k in KEYS:
            self.p[k] = torch.cat((self.p[k], new[k]), dim=0)
            self.m[k] = torch.cat((self.m[k], torch.zeros_like(new[k])), dim=0)
            self.v[k] = torch.cat((self.v[k], torch.zeros_like(new[k])), dim=0)
        n = self.p["xyz"].shape[0]
        self.xyz_gradient_accum = torch.zeros((n, 1))
        self.denom = torch.zeros((n, 1))
        self.max_radii2D = torch.zeros((n))

    def densify_and_clone(self, grads, thr, extent):                # :628-661
        sel = torch.where(torch.norm(grads, dim=-1) >= thr, True, False)
        sel = torch.logical_and(sel, torch.max(self.scaling, dim=1).values <= self.percent_dense * extent)
        self.postfix({k: self.p[k][sel] for k in KEYS})
        return int(sel.sum())

    def densify_and_split(self, grads, thr, extent, unit_samples, N=2):   # :593-626
        n_init = self.p["xyz"].shape[0]
        padded = torch.zeros((n_init))
        padded[: grads.shape[0]] = grads.squeeze()
        sel = torch.where(padded >= thr, True, False)
        sel = torch.logical_and(sel, torch.max(self.scaling, dim=1).values > self.percent_dense * extent)
        stds = self.scaling[sel].repeat(N, 1)
        samples = unit_samples * stds                                # torch.normal(mean=0, std=stds)
        rots = build_rotation(self.p["rotation"][sel]).repeat(N, 1, 1)
        new = {k: self.p[k][sel].repeat(N, *([1] * (self.p[k].dim() - 1))) for k in KEYS}
        new["xyz"] = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + self.p["xyz"][sel].repeat(N, 1)
        new["scaling"] = torch.log(self.scaling[sel].repeat(N, 1) / (0.8 * N))
        self.postfix(new)
        self.prune_points(torch.cat((sel, torch.zeros(N * int(sel.sum()), dtype=bool))))
        return int(sel.sum())

    def densify_and_prune(self, max_grad, min_opacity, extent, max_screen_size, unit_samples):   # :663-689
        grads = self.xyz_gradient_accum / self.denom
        grads[grads.isnan()] = 0.0
        nc = self.densify_and_clone(grads, max_grad, extent)
        ns = self.densify_and_split(grads, max_grad, extent, unit_samples)
        prune_mask = (self.opacity < min_opacity).squeeze()
        if max_screen_size:
            big_vs = self.max_radii2D > max_screen_size
            big_ws = self.scaling.max(dim=1).values > 0.1 * extent
            prune_mask = torch.logical_or(torch.logical_or(prune_mask, big_vs), big_ws)
        n_pruned = int(prune_mask.sum())
        self.prune_points(prune_mask)
        return nc, ns, n_pruned
