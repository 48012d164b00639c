"""Densify / prune on the packed parameter arena (BASELINE config 4).

Same decisions and the same resulting point order as the reference's
  densify_and_prune      gaussian_splatting/scene/gaussian_model.py:663-689
  densify_and_clone      :628-661
  densify_and_split      :593-626   (N = 2, scale / (0.8 N), samples ~ N(0, scale) rotated by R(q))
  prune_points           :500-517
  cat_tensors_to_optimizer / _prune_optimizer  :468-555 (new rows get zero Adam moments,
                         kept rows keep theirs)
but as ONE plan: the reference materialises the whole model four times (clone cat, split cat,
split prune, final prune — each re-wrapping 6 Parameters + 12 moment tensors and bouncing masks
through the CPU for unique_kfIDs / n_obs); here all masks are evaluated on the device, the final
row list [kept originals | kept clones | kept split children] is computed once, and every arena is
gathered once.  The only host sync is the size of the new arena.

Reference quirk reproduced (q9): densification_postfix zeroes max_radii2D before the final prune
(:583-585), so the `max_screen_size` test never fires after a densify; it is kept for fidelity."""
import torch

from .mapping import _GROUPS


def build_rotation(r):
    """gaussian_splatting/utils/general_utils.py:113-135 (normalises the quaternion)."""
    q = r / torch.sqrt((r * r).sum(1))[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - w * z), 2 * (x * z + w * y),
                     2 * (x * y + w * z), 1 - 2 * (x * x + z * z), 2 * (y * z - w * x),
                     2 * (x * z - w * y), 2 * (y * z + w * x), 1 - 2 * (x * x + y * y)], dim=1)
    return R.view(-1, 3, 3)


def _state(g):
    par = {k: g._params[k].detach() for k in _GROUPS}
    mom = {k: (g.optimizer.state[g._params[k]]["exp_avg"], g.optimizer.state[g._params[k]]["exp_avg_sq"])
           for k in _GROUPS}
    steps = {k: g.optimizer.state[g._params[k]]["step"] for k in _GROUPS}
    return par, mom, steps


def _extra(g, idx_fn):
    """per-Gaussian bookkeeping that rides along (kept on the DEVICE here; the reference keeps
    unique_kfIDs / n_obs on the CPU, gaussian_model.py:48-49,520-521)."""
    for a in ("unique_kfIDs", "n_obs"):
        if hasattr(g, a):
            setattr(g, a, idx_fn(getattr(g, a)))


def prune_points(g, mask):
    """keep rows where mask is False (gaussian_model.py:500-517)."""
    keep = torch.nonzero(~mask.reshape(-1), as_tuple=False).squeeze(1)
    par, mom, steps = _state(g)
    stats = [getattr(g, a)[keep].clone() for a in ("xyz_gradient_accum", "denom", "max_radii2D")] \
        if hasattr(g, "xyz_gradient_accum") else None
    g._allocate({k: par[k][keep] for k in _GROUPS},
                {k: (mom[k][0][keep], mom[k][1][keep]) for k in _GROUPS}, steps)
    if stats is not None:
        g.xyz_gradient_accum, g.denom, g.max_radii2D = stats
    _extra(g, lambda t: t[keep.to(t.device)])
    return int(keep.numel())


@torch.no_grad()
def _densify_and_prune_kernels(g, max_grad, min_opacity, extent, max_screen_size, N, samples, generator):
    """densify_and_prune as libb2gs kernels (csrc/densify.cu): plan -> one host read of the totals ->
    the surviving rows of every group of the parameter / exp_avg / exp_avg_sq arenas written straight
    to their final position in freshly allocated arenas."""
    import ctypes as C
    from . import _lib
    from ._lib import lib, check, ptr
    acc, den, _ = g._stats()
    dev = g.flat.device
    P = g._xyz.shape[0]
    par, mom, steps = _state(g)
    a = _lib.B2gsDensifyArgs()
    a.P, a.n_groups = P, len(_GROUPS)
    for i, k in enumerate(_GROUPS):
        a.param[i], a.exp_avg[i], a.exp_avg_sq[i] = ptr(par[k]), ptr(mom[k][0]), ptr(mom[k][1])
        a.width[i] = par[k].numel() // max(P, 1)
    a.xyz, a.opacity, a.scaling, a.rotation = (_GROUPS.index(k) for k in ("xyz", "opacity", "scaling", "rotation"))
    a.grad_accum, a.denom = acc.data_ptr(), den.data_ptr()
    a.max_grad, a.min_opacity = float(max_grad), float(min_opacity)
    a.dense_threshold = g.percent_dense * extent              # rounded to fp32 once, as torch rounds the scalar
    a.world_size_threshold = 0.1 * extent
    a.max_screen_size_on, a.N, a.split_divisor = int(bool(max_screen_size)), int(N), 0.8 * N
    aux_old = [getattr(g, n, None) for n in ("unique_kfIDs", "n_obs")]
    aux_old = [t.to(dev).int().contiguous() if t is not None else None for t in aux_old]
    for i, t in enumerate(aux_old):
        a.aux[i] = ptr(t)
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        ws = torch.empty(lib.b2gs_densify_bytes(P), dtype=torch.uint8, device=dev)
        totals = torch.zeros(8, dtype=torch.int64, device=dev)
        check(lib.b2gs_densify_plan(C.byref(a), ws.data_ptr(), ws.numel(), totals.data_ptr(), st))
        tot = totals.cpu().tolist()                           # the one host sync: sizes the new arenas
        n_o, n_c, n_s, nc, ns = tot[:5]
        P_new = n_o + n_c + N * n_s
        if ns and samples is None:
            samples = torch.randn(N * ns, 3, device=dev, generator=generator)
        samples = samples.to(dev).float().contiguous() if ns else None
        old = (g.flat, g.flat_m, g.flat_v)                    # the kernels read these: keep them alive
        shapes = {k: (P_new,) + tuple(par[k].shape[1:]) for k in _GROUPS}
        g._allocate(None, None, steps, shapes=shapes, device=dev)
        new_p = (C.c_void_p * 8)(*[ptr(g._params[k]) for k in _GROUPS])
        new_m = (C.c_void_p * 8)(*[ptr(g.optimizer.state[g._params[k]]["exp_avg"]) for k in _GROUPS])
        new_v = (C.c_void_p * 8)(*[ptr(g.optimizer.state[g._params[k]]["exp_avg_sq"]) for k in _GROUPS])
        aux_new = [torch.empty(P_new, dtype=torch.int32, device=dev) if t is not None else None for t in aux_old]
        new_a = (C.c_void_p * 2)(*[ptr(t) for t in aux_new])
        th = (C.c_int64 * 8)(*tot)
        check(lib.b2gs_densify_apply(C.byref(a), ws.data_ptr(), th, ptr(samples), new_p, new_m, new_v, new_a, st))
    for n, t in zip(("unique_kfIDs", "n_obs"), aux_new):
        if t is not None:
            setattr(g, n, t)
    g._stats()                                                # zeros of the new size
    del old
    return nc, ns, (P - ns + nc + N * ns) - P_new


@torch.no_grad()
def densify_and_prune(g, max_grad, min_opacity, extent, max_screen_size, N=2, samples=None, generator=None):
    """Returns (n_clone, n_split, n_pruned_final).  `samples` ([n_split * N, 3] unit normals) can be
    injected for parity tests; otherwise drawn on the device.  On the GPU the whole sequence runs as
    three libb2gs kernels; the torch index plan below is the host-logic (CPU) path."""
    if g.flat.is_cuda and getattr(g, "peer", None) is None:
        return _densify_and_prune_kernels(g, max_grad, min_opacity, extent, max_screen_size, N, samples, generator)
    acc, den, max_radii = g._stats()
    P = g._xyz.shape[0]
    dev = g.flat.device
    grads = acc / den
    grads[grads.isnan()] = 0.0
    grads = grads.reshape(-1)
    par, mom, steps = _state(g)
    scaling = torch.exp(par["scaling"])
    smax = scaling.max(dim=1).values
    big = smax > g.percent_dense * extent
    sel = grads >= max_grad                       # torch.norm of a [P,1] row == |value| == value (>= 0)
    clone_idx = torch.nonzero(sel & ~big).squeeze(1)
    split_idx = torch.nonzero(sel & big).squeeze(1)     # clones carry zero padded_grad: never split
    nc, ns = int(clone_idx.numel()), int(split_idx.numel())

    # ---- split children (gaussian_model.py:606-619) ----
    stds = scaling[split_idx].repeat(N, 1)
    if samples is None:
        samples = torch.randn(stds.shape, device=dev, generator=generator)
    samples = samples.to(dev) * stds                                   # torch.normal(mean=0, std=stds)
    rots = build_rotation(par["rotation"][split_idx]).repeat(N, 1, 1)
    child_xyz = torch.bmm(rots, samples.unsqueeze(-1)).squeeze(-1) + par["xyz"][split_idx].repeat(N, 1)
    child_scaling = torch.log(scaling[split_idx].repeat(N, 1) / (0.8 * N))
    child_src = split_idx.repeat(N)

    # ---- candidate rows in the reference's final order: originals minus split parents | clones | children
    orig_keep = torch.ones(P, dtype=torch.bool, device=dev)
    orig_keep[split_idx] = False
    src = torch.cat([torch.nonzero(orig_keep).squeeze(1), clone_idx, child_src])
    n_orig = P - ns
    is_child = torch.zeros(src.numel(), dtype=torch.bool, device=dev)
    is_child[n_orig + nc:] = True
    new = {k: par[k][src] for k in _GROUPS}
    new["xyz"][is_child] = child_xyz
    new["scaling"][is_child] = child_scaling

    # ---- final prune (gaussian_model.py:676-689); max_radii2D was reset by the postfix (quirk q9)
    prune = (torch.sigmoid(new["opacity"]) < min_opacity).reshape(-1)
    if max_screen_size:
        big_vs = torch.zeros_like(prune)                               # max_radii2D == 0 > max_screen_size: False
        big_ws = torch.exp(new["scaling"]).max(dim=1).values > 0.1 * extent
        prune = prune | big_vs | big_ws
    keep = torch.nonzero(~prune).squeeze(1)
    is_new = torch.ones(src.numel(), dtype=torch.bool, device=dev)
    is_new[:n_orig] = False
    fresh = is_new[keep]
    src_k = src[keep]
    moments = {}
    for k in _GROUPS:
        m, v = mom[k][0][src_k], mom[k][1][src_k]
        m[fresh] = 0                                                    # cat_tensors_to_optimizer: zeros
        v[fresh] = 0
        moments[k] = (m, v)
    g._allocate({k: new[k][keep] for k in _GROUPS}, moments, steps)
    g._stats()                                                          # zeros of the new size
    _extra(g, lambda t: t[src_k.to(t.device)])
    return nc, ns, int(src.numel() - keep.numel())
