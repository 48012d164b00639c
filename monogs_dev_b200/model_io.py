"""GaussianModel bookkeeping around the mapping loop, on the packed arena: learning-rate schedule,
opacity resets, SH degree, PLY export / import (SURVEY.md §8a a20 callers, §8f rank 4 "PLY writer").

Mirrors, with the reference's names and argument meaning,
  helper / get_expon_lr_func          gaussian_splatting/utils/general_utils.py:42-94
  GaussianModel.training_setup        gaussian_splatting/scene/gaussian_model.py:245-294
  GaussianModel.update_learning_rate  :296-310    (utils/slam_backend.py:309,352)
  GaussianModel.reset_opacity         :362-365    (utils/slam_backend.py:133)
  GaussianModel.reset_opacity_nonvisible :367-375 (utils/slam_backend.py:304)
  GaussianModel.save_ply / load_ply   :312-360, 377-466 (utils/eval_utils.py:192)

The reference replaces the opacity nn.Parameter and its two Adam moment tensors by new
allocations (replace_tensor_to_optimizer, :468-481); here the three slices of the packed arenas
are overwritten in place, so a captured mapping graph stays valid across a reset.  The PLY file is
written / parsed directly (binary little-endian, one float property per attribute, the layout
plyfile produces for the reference's structured array); plyfile itself is not a dependency.
"""
import math
import os

import numpy as np
import torch

RESET_OPACITY = 0.01              # gaussian_model.py:363
RESET_OPACITY_NONVISIBLE = 0.4    # gaussian_model.py:370


def helper(step, lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """log-linear decay lr_init -> lr_final over max_steps, optional delayed start
    (gaussian_splatting/utils/general_utils.py:79-94)."""
    if step < 0 or (lr_init == 0.0 and lr_final == 0.0):
        return 0.0
    if lr_delay_steps > 0:
        ramp = min(max(step / lr_delay_steps, 0.0), 1.0)
        delay_rate = lr_delay_mult + (1 - lr_delay_mult) * math.sin(0.5 * math.pi * ramp)
    else:
        delay_rate = 1.0
    t = min(max(step / max_steps, 0.0), 1.0)
    return delay_rate * math.exp(math.log(lr_init) * (1 - t) + math.log(lr_final) * t)


def get_expon_lr_func(lr_init, lr_final, lr_delay_steps=0, lr_delay_mult=1.0, max_steps=1000000):
    """general_utils.py:42-74 returns the bare `helper`; kept callable the same way."""
    return helper


def training_setup(model, training_args):
    """learning rates of the six groups + the xyz schedule from an opt_params-like object
    (attributes as in gaussian_splatting/arguments/__init__.py: OptimizationParams)."""
    s = model.spatial_lr_scale
    model.percent_dense = training_args.percent_dense
    lr = dict(xyz=training_args.position_lr_init * s, f_dc=training_args.feature_lr,
              f_rest=training_args.feature_lr / 20.0, opacity=training_args.opacity_lr,
              scaling=training_args.scaling_lr * s, rotation=training_args.rotation_lr)
    for grp in model.optimizer.param_groups:
        grp["lr"] = lr[grp["name"]]
    model._lr = lr
    model.xyz_scheduler_args = get_expon_lr_func(
        lr_init=lr["xyz"], lr_final=training_args.position_lr_final * s,
        lr_delay_mult=training_args.position_lr_delay_mult, max_steps=training_args.position_lr_max_steps)
    model.lr_init = lr["xyz"]
    model.lr_final = training_args.position_lr_final * s
    model.lr_delay_mult = training_args.position_lr_delay_mult
    model.max_steps = training_args.position_lr_max_steps
    model._stats()


def update_learning_rate(model, iteration):
    """xyz group only, as in the reference (:296-310).  With a capturable FusedAdam the new rate
    reaches the captured kernel through the device lr array (FusedAdam.sync_lr before a replay)."""
    for grp in model.optimizer.param_groups:
        if grp["name"] == "xyz":
            lr = helper(iteration, lr_init=model.lr_init, lr_final=model.lr_final,
                        lr_delay_mult=model.lr_delay_mult, max_steps=model.max_steps)
            grp["lr"] = lr
            return lr


def inverse_sigmoid(x):
    return torch.log(x / (1 - x))            # gaussian_splatting/utils/general_utils.py:22-23


def _replace_opacity(model, new):
    """replace_tensor_to_optimizer(new, "opacity") on the arena: value in place, both moments
    zeroed, step count kept (:468-481 keeps stored_state["step"])."""
    p = model._params["opacity"]
    st = model.optimizer.state[p]
    with torch.no_grad():
        p.copy_(new)
        st["exp_avg"].zero_()
        st["exp_avg_sq"].zero_()


def reset_opacity(model):
    _replace_opacity(model, inverse_sigmoid(torch.ones_like(model.get_opacity) * RESET_OPACITY))


def reset_opacity_nonvisible(model, visibility_filters):
    """Gaussians seen by none of the filters get opacity 0.4; the visible ones are assigned their
    ACTIVATED opacity as the new raw value — the reference's own behaviour (:373 copies
    get_opacity, not _opacity, into the pre-sigmoid tensor), kept as is."""
    with torch.no_grad():
        act = model.get_opacity
        new = inverse_sigmoid(torch.ones_like(act) * RESET_OPACITY_NONVISIBLE)
        for f in visibility_filters:
            new[f] = act[f]
    _replace_opacity(model, new)


def oneupSHdegree(model):
    if model.active_sh_degree < model.max_sh_degree:
        model.active_sh_degree += 1


# -- PLY --------------------------------------------------------------------------------------

def construct_list_of_attributes(model):
    names = ["x", "y", "z", "nx", "ny", "nz"]
    names += ["f_dc_%d" % i for i in range(model._features_dc.shape[1] * model._features_dc.shape[2])]
    names += ["f_rest_%d" % i for i in range(model._features_rest.shape[1] * model._features_rest.shape[2])]
    names.append("opacity")
    names += ["scale_%d" % i for i in range(model._scaling.shape[1])]
    names += ["rot_%d" % i for i in range(model._rotation.shape[1])]
    return names


def save_ply(model, path):
    """one `vertex` element, every attribute a little-endian float32 property; SH coefficients are
    stored channel-major ([P,M,3] -> transpose(1,2) -> flatten), normals are zeros (:326-360)."""
    d = os.path.dirname(path)
    if d:
        os.makedirs(d, exist_ok=True)
    host = lambda t: t.detach().float().cpu().numpy()
    xyz = host(model._xyz)
    P = xyz.shape[0]
    cols = [xyz, np.zeros_like(xyz),
            host(model._features_dc.detach().transpose(1, 2).flatten(start_dim=1).contiguous()).reshape(P, -1),
            host(model._features_rest.detach().transpose(1, 2).flatten(start_dim=1).contiguous()).reshape(P, -1),
            host(model._opacity), host(model._scaling), host(model._rotation)]
    names = construct_list_of_attributes(model)
    table = np.ascontiguousarray(np.concatenate(cols, axis=1), dtype="<f4")
    assert table.shape == (P, len(names))
    header = ["ply", "format binary_little_endian 1.0", "element vertex %d" % P]
    header += ["property float %s" % n for n in names]
    header.append("end_header")
    with open(path, "wb") as f:
        f.write(("\n".join(header) + "\n").encode("ascii"))
        f.write(table.tobytes())


_PLY_TYPES = {"char": "i1", "int8": "i1", "uchar": "u1", "uint8": "u1", "short": "i2", "int16": "i2",
              "ushort": "u2", "uint16": "u2", "int": "i4", "int32": "i4", "uint": "u4", "uint32": "u4",
              "float": "f4", "float32": "f4", "double": "f8", "float64": "f8"}


def read_ply_vertices(path):
    """-> structured numpy array of the FIRST element of a PLY file (ascii or binary, scalar
    properties only — what plydata.elements[0] is for the files save_ply writes)."""
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise ValueError("%s: not a PLY file" % path)
        fmt, count, props, seen = None, None, [], 0
        while True:
            line = f.readline()
            if not line:
                raise ValueError("%s: header without end_header" % path)
            tok = line.decode("ascii").split()
            if not tok or tok[0] == "comment" or tok[0] == "obj_info":
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                seen += 1
                if seen == 1:
                    count = int(tok[2])
            elif tok[0] == "property" and seen == 1:
                if tok[1] == "list":
                    raise ValueError("%s: list properties are not supported" % path)
                props.append((tok[2], _PLY_TYPES[tok[1]]))
            elif tok[0] == "end_header":
                break
        if fmt is None or count is None:
            raise ValueError("%s: incomplete PLY header" % path)
        if fmt == "ascii":
            rows = np.loadtxt(f, max_rows=count, ndmin=2, dtype=np.float64)
            out = np.empty(count, dtype=[(n, t) for n, t in props])
            for i, (n, _) in enumerate(props):
                out[n] = rows[:, i]
            return out
        order = {"binary_little_endian": "<", "binary_big_endian": ">"}[fmt]
        dt = np.dtype([(n, order + t) for n, t in props])
        raw = f.read(count * dt.itemsize)
        if len(raw) != count * dt.itemsize:
            raise ValueError("%s: truncated vertex data" % path)
        return np.frombuffer(raw, dtype=dt, count=count)


def load_ply_tensors(path, max_sh_degree, device):
    """per-group tensors in the arena's shapes from a PLY written by save_ply (:377-466)."""
    v = read_ply_vertices(path)
    col = lambda names: np.stack([np.asarray(v[n], dtype=np.float32) for n in names], axis=1)
    by_index = lambda prefix: sorted((n for n in v.dtype.names if n.startswith(prefix)),
                                     key=lambda s: int(s.split("_")[-1]))
    rest_names = by_index("f_rest_")
    M = (max_sh_degree + 1) ** 2
    if len(rest_names) != 3 * M - 3:
        raise ValueError("PLY holds %d f_rest properties, max_sh_degree=%d needs %d"
                         % (len(rest_names), max_sh_degree, 3 * M - 3))
    P = v.shape[0]
    dc = col(["f_dc_0", "f_dc_1", "f_dc_2"]).reshape(P, 3, 1)
    rest = (col(rest_names) if rest_names else np.zeros((P, 0), np.float32)).reshape(P, 3, M - 1)
    t = lambda a: torch.tensor(np.ascontiguousarray(a), dtype=torch.float, device=device)
    return dict(xyz=t(col(["x", "y", "z"])),
                f_dc=t(dc).transpose(1, 2).contiguous(), f_rest=t(rest).transpose(1, 2).contiguous(),
                opacity=t(col(["opacity"])), scaling=t(col(by_index("scale_"))),
                rotation=t(col(by_index("rot"))))
