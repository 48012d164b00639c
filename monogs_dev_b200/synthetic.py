"""Synthetic scenes wrapped in the product's camera class — test / bench plumbing.

The generator itself lives in the neutral top-level `synthdata` package (it feeds the reference
arm and the oracle too, which must not import this package); here its poses become
`monogs_dev_b200.camera.Camera` objects on the requested device."""
from types import SimpleNamespace

import torch

import synthdata
from synthdata import CONFIGS, se3_exp, skew  # noqa: F401  (re-exported for the tests)
from .camera import Camera, projection_matrix  # noqa: F401


def make_scene(name=None, *, device="cpu", n_views=1, **kw):
    """Raw (pre-activation) Gaussian parameters + cameras, per SURVEY.md §8d."""
    t = synthdata.scene_tensors(name, n_views=n_views, **kw)
    to = lambda x: x.to(device).contiguous()
    cams = [Camera(t.W, t.H, t.fx, t.fy, t.cx, t.cy, w2c=m, device=device) for m in t.w2c]
    return SimpleNamespace(P=t.P, W=t.W, H=t.H, sh_degree=t.sh_degree, xyz=to(t.xyz),
                           scaling=to(t.scaling), rotation=to(t.rotation), opacity=to(t.opacity),
                           f_dc=to(t.f_dc), f_rest=to(t.f_rest), cameras=cams, generator=t.generator,
                           bg=torch.zeros(3, device=device))


def activated(scene):
    """the activations render() applies before the rasterizer
    (gaussian_splatting/scene/gaussian_model.py:76-101)."""
    return dict(means3D=scene.xyz, opacities=torch.sigmoid(scene.opacity),
                scales=torch.exp(scene.scaling),
                rotations=torch.nn.functional.normalize(scene.rotation),
                shs=torch.cat([scene.f_dc, scene.f_rest], dim=1).contiguous())


def upstream_grads(scene, device=None):
    """kernel-level upstream gradients: dL/dcolor ~ N(0,1)/(3N), dL/ddepth ~ N(0,1)/N."""
    return synthdata.upstream_grads(scene.W, scene.H, device or scene.xyz.device)
