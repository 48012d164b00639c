"""distCUDA2: mean squared distance to the 3 nearest neighbours.

Mirror of `simple_knn._C.distCUDA2` (submodules/simple-knn/spatial.cu:15-26, called at
gaussian_splatting/scene/gaussian_model.py:185): float32 CUDA [P,3] -> float32 CUDA [P].
"""
import torch

from ._lib import lib, check


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    if not points.is_cuda:
        raise RuntimeError("distCUDA2 needs a CUDA tensor: libb2gs has no CPU path")
    if points.dim() != 2 or points.size(1) != 3:
        raise RuntimeError("points must have dimensions (num_points, 3)")
    pts = points.detach().to(torch.float32).contiguous()
    P = pts.size(0)
    out = torch.zeros(P, dtype=torch.float32, device=pts.device)
    if P == 0:
        return out
    with torch.cuda.device(pts.device):
        ws = torch.empty(lib.b2gs_knn_bytes(P), dtype=torch.uint8, device=pts.device)
        check(lib.b2gs_knn3_dist2(pts.data_ptr(), P, out.data_ptr(), ws.data_ptr(), ws.numel(),
                                  torch.cuda.current_stream().cuda_stream))
    return out
