"""monogs_dev_b200 — B200-native (sm_100a) hot path of MonoGS behind the reference's own
operator API: the differentiable Gaussian rasterizer with pose gradients, distCUDA2 and a
fused Adam.  Everything computes in libb2gs.so (hand-written CUDA, C ABI in include/b2gs.h);
importing this package fails if the library is not built."""
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians,
                         _RasterizeGaussians)
from .knn import distCUDA2
from .adam import FusedAdam

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians",
           "distCUDA2", "FusedAdam"]
