"""Drop-in import name: `from diff_gaussian_rasterization import GaussianRasterizationSettings,
GaussianRasterizer` (gaussian_splatting/gaussian_renderer/__init__.py:15-18 of the reference)."""
from monogs_dev_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,
                                        rasterize_gaussians, _RasterizeGaussians)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians"]
