"""Drop-in import name: `from simple_knn._C import distCUDA2`
(gaussian_splatting/scene/gaussian_model.py:18 of the reference)."""
from monogs_dev_b200.knn import distCUDA2

__all__ = ["distCUDA2"]
