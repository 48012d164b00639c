"""`Camera`: the part of the reference's utils/camera_utils.py:Camera the rasterizer path reads, with
its pose-derived matrices in one device block (see synthetic.Camera); `Camera.from_reference(vp)`
wraps a MonoGS viewpoint for MappingWindow / TrackingLoop / BackEnd / FrontEnd."""
from .synthetic import Camera, projection_matrix, se3_exp  # noqa: F401
