"""Drop-in `simple_knn`: `from simple_knn._C import distCUDA2` as the reference callers do
(gs3dgs/scene/gaussian_model.py:22,177; gs2dgs/scene/gaussian_model.py:20,135)."""
from . import _C  # noqa: F401
