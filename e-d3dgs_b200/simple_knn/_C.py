"""Drop-in for ``from simple_knn._C import distCUDA2`` (scene/gaussian_model.py:21)."""
from ed3dgs_b200.knn import distCUDA2  # noqa: F401
