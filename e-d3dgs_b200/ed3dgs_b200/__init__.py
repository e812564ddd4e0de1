"""ed3dgs_b200 -- B200-native (sm_100a) implementation of the E-D3DGS per-step training hot path.

Public surface (mirrors the reference's names):
  GaussianRasterizationSettings, GaussianRasterizer   (diff_gaussian_rasterization)
  distCUDA2                                          (simple_knn._C)
  deform_network                                     (scene.deformation)
"""
from ._lib import Ed3dgsError, LIB_PATH  # noqa: F401
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                         rasterize_gaussians)
from .knn import distCUDA2  # noqa: F401
from .dropin import install, uninstall  # noqa: F401  (import hook for `scene.deformation`, see dropin.py)

__version__ = "0.1.0"
