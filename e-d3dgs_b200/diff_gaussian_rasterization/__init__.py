"""Drop-in module name: with ``e-d3dgs_b200/`` on PYTHONPATH, the reference's
``from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer``
(gaussian_renderer/__init__.py:3) resolves here."""
from ed3dgs_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                    _RasterizeGaussians, rasterize_gaussians, cpu_deep_copy_tuple)
