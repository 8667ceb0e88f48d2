"""Drop-in for the reference package `diff_surfel_rasterization`
(submodules/diff-surfel-rasterization/diff_surfel_rasterization/__init__.py)."""
from gaussmart_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                       rasterize_gaussians, _RasterizeGaussians, cpu_deep_copy_tuple)
