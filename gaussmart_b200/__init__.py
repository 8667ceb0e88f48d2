"""gaussmart_b200: B200-native (sm_100a) 2D Gaussian surfel rasterizer + kNN scale initialiser,
a drop-in for GauSSmart's `diff_surfel_rasterization` and `simple_knn._C` (and nothing else).

    import gaussmart_b200; gaussmart_b200.install()   # makes `import diff_surfel_rasterization`
                                                      # and `from simple_knn._C import distCUDA2`
                                                      # resolve to this implementation
"""
import os
import sys

from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer, rasterize_gaussians,  # noqa: F401
                         mark_visible)
from .knn import distCUDA2  # noqa: F401

COMPAT_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "compat")


def install():
    """Put the drop-in packages `diff_surfel_rasterization` and `simple_knn` on sys.path."""
    if COMPAT_DIR not in sys.path:
        sys.path.insert(0, COMPAT_DIR)
