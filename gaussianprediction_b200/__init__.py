"""gaussianprediction_b200 -- B200 (sm_100a) implementation of GaussianPrediction's per-frame render hot path.

Keypoint-driven deformation followed by depth-aware tile rasterization, forward and backward, behind
the reference's Python API.  All compute runs in hand-written CUDA kernels inside libgp_b200.so
(C ABI: include/gp_b200.h); there is no CPU or PyTorch fallback.
"""
from ._lib import GpError, lib  # noqa: F401
from .rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                         mark_visible, rasterize_gaussians)

__all__ = ["GaussianRasterizationSettings", "GaussianRasterizer", "rasterize_gaussians", "mark_visible", "GpError",
           "lib", "install_dropin"]


def install_dropin():
    """Put the reference's module names (``diff_gaussian_rasterization``, ``simple_knn``, ``frnn``) on
    ``sys.path`` so that unmodified reference code (``gaussian_renderer/__init__.py:14``,
    ``scene/gaussian_model.py:20-27``) imports this implementation."""
    import os
    import sys
    d = os.path.join(os.path.dirname(os.path.abspath(__file__)), "dropin")
    if d not in sys.path:
        sys.path.insert(0, d)
    return d
