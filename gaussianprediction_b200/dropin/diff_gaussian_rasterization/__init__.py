"""Import-name shim: ``from diff_gaussian_rasterization import GaussianRasterizationSettings, GaussianRasterizer``
(reference gaussian_renderer/__init__.py:14) resolves to the B200 implementation."""
from gaussianprediction_b200.rasterizer import (GaussianRasterizationSettings, GaussianRasterizer,  # noqa: F401
                                                _RasterizeGaussians, mark_visible, rasterize_gaussians)
