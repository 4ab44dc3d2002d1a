"""Import-name shim: ``import frnn; frnn.frnn_grid_points(...)`` (reference scene/gaussian_model.py:27,113,117)."""
from gaussianprediction_b200.frnn import frnn_gather, frnn_grid_points  # noqa: F401
