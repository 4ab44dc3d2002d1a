"""Import-name shim: ``from simple_knn._C import distCUDA2`` (reference scene/gaussian_model.py:20)."""
from gaussianprediction_b200.simple_knn import distCUDA2  # noqa: F401
