"""Host-side mirror of the reference's second native submodule: ``from simple_knn._C import
distCUDA2`` (scene/gaussian_model.py:21,301). Same name, argument and result."""
from __future__ import annotations

import torch

from . import cabi

__all__ = ["distCUDA2"]


def distCUDA2(points: torch.Tensor) -> torch.Tensor:
    """[P] float32: mean squared distance of each point to its three nearest neighbours
    (submodules/simple-knn/spatial.cu:15-27). ``points`` is a [P,3] CUDA tensor; there is no CPU path."""
    if not points.is_cuda:
        raise ValueError("distCUDA2: points must be a CUDA tensor (there is no CPU path)")
    return cabi.knn_mean_dist2(points)
