"""Covariance codebook -> (rotation, scaling): the reference's utils/splats.py:7-36, the step that follows
join_features in compress_covariance (compression/vq.py:186).

Same names and argument meaning as the reference: `to_full_cov(cov [n,6]) -> [n,3,3]`,
`extract_rot_scale(cov [n,3,3]) -> (rot [n,4], scaling [n,3])`.  extract_rot_scale also accepts the [n,6] stored
form directly, which skips materialising the 3x3 (one kernel for the whole step, c3vq_extract_rot_scale).
"""
from __future__ import annotations

from typing import Tuple

import torch

from . import _lib


def to_full_cov(cov: torch.Tensor) -> torch.Tensor:
    """utils/splats.py:7-25: (xx, xy, xz, yy, yz, zz) -> symmetric [n,3,3] (layout only: one gather)."""
    if cov.dim() != 2 or cov.size(1) != 6:
        raise RuntimeError(f"cov must be [n,6], got {tuple(cov.shape)}")
    sel = torch.tensor([0, 1, 2, 1, 3, 4, 2, 4, 5], device=cov.device)
    return cov.float().index_select(1, sel).reshape(-1, 3, 3)


def extract_rot_scale(cov: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """utils/splats.py:28-36.  Eigenvalues (ascending) and R diag(s^2) R^T match the reference to float32 rounding;
    eigenvector signs (hence the quaternion among its equivalent choices) are not LAPACK's."""
    return _lib.extract_rot_scale(cov)


def morton_order(xyz: torch.Tensor) -> torch.Tensor:
    """The permutation GaussianModel._sort_morton computes (scene/gaussian_model.py:997-1004):
    `mortonEncode(xyz_q, pp_diap.argsort()).sort().indices` for xyz [n,3] float32, as one key kernel and a stable
    radix sort (c3vq_morton_order).  The caller applies it to its per-Gaussian tensors (`t[order]`)."""
    return _lib.morton_order(xyz)
