"""Indexed-codebook decode (scene/gaussian_model.py:213-257, :889-899; compression/vq.py:40-42).

`indexed_decode(table, idx)` is `table[idx]` for a float32 table [R, ...] and int64 idx [N],
as one vectorised gather kernel (c3vq_gather) with an autograd backward that is the same
segmented sum as the centroid update with unit weights (c3vq_scatter_add) — so finetune.py's
QAT loop can train codebook entries through it exactly as it does through torch indexing.
"""
from __future__ import annotations

import torch

from . import _lib


class _IndexedDecode(torch.autograd.Function):
    @staticmethod
    def forward(ctx, table: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
        ctx.save_for_backward(idx)
        ctx.table_rows = table.size(0)
        return _lib.gather(table, idx)

    @staticmethod
    def backward(ctx, grad_out: torch.Tensor):
        (idx,) = ctx.saved_tensors
        return _lib.scatter_add(grad_out, idx, ctx.table_rows), None


def indexed_decode(table: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """table[idx] with gradient to `table` (rows of any trailing shape)."""
    if table.requires_grad and torch.is_grad_enabled():
        return _IndexedDecode.apply(table, idx)
    return _lib.gather(table.detach(), idx)


def decode_gaussians(scaling_n, rotation, features, gaussian_indices, feature_indices, scaling_factor=None):
    """The three decodes `GaussianModel.get_scaling / get_rotation / get_features` perform on an
    indexed model (gaussian_model.py:220, :240, :257): returns (scaling [N,3], rotation [N,4],
    features [N,16,3])."""
    s = indexed_decode(scaling_n, gaussian_indices)
    if scaling_factor is not None:
        s = scaling_factor * s
    r = indexed_decode(rotation, gaussian_indices)
    f = indexed_decode(features, feature_indices)
    return s, r, f
