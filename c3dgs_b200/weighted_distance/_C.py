"""`weighted_distance._C` — same single symbol as the reference's pybind module (ext.cpp:4-6)."""
from __future__ import annotations

from typing import Tuple

import torch

from .. import _lib


def weightedDistance(coefs: torch.Tensor, codebook: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Nearest codeword of every row of `coefs` [N, D] among `codebook` [C, D].

    Mirrors weightedDistanceCUDA (weighted_distance.cu:60-93): RuntimeError when an argument is
    not 2-D or the channel counts differ; returns freshly allocated (min_distances f32 [N],
    min_idx i64 [N]) on coefs' device.  Results are bit-identical to the reference kernel.
    Differences, all deliberate: runs on the current torch stream of coefs' device (the
    reference uses the legacy default stream, :52), and non-CUDA / non-fp32 inputs raise
    instead of reading through a bad pointer.
    """
    return _lib.assign(coefs, codebook, exact=False, return_dist=True)


def weightedDistanceExact(coefs: torch.Tensor, codebook: torch.Tensor) -> Tuple[torch.Tensor, torch.Tensor]:
    """Same contract through the SIMT in-order kernel (parity anchor)."""
    return _lib.assign(coefs, codebook, exact=True, return_dist=True)
