"""The callers of the clustering path: compress_color / compress_covariance / compress_gaussians and their
settings record (compression/vq.py:106-223), with the reference's names, argument order, defaults and printed
messages, on top of this package's kernels:

    prune mask  -> GaussianModel.mask_splats            (SplatScene: c3vq_join_count + c3vq_compact_rows)
    thresholds  -> torch.quantile                        (library call, as in the reference)
    covariance  -> get_normalized_covariance             (SplatScene: c3vq_build_covariance)
    codebooks   -> vq_features                           (c3vq_update ... : the hot path)
    join        -> join_features                         (c3vq_join_count / c3vq_join_write)
    rot, scale  -> extract_rot_scale(to_full_cov(.))     (c3vq_extract_rot_scale on the 6-value rows)

`gaussians` is anything with the GaussianModel members used below — the reference's own GaussianModel or
c3dgs_b200.scene.SplatScene.
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Optional

import torch

from . import _lib
from . import vq as _vq
from . import splats as _splats


@dataclass
class CompressionSettings:
    """compression/vq.py:106-114."""
    codebook_size: int
    importance_prune: float
    importance_include: float
    importance_include_relative: float
    steps: int
    decay: float
    batch_size: int


def _quantile(values: torch.Tensor, q: float) -> float:
    """torch.quantile (vq.py:212, :216); torch caps its input at 2^24 elements, larger inputs use the same
    linear-interpolation rule on a full sort."""
    if values.numel() <= (1 << 24):
        return torch.quantile(values, q).item()
    ordered = values.float().sort().values
    pos = q * (ordered.numel() - 1)
    lo = int(pos)
    hi = min(lo + 1, ordered.numel() - 1)
    return torch.lerp(ordered[lo], ordered[hi], pos - lo).item()


def compress_color(gaussians, color_importance: torch.Tensor, color_comp: CompressionSettings,
                   color_compress_non_dir: bool, silent: bool):
    """compression/vq.py:117-147: SH-colour codebook over the Gaussians at or below the importance threshold; the
    others keep their own row behind the codebook."""
    keep_mask = color_importance > color_comp.importance_include
    if not silent:
        print(f"color keep: {keep_mask.float().mean() * 100:.2f}%")
    feats = gaussians.get_features.detach()
    if color_compress_non_dir:
        n_sh_coefs = feats.shape[1]
    else:                                        # leave the view-independent (DC) coefficient out of the codebook
        n_sh_coefs = feats.shape[1] - 1
        feats = feats[:, 1:]
    color_features = feats.flatten(-2).contiguous()
    vq_mask = ~keep_mask
    if vq_mask.any():
        if not silent:
            print("compressing color...")
        vq_rows, vq_weights = _lib.compact([color_features, color_importance.float()], vq_mask)
        codebook, vq_indices = _vq.vq_features(vq_rows, vq_weights, color_comp.codebook_size,
                                               color_comp.batch_size, color_comp.steps, silent=silent)
    else:
        codebook = torch.empty((0, color_features.shape[-1]), device=color_features.device)
        vq_indices = torch.empty((0,), device=color_features.device, dtype=torch.long)
    compressed, indices = _vq.join_features(color_features, keep_mask, codebook, vq_indices)
    gaussians.set_color_indexed(compressed.reshape(-1, n_sh_coefs, 3), indices)


def compress_covariance(gaussians, gaussian_importance: torch.Tensor, gaussian_comp: CompressionSettings,
                        silent: bool):
    """compression/vq.py:149-191: shape codebook on the normalised covariance (trace-normalised codewords), then
    rotation + scaling of every table row."""
    keep_mask = gaussian_importance > gaussian_comp.importance_include
    if not silent:
        print(f"gaussians keep: {keep_mask.float().mean() * 100:.2f}%")
    covariance = gaussians.get_normalized_covariance(strip_sym=True).detach()
    vq_mask = ~keep_mask
    if vq_mask.any():
        if not silent:
            print("compressing gaussian splats...")
        vq_rows, vq_weights = _lib.compact([covariance, gaussian_importance.float()], vq_mask)
        codebook, vq_indices = _vq.vq_features(vq_rows, vq_weights, gaussian_comp.codebook_size,
                                               gaussian_comp.batch_size, gaussian_comp.steps, scale_normalize=True,
                                               silent=silent)
    else:
        codebook = torch.empty((0, covariance.shape[1]), device=covariance.device)
        vq_indices = torch.empty((0,), device=covariance.device, dtype=torch.long)
    compressed_cov, cov_indices = _vq.join_features(covariance, keep_mask, codebook, vq_indices)
    rot_vq, scale_vq = _splats.extract_rot_scale(compressed_cov)       # [n,6] rows: to_full_cov fused away
    gaussians.set_gaussian_indexed(rot_vq.to(compressed_cov.device), scale_vq.to(compressed_cov.device), cov_indices)


def compress_gaussians(gaussians, color_importance: torch.Tensor, gaussian_importance: torch.Tensor,
                       color_comp: Optional[CompressionSettings], gaussian_comp: Optional[CompressionSettings],
                       color_compress_non_dir: bool, prune_threshold: float = 0., silent: bool = False):
    """compression/vq.py:194-223: prune, fix the two thresholds, cluster colours, cluster shapes."""
    with torch.no_grad():
        if prune_threshold >= 0:
            non_prune_mask = color_importance > prune_threshold
            if not silent:
                print(f"prune: {(1 - non_prune_mask.float().mean()) * 100:.2f}%")
            gaussians.mask_splats(non_prune_mask)
            gaussian_importance, color_importance = _lib.compact(
                [gaussian_importance.float(), color_importance.float()], non_prune_mask)
        # like the reference, the thresholds are read before the None checks below (vq.py:211-218)
        if color_comp.importance_include is None:
            color_comp.importance_include = _quantile(color_importance, color_comp.importance_include_relative)
            if not silent:
                print(f"Setting color threshold to {color_comp.importance_include}")
        if gaussian_comp.importance_include is None:
            gaussian_comp.importance_include = _quantile(gaussian_importance,
                                                         gaussian_comp.importance_include_relative)
            if not silent:
                print(f"Setting gaussian threshold to {gaussian_comp.importance_include}")
        if color_comp is not None:
            compress_color(gaussians, color_importance, color_comp, color_compress_non_dir, silent=silent)
        if gaussian_comp is not None:
            compress_covariance(gaussians, gaussian_importance, gaussian_comp, silent=silent)
