"""c3dgs_b200 — B200 (sm_100a) implementation of c3dgs's sensitivity-aware vector clustering.

Public surface (mirrors the reference, compression/vq.py and submodules/weighted_distance):
    VectorQuantize, vq_features, ema_inplace, join_features      (c3dgs_b200.vq)
    weighted_distance._C.weightedDistance                        (c3dgs_b200.weighted_distance)
    indexed_decode                                               (c3dgs_b200.decode)
    to_full_cov, extract_rot_scale, morton_order                 (c3dgs_b200.splats)
    CompressionSettings, compress_color, compress_covariance, compress_gaussians   (c3dgs_b200.compress)
    SplatScene (mask_splats, get_normalized_covariance, save_npz / load_npz, _sort_morton)   (c3dgs_b200.scene)
    install()   register the drop-ins under the reference's module names
"""
from __future__ import annotations

import sys

from . import _lib  # noqa: F401
from .decode import indexed_decode, decode_gaussians  # noqa: F401
from .splats import extract_rot_scale, morton_order, to_full_cov  # noqa: F401
from .vq import VectorQuantize, ema_inplace, join_features, vq_features  # noqa: F401
from .compress import CompressionSettings, compress_color, compress_covariance, compress_gaussians  # noqa: F401
from .scene import SplatScene, affine_qparams  # noqa: F401

__all__ = ["VectorQuantize", "vq_features", "ema_inplace", "join_features", "indexed_decode",
           "decode_gaussians", "to_full_cov", "extract_rot_scale", "morton_order", "install",
           "CompressionSettings", "compress_color", "compress_covariance", "compress_gaussians", "SplatScene",
           "affine_qparams"]


def install(patch_vq_module: bool = True) -> None:
    """Make the reference pick this implementation up without editing it.

    1. `weighted_distance` / `weighted_distance._C` in sys.modules -> this package's drop-in, so
       `from weighted_distance._C import weightedDistance` (compression/vq.py:12) binds to libc3vq.
    2. If the reference's `compression.vq` is (or later gets) imported and `patch_vq_module` is
       true, its `VectorQuantize`, `vq_features`, `ema_inplace`, `join_features`, `to_full_cov` and
       `extract_rot_scale` are rebound to the versions here, so compress_color / compress_covariance (vq.py:139, :164) call them.
    """
    from . import weighted_distance as wd

    sys.modules["weighted_distance"] = wd
    sys.modules["weighted_distance._C"] = wd._C
    if patch_vq_module:
        mod = sys.modules.get("compression.vq")
        if mod is not None:
            patch_reference_module(mod)


def patch_reference_module(mod) -> None:
    """Rebind the hot-path symbols of an imported reference `compression.vq` module."""
    from . import vq as _vq
    from .weighted_distance._C import weightedDistance

    mod.weightedDistance = weightedDistance
    mod.VectorQuantize = _vq.VectorQuantize
    mod.vq_features = _vq.vq_features
    mod.ema_inplace = _vq.ema_inplace
    mod.join_features = _vq.join_features
    if hasattr(mod, "extract_rot_scale"):                     # compression/vq.py:11 imports both by name
        from . import splats as _splats
        mod.extract_rot_scale = _splats.extract_rot_scale
        mod.to_full_cov = _splats.to_full_cov
