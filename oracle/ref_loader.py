"""Import the reference's own compression/vq.py, unchanged, in the build container.

TEST INFRASTRUCTURE ONLY.  /root/reference does not exist on the GPU box, so this
module is used only by oracle/make_golden.py (run here, output committed under
tests/golden/) and by CPU tests that skip when the checkout is absent.

The reference module needs two native dependencies that are not importable here:
  * `weighted_distance._C.weightedDistance`  (compression/vq.py:12)  -> `distance_impl`
  * `torch_scatter.scatter`                  (compression/vq.py:6)   -> zeros + index_add_
    (torch_scatter is un-vendored and unpinned, environment.yml:15; its documented
     semantics for reduce="sum" are zero-init + scatter_add_ along `dim`).
They are injected into sys.modules before the import; nothing in the reference file
is edited or copied.
"""
from __future__ import annotations

import importlib
import os
import sys
import types

REFERENCE_ROOT = os.environ.get("C3DGS_REFERENCE", "/root/reference")


def available() -> bool:
    return os.path.isfile(os.path.join(REFERENCE_ROOT, "compression", "vq.py"))


def _oracle_weighted_distance(coefs, codebook):
    """CPU stand-in for weightedDistanceCUDA (weighted_distance.cu:60-93) backed by the C oracle."""
    import torch
    from . import oracle

    if coefs.ndimension() != 2 or codebook.ndimension() != 2:
        raise RuntimeError("ceofs and codebook must have dimension 2")
    if codebook.size(1) != coefs.size(1):
        raise RuntimeError("coefs and codebook must have same number of channels")
    dist, idx = oracle.assign(coefs.detach().cpu().numpy(), codebook.detach().cpu().numpy())
    return torch.from_numpy(dist), torch.from_numpy(idx)


def _scatter(src, index, dim=0, reduce="sum", dim_size=None):
    """torch_scatter.scatter(reduce='sum') on dim 0 — the only form the path uses (vq.py:31,33)."""
    import torch

    assert dim == 0 and reduce == "sum"
    out = torch.zeros((dim_size,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    return out.index_add_(0, index, src)


def load(distance_impl=None):
    """Return the reference `compression.vq` module object (fresh import each call)."""
    if not available():
        raise FileNotFoundError(f"reference checkout not found at {REFERENCE_ROOT}")
    wd_pkg = types.ModuleType("weighted_distance")
    wd_c = types.ModuleType("weighted_distance._C")
    wd_c.weightedDistance = distance_impl or _oracle_weighted_distance
    wd_pkg._C = wd_c
    ts = types.ModuleType("torch_scatter")
    ts.scatter = _scatter

    saved = {k: sys.modules.get(k) for k in
             ("weighted_distance", "weighted_distance._C", "torch_scatter", "compression", "compression.vq",
              "utils", "utils.splats")}
    sys.modules["weighted_distance"] = wd_pkg
    sys.modules["weighted_distance._C"] = wd_c
    sys.modules["torch_scatter"] = ts
    for k in ("compression", "compression.vq", "utils", "utils.splats"):
        sys.modules.pop(k, None)
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        mod = importlib.import_module("compression.vq")
    finally:
        sys.path.remove(REFERENCE_ROOT)
        # leave no trace: the reference's `utils`/`compression` packages must not shadow ours
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v
    return mod
