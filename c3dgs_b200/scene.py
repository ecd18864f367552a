"""The slice of the reference's GaussianModel (scene/gaussian_model.py) that sits either side of the clustering
path: the per-Gaussian tensors, the indexed (codebook) representation the compression produces, pruning, Morton
ordering and the compressed `.npz` scene file.

`SplatScene` keeps the reference's attribute and method names for exactly the members compress_gaussians
(compression/vq.py:117-223) and the scene file touch, so the pipeline code in c3dgs_b200.compress runs on either a
SplatScene or the reference's own GaussianModel:

    get_features, get_normalized_covariance, mask_splats, set_color_indexed, set_gaussian_indexed,
    _sort_morton, save_npz, load_npz, is_color_indexed, is_gaussian_indexed

Training state (optimizer, densification statistics, fake-quantisation observers, rasteriser hooks) is out of scope
(not on the path).  The int8 scale / zero-point of each tensor, which the reference takes from its QAT
observers (gaussian_model.py:109-117), are inputs here (`qparams`); `affine_qparams` restates the observers' min/max
rule for callers that have no finetuned values.

Files written here load in the reference's `load_npz` and vice versa: same keys, dtypes and shapes
(gaussian_model.py:505-740).  All device work is libc3vq kernels (c3dgs_b200._lib); numpy does the file IO.
"""
from __future__ import annotations

import os
from typing import Dict, Optional, Tuple

import numpy as np
import torch

from . import _lib

_QUANTIZED_FIELDS = ("features_dc", "features_rest", "opacity", "scaling", "scaling_factor", "rotation")


def affine_qparams(t: torch.Tensor) -> Tuple[float, int]:
    """(scale, zero_point) a min/max observer reports for a qint8 per-tensor-affine FakeQuantize
    (torch.ao.quantization.observer `_calculate_qparams`, the module the reference instantiates at
    gaussian_model.py:109-117): range widened to include 0, scale = range / 255 (at least float32 eps),
    zero_point = -128 - round(min / scale) clamped to [-128, 127]."""
    lo, hi = (float(v) for v in torch.aminmax(t.detach().float()))
    lo, hi = min(lo, 0.0), max(hi, 0.0)
    scale = max((hi - lo) / 255.0, float(torch.finfo(torch.float32).eps))
    scale = float(np.float32(scale))
    zp = -128 - int(round(lo / scale))
    return scale, max(-128, min(127, zp))


class SplatScene:
    """Per-Gaussian tensors + optional codebook indices (float32 / int64, one CUDA device)."""

    def __init__(self, xyz, features_dc, features_rest, opacity, scaling, rotation, scaling_factor=None,
                 quantization: bool = True, qparams: Optional[Dict[str, Tuple[float, int]]] = None):
        self._xyz = xyz                          # [N,3]
        self._features_dc = features_dc          # [N,1,3]   (or [K+kept,1,3] once colour-indexed)
        self._features_rest = features_rest      # [N,15,3]
        self._opacity = opacity                  # [N,1], pre-sigmoid
        self._scaling = scaling                  # [N,3]     (or [K+kept,3] once gaussian-indexed)
        self._rotation = rotation                # [N,4]
        self._scaling_factor = scaling_factor    # [N,1] log of the per-Gaussian size, or None (no factorisation)
        self._feature_indices = None
        self._gaussian_indices = None
        self.quantization = quantization
        self.qparams = dict(qparams or {})
        self.device = xyz.device

    # -- activations (gaussian_model.py:65-77) ------------------------------------------------------------------
    @property
    def use_factor_scaling(self) -> bool:
        return self._scaling_factor is not None

    def scaling_activation(self, s):
        if self.use_factor_scaling:
            return torch.nn.functional.normalize(torch.nn.functional.relu(s))
        return torch.exp(s)

    def scaling_inverse_activation(self, s):
        return s if self.use_factor_scaling else torch.log(s)

    # -- what compress_gaussians reads ----------------------------------------------------------------------------
    @property
    def is_color_indexed(self) -> bool:
        return self._feature_indices is not None

    @property
    def is_gaussian_indexed(self) -> bool:
        return self._gaussian_indices is not None

    @property
    def get_xyz(self):
        return self._xyz

    @property
    def get_features(self) -> torch.Tensor:
        """[N,16,3]: cat(dc, rest), decoded through the colour indices when indexed (gaussian_model.py:252-258)."""
        feats = torch.cat((self._features_dc, self._features_rest), dim=1)
        if self.is_color_indexed:
            return _lib.gather(feats.detach(), self._feature_indices)
        return feats

    @property
    def get_rotation(self) -> torch.Tensor:
        rot = torch.nn.functional.normalize(self._rotation)
        return _lib.gather(rot, self._gaussian_indices) if self.is_gaussian_indexed else rot

    @property
    def get_scaling_normalized(self) -> torch.Tensor:
        s = self.scaling_activation(self._scaling)
        return _lib.gather(s, self._gaussian_indices) if self.is_gaussian_indexed else s

    def get_normalized_covariance(self, scaling_modifier=1, strip_sym: bool = True) -> torch.Tensor:
        """gaussian_model.py:274-278: covariance of the unit-size Gaussian, [N,6] — one kernel that also applies the
        two activations when the model is not indexed."""
        if not strip_sym:
            raise RuntimeError("only the stripped (6-value) form is on the compression path")
        if not self.is_gaussian_indexed and self.use_factor_scaling:
            return _lib.build_covariance(self._scaling.detach(), self._rotation.detach(), scaling_modifier,
                                         activate=True)
        return _lib.build_covariance(self.get_scaling_normalized.detach(), self.get_rotation.detach(),
                                     scaling_modifier, activate=False)

    # -- what compress_gaussians writes ---------------------------------------------------------------------------
    def mask_splats(self, mask: torch.Tensor) -> None:
        """gaussian_model.py:1027-1046: keep the Gaussians where mask is true (one mask scan, one move per tensor)."""
        names = ["_xyz", "_opacity", "_scaling_factor"]
        names += ["_feature_indices"] if self.is_color_indexed else ["_features_dc", "_features_rest"]
        names += ["_gaussian_indices"] if self.is_gaussian_indexed else ["_scaling", "_rotation"]
        outs = _lib.compact([getattr(self, n) for n in names], mask)
        for n, t in zip(names, outs):
            setattr(self, n, t)

    @staticmethod
    def _check_indices(indices: torch.Tensor, rows: int, what: str) -> None:
        """The decode gathers read table rows unchecked (an out-of-range index decodes to a row of zeros where torch
        indexing would raise), so index arrays are validated once, where they enter the scene."""
        if indices.numel():
            lo, hi = (int(v) for v in torch.aminmax(indices))
            if lo < 0 or hi >= rows:
                raise IndexError(f"{what} must lie in [0, {rows}), got [{lo}, {hi}]")

    def set_color_indexed(self, features: torch.Tensor, indices: torch.Tensor) -> None:
        self._check_indices(indices, features.shape[0], "feature indices")
        self._feature_indices = indices
        self._features_dc = features[:, :1].detach()
        self._features_rest = features[:, 1:].detach()

    def set_gaussian_indexed(self, rotation: torch.Tensor, scaling: torch.Tensor, indices: torch.Tensor) -> None:
        self._check_indices(indices, rotation.shape[0], "gaussian indices")
        self._gaussian_indices = indices.detach()
        self._rotation = rotation.detach()
        self._scaling = scaling.detach()

    def _sort_morton(self) -> None:
        """gaussian_model.py:997-1023: reorder the per-Gaussian tensors along the Z-order curve of the positions."""
        order = _lib.morton_order(self._xyz.detach())
        per_gaussian = ["_xyz", "_opacity", "_scaling_factor"]
        per_gaussian += ["_feature_indices"] if self.is_color_indexed else ["_features_dc", "_features_rest"]
        per_gaussian += ["_gaussian_indices"] if self.is_gaussian_indexed else ["_scaling", "_rotation"]
        for n in per_gaussian:
            t = getattr(self, n)
            if t is None:
                continue
            if t.dtype == torch.float32:
                flat = t.detach().reshape(t.size(0), -1)
                setattr(self, n, _lib.gather(flat, order).reshape(t.shape))
            else:                                               # int64 index arrays: gather as pairs of words
                pairs = t.detach().contiguous().view(torch.float32).reshape(t.size(0), 2)
                setattr(self, n, _lib.gather(pairs, order).view(torch.int64).reshape(t.shape))

    # -- scene file -------------------------------------------------------------------------------------------------
    def _stored(self, name: str) -> torch.Tensor:
        """The float32 tensor the file stores for `name` (activation applied where save_npz does)."""
        if name == "opacity":
            return torch.sigmoid(self._opacity.detach())
        if name == "scaling":
            return self.scaling_activation(self._scaling.detach())
        if name == "rotation":
            return torch.nn.functional.normalize(self._rotation.detach())
        return getattr(self, "_" + name).detach()

    def save_npz(self, path, compress: bool = True, half_precision: bool = False, sort_morton: bool = False) -> None:
        """gaussian_model.py:505-620, same keys / dtypes / shapes."""
        with torch.no_grad():
            if sort_morton:
                self._sort_morton()
            if isinstance(path, str):
                os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            out = {"quantization": self.quantization}
            if self.quantization:
                out["xyz"] = _lib.convert(self._xyz.detach(), torch.float16).cpu().numpy()
                for name in _QUANTIZED_FIELDS:
                    if name == "scaling_factor" and self._scaling_factor is None:
                        continue
                    t = self._stored(name)
                    if name not in self.qparams:
                        self.qparams[name] = affine_qparams(t)
                    scale, zp = self.qparams[name]
                    out[name] = _lib.quantize_i8(t, scale, zp).cpu().numpy()
                    out[name + "_scale"] = np.array([scale], dtype=np.float32)
                    out[name + "_zero_point"] = np.array([zp], dtype=np.int32)
            else:
                dtype = torch.float16 if half_precision else torch.float32
                out["xyz"] = self._xyz.detach().cpu().numpy()
                out["features_dc"] = self._features_dc.detach().cpu().numpy()
                out["features_rest"] = self._features_rest.detach().cpu().numpy()
                for name in ("opacity", "scaling", "scaling_factor", "rotation"):
                    t = getattr(self, "_" + name)
                    if t is None:
                        continue
                    t = t.detach()
                    out[name] = (_lib.convert(t, dtype) if dtype != t.dtype else t).cpu().numpy()
            if self.is_color_indexed:
                out["feature_indices"] = _lib.convert(self._feature_indices.detach(), torch.int32).cpu().numpy()
            if self.is_gaussian_indexed:
                out["gaussian_indices"] = _lib.convert(self._gaussian_indices.detach(), torch.int32).cpu().numpy()
            (np.savez_compressed if compress else np.savez)(path, **out)

    @classmethod
    def load_npz(cls, path, device="cuda") -> "SplatScene":
        """gaussian_model.py:625-740: the inverse of save_npz."""
        sd = np.load(path)
        dev = torch.device(device)
        quantization = bool(sd["quantization"])
        qparams = {}

        def dev_t(key):
            return torch.from_numpy(sd[key]).to(dev)

        def field(name):
            if name not in sd.files:
                return None
            if quantization:
                scale = float(sd[name + "_scale"].reshape(-1)[0])
                zp = int(sd[name + "_zero_point"].reshape(-1)[0])
                qparams[name] = (scale, zp)
                return _lib.dequantize_i8(dev_t(name), scale, zp)
            t = dev_t(name)
            return _lib.convert(t, torch.float32) if t.dtype == torch.float16 else t.float()

        xyz = dev_t("xyz")
        xyz = _lib.convert(xyz, torch.float32) if xyz.dtype == torch.float16 else xyz.float()
        scene = cls(xyz, field("features_dc"), field("features_rest"), None, None, None,
                    scaling_factor=field("scaling_factor"), quantization=quantization)
        opacity, scaling, rotation = field("opacity"), field("scaling"), field("rotation")
        if quantization:                                        # the file holds activated values (:674, :688)
            opacity = torch.log(opacity / (1 - opacity))
            scaling = scene.scaling_inverse_activation(scaling)
        scene._opacity, scene._scaling, scene._rotation = opacity, scaling, rotation
        scene.qparams = qparams
        if "gaussian_indices" in sd.files:
            scene._gaussian_indices = _lib.convert(dev_t("gaussian_indices"), torch.int64)
            cls._check_indices(scene._gaussian_indices, rotation.shape[0], "gaussian_indices of " + str(path))
        if "feature_indices" in sd.files:
            scene._feature_indices = _lib.convert(dev_t("feature_indices"), torch.int64)
            cls._check_indices(scene._feature_indices, scene._features_dc.shape[0], "feature_indices of " + str(path))
        return scene
