"""Golden fixture for the reference-side interop of the compressed scene file and the indexed decode sites.

TEST INFRASTRUCTURE ONLY.  Run in the build container (the reference checkout is absent on the GPU box):

    python oracle/make_scene_golden.py        ->  tests/golden/scene_interop.npz

Imports the reference's own scene/gaussian_model.py UNCHANGED on CPU (device="cpu"), with stub modules for the
native / graphics dependencies that are not installed here and that the file format does not touch
(matplotlib, plyfile, simple_knn._C, diff_gaussian_rasterization_no_camera; compression.vq's torch_scatter and
weighted_distance._C as in oracle/ref_loader.py), builds a small indexed + quantisation-aware model and records

  in/*       the raw tensors and (scale, zero_point) pairs the model was built from
  fileA/*    every array of the file the reference's save_npz (gaussian_model.py:505-623) wrote
  fileB/*    the same with sort_morton=True (gaussian_model.py:997-1023, 1415-1432)
  loaded/*   the tensors the reference's load_npz (gaussian_model.py:625-740) produced from file A
  decoded/*  get_features / get_rotation / get_scaling of the loaded model with the fake quantisers switched off
             (gaussian_model.py:213-257: the indexed decode sites)

tests/test_gpu_pipeline.py checks SplatScene against it in both directions: files written by the reference load
into the same tensors, and files written by SplatScene from the same inputs hold the same arrays.
"""
from __future__ import annotations

import os
import sys
import tempfile
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
REFERENCE_ROOT = os.environ.get("C3DGS_REFERENCE", "/root/reference")


def load_reference_model_module():
    """scene.gaussian_model of the reference, imported from where it lies."""
    from oracle import ref_loader

    def stub(name, **attrs):
        m = types.ModuleType(name)
        for k, v in attrs.items():
            setattr(m, k, v)
        sys.modules[name] = m
        return m

    def unavailable(*a, **k):
        raise RuntimeError("native dependency stubbed out in oracle/make_scene_golden.py")

    mpl = stub("matplotlib")
    mpl.pyplot = stub("matplotlib.pyplot")
    stub("plyfile", PlyData=object, PlyElement=object)
    knn = stub("simple_knn")
    knn._C = stub("simple_knn._C", distCUDA2=unavailable)
    stub("diff_gaussian_rasterization_no_camera", GaussianRasterizationSettings=object, GaussianRasterizer=object,
         GaussianRasterizerIndexed=object, getProjectionMatrix=unavailable, quat_to_mat=unavailable)
    wd = stub("weighted_distance")
    wd._C = stub("weighted_distance._C", weightedDistance=ref_loader._oracle_weighted_distance)
    stub("torch_scatter", scatter=ref_loader._scatter)
    sys.path.insert(0, REFERENCE_ROOT)
    try:
        import importlib
        return importlib.import_module("scene.gaussian_model")
    finally:
        sys.path.remove(REFERENCE_ROOT)


def qparams(t: torch.Tensor):
    """What the reference's FakeQuantize observers would hold for `t` (min/max rule, qint8)."""
    obs = torch.ao.quantization.MinMaxObserver(dtype=torch.qint8)
    obs(t)
    scale, zp = obs.calculate_qparams()
    return scale.float().reshape(1), zp.int().reshape(1)


def main():
    gm = load_reference_model_module()
    g = torch.Generator().manual_seed(11)
    n, f_rows, g_rows = 700, 96, 64
    raw = {
        "xyz": torch.randn(n, 3, generator=g) * torch.tensor([4.0, 1.5, 2.5]),
        "features_dc": torch.randn(f_rows, 1, 3, generator=g) * 0.6,
        "features_rest": torch.randn(f_rows, 15, 3, generator=g) * 0.15,
        "opacity": torch.randn(n, 1, generator=g) * 2.0,
        "scaling": torch.rand(g_rows, 3, generator=g) + 0.02,
        "scaling_factor": torch.randn(n, 1, generator=g) * 0.7 - 3.5,
        "rotation": torch.randn(g_rows, 4, generator=g),
        "feature_indices": torch.randint(0, f_rows, (n,), generator=g),
        "gaussian_indices": torch.randint(0, g_rows, (n,), generator=g),
    }

    def build():
        m = gm.GaussianModel(3, quantization=True, use_factor_scaling=True, device="cpu")
        P = torch.nn.Parameter
        m._xyz = P(raw["xyz"].clone())
        m._features_dc, m._features_rest = P(raw["features_dc"].clone()), P(raw["features_rest"].clone())
        m._opacity, m._scaling = P(raw["opacity"].clone()), P(raw["scaling"].clone())
        m._scaling_factor, m._rotation = P(raw["scaling_factor"].clone()), P(raw["rotation"].clone())
        m._feature_indices = P(raw["feature_indices"].clone(), requires_grad=False)
        m._gaussian_indices = P(raw["gaussian_indices"].clone(), requires_grad=False)
        m.color_index_mode = gm.ColorMode.ALL_INDEXED
        stored = {"features_dc": m._features_dc, "features_rest": m._features_rest,
                  "opacity": m.opacity_activation(m._opacity), "scaling": m.scaling_activation(m._scaling),
                  "scaling_factor": m._scaling_factor, "rotation": m.rotation_activation(m._rotation)}
        qp = {}
        for name, t in stored.items():
            scale, zp = qparams(t.detach())
            qa = getattr(m, name + "_qa")
            qa.scale, qa.zero_point = scale, zp
            qp[name] = (scale, zp)
        return m, qp

    out = {}
    for k, v in raw.items():
        out["in/" + k] = v.numpy()
    with tempfile.TemporaryDirectory() as tmp:
        m, qp = build()
        for name, (scale, zp) in qp.items():
            out[f"in/{name}_scale"], out[f"in/{name}_zero_point"] = scale.numpy(), zp.numpy()
        path_a = os.path.join(tmp, "a.npz")
        m.save_npz(path_a)                                                   # gaussian_model.py:505-623
        for k, v in np.load(path_a).items():
            out["fileA/" + k] = v
        m_b, _ = build()
        path_b = os.path.join(tmp, "b.npz")
        m_b.save_npz(path_b, sort_morton=True)
        for k, v in np.load(path_b).items():
            out["fileB/" + k] = v
        m2 = gm.GaussianModel(3, quantization=True, use_factor_scaling=True, device="cpu")
        m2.load_npz(path_a)                                                  # gaussian_model.py:625-740
        for name in ("xyz", "features_dc", "features_rest", "opacity", "scaling", "scaling_factor", "rotation",
                     "feature_indices", "gaussian_indices"):
            out["loaded/" + name] = getattr(m2, "_" + name).detach().numpy()
        for qa in (m2.features_dc_qa, m2.features_rest_qa, m2.opacity_qa, m2.scaling_qa, m2.scaling_factor_qa,
                   m2.rotation_qa):
            qa.disable_fake_quant()
            qa.disable_observer()
        with torch.no_grad():                                                # the indexed decode sites, :213-257
            out["decoded/features"] = m2.get_features.numpy()
            out["decoded/rotation"] = m2.get_rotation.numpy()
            out["decoded/scaling"] = m2.get_scaling.numpy()
    dst = os.path.join(ROOT, "tests", "golden", "scene_interop.npz")
    np.savez_compressed(dst, **out)
    print(f"wrote {dst}: {len(out)} arrays, {os.path.getsize(dst)} bytes")
    for k in sorted(out):
        if k.startswith("fileA/"):
            print(f"  {k:36s} {str(out[k].dtype):8s} {out[k].shape}")


if __name__ == "__main__":
    main()
