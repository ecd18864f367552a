"""Small-case run of every kernel added in r1p-r1t (wire formats, compaction, covariance, the compress pipeline incl.
the outlier-codeword path, scene file, indexed preprocess with buffer reuse) — meant for compute-sanitizer where that
is available (it is closed on this pool), otherwise a quick end-to-end smoke run: python tools/sanitize_run.py"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from c3dgs_b200 import CompressionSettings, SplatScene, _lib, compress_gaussians  # noqa: E402
from preprocess_scene import camera, scene  # noqa: E402

dev = "cuda:0"
g = torch.Generator().manual_seed(0)
for n in (1, 1023, 5003):
    x = torch.randn(n, 7, generator=g).to(dev)
    q = _lib.quantize_i8(x, 0.0123, -7)
    _lib.dequantize_i8(q, 0.0123, -7)
    _lib.convert(_lib.convert(x, torch.float16), torch.float32)
    idx = torch.randint(0, 1000, (n,), generator=g).to(dev)
    _lib.convert(_lib.convert(idx, torch.int32), torch.int64)
    _lib.build_covariance(x[:, :3].contiguous(), x[:, 3:].contiguous(), activate=True)
    _lib.compact([x, idx], (torch.rand(n, generator=g) < 0.5).to(dev))
n = 6000
feats = torch.randn(n, 16, 3, generator=g)
sc = SplatScene(torch.randn(n, 3, generator=g).to(dev), feats[:, :1].contiguous().to(dev), feats[:, 1:].contiguous().to(dev),
                torch.randn(n, 1, generator=g).to(dev), (torch.rand(n, 3, generator=g) + 0.05).to(dev),
                torch.randn(n, 4, generator=g).to(dev), (torch.randn(n, 1, generator=g) - 3).to(dev))
ci, gi = (torch.rand(n, generator=g) ** 4).to(dev), (torch.rand(n, generator=g) ** 4).to(dev)
cc = CompressionSettings(128, 0.0, None, 0.8, 6, 0.8, 1024)
gc = CompressionSettings(128, 0.0, None, 0.7, 6, 0.8, 1024)
compress_gaussians(sc, ci, gi, cc, gc, True, prune_threshold=1e-6, silent=True)     # incl. the outlier-codeword path
import io
buf = io.BytesIO()
sc.save_npz(buf, sort_morton=True)
buf.seek(0)
cam = camera(width=320, height=200)
camt = {k: torch.from_numpy(v).to(dev) for k, v in cam.items() if isinstance(v, np.ndarray)}
out = _lib.preprocess_indexed(sc._xyz, sc._scaling, torch.exp(sc._scaling_factor), sc._rotation, torch.sigmoid(sc._opacity),
                              torch.cat((sc._features_dc, sc._features_rest), 1).contiguous(), sc._feature_indices,
                              sc._gaussian_indices, camt["viewmatrix"], camt["projmatrix"], camt["cam_pos"], 320, 200,
                              cam["tan_fovx"], cam["tan_fovy"])
out = _lib.preprocess_indexed(sc._xyz, sc._scaling, torch.exp(sc._scaling_factor), sc._rotation, torch.sigmoid(sc._opacity),
                              torch.cat((sc._features_dc, sc._features_rest), 1).contiguous(), sc._feature_indices,
                              sc._gaussian_indices, camt["viewmatrix"], camt["projmatrix"], camt["cam_pos"], 320, 200,
                              cam["tan_fovx"], cam["tan_fovy"], out=out)
torch.cuda.synchronize()
print("visible", int((out["radii"] > 0).sum()), "of", out["radii"].numel(), "- done")
