"""Stage timing of the full compress pipeline (BASELINE configs[4]) on a synthetic 6M-Gaussian scene: every stage
wrapped with a device synchronise + host clock, two runs (the first pays the allocator's cudaMallocs).

    python tools/pipeline_profile.py [n_gaussians] > profiles/<round>_pipeline_stages.txt
"""
import os
import sys
import time
from collections import defaultdict

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import c3dgs_b200  # noqa: E402,F401
from c3dgs_b200 import CompressionSettings, SplatScene, _lib, compress_gaussians, splats, vq  # noqa: E402

dev = torch.device("cuda:0")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 6_000_000
K = 4096
stages = defaultdict(float)


def wrap(mod, name, label=None):
    fn = getattr(mod, name)

    def timed(*a, **kw):
        torch.cuda.synchronize()
        t0 = time.time()
        out = fn(*a, **kw)
        torch.cuda.synchronize()
        stages[label or name] += (time.time() - t0) * 1e3
        return out

    setattr(mod, name, timed)


wrap(vq, "vq_features")
wrap(vq.VectorQuantize, "update_indexed", "  vq: update_indexed (all steps, synchronised)")
wrap(vq.VectorQuantize, "assign", "  vq: final assign")
wrap(vq.VectorQuantize, "uniform_init", "  vq: uniform_init")
wrap(torch, "randint", "  vq: host torch.randint")
wrap(vq, "join_features")
wrap(_lib, "compact")
wrap(_lib, "build_covariance")
wrap(splats, "extract_rot_scale")
wrap(_lib, "morton_order")
wrap(_lib, "quantize_i8")
wrap(_lib, "convert")


def make_scene(seed):
    gen = torch.Generator(device=dev).manual_seed(seed)
    feats = torch.randn(n, 16, 3, device=dev, generator=gen)
    scene = SplatScene(torch.randn(n, 3, device=dev, generator=gen), feats[:, :1].contiguous(),
                       feats[:, 1:].contiguous(), torch.randn(n, 1, device=dev, generator=gen),
                       torch.rand(n, 3, device=dev, generator=gen) + 0.05,
                       torch.randn(n, 4, device=dev, generator=gen), torch.randn(n, 1, device=dev, generator=gen) - 3)
    return scene, torch.rand(n, device=dev, generator=gen) ** 4, torch.rand(n, device=dev, generator=gen) ** 4


for run in range(2):
    scene, cimp, gimp = make_scene(run)
    cc = CompressionSettings(K, 0.0, None, 0.6, 100, 0.8, 2 ** 18)        # thresholds by torch.quantile, 60 % clustered
    gc = CompressionSettings(K, 0.0, None, 0.3, 100, 0.8, 2 ** 18)
    stages.clear()
    torch.cuda.synchronize()
    t0 = time.time()
    compress_gaussians(scene, cimp, gimp, cc, gc, color_compress_non_dir=True, prune_threshold=1e-8, silent=True)
    torch.cuda.synchronize()
    t1 = time.time()
    import io
    buf = io.BytesIO()
    t2 = time.time()
    scene.save_npz(buf, compress=False, sort_morton=True)
    torch.cuda.synchronize()
    t3 = time.time()
    print(f"run {run}: {n} Gaussians, K={K}, 2 x (100 mini-batch steps of 2^18 rows + full assignment)")
    print(f"  compress_gaussians total {1e3 * (t1 - t0):9.1f} ms")
    print(f"  save_npz (uncompressed, in memory; incl. D2H + numpy) {1e3 * (t3 - t2):9.1f} ms")
    for k, v in sorted(stages.items(), key=lambda kv: -kv[1]):
        print(f"    {k:22s} {v:9.2f} ms")
    print(f"  file bytes {buf.getbuffer().nbytes}, colour table rows {scene._features_dc.shape[0]}, "
          f"shape table rows {scene._rotation.shape[0]}")
    del scene, cimp, gimp, buf
