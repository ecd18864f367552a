"""CPU: host-side logic around the kernels that needs no GPU — the observer rule behind the scene file's int8
parameters, the quantile fallback for inputs torch.quantile rejects, the settings record, and that the pipeline
modules refuse CPU tensors (no fallback)."""
import numpy as np
import pytest
import torch


def test_affine_qparams_matches_torch_observer():
    from c3dgs_b200.scene import affine_qparams
    g = torch.Generator().manual_seed(0)
    for shift, spread in [(0.0, 1.0), (3.0, 0.5), (-2.0, 0.1), (0.5, 1e-9)]:
        t = torch.randn(5000, generator=g) * spread + shift
        obs = torch.ao.quantization.MinMaxObserver(dtype=torch.qint8)     # the rule FakeQuantize(dtype=qint8) applies
        obs(t)
        scale, zp = obs.calculate_qparams()
        got_scale, got_zp = affine_qparams(t)
        assert got_scale == pytest.approx(float(scale), rel=1e-6)
        assert got_zp == int(zp)


def test_quantile_fallback_equals_torch_quantile():
    from c3dgs_b200 import compress
    g = torch.Generator().manual_seed(1)
    v = torch.rand(10_001, generator=g) ** 4
    for q in (0.0, 0.3, 0.6, 0.999, 1.0):
        ordered = v.sort().values
        pos = q * (v.numel() - 1)
        lo = int(pos)
        hi = min(lo + 1, v.numel() - 1)
        manual = torch.lerp(ordered[lo], ordered[hi], pos - lo).item()
        assert manual == pytest.approx(torch.quantile(v, q).item(), rel=1e-6, abs=1e-12)
        assert compress._quantile(v, q) == torch.quantile(v, q).item()


def test_settings_record_and_cpu_inputs_raise():
    from c3dgs_b200 import CompressionSettings, SplatScene, compress_gaussians
    cs = CompressionSettings(4096, 0.0, None, 0.6, 1000, 0.8, 2 ** 18)
    assert (cs.codebook_size, cs.importance_include, cs.batch_size) == (4096, None, 262144)
    n = 16
    scene = SplatScene(torch.zeros(n, 3), torch.zeros(n, 1, 3), torch.zeros(n, 15, 3), torch.zeros(n, 1),
                       torch.ones(n, 3), torch.ones(n, 4), torch.zeros(n, 1))
    imp = torch.rand(n)
    with pytest.raises(RuntimeError, match="CUDA"):
        compress_gaussians(scene, imp, imp, cs, cs, True, prune_threshold=0.0, silent=True)
    with pytest.raises(RuntimeError, match="CUDA"):
        scene.get_normalized_covariance()
    with pytest.raises(RuntimeError, match="CUDA"):
        scene.save_npz("/tmp/never_written.npz")
