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


class _RecordingModel:
    """Stand-in for VectorQuantize that records the batches vq_features hands to the kernels."""
    fail_at = None

    def __init__(self, *a, **k):
        self.batches = []
        self.codebook = torch.nn.Parameter(torch.zeros(4, 3))

    def to(self, **k):
        return self

    def uniform_init(self, x):
        pass

    def update_indexed(self, features, importance, batch, scale_normalize=False, batch_checked=False):
        assert batch_checked                    # the loop vouches for its own randint(0, n) draws: no per-step host sync
        if self.fail_at is not None and len(self.batches) == self.fail_at:
            raise ValueError("boom")
        self.batches.append(batch)

    def assign(self, features):
        return torch.zeros(features.shape[0], dtype=torch.int64)


def _host_only():
    """vq_features with every CUDA call stubbed: what is left is the batch-draw thread and its ring of buffers."""
    import contextlib
    import unittest.mock as um
    stack = contextlib.ExitStack()
    stack.enter_context(um.patch.object(torch.Tensor, "is_cuda", new=property(lambda self: True)))
    stack.enter_context(um.patch.object(torch.Tensor, "pin_memory", lambda self: self))
    stack.enter_context(um.patch.object(torch.Tensor, "to", lambda self, *a, **k: self.clone()))
    for name in ("torch.cuda.Event", "torch.cuda.current_stream", "torch.cuda.synchronize"):
        stack.enter_context(um.patch(name))
    return stack


def test_vq_features_batch_draws_keep_the_reference_random_stream(monkeypatch):
    """The draws of vq.py:69 run ahead on a helper thread; the batches must still be the sequence a plain loop of
    torch.randint calls produces from the same CPU seed."""
    from c3dgs_b200 import vq
    made = []
    monkeypatch.setattr(vq, "VectorQuantize", lambda *a, **k: made.append(_RecordingModel()) or made[-1])
    x, w = torch.randn(1000, 3), torch.rand(1000)
    for steps in (0, 1, 3, 250):
        made.clear()
        torch.manual_seed(5)
        with _host_only():
            vq.vq_features(x, w, 4, vq_chunk=64, steps=steps, silent=True)
        torch.manual_seed(5)
        want = [torch.randint(low=0, high=1000, size=[64]) for _ in range(steps)]
        assert len(made[0].batches) == steps
        assert all(torch.equal(a, b) for a, b in zip(made[0].batches, want))


@pytest.mark.timeout(60)
def test_vq_features_error_in_a_step_does_not_hang_the_draw_thread(monkeypatch):
    import threading
    from c3dgs_b200 import vq

    class Failing(_RecordingModel):
        fail_at = 3

    monkeypatch.setattr(vq, "VectorQuantize", lambda *a, **k: Failing())
    with _host_only(), pytest.raises(ValueError, match="boom"):
        vq.vq_features(torch.randn(100, 3), torch.rand(100), 4, vq_chunk=16, steps=200, silent=True)
    assert not [t for t in threading.enumerate() if t.name == "c3vq-batch-draw"]


def test_gpu_local_host_memory_restores_the_affinity():
    """The NUMA helper must leave the process affinity as it found it (and be a no-op without NVML / a GPU)."""
    import os
    from c3dgs_b200.sharded import gpu_local_host_memory
    before = os.sched_getaffinity(0)
    with gpu_local_host_memory(0):
        inside = os.sched_getaffinity(0)
        assert inside and inside <= before
    assert os.sched_getaffinity(0) == before
