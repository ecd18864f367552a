"""GPU (-m gpu): the CUDA path through the C ABI against the oracle, the golden fixtures, and — when
oracle/_ref holds it — the reference's own CUDA kernel compiled unmodified.

Bar: bit-exact for assignments / indices / min distances / gathers; codebooks within 1e-5 relative
(row-normalised) because fp32 atomics have no defined order in the reference either."""
import os
import sys

import numpy as np
import pytest
import torch

from conftest import GOLDEN_CASES, ROOT, golden_path

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
CB_RTOL = 1e-5      # north_star: codebooks match within 1e-5 relative


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


def _synth(n, d, k, seed, clustered=True):
    g = torch.Generator().manual_seed(seed)
    if clustered:
        centres = torch.randn(max(k // 2, 1), d, generator=g)
        x = centres[torch.randint(0, centres.shape[0], (n,), generator=g)] + 0.1 * torch.randn(n, d, generator=g)
    else:
        x = torch.randn(n, d, generator=g)
    w = torch.rand(n, generator=g) ** 4
    amin, amax = x.aminmax()
    cb = torch.rand(k, d, generator=g) * (amax - amin) + amin
    return x.contiguous(), w.contiguous(), cb.contiguous()


def _cb_close(a, b, rtol=CB_RTOL):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    fin = np.isfinite(b).all(1)
    assert np.array_equal(np.isfinite(a).all(1), fin)
    scale = np.abs(b[fin]).max(1, keepdims=True) + 1e-30
    err = np.abs(a[fin] - b[fin]) / scale
    assert err.max() <= rtol, f"codebook rel err {err.max():.3e} > {rtol}"


@pytest.fixture(scope="module")
def L(lib):
    from c3dgs_b200 import _lib
    assert torch.cuda.is_available()
    return _lib


# ---------------------------------------------------------------- assignment ----------------
@pytest.mark.parametrize("d", [6, 7, 45, 48])
@pytest.mark.parametrize("n,k", [(1, 1), (127, 64), (129, 100), (1000, 256), (5003, 4096)])
@pytest.mark.parametrize("exact", [True, False])
def test_assign_bit_exact_vs_oracle(L, oracle, d, n, k, exact):
    x, _, cb = _synth(n, d, k, seed=d * 1000 + n)
    dist_o, idx_o = oracle.assign(x.numpy(), cb.numpy())
    dist, idx = L.assign(x.to(DEV), cb.to(DEV), exact=exact)
    assert idx.dtype == torch.int64 and dist.dtype == torch.float32
    assert np.array_equal(idx.cpu().numpy(), idx_o)
    assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))


@pytest.mark.parametrize("screen", ["f16", "tf32"])
@pytest.mark.parametrize("d", [1, 2, 3, 5, 13, 14, 16, 20, 29, 33, 44, 46, 50, 61, 62, 64])
def test_assign_generic_widths_bit_exact(L, oracle, d, screen, monkeypatch):
    """Widths without a specialised kernel (runtime d inside the DP = 16 ... 80 instantiations, both screenings): the
    assignment twice in a row (unseeded, then seeded) and the fused sums against the oracle."""
    monkeypatch.setenv("C3VQ_SCREEN", screen)
    n, k = 3001, 300
    x, w, cb = _synth(n, d, k, seed=500 + d)
    dist_o, idx_o = oracle.assign(x.numpy(), cb.numpy())
    xd, cbd = x.to(DEV), cb.to(DEV)
    for _ in range(2):
        dist, idx = L.assign(xd, cbd)
        assert np.array_equal(idx.cpu().numpy(), idx_o)
        assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))
    wn = (w / w.max()).contiguous().to(DEV)
    acc = torch.zeros(k, L.acc_stride(d), device=DEV)
    _, _, idx2 = L.assign_accumulate(xd, wn, cbd, acc=acc, tensor_cores="fused")
    assert np.array_equal(idx2.cpu().numpy(), idx_o)
    sx64, sw64 = oracle.scatter_sum_f64(x.numpy(), wn.cpu().numpy(), idx_o, k)
    a = acc.cpu().numpy().astype(np.float64)
    scale = np.abs(sx64).max(1, keepdims=True) + 1e-30
    assert (np.abs(a[:, :d] - sx64) / scale).max() <= 1e-5
    assert np.abs(a[:, d] - sw64).max() <= 1e-5 * (np.abs(sw64).max() + 1e-30)


@pytest.mark.parametrize("d", [6, 7, 45])
@pytest.mark.parametrize("k", [1, 64, 97, 150, 193, 300])
def test_assign_many_passes_few_steps(L, d, k):
    """Several 256-row passes per CTA against codebooks of one to four 96-code steps (odd and even step counts: the
    accumulator slots alternate their role from pass to pass when the count is odd): tensor-core path == exact kernel,
    twice (unseeded, seeded)."""
    n = 300_007
    x, _, cb = _synth(n, d, k, seed=7000 + 10 * d + k)
    xd, cbd = x.to(DEV), cb.to(DEV)
    dist_e, idx_e = L.assign(xd, cbd, exact=True)
    for _ in range(2):
        dist, idx = L.assign(xd, cbd)
        assert torch.equal(idx, idx_e)
        assert torch.equal(dist.view(torch.int32), dist_e.view(torch.int32))


@pytest.mark.parametrize("d", [45, 48])
def test_assign_seeded_running_minima_are_harmless(L, oracle, d, monkeypatch):
    """The tensor-core kernel seeds every row's running minimum with the row's winner of the previous call on the same
    workspace (evaluated against the current codebook).  Whatever that previous call was, the results are the oracle's:
    (1) the same rows and codebook again (perfect seeds), (2) a moved codebook (the k-means case), (3) unrelated rows of
    the same shape (stale seeds that point anywhere), (4) a smaller codebook (seeds beyond k), (5) outlier / non-finite
    codewords as previous winners, (6) seeds switched off (C3VQ_SEEDS=0)."""
    n, k = 20011, 1024
    x1, _, cb1 = _synth(n, d, k, seed=11)
    x2, _, cb2 = _synth(n, d, k, seed=12, clustered=False)

    def check(x, cb):
        dist_o, idx_o = oracle.assign(x.numpy(), cb.numpy())
        dist, idx = L.assign(x.to(DEV), cb.to(DEV))
        assert np.array_equal(idx.cpu().numpy(), idx_o)
        assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))
        return idx_o

    idx1 = check(x1, cb1)
    check(x1, cb1)                                             # (1)
    g = torch.Generator().manual_seed(3)
    check(x1, (cb1 + 0.05 * torch.randn(cb1.shape, generator=g)).contiguous())     # (2)
    check(x2, cb2)                                             # (3)
    check(x1, cb1[:100].contiguous())                          # (4)
    check(x1, cb1)
    cb_bad = cb1.clone()                                       # (5): the most popular previous winners leave the screening
    top = np.bincount(idx1, minlength=k).argsort()[::-1][:8].copy()
    cb_bad[torch.from_numpy(top[:4])] *= 1.0e4                 # finite outliers
    cb_bad[int(top[4])] = float("nan")
    cb_bad[int(top[5])] = float("inf")
    check(x1, cb_bad)
    check(x1, cb1)
    monkeypatch.setenv("C3VQ_SEEDS", "0")                      # (6)
    check(x1, cb1)
    check(x2, cb2)


@pytest.mark.parametrize("exact", [True, False])
def test_assign_kats(L, oracle, exact):
    g = torch.Generator().manual_seed(5)
    cb = torch.randn(512, 45, generator=g)
    # (1) codebook rows map to themselves with distance 0
    dist, idx = L.assign(cb.to(DEV), cb.to(DEV), exact=exact)
    assert torch.equal(idx.cpu(), torch.arange(512)) and float(dist.abs().max()) == 0.0
    # (2) duplicated codewords: lowest index wins
    cb3 = torch.cat([cb, cb, cb])
    x = cb + 1e-3 * torch.randn(cb.shape, generator=g)
    _, idx = L.assign(x.to(DEV), cb3.to(DEV), exact=exact)
    _, idx_o = oracle.assign(x.numpy(), cb3.numpy())
    assert np.array_equal(idx.cpu().numpy(), idx_o) and int(idx.max()) < 512
    # (3) one-hot +-1 data (exact in TF32)
    eye = torch.eye(45)
    cbh = torch.cat([eye, -eye])
    xh = cbh[torch.randint(0, 90, (777,), generator=g)]
    dist, idx = L.assign(xh.to(DEV), cbh.to(DEV), exact=exact)
    dist_o, idx_o = oracle.assign(xh.numpy(), cbh.numpy())
    assert np.array_equal(idx.cpu().numpy(), idx_o) and np.array_equal(dist.cpu().numpy(), dist_o)


@pytest.mark.parametrize("exact", [True, False])
def test_assign_nan_inf_codewords_never_win(L, oracle, exact):
    g = torch.Generator().manual_seed(6)
    cb = torch.randn(64, 6, generator=g)
    cb[0, :] = float("nan")
    cb[5, 2] = float("inf")
    cb[9, :] = float("-inf")
    x = torch.randn(300, 6, generator=g)
    x[7, :] = float("nan")                       # all distances NaN -> index 0 / FLT_MAX (defined fallback)
    dist, idx = L.assign(x.to(DEV), cb.to(DEV), exact=exact)
    dist_o, idx_o = oracle.assign(x.numpy(), cb.numpy())
    assert np.array_equal(idx.cpu().numpy(), idx_o)
    assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))
    assert idx_o[7] == 0 and dist_o[7] == np.finfo(np.float32).max


def test_assign_rows_indirection_and_noncontiguous(L, oracle):
    x, _, cb = _synth(3000, 45, 128, seed=7)
    rows = torch.randint(0, 3000, (1111,), generator=torch.Generator().manual_seed(8))
    for exact in (True, False):
        dist, idx = L.assign(x.to(DEV), cb.to(DEV), rows=rows.to(DEV), exact=exact)
        dist_o, idx_o = oracle.assign(x[rows].numpy(), cb.numpy())
        assert np.array_equal(idx.cpu().numpy(), idx_o) and np.array_equal(dist.cpu().numpy(), dist_o)
    # a strided view goes through .contiguous() like the reference (weighted_distance.cu:86-87)
    xs = x.to(DEV)[::2]
    dist, idx = L.assign(xs, cb.to(DEV))
    _, idx_o = oracle.assign(x[::2].numpy(), cb.numpy())
    assert np.array_equal(idx.cpu().numpy(), idx_o)


def test_weighted_distance_dropin_errors_and_empty(L):
    from c3dgs_b200.weighted_distance._C import weightedDistance
    with pytest.raises(RuntimeError, match="dimension 2"):
        weightedDistance(torch.zeros(5, device=DEV), torch.zeros(3, 5, device=DEV))
    with pytest.raises(RuntimeError, match="same number of channels"):
        weightedDistance(torch.zeros(5, 4, device=DEV), torch.zeros(3, 5, device=DEV))
    dist, idx = weightedDistance(torch.zeros(0, 5, device=DEV), torch.zeros(3, 5, device=DEV))
    assert dist.shape == (0,) and idx.shape == (0,) and idx.dtype == torch.int64


def _load_reference_kernel():
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if not os.path.exists(os.path.join(ref_dir, "weighted_distance_ref", "_C.so")):
        return None
    sys.path.insert(0, ref_dir)
    try:
        from weighted_distance_ref import _C
        return _C
    finally:
        sys.path.remove(ref_dir)


@pytest.mark.parametrize("d,k", [(45, 4096), (48, 1024), (6, 4096), (7, 512)])
def test_reference_cuda_kernel_pins_oracle_and_product(L, oracle, d, k):
    """The reference's own weighted_distance.cu, compiled unmodified (oracle/Makefile `ref`), run here."""
    ref = _load_reference_kernel()
    if ref is None:
        pytest.skip("oracle/_ref not built (reference checkout was absent at build time)")
    x, _, cb = _synth(4099, d, k, seed=31 + d)
    torch.cuda.synchronize()
    dist_r, idx_r = ref.weightedDistance(x.to(DEV), cb.to(DEV))     # legacy default stream (:52)
    torch.cuda.synchronize()
    dist_o, idx_o = oracle.assign(x.numpy(), cb.numpy())
    assert np.array_equal(idx_r.cpu().numpy(), idx_o), "oracle disagrees with the reference kernel"
    assert np.array_equal(dist_r.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))
    for exact in (True, False):
        dist, idx = L.assign(x.to(DEV), cb.to(DEV), exact=exact)
        assert torch.equal(idx, idx_r) and torch.equal(dist, dist_r)


# ---------------------------------------------------------------- update ---------------------
@pytest.mark.parametrize("name", GOLDEN_CASES)
@pytest.mark.parametrize("tc", [False, True])
def test_update_teacher_forced_against_reference_golden(L, name, tc):
    """Every recorded reference step, replayed from the recorded previous state (SURVEY §7.2 protocol)."""
    g = np.load(golden_path(name))
    n, d, k, chunk, steps, sn, seed = [int(v) for v in g["meta"]]
    f = _t(g["features"])
    w = _t(g["importance"])
    wn = (w / w.max()).contiguous()
    for s in range(steps):
        if s == 0:
            cb_prev, ei_prev = g["codebook0"], np.zeros(k, np.float32)
        else:
            cb_prev, ei_prev = g["cb_after_update"][s - 1].copy(), g["ei_after_update"][s - 1]
            if sn:
                from oracle import oracle as o
                o.trace_normalize(cb_prev)
        cb, ei = _t(cb_prev), _t(ei_prev)
        batch = _t(g["batches"][s])
        dist, idx = L.update(f, wn, cb, ei, float(g["decay"]), float(g["eps"]), rows=batch,
                             scale_normalize=False, tensor_cores=tc, return_dist=True, return_idx=True)
        assert np.array_equal(idx.cpu().numpy(), g["upd_idx"][s])
        assert np.array_equal(dist.cpu().numpy().view(np.uint32), g["upd_dist"][s].view(np.uint32))
        _cb_close(cb.cpu().numpy(), g["cb_after_update"][s])
        assert np.allclose(ei.cpu().numpy(), g["ei_after_update"][s], rtol=1e-5, atol=1e-7)


@pytest.mark.parametrize("d", [6, 7, 45, 48])
def test_accumulate_and_ema_vs_oracle(L, oracle, d):
    n, k = 20000, 300
    x, w, cb = _synth(n, d, k, seed=50 + d)
    _, idx = oracle.assign(x.numpy(), cb.numpy())
    sx64, sw64 = oracle.scatter_sum_f64(x.numpy(), w.numpy(), idx, k)
    acc = L.accumulate(x.to(DEV), w.to(DEV), _t(idx), k).cpu().numpy()
    assert acc.shape == (k, L.acc_stride(d))
    scale = np.abs(sx64).max(1, keepdims=True) + 1e-30
    assert np.max(np.abs(acc[:, :d] - sx64) / scale) < 1e-5
    assert np.allclose(acc[:, d], sw64, rtol=1e-5)
    assert not acc[:, d + 1:].any()
    # EMA epilogue is deterministic: bit-exact against the oracle on the SAME sums
    for sn in (False, True):
        cb_g, ei_g = cb.to(DEV).clone(), torch.rand(k, generator=torch.Generator().manual_seed(1)).to(DEV)
        cb_o, ei_o = cb.numpy().copy(), ei_g.cpu().numpy().copy()
        L.ema(cb_g, ei_g, _t(acc), 0.8, 1e-5, scale_normalize=sn)
        oracle.ema(cb_o, ei_o, np.ascontiguousarray(acc[:, :d]), np.ascontiguousarray(acc[:, d]), 0.8, 1e-5)
        if sn:
            oracle.trace_normalize(cb_o)
        assert np.array_equal(cb_g.cpu().numpy().view(np.uint32), cb_o.view(np.uint32))
        assert np.array_equal(ei_g.cpu().numpy().view(np.uint32), ei_o.view(np.uint32))


@pytest.mark.parametrize("d", [3, 7, 45, 48, 64])
@pytest.mark.parametrize("skew", [False, True])
def test_accumulate_sorted_vs_oracle(L, oracle, d, skew):
    """Sorted segmented sums (counting sort by codeword, no per-row atomics): same sums as the fp64 scatter,
    incl. out-of-range codes ignored, empty codewords, a giant segment, the rows indirection and unit weights."""
    n, k = 70001, 517
    x, w, cb = _synth(n, d, k, seed=90 + d)
    g = torch.Generator().manual_seed(d)
    idx = torch.randint(0, k, (n,), generator=g)
    if skew:
        idx[torch.rand(n, generator=g) < 0.7] = 5             # one codeword owns most rows
        idx[idx == 6] = 7                                     # and its neighbour is empty
    sx64, sw64 = oracle.scatter_sum_f64(x.numpy(), w.numpy(), idx.numpy(), k)
    bad = idx.clone()
    extra_rows = torch.randint(0, n, (1000,), generator=g)
    acc = L.accumulate(x.to(DEV), w.to(DEV), bad.to(DEV), k, sorted_path=True).cpu().numpy()
    scale = np.abs(sx64).max(1, keepdims=True) + 1e-30
    assert np.max(np.abs(acc[:, :d] - sx64) / scale) < 1e-5
    assert np.allclose(acc[:, d], sw64, rtol=1e-5)
    assert not acc[:, d + 1:].any()
    # out-of-range codes are ignored (as in the atomic kernel); rows indirection; unit weights
    idx2 = idx[extra_rows].clone()
    idx2[::7] = k + 3
    idx2[::11] = -1
    keep = (idx2 >= 0) & (idx2 < k)
    sx2, sw2 = oracle.scatter_sum_f64(x[extra_rows][keep].numpy(), np.ones(int(keep.sum()), np.float32),
                                      idx2[keep].numpy(), k)
    acc2 = L.accumulate(x.to(DEV), None, idx2.to(DEV), k, rows=extra_rows.to(DEV), sorted_path=True).cpu().numpy()
    assert np.max(np.abs(acc2[:, :d] - sx2) / (np.abs(sx2).max(1, keepdims=True) + 1e-30)) < 1e-5
    assert np.allclose(acc2[:, d], sw2, rtol=1e-6)


@pytest.mark.parametrize("d,k", [(45, 4096), (6, 33), (48, 1)])
def test_accumulate_sorted_large_and_invalid_codes(L, oracle, d, k):
    """The sorted sums on 200 k rows in storage order: rows with out-of-range codes (ignored), a codeword that owns
    most of one stretch of the input only, a one-codeword table, unit weights."""
    n = 3 * 65536 + 777
    x, w, _ = _synth(n, d, max(k, 2), seed=300 + d)
    g = torch.Generator().manual_seed(k)
    idx = torch.randint(0, k, (n,), generator=g)
    idx[65536:2 * 65536][torch.rand(65536, generator=g) < 0.8] = k - 1      # hot codeword inside the second window
    idx[::13] = -1
    idx[5::17] = k + 7
    keep = (idx >= 0) & (idx < k)
    for weights in (w, None):
        wk = (weights[keep] if weights is not None else torch.ones(int(keep.sum()))).numpy()
        sx64, sw64 = oracle.scatter_sum_f64(x[keep].numpy(), wk, idx[keep].numpy(), k)
        acc = L.accumulate(x.to(DEV), None if weights is None else weights.to(DEV), idx.to(DEV), k,
                           sorted_path=True).cpu().numpy()
        assert np.max(np.abs(acc[:, :d] - sx64) / (np.abs(sx64).max(1, keepdims=True) + 1e-30)) < 1e-5
        assert np.allclose(acc[:, d], sw64, rtol=1e-5)
        assert not acc[:, d + 1:].any()


@pytest.mark.parametrize("mode", [True, "fused", False])
def test_assign_accumulate_modes_agree(L, oracle, mode):
    """tensor-core assignment + sorted sums (default), fused bulk-reduce sums, SIMT + atomic sums."""
    n, d, k = 30011, 45, 777
    x, w, cb = _synth(n, d, k, seed=7)
    dist_o, idx_o = oracle.assign(x.numpy(), cb.numpy())
    sx64, sw64 = oracle.scatter_sum_f64(x.numpy(), w.numpy(), idx_o, k)
    acc, dist, idx = L.assign_accumulate(x.to(DEV), w.to(DEV), cb.to(DEV), tensor_cores=mode, return_dist=True)
    assert np.array_equal(idx.cpu().numpy(), idx_o)
    assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))
    acc = acc.cpu().numpy()
    assert np.max(np.abs(acc[:, :d] - sx64) / (np.abs(sx64).max(1, keepdims=True) + 1e-30)) < 1e-5
    assert np.allclose(acc[:, d], sw64, rtol=1e-5)


def test_update_kats_empty_cluster_and_single_cluster(L):
    cb = torch.tensor([[1.0, 2.0], [100.0, 100.0]], device=DEV)
    ei = torch.tensor([0.5, 0.25], device=DEV)
    x = torch.tensor([[1.0, 2.0], [1.5, 2.5]], device=DEV)
    w = torch.ones(2, device=DEV)
    for tc in (False, True):
        c, e = cb.clone(), ei.clone()
        L.update(x, w, c, e, 0.8, 1e-5, tensor_cores=tc)
        assert torch.equal(c[1], cb[1] * 0.8) and float(e[1]) == float(torch.tensor(0.25) * torch.tensor(0.8))
    g = torch.Generator().manual_seed(3)
    x = torch.randn(5000, 7, generator=g)
    w = torch.rand(5000, generator=g)
    c, e = torch.zeros(1, 7, device=DEV), torch.zeros(1, device=DEV)
    L.update(x.to(DEV), w.to(DEV), c, e, 0.0, 1e-5)
    expect = (x.double() * w.double()[:, None]).sum(0) / (w.double().sum() + 1e-5)
    assert torch.allclose(c[0].cpu().double(), expect, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_vq_loop_end_to_end_vs_oracle(L, oracle, name):
    """The whole mini-batch loop with the reference's recorded draws: final codebook within tolerance,
    final assignments equal wherever the two codebooks agree on the winner by a clear margin."""
    from c3dgs_b200 import VectorQuantize
    g = np.load(golden_path(name))
    n, d, k, chunk, steps, sn, seed = [int(v) for v in g["meta"]]
    f, w = _t(g["features"]), _t(g["importance"])
    wn = (w / w.max()).contiguous()
    vq = VectorQuantize(d, k, decay=0.8).to(DEV)
    vq.codebook.data = _t(g["codebook0"])
    for s in range(steps):
        vq.update_indexed(f, wn, _t(g["batches"][s]), scale_normalize=bool(sn))
    _cb_close(vq.codebook.data.cpu().numpy(), g["codebook"], rtol=5e-5)
    idx = vq.assign(f).cpu().numpy()
    assert (idx == g["indices"]).mean() > 0.99
    # and exactly the oracle's answer for the codebook we actually hold
    _, idx_o = oracle.assign(g["features"], vq.codebook.data.cpu().numpy())
    assert np.array_equal(idx, idx_o)


def test_vq_features_dropin_contract(L):
    from c3dgs_b200 import vq_features
    x, w, _ = _synth(20000, 6, 64, seed=77)
    torch.manual_seed(0)
    cb, idx = vq_features(x.to(DEV), w.to(DEV), 64, vq_chunk=4096, steps=5, scale_normalize=True, silent=True)
    assert cb.shape == (64, 6) and cb.dtype == torch.float32 and not cb.requires_grad
    assert idx.shape == (20000,) and idx.dtype == torch.int64
    assert int(idx.min()) >= 0 and int(idx.max()) < 64
    torch.manual_seed(0)
    cb2, idx2 = vq_features(x.to(DEV), w.to(DEV), 64, vq_chunk=4096, steps=5, scale_normalize=True, silent=True)
    assert torch.allclose(cb, cb2, rtol=1e-4, atol=1e-6, equal_nan=True)


def test_reference_vq_module_runs_on_the_dropin(L):
    """compression/vq.py's own import line resolves to libc3vq after install()."""
    import c3dgs_b200
    saved = {k: sys.modules.get(k) for k in ("weighted_distance", "weighted_distance._C")}
    try:
        c3dgs_b200.install()
        from weighted_distance._C import weightedDistance
        x, _, cb = _synth(500, 48, 32, seed=9)
        dist, idx = weightedDistance(x.to(DEV), cb.to(DEV))
        assert dist.device.type == "cuda" and idx.dtype == torch.int64
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


# ---------------------------------------------------------------- decode ---------------------
@pytest.mark.parametrize("shape", [(100, 3), (100, 4), (5000, 16, 3), (300, 45), (64, 7), (10, 1)])
def test_gather_bit_exact(L, oracle, shape):
    g = torch.Generator().manual_seed(sum(shape))
    table = torch.randn(*shape, generator=g)
    idx = torch.randint(0, shape[0], (12345,), generator=g)
    out = L.gather(table.to(DEV), idx.to(DEV))
    assert out.shape == (12345,) + tuple(shape[1:])
    assert np.array_equal(out.cpu().numpy(), oracle.gather(table.numpy(), idx.numpy()))
    assert torch.equal(out.cpu(), table[idx])
    empty = L.gather(table.to(DEV), torch.zeros(0, dtype=torch.int64, device=DEV))
    assert empty.shape == (0,) + tuple(shape[1:])


@pytest.mark.parametrize("d", [1, 3, 5, 6, 7, 9, 13, 45, 47, 48, 63, 101, 179, 181])
def test_gather_unaligned_rows_every_alignment(L, d):
    """Rows that are not whole float4, or whose table is not 16-byte aligned (k_gather_span, and k_gather_padded from
    the four padded copies once n >= 8 table rows): every residue of d mod 4, the table starting at each of the four
    4-byte alignments, ragged row counts (the partial last span), the last table row, and out-of-range indices, which
    decode to zeros (the unchecked decode's contract)."""
    g = torch.Generator().manual_seed(1000 + d)
    rows = 517
    store = torch.randn(rows * d + 8, generator=g).to(DEV)
    for shift in range(4):
        table = store[shift:shift + rows * d].view(rows, d)
        assert table.data_ptr() % 16 == (store.data_ptr() + 4 * shift) % 16
        for n in (1, 31, 32, 33, 64, 4099, 4136, 4137, 100_003):
            idx = torch.randint(0, rows, (n,), generator=g)
            idx[::97] = rows - 1
            idx[5::389] = 0
            want = table.cpu()[idx]
            if n > 40:
                idx[7], idx[40] = -1, rows
                want[7], want[40] = 0.0, 0.0
            dev_idx = idx.to(DEV)
            out = L.gather(table, dev_idx)
            assert torch.equal(out.cpu(), want), (d, shift, n)
            # the C entry points side by side: without scratch (c3vq_gather), and with scratch refused (null / short)
            lib = L.load()
            need = lib.c3vq_gather_workspace_bytes(rows, d, n)
            assert (need > 0) == (d % 4 != 0 and 12 <= d < 181 and n >= 8 * rows), (d, n, need)
            plain = torch.empty_like(out)
            assert lib.c3vq_gather(table.data_ptr(), rows, d, dev_idx.data_ptr(), n, plain.data_ptr(), None) == 0
            assert torch.equal(plain, out)
            if need:
                short = torch.empty(need - 16, dtype=torch.uint8, device=DEV)
                for ws, nb in ((None, 0), (short, short.numel())):
                    o2 = torch.full_like(out, float("nan"))
                    assert lib.c3vq_gather_ws(table.data_ptr(), rows, d, dev_idx.data_ptr(), n, o2.data_ptr(),
                                              ws.data_ptr() if ws is not None else None, nb, None) == 0
                    assert torch.equal(o2, out)


def test_indexed_decode_autograd_matches_torch_indexing(L):
    from c3dgs_b200 import indexed_decode
    g = torch.Generator().manual_seed(12)
    table = torch.randn(200, 16, 3, generator=g).to(DEV).requires_grad_(True)
    idx = torch.randint(0, 200, (5000,), generator=g).to(DEV)
    upstream = torch.randn(5000, 16, 3, generator=g).to(DEV)
    out = indexed_decode(table, idx)
    out.backward(upstream)
    grad = table.grad.clone()
    table.grad = None
    ref = table[idx]
    ref.backward(upstream)
    assert torch.equal(out, ref)
    assert torch.allclose(grad, table.grad, rtol=1e-5, atol=1e-5)


def test_gather_rows_matches_indexing(L):
    x, w, _ = _synth(5000, 45, 8, seed=13)
    rows = torch.randint(0, 5000, (3000,), generator=torch.Generator().manual_seed(14))
    ox, ow = L.gather_rows(x.to(DEV), w.to(DEV), rows.to(DEV))
    assert torch.equal(ox.cpu(), x[rows]) and torch.equal(ow.cpu(), w[rows])


# ---------------------------------------------------------------- join_features -------
def test_join_features_bit_exact_vs_reference_golden(L, oracle):
    g = np.load(golden_path("join_features").replace("vq_join_features", "join_features"))
    import c3dgs_b200
    for name in g["names"]:
        comp, idx = c3dgs_b200.join_features(_t(g[f"{name}_feats"]), _t(g[f"{name}_keep"]), _t(g[f"{name}_cb"]),
                                             _t(g[f"{name}_cbidx"]))
        assert idx.dtype == torch.int64 and comp.dtype == torch.float32
        assert np.array_equal(idx.cpu().numpy(), g[f"{name}_idx"]), name
        assert tuple(comp.shape) == g[f"{name}_out"].shape, name
        assert np.array_equal(comp.cpu().numpy().view(np.uint32), g[f"{name}_out"].view(np.uint32)), name


def test_join_features_large_and_errors(L, oracle):
    n, d, k = 1_000_003, 45, 4096
    gen = torch.Generator(device=DEV).manual_seed(4)
    feats = torch.randn(n, d, device=DEV, generator=gen)
    keep = torch.rand(n, device=DEV, generator=gen) < 0.04
    cb = torch.randn(k, d, device=DEV, generator=gen)
    cb_idx = torch.randint(0, k, (int((~keep).sum()),), device=DEV, generator=gen)
    comp, idx = L.join_features(feats, keep, cb, cb_idx)
    comp_o, idx_o = oracle.join_features(feats.cpu().numpy(), keep.cpu().numpy(), cb.cpu().numpy(), cb_idx.cpu().numpy())
    assert np.array_equal(idx.cpu().numpy(), idx_o) and np.array_equal(comp.cpu().numpy(), comp_o)
    assert torch.equal(comp[idx], torch.where(keep[:, None], feats, cb[idx.clamp(max=k - 1)]))   # decode round trip
    with pytest.raises(RuntimeError):
        L.join_features(feats, keep, cb, cb_idx[:-1])
    with pytest.raises(RuntimeError):
        L.join_features(feats, keep.to(torch.uint8), cb, cb_idx)


def test_update_from_host_equals_update(L):
    """The streamed first iteration (chunked upload on a copy stream, chunk-wise assignment) gives the same
    assignments and the same codebook as a plain copy followed by update()."""
    from c3dgs_b200.sharded import ShardedVectorQuantize
    n, d, k = 300_001, 45, 512
    g = torch.Generator().manual_seed(21)
    xh = torch.randn(n, d, generator=g).pin_memory()
    wh = (torch.rand(n, generator=g) ** 4).pin_memory()
    cb0 = xh[torch.randint(0, n, (k,), generator=g)].to(DEV)
    a = ShardedVectorQuantize(d, k, 0.8, device=DEV)
    b = ShardedVectorQuantize(d, k, 0.8, device=DEV)
    a.codebook.copy_(cb0)
    b.codebook.copy_(cb0)
    _, idx_a = a.update(xh.to(DEV), wh.to(DEV))
    xd, wd = torch.empty(n, d, device=DEV), torch.empty(n, device=DEV)
    idx_b = b.update_from_host(xh, wh, xd, wd, chunk_rows=70_000)
    torch.cuda.synchronize()
    assert torch.equal(idx_a, idx_b)
    assert torch.equal(xd.cpu(), xh) and torch.equal(wd.cpu(), wh)
    # the sums are fp32 reductions in whatever order the chunks and warps deliver them: north-star tolerance
    scale = a.codebook.abs().max(1, keepdim=True).values + 1e-30
    assert float(((a.codebook - b.codebook).abs() / scale).max()) < CB_RTOL
    assert torch.allclose(a.entry_importance, b.entry_importance, rtol=1e-5, atol=1e-7)


# ---------------------------------------------------------------- extract_rot_scale --
def test_extract_rot_scale_vs_reference_golden(L, oracle):
    """c3vq_extract_rot_scale on the fixture the reference's extract_rot_scale(to_full_cov(.)) produced
    (utils/splats.py:7-36): squared scalings and R diag(s^2) R^T within float32 rounding of the float64 oracle
    (tolerance 4e-6 of the row's largest entry; the reference itself sits at ~1e-6), unit quaternions, ascending
    order, and eigenvectors equal to the reference's up to sign where the eigenvalues are separated."""
    from test_oracle_golden import ROT_SCALE_TOL, rot_scale_errors
    import c3dgs_b200
    g = np.load(golden_path("rot_scale").replace("vq_rot_scale", "rot_scale"))
    cov6 = g["cov6"]
    rot, scale = c3dgs_b200.extract_rot_scale(_t(cov6))                              # [n,6] form
    rot_f, scale_f = c3dgs_b200.extract_rot_scale(c3dgs_b200.to_full_cov(_t(cov6)))  # the reference's [n,3,3] form
    assert torch.equal(rot, rot_f) and torch.equal(scale, scale_f)
    assert np.array_equal(c3dgs_b200.to_full_cov(_t(cov6)).cpu().numpy(), oracle.to_full_cov(cov6))
    rot, scale = rot.cpu().numpy(), scale.cpu().numpy()
    assert rot.shape == g["rot"].shape and scale.shape == g["scale"].shape
    e_s, e_c, e_q = rot_scale_errors(oracle, cov6, rot, scale)
    assert e_s < ROT_SCALE_TOL and e_c < ROT_SCALE_TOL and e_q < 1e-6, (e_s, e_c, e_q)
    assert np.all(np.diff(scale, axis=1) >= 0) and np.all(np.isfinite(rot)) and np.all(np.isfinite(scale))
    assert np.array_equal(scale[5], np.full(3, 1e-6, np.float32))                    # negative definite: NaN -> 1e-6
    lam = g["scale"].astype(np.float64) ** 2
    sep = (np.diff(lam, axis=1).min(1) > 0.05 * lam[:, 2]) & (lam[:, 0] > 0)
    dots = np.abs(np.einsum("nij,nij->nj", oracle.quaternion_to_matrix(g["rot"][sep]),
                            oracle.quaternion_to_matrix(rot[sep])))
    assert dots.min() > 1 - 1e-4


def test_extract_rot_scale_large_and_errors(L, oracle):
    """1.5 M rows (25 % of 6 M, the joined covariance table of SURVEY §8f): round trip through build_covariance
    (utils/splats.py:39-42) in float64 on a sample, plus empty input and shape errors."""
    n = 1_500_000
    gen = torch.Generator(device=DEV).manual_seed(9)
    q = torch.nn.functional.normalize(torch.randn(n, 4, device=DEV, generator=gen))
    s = torch.exp(torch.randn(n, 3, device=DEV, generator=gen) - 2.0)
    r, x, y, z = q.unbind(-1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(n, 3, 3)
    cov = R @ torch.diag_embed(s * s) @ R.transpose(-1, -2)
    cov = cov / cov.diagonal(dim1=-1, dim2=-2).sum(-1)[:, None, None]                # trace-normalised
    cov6 = torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], -1)
    rot, scale = L.extract_rot_scale(cov6.contiguous())
    from test_oracle_golden import ROT_SCALE_TOL, rot_scale_errors
    sel = torch.arange(0, n, 37, device=DEV)
    e_s, e_c, e_q = rot_scale_errors(oracle, cov6[sel].cpu().numpy(), rot[sel].cpu().numpy(), scale[sel].cpu().numpy())
    assert e_s < ROT_SCALE_TOL and e_c < ROT_SCALE_TOL and e_q < 1e-6, (e_s, e_c, e_q)
    assert bool((scale[:, 1:] >= scale[:, :-1]).all()) and bool(torch.isfinite(rot).all())
    assert torch.allclose((scale.double() ** 2).sum(1), torch.ones(n, device=DEV, dtype=torch.double) + 3e-8, atol=1e-5)
    r0, s0 = L.extract_rot_scale(torch.empty(0, 6, device=DEV))
    assert r0.shape == (0, 4) and s0.shape == (0, 3)
    with pytest.raises(RuntimeError):
        L.extract_rot_scale(torch.zeros(4, 5, device=DEV))
    with pytest.raises(RuntimeError):
        L.extract_rot_scale(torch.zeros(4, 6, device=DEV, dtype=torch.float64))


# ---------------------------------------------------------------- Morton order ------
def test_morton_order_bit_exact_vs_reference_golden(L, oracle):
    """c3vq_morton_order on the fixture made with the reference's own mortonEncode (scene/gaussian_model.py:997-1004,
    :1415-1432): keys bit-exact, permutation identical to the stable argsort (= the reference's wherever its keys
    are unique)."""
    g = np.load(golden_path("morton").replace("vq_morton", "morton"))
    for name in g["names"]:
        xyz, keys, order = g[f"{name}_xyz"], g[f"{name}_keys"], g[f"{name}_order"]
        o, ks = L.morton_order(_t(xyz), return_keys=True)
        o, ks = o.cpu().numpy(), ks.cpu().numpy()
        assert o.dtype == np.int64 and np.array_equal(ks, keys[order]), name            # the sorted key sequence
        assert np.array_equal(o, np.argsort(keys, kind="stable")), name                  # stable permutation
        if len(np.unique(keys)) == len(keys):
            assert np.array_equal(o, order), name


def test_morton_order_large_and_errors(L, oracle):
    """6 M points (config 5's Gaussian count): against the numpy oracle on the full array (keys and permutation),
    plus empty input and shape errors."""
    n = 6_000_000
    gen = torch.Generator(device=DEV).manual_seed(31)
    xyz = torch.randn(n, 3, device=DEV, generator=gen) * torch.tensor([4.0, 9.0, 1.5], device=DEV)
    xyz[1000:2000] = xyz[:1000]                                   # exact duplicates: ties must keep row order
    o, ks = L.morton_order(xyz, return_keys=True)
    k_or = oracle.morton_keys(xyz.cpu().numpy())
    o_or = np.argsort(k_or, kind="stable")
    assert np.array_equal(ks.cpu().numpy(), k_or[o_or])
    assert np.array_equal(o.cpu().numpy(), o_or)
    # sizes around the 4096-key tile of the radix sort, an unaligned view (the vectorised min/max needs 16-byte alignment
    # and must fall back), heavy ties: the permutation equals torch's stable sort of the same keys
    for m in (1, 31, 4095, 4096, 4097, 12289, 70001):
        g2 = torch.Generator().manual_seed(m)
        base = torch.randn(m + 1, 3, generator=g2).to(DEV)
        p3 = base.reshape(-1)[3:].reshape(m, 3)                                # starts 12 bytes into the storage
        p3[m // 3:m // 3 + m // 4] = p3[:m // 4].clone()
        o2, k2 = L.morton_order(p3, return_keys=True)
        krow = torch.empty_like(k2)
        krow[o2] = k2
        assert np.array_equal(krow.cpu().numpy().astype(np.uint64), oracle.morton_keys(p3.cpu().numpy()))
        assert torch.equal(o2, torch.sort(krow, stable=True).indices)
    assert L.morton_order(torch.empty(0, 3, device=DEV)).shape == (0,)
    with pytest.raises(RuntimeError):
        L.morton_order(torch.zeros(4, 2, device=DEV))
    with pytest.raises(RuntimeError):
        L.morton_order(torch.zeros(4, 3, device=DEV, dtype=torch.float64))


# ---------------------------------------------------------------- full size ------------------
@pytest.mark.parametrize("d", [45, 7])
def test_full_size_properties(L, d):
    """BASELINE config 2/3 sizes (5.8M rows, K=4096): properties that need no CPU oracle."""
    n, k = 5_800_000, 4096
    g = torch.Generator(device=DEV).manual_seed(d)
    x = torch.randn(n, d, device=DEV, generator=g)
    w = torch.rand(n, device=DEV, generator=g) ** 4
    cb = x[torch.randint(0, n, (k,), device=DEV, generator=g)].clone()
    cb[1] = cb[0]                                             # a duplicated codeword: index 0 must win
    dist_tc, idx_tc = L.assign(x, cb, exact=False)
    # tensor-core path == SIMT exact path on every row (bit-exact indices and distances)
    dist_ex, idx_ex = L.assign(x, cb, exact=True)
    assert torch.equal(idx_tc, idx_ex) and torch.equal(dist_tc, dist_ex)
    assert int((idx_tc == 1).sum()) == 0
    # codeword rows map to themselves
    _, self_idx = L.assign(cb, cb, exact=False)
    assert torch.equal(cb[self_idx], cb)                      # sampled rows may repeat: the copy with the
    assert bool((self_idx <= torch.arange(k, device=DEV)).all())   # lowest index wins, never a later one
    # sums: total weight is conserved; decode of the assignment is idempotent under re-assignment
    acc = L.accumulate(x, w, idx_tc, k)
    assert abs(float(acc[:, d].double().sum()) - float(w.double().sum())) < 1e-4 * float(w.double().sum())
    acc_s = L.accumulate(x, w, idx_tc, k, sorted_path=True)
    assert abs(float(acc_s[:, d].double().sum()) - float(w.double().sum())) < 1e-5 * float(w.double().sum())
    ref = torch.zeros(k, d, device=DEV, dtype=torch.float64).index_add_(0, idx_tc, x.double() * w.double()[:, None])
    scale = ref.abs().max(1, keepdim=True).values + 1e-30
    assert float(((acc_s[:, :d].double() - ref).abs() / scale).max()) < 1e-5
    q = L.gather(cb, idx_tc[:1_000_000])
    _, idx_q = L.assign(q, cb, exact=False)
    assert torch.equal(cb[idx_q], q)


def test_full_size_reference_kernel_bit_exact(L):
    """BASELINE configs[1] at full size (5.8 M x 45-D, K = 4096, the bench's own clustered rows): the reference's
    weighted_distance.cu:20-44, compiled unmodified (oracle/_ref), against c3vq_assign — every index and every
    distance bit for bit.  Two codebooks: the uniform-box start of the bench (a handful of codewords take everything)
    and rows sampled from the data (every codeword in use).  The reference kernel needs ~1.6 s per codebook here."""
    ref = _load_reference_kernel()
    if ref is None:
        pytest.skip("oracle/_ref not built (reference checkout was absent at build time)")
    sys.path.insert(0, ROOT)
    from bench import init_codebook, synth
    n, d, k = 5_800_000, 45, 4096
    x, _ = synth(n, d, 1234, device=torch.device(DEV))
    g = torch.Generator(device=DEV).manual_seed(3)
    books = (init_codebook(x, k, 99), x[torch.randint(0, n, (k,), device=DEV, generator=g)].clone())
    for cb in books:
        torch.cuda.synchronize()
        dist_r, idx_r = ref.weightedDistance(x, cb)                  # legacy default stream (weighted_distance.cu:52)
        torch.cuda.synchronize()
        dist, idx = L.assign(x, cb, exact=False)
        assert idx.dtype == idx_r.dtype == torch.int64
        assert torch.equal(idx, idx_r), f"{int((idx != idx_r).sum())} assignments differ from the reference kernel"
        assert torch.equal(dist.view(torch.int32), dist_r.view(torch.int32))
        del dist_r, idx_r, dist, idx


@pytest.mark.parametrize("book", ["uniform-box", "sampled"])
def test_full_size_fused_sums_within_tolerance(L, book):
    """BASELINE configs[1] at full size: the centroid sums the assignment kernel forms itself (default mode; warp
    combining + 32 partial tables folded pairwise) against float64 accumulation of the same float32 products, as
    centroids sum_x / (sum_w + eps), within the north star's 1e-5 (row-normalised).  The uniform-box codebook is the
    hard case: four codewords own 88 % of the rows (a plain fp32 atomic scatter, the reference's torch_scatter, is off
    by 2e-3 there: profiles/r2e_sums_probe.txt)."""
    sys.path.insert(0, ROOT)
    from bench import init_codebook, synth
    n, d, k = 5_800_000, 45, 4096
    x, w = synth(n, d, 1234, device=torch.device(DEV))
    wn = (w / w.max()).contiguous()
    g = torch.Generator(device=DEV).manual_seed(3)
    cb = init_codebook(x, k, 99) if book == "uniform-box" else x[torch.randint(0, n, (k,), device=DEV, generator=g)].clone()
    acc, _, idx = L.assign_accumulate(x, wn, cb, tensor_cores="fused")
    ref_x = torch.zeros(k, d, device=DEV, dtype=torch.float64).index_add_(0, idx, (x * wn[:, None]).double())
    ref_w = torch.zeros(k, device=DEV, dtype=torch.float64).index_add_(0, idx, wn.double())
    ref_c = ref_x / (ref_w[:, None] + 1e-5)
    got_c = acc[:, :d].double() / (acc[:, d].double()[:, None] + 1e-5)
    err = (got_c - ref_c).abs().max(1).values / (ref_c.abs().max(1).values + 1e-30)
    assert float(err[ref_w > 0].max()) < CB_RTOL, f"centroid error {float(err[ref_w > 0].max()):.2e}"
    assert torch.allclose(acc[:, d].double(), ref_w, rtol=1e-5, atol=1e-9)
    assert not acc[:, d + 1:].any()


# ---------------------------------------------------------------- outlier codewords ----------------
@pytest.mark.parametrize("d", [6, 45, 48])
def test_assign_outlier_codewords_stay_exact_and_off_the_fallback(L, oracle, d):
    """The trace normalisation of the shape codebook (vq.py:76-77) turns unused codewords into far outliers (their
    random initial trace can be ~0).  They must not widen every row's error band: results stay bit-exact, and only the
    rows that are themselves far out may take the exact scan."""
    n, k = 20_000, 512
    x, _, cb = _synth(n, d, k, seed=77 + d)
    g = torch.Generator().manual_seed(d)
    cb[5] *= 3.0e3
    cb[100] *= 1.0e6
    cb[101] = cb[101] * 1.0e19
    cb[300] = -cb[300] * 4.0e4
    cb[301, 0] = 3.0e38
    cb[400] = float("nan")
    far = torch.randint(0, n, (40,), generator=g)
    x[far[:20]] = cb[5] + 0.5 * torch.randn(20, d, generator=g)           # rows whose nearest code IS an outlier
    x[far[20:]] = cb[300] * (1 + 1e-3 * torch.randn(20, 1, generator=g))
    dist_o, idx_o = oracle.assign(x.numpy(), cb.numpy())
    assert (idx_o[far.numpy()[:20]] == 5).all()
    dist, idx = L.assign(x.to(DEV), cb.to(DEV), exact=False)
    rows, cands, fb = L.assign_stats(DEV)
    assert np.array_equal(idx.cpu().numpy(), idx_o)
    assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))
    assert rows == n and fb <= 200, f"{fb} of {n} rows took the exact scan"


def test_trace_normalised_shape_codebook_keeps_the_tensor_core_path(L, oracle):
    """vq_features(..., scale_normalize=True) on covariance rows: after the first step the codebook holds dead codewords
    divided by an arbitrary trace; the mini-batch steps must keep screening on the tensor cores."""
    n, k, b = 200_000, 4096, 65_536
    g = torch.Generator().manual_seed(9)
    s = torch.nn.functional.normalize(torch.rand(n, 3, generator=g) + 0.05)
    r = torch.randn(n, 4, generator=g)
    cov = L.build_covariance(s.to(DEV), r.to(DEV))
    w = (torch.rand(n, generator=g) ** 4).to(DEV)
    from c3dgs_b200.vq import VectorQuantize
    m = VectorQuantize(6, k, 0.8).to(DEV)
    amin, amax = cov.aminmax()
    m.codebook.data = (torch.rand(k, 6, generator=g).to(DEV) * (amax - amin) + amin)
    for step in range(4):
        batch = torch.randint(0, n, (b,), generator=g).to(DEV)
        cb_before = m.codebook.data.clone()
        dist = m.update_indexed(cov, w, batch, scale_normalize=True, return_dist=True)
        rows, cands, fb = L.assign_stats(DEV)
        assert rows == b and fb <= b // 20, f"step {step}: {fb} of {b} rows took the exact scan"
        dist_o, _ = oracle.assign(cov[batch].cpu().numpy(), cb_before.cpu().numpy())
        assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32))


def test_full_size_shape_codebook_trace_normalised(L):
    """BASELINE configs[2] at full size on the fork's real shape width: 5.8 M normalised covariances x 6-D, K = 4096,
    full-batch updates with the trace normalisation (vq.py:76-77).  Every iteration's tensor-core result must equal the
    SIMT exact kernel's on the same codebook bit for bit, and the outlier codewords the normalisation creates must not
    push more than 2 % of the rows to the exact scan."""
    n, k = 5_800_000, 4096
    g = torch.Generator(device=DEV).manual_seed(3)
    s = torch.nn.functional.normalize(torch.rand(n, 3, device=DEV, generator=g) + 0.05)
    r = torch.randn(n, 4, device=DEV, generator=g)
    cov = L.build_covariance(s, r)
    w = torch.rand(n, device=DEV, generator=g) ** 4
    amin, amax = cov.aminmax()
    cb = (torch.rand(k, 6, device=DEV, generator=g) * (amax - amin) + amin).contiguous()
    ei = torch.zeros(k, device=DEV)
    for it in range(4):
        d_exact, i_exact = L.assign(cov, cb, exact=True)
        d_tc, i_tc = L.update(cov, w, cb, ei, 0.8, 1e-5, scale_normalize=True, tensor_cores=True, return_dist=True,
                              return_idx=True)
        rows, cands, fb = L.assign_stats(DEV)
        assert torch.equal(d_tc, d_exact), f"iteration {it}: min distances differ"
        assert torch.equal(i_tc, i_exact), f"iteration {it}: assignments differ"
        assert rows == n and fb <= n // 50, f"iteration {it}: {fb} of {n} rows took the exact scan"
        tr = cb[:, 0] + cb[:, 3] + cb[:, 5]
        fin = torch.isfinite(tr)
        assert (tr[fin] - 1).abs().max() < 1e-3


def test_outlier_rows_keep_their_sums_in_every_mode(L, oracle):
    """Rows that the outlier check sends to the exact scan after the re-rank must still contribute their weighted row
    to the centroid sums, in the sorted, the fused and the SIMT mode alike (codebooks within 1e-5 of the oracle)."""
    n, k, d = 30_000, 512, 6
    g = torch.Generator().manual_seed(21)
    s = torch.nn.functional.normalize(torch.rand(n, 3, generator=g) + 0.05)
    r = torch.randn(n, 4, generator=g)
    cov = L.build_covariance(s.to(DEV), r.to(DEV)).cpu()
    w = torch.rand(n, generator=g) ** 4
    amin, amax = cov.aminmax()
    cb0 = torch.rand(k, d, generator=g) * (amax - amin) + amin
    cb0[7] *= 2.0e3                                              # planted outliers, one of them with members
    cb0[9] = cov[5] * 300.0
    cov[100:140] = cb0[9] + 0.01 * torch.randn(40, d, generator=g)
    cb_o, ei_o = cb0.numpy().copy(), np.zeros(k, np.float32)
    dist_o, idx_o = oracle.assign(cov.numpy(), cb_o)
    assert (idx_o[100:140] == 9).all()
    oracle.update(cov.numpy(), w.numpy(), cb_o, ei_o, 0.8)
    for mode in (True, "fused", False):
        cb, ei = cb0.clone().to(DEV), torch.zeros(k, device=DEV)
        dist, idx = L.update(cov.to(DEV), w.to(DEV), cb, ei, 0.8, 1e-5, tensor_cores=mode, return_dist=True,
                             return_idx=True)
        assert np.array_equal(idx.cpu().numpy(), idx_o), mode
        assert np.array_equal(dist.cpu().numpy().view(np.uint32), dist_o.view(np.uint32)), mode
        if mode:
            rows, _, fb = L.assign_stats(DEV)
            assert 40 <= fb <= 2000, f"{mode}: {fb} rows in the exact scan"
        _cb_close(cb.cpu().numpy(), cb_o)
        assert np.allclose(ei.cpu().numpy(), ei_o, rtol=1e-5, atol=1e-7)
