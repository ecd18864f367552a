"""GPU (-m gpu), collected last: c3vq_assign_static, the allocation-free variant of c3vq_assign whose scratch is a
fixed __device__ array inside the library (include/c3vq.h; no counterpart in the reference).

Called through the C ABI with raw pointers, checked against the existing entry point c3vq_assign on the same
inputs.  Bar: the one the assignment tests use, bit-exact indices and minimum distances.  Two calls in a row with
different shapes, so the second one runs on the array as the first left it."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

DEV = "cuda:0"
STATIC_BYTES = 4 << 20          # C3VQ_STATIC_SCRATCH_BYTES


def _inputs(n, d, k, seed):
    g = torch.Generator().manual_seed(seed)
    centres = torch.randn(max(k // 2, 1), d, generator=g)
    x = centres[torch.randint(0, centres.shape[0], (n,), generator=g)] + 0.1 * torch.randn(n, d, generator=g)
    lo, hi = x.aminmax()
    cb = torch.rand(k, d, generator=g) * (hi - lo) + lo
    cb[k // 3] = cb[k // 3 + 1]                     # an exact tie: the lower index must win in both entry points
    return x.contiguous().to(DEV), cb.contiguous().to(DEV)


def _p(t):
    return ctypes.c_void_p(t.data_ptr() if t is not None else None)


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream(torch.device(DEV)).cuda_stream)


def _both(L, x, cb, rows=None):
    """(dist, idx) of c3vq_assign with a caller workspace and of c3vq_assign_static, raw C-ABI calls."""
    n = rows.numel() if rows is not None else x.shape[0]
    d, k = x.shape[1], cb.shape[0]
    ws_bytes = L.c3vq_assign_workspace_bytes(n, d, k)
    assert ws_bytes <= STATIC_BYTES
    ws = torch.empty(ws_bytes, dtype=torch.uint8, device=DEV)
    out = []
    for static in (False, True):
        dist = torch.full((n,), -1.0, dtype=torch.float32, device=DEV)
        idx = torch.full((n,), -1, dtype=torch.int64, device=DEV)
        torch.cuda.synchronize()
        before = torch.cuda.memory_allocated(DEV)
        if static:
            rc = L.c3vq_assign_static(_p(x), _p(rows), n, d, _p(cb), k, _p(dist), _p(idx), _stream())
        else:
            rc = L.c3vq_assign(_p(x), _p(rows), n, d, _p(cb), k, _p(dist), _p(idx), _p(ws), ws_bytes, _stream())
        assert rc == 0, L.c3vq_last_error().decode()
        torch.cuda.synchronize()
        assert torch.cuda.memory_allocated(DEV) == before
        out.append((dist.cpu(), idx.cpu()))
    return out


def test_assign_static_matches_assign_twice_in_a_row(lib):
    assert torch.cuda.is_available()
    L = lib
    # first call: the largest advertised shape (65536 rows x 64 against 4096 codewords) fills most of the array;
    # second call: a small ragged one through a row list, on whatever the first left behind; third: the first again
    x1, cb1 = _inputs(65536, 64, 4096, seed=1)
    x2, cb2 = _inputs(5003, 45, 1000, seed=2)
    rows2 = torch.randperm(5003, generator=torch.Generator().manual_seed(3))[:3001].to(DEV)
    for x, cb, rows in ((x1, cb1, None), (x2, cb2, rows2), (x1, cb1, None)):
        (dist_a, idx_a), (dist_s, idx_s) = _both(L, x, cb, rows)
        assert int(idx_a.min()) >= 0 and int(idx_a.max()) < cb.shape[0]
        assert torch.equal(idx_s, idx_a)
        assert torch.equal(dist_s, dist_a)            # bit-exact, like the assignment parity tests
    # against plain fp32 torch arithmetic as a sanity check that both are a real nearest-codeword search
    sel = torch.arange(0, 5003, 37, device=DEV)
    ref = torch.cdist(x2[sel].double(), cb2.double()).argmin(1).cpu()
    _, (_, idx_s) = _both(L, x2, cb2, sel)
    assert (idx_s == ref).float().mean() > 0.99       # fp32 vs float64 may differ on near ties only


def test_assign_static_rejects_sizes_beyond_the_array(lib):
    L = lib
    n, d, k = 400_000, 64, 4096
    assert L.c3vq_assign_workspace_bytes(n, d, k) > STATIC_BYTES
    x = torch.zeros(n, d, device=DEV)
    cb = torch.zeros(k, d, device=DEV)
    idx = torch.full((n,), -7, dtype=torch.int64, device=DEV)
    rc = L.c3vq_assign_static(_p(x), _p(None), n, d, _p(cb), k, _p(None), _p(idx), _stream())
    assert rc == -2                                                       # C3VQ_EWORKSPACE
    msg = L.c3vq_last_error().decode()
    assert "c3vq_assign_static" in msg and "400000" in msg and str(STATIC_BYTES) in msg
    torch.cuda.synchronize()
    assert int(idx.min()) == -7 and int(idx.max()) == -7                  # nothing was launched
    # n == 0 is a no-op like c3vq_assign
    assert L.c3vq_assign_static(_p(x), _p(None), 0, d, _p(cb), k, _p(None), _p(None), _stream()) == 0


def test_assign_static_host_wrapper(lib):
    from c3dgs_b200 import _lib
    x, cb = _inputs(2049, 7, 256, seed=9)
    idx = torch.empty(2049, dtype=torch.int64, device=DEV)
    dist = torch.empty(2049, dtype=torch.float32, device=DEV)
    _lib.assign_static(x, cb, dist, idx)
    d0, i0 = _lib.assign(x, cb)
    assert torch.equal(idx, i0) and torch.equal(dist, d0)
