"""ctypes binding of libc3vq.so (include/c3vq.h) for torch CUDA tensors.

PyTorch is plumbing here: it owns device memory and streams; every operation below is
one call through the C ABI with raw device pointers and the current stream.  There is
no CPU fallback and no alternative backend: if the library is missing or a tensor is
not a CUDA tensor the call raises.
"""
from __future__ import annotations

import ctypes
import os
import threading

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("C3VQ_LIB") or os.path.join(_HERE, "libc3vq.so")   # C3VQ_LIB: developer builds

_lib = None
_lock = threading.Lock()
DEFAULT_SUMS = "fused"


def default_tensor_cores() -> bool:
    """Assignment kernel used by default: the tcgen05 path; C3VQ_ASSIGN=simt selects the SIMT exact
    kernel (same results, ~10x slower) for debugging.  Both are CUDA kernels of libc3vq.so."""
    return os.environ.get("C3VQ_ASSIGN", "tc").lower() != "simt"


def default_sums() -> str:
    """Centroid sums after / inside the tensor-core assignment: "sorted" (counting sort by codeword + segmented sum,
    second pass over x) or "fused" (one bulk reduce-add per distinct codeword of a warp from the rows staged in
    shared memory; x read once).  C3VQ_SUMS overrides the default."""
    v = os.environ.get("C3VQ_SUMS", DEFAULT_SUMS).lower()
    return "fused" if v == "fused" else "sorted"


def _tc_mode(tensor_cores) -> int:
    """use_tensor_cores argument of c3vq_update / c3vq_assign_accumulate: 0 = SIMT exact assignment + atomic sums,
    1 = tcgen05 assignment + sorted segmented sums (second pass over x), 2 = tcgen05 assignment with the sums fused
    into the kernel (one bulk reduce-add per distinct codeword of a warp; x read once: the default).  True selects the
    default (C3VQ_SUMS=sorted|fused overrides it), "sorted" / "fused" select explicitly."""
    if tensor_cores == "fused" or (tensor_cores == 2 and tensor_cores is not True):
        return 2
    if tensor_cores == "sorted":
        return 1
    if not tensor_cores:
        return 0
    return 2 if default_sums() == "fused" else 1


_c_f32p = ctypes.c_void_p
_c_i64 = ctypes.c_int64
_c_int = ctypes.c_int
_c_dbl = ctypes.c_double
_c_sz = ctypes.c_size_t
_c_vp = ctypes.c_void_p

# name -> (restype, argtypes); mirrors include/c3vq.h one to one (tests check every symbol resolves)
SIGNATURES = {
    "c3vq_version": (_c_int, []),
    "c3vq_last_error": (ctypes.c_char_p, []),
    "c3vq_launch_count": (_c_i64, []),
    "c3vq_assign_workspace_bytes": (_c_sz, [_c_i64, _c_int, _c_int]),
    "c3vq_assign": (_c_int, [_c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_int, _c_vp, _c_vp, _c_vp, _c_sz, _c_vp]),
    "c3vq_assign_exact": (_c_int, [_c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_int, _c_vp, _c_vp, _c_vp]),
    "c3vq_assign_static": (_c_int, [_c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_int, _c_vp, _c_vp, _c_vp]),
    "c3vq_acc_stride": (_c_int, [_c_int]),
    "c3vq_accumulate": (_c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_int, _c_int, _c_vp, _c_vp]),
    "c3vq_accumulate_sorted_workspace_bytes": (_c_sz, [_c_i64, _c_int]),
    "c3vq_accumulate_sorted": (_c_int, [_c_vp, _c_vp, _c_vp, _c_vp, _c_i64, _c_int, _c_int, _c_vp, _c_vp, _c_sz, _c_vp]),
    "c3vq_ema": (_c_int, [_c_vp, _c_vp, _c_vp, _c_int, _c_int, _c_dbl, _c_dbl, _c_int, _c_vp]),
    "c3vq_update_workspace_bytes": (_c_sz, [_c_i64, _c_int, _c_int]),
    "c3vq_update": (_c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_vp, _c_int, _c_dbl, _c_dbl, _c_int,
                             _c_int, _c_vp, _c_vp, _c_vp, _c_sz, _c_vp]),
    "c3vq_assign_accumulate": (_c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_int, _c_int, _c_vp, _c_vp,
                                        _c_vp, _c_vp, _c_sz, _c_vp]),
    "c3vq_gather": (_c_int, [_c_vp, _c_i64, _c_int, _c_vp, _c_i64, _c_vp, _c_vp]),
    "c3vq_gather_workspace_bytes": (_c_sz, [_c_i64, _c_int, _c_i64]),
    "c3vq_gather_ws": (_c_int, [_c_vp, _c_i64, _c_int, _c_vp, _c_i64, _c_vp, _c_vp, _c_sz, _c_vp]),
    "c3vq_scatter_add": (_c_int, [_c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_i64, _c_vp]),
    "c3vq_join_workspace_bytes": (_c_sz, [_c_i64]),
    "c3vq_join_count": (_c_int, [_c_vp, _c_i64, ctypes.POINTER(_c_i64), _c_vp, _c_sz, _c_vp]),
    "c3vq_join_write": (_c_int, [_c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_i64, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp]),
    "c3vq_extract_rot_scale": (_c_int, [_c_vp, _c_i64, _c_int, _c_vp, _c_vp, _c_vp]),
    "c3vq_morton_workspace_bytes": (_c_sz, [_c_i64]),
    "c3vq_morton_order": (_c_int, [_c_vp, _c_i64, _c_vp, _c_vp, _c_vp, _c_sz, _c_vp]),
    "c3vq_quantize_i8": (_c_int, [_c_vp, _c_i64, _c_dbl, _c_i64, _c_vp, _c_vp]),
    "c3vq_dequantize_i8": (_c_int, [_c_vp, _c_i64, ctypes.c_float, ctypes.c_int32, _c_vp, _c_vp]),
    "c3vq_convert": (_c_int, [_c_int, _c_vp, _c_i64, _c_vp, _c_vp]),
    "c3vq_build_covariance": (_c_int, [_c_vp, _c_vp, _c_i64, ctypes.c_float, _c_int, _c_vp, _c_vp]),
    "c3vq_compact_rows": (_c_int, [_c_vp, _c_vp, _c_i64, _c_i64, _c_vp, _c_vp, _c_vp]),
    "c3vq_preprocess_indexed": (_c_int, [_c_i64, _c_int, _c_int, _c_vp, _c_vp, _c_vp, ctypes.c_float, _c_vp, _c_vp, _c_vp,
                                         _c_vp, _c_vp, _c_vp, _c_int, _c_int, ctypes.c_float, ctypes.c_float,
                                         ctypes.c_float, ctypes.c_float, _c_int, _c_vp, _c_vp, _c_int, _c_vp, _c_vp,
                                         _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp]),
    "c3vq_preprocess_indexed_backward": (_c_int, [_c_i64, _c_int, _c_int, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp,
                                                  ctypes.c_float, _c_vp, _c_vp, _c_vp, ctypes.c_float, ctypes.c_float,
                                                  ctypes.c_float, ctypes.c_float, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp,
                                                  _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp, _c_vp]),
    "c3vq_gather_rows": (_c_int, [_c_vp, _c_vp, _c_vp, _c_i64, _c_int, _c_vp, _c_vp, _c_vp]),
    "c3vq_debug_scores": (_c_int, [_c_vp, _c_i64, _c_int, _c_vp, _c_int, _c_vp, _c_vp, _c_vp, _c_vp, _c_sz, _c_vp]),
    "c3vq_assign_stats": (_c_int, [_c_vp, ctypes.POINTER(_c_i64), _c_vp]),
}


def load() -> ctypes.CDLL:
    """Load libc3vq.so; raises if it has not been built (python -m c3dgs_b200.build)."""
    global _lib
    with _lock:
        if _lib is None:
            if not os.path.exists(LIB_PATH):
                raise RuntimeError(
                    f"{LIB_PATH} is missing: build it with `python -m c3dgs_b200.build` "
                    "(nvcc, sm_100a). There is no fallback implementation.")
            L = ctypes.CDLL(LIB_PATH)
            for name, (res, args) in SIGNATURES.items():
                fn = getattr(L, name)  # AttributeError if the symbol is not exported
                fn.restype = res
                fn.argtypes = args
            _lib = L
    return _lib


def _err(rc: int) -> RuntimeError:
    msg = load().c3vq_last_error()
    return RuntimeError(f"libc3vq error {rc}: {msg.decode() if msg else '?'}")


def _stream(t: torch.Tensor) -> int:
    return torch.cuda.current_stream(t.device).cuda_stream


def _ptr(t):
    return None if t is None else t.data_ptr()


def _check(t: torch.Tensor, dtype, name: str, ndim=None):
    if not isinstance(t, torch.Tensor):
        raise RuntimeError(f"{name} must be a torch.Tensor")
    if not t.is_cuda:
        raise RuntimeError(f"{name} must be a CUDA tensor (libc3vq has no CPU path)")
    if t.dtype != dtype:
        raise RuntimeError(f"{name} must be {dtype}, got {t.dtype}")
    if ndim is not None and t.dim() != ndim:
        raise RuntimeError(f"{name} must have {ndim} dimensions, got {t.dim()}")


def _check_rows(rows: torch.Tensor, n_rows: int, trusted: bool = False) -> torch.Tensor:
    """Mini-batch indirection: int64 row numbers into x.  The kernels read x[rows[i]] unchecked, so the range is
    verified here (the reference's features[batch] raises a device-side assert for the same input) unless the
    caller vouches for it (`rows_checked=True`: indices drawn by torch.randint(0, n), as vq_features does — the
    check costs a reduction and a host sync per call)."""
    _check(rows, torch.int64, "rows", 1)
    rows = rows.contiguous()
    if rows.numel() and not trusted:
        lo, hi = torch.aminmax(rows)
        if int(lo) < 0 or int(hi) >= n_rows:
            raise IndexError(f"rows must lie in [0, {n_rows}), got [{int(lo)}, {int(hi)}]")
    return rows


class Workspace:
    """Grow-only device scratch owned by the Python side (the C side never allocates), one buffer per
    (device, stream).  The scratch holds state a running kernel depends on (stats header, TF32 codebook image,
    fallback list, sorted-sum tables), so two streams of one device must not share it; a buffer is allocated while
    its stream is current, which is also what makes dropping it on a grow safe (the caching allocator hands a freed
    block to later work of the same stream only).  Callers on one stream are serialised by the stream itself;
    Python threads are serialised by the lock around the lookup."""

    def __init__(self):
        self._buf = {}
        self._lock = threading.Lock()

    def get(self, device: torch.device, nbytes: int, tag: str = "assign") -> torch.Tensor:
        """`tag` separates scratch areas with different layouts: "assign" starts with the assignment's stats
        header (read back by assign_stats), "sums" is the sorted-sum tables of a stand-alone accumulate."""
        index = device.index if device.index is not None else torch.cuda.current_device()
        dev = torch.device("cuda", index)
        key = (index, torch.cuda.current_stream(dev).cuda_stream, tag)
        with self._lock:
            buf = self._buf.get(key)
            if buf is None or buf.numel() < nbytes:
                with torch.cuda.device(dev):
                    buf = torch.empty(max(int(nbytes), 256), dtype=torch.uint8, device=dev)
                self._buf[key] = buf
        return buf


_workspace = Workspace()


def assign(x: torch.Tensor, codebook: torch.Tensor, rows: torch.Tensor | None = None, exact: bool = False,
           return_dist: bool = True, out_idx: torch.Tensor | None = None):
    """Nearest codeword per row -> (min_dist f32[n] | None, idx i64[n]).  c3vq_assign / c3vq_assign_exact.
    `out_idx`: optional contiguous int64 [n] destination (e.g. a slice of a larger index array)."""
    _check(x, torch.float32, "coefs")
    _check(codebook, torch.float32, "codebook")
    if x.dim() != 2 or codebook.dim() != 2:
        raise RuntimeError("ceofs and codebook must have dimension 2")          # weighted_distance.cu:64-67
    if codebook.size(1) != x.size(1):
        raise RuntimeError("coefs and codebook must have same number of channels")  # weighted_distance.cu:69-72
    if codebook.device != x.device:
        raise RuntimeError("coefs and codebook must be on the same device")
    x = x.contiguous()
    codebook = codebook.contiguous()
    if rows is not None:
        _check(rows, torch.int64, "rows", 1)
        rows = rows.contiguous()
    n = rows.numel() if rows is not None else x.size(0)
    d, k = x.size(1), codebook.size(0)
    L = load()
    with torch.cuda.device(x.device):
        dist = torch.empty(n, dtype=torch.float32, device=x.device) if return_dist else None
        if out_idx is not None:
            _check(out_idx, torch.int64, "out_idx", 1)
            if out_idx.numel() != n or not out_idx.is_contiguous() or out_idx.device != x.device:
                raise RuntimeError("out_idx must be a contiguous int64 tensor with one entry per row on x's device")
            idx = out_idx
        else:
            idx = torch.empty(n, dtype=torch.int64, device=x.device)
        if exact:
            rc = L.c3vq_assign_exact(_ptr(x), _ptr(rows), n, d, _ptr(codebook), k, _ptr(dist), _ptr(idx),
                                     _stream(x))
        else:
            nbytes = L.c3vq_assign_workspace_bytes(n, d, k)
            ws = _workspace.get(x.device, nbytes)
            rc = L.c3vq_assign(_ptr(x), _ptr(rows), n, d, _ptr(codebook), k, _ptr(dist), _ptr(idx),
                               _ptr(ws), ws.numel(), _stream(x))
    if rc:
        raise _err(rc)
    return dist, idx


STATIC_SCRATCH_BYTES = 4 << 20      # C3VQ_STATIC_SCRATCH_BYTES of include/c3vq.h


def assign_static(x: torch.Tensor, codebook: torch.Tensor, dist: torch.Tensor | None, idx: torch.Tensor,
                  rows: torch.Tensor | None = None) -> None:
    """c3vq_assign_static: nearest codeword per row into the caller's `idx` (int64 [n]) and optional `dist`
    (float32 [n]); the scratch is the library's fixed device array, nothing is allocated here or in the library.
    No counterpart in the reference (added for callers that cannot allocate); tensors must already be contiguous."""
    for t, dt, nm in ((x, torch.float32, "coefs"), (codebook, torch.float32, "codebook"), (idx, torch.int64, "idx")):
        _check(t, dt, nm)
        if not t.is_contiguous():
            raise RuntimeError(f"{nm} must be contiguous (assign_static makes no copies)")
    n = rows.numel() if rows is not None else x.size(0)
    if idx.numel() != n or (dist is not None and (dist.numel() != n or dist.dtype != torch.float32)):
        raise RuntimeError("idx / dist must hold one entry per row")
    with torch.cuda.device(x.device):
        rc = load().c3vq_assign_static(_ptr(x), _ptr(rows), n, x.size(1), _ptr(codebook), codebook.size(0),
                                       _ptr(dist), _ptr(idx), _stream(x))
    if rc:
        raise _err(rc)


def update(x: torch.Tensor, w: torch.Tensor, codebook: torch.Tensor, entry_importance: torch.Tensor,
           decay: float, eps: float, rows: torch.Tensor | None = None, scale_normalize: bool = False,
           tensor_cores: bool = True, return_dist: bool = True, return_idx: bool = False, rows_checked: bool = False):
    """One VectorQuantize.update (vq.py:28-35) in place on codebook / entry_importance.  c3vq_update."""
    _check(x, torch.float32, "x", 2)
    _check(w, torch.float32, "importance", 1)
    _check(codebook, torch.float32, "codebook", 2)
    _check(entry_importance, torch.float32, "entry_importance", 1)
    if not (codebook.is_contiguous() and entry_importance.is_contiguous()):
        raise RuntimeError("codebook and entry_importance must be contiguous (they are updated in place)")
    if codebook.size(1) != x.size(1):
        raise RuntimeError("coefs and codebook must have same number of channels")
    x = x.contiguous()
    w = w.contiguous()
    if w.numel() != x.size(0):
        # with `rows` the kernels read w[rows[i]]: importance is indexed like x, not like the batch
        raise RuntimeError("importance must have one entry per row of x")
    if rows is not None:
        rows = _check_rows(rows, x.size(0), rows_checked)
        n = rows.numel()
    else:
        n = x.size(0)
    d, k = x.size(1), codebook.size(0)
    if scale_normalize and d < 6:
        raise RuntimeError("scale_normalize needs the 6 stored covariance entries (columns 0, 3, 5 are the trace)")
    L = load()
    with torch.cuda.device(x.device):
        dist = torch.empty(n, dtype=torch.float32, device=x.device) if return_dist else None
        idx = torch.empty(n, dtype=torch.int64, device=x.device) if return_idx else None
        nbytes = L.c3vq_update_workspace_bytes(n, d, k)
        ws = _workspace.get(x.device, nbytes)
        rc = L.c3vq_update(_ptr(x), _ptr(w), _ptr(rows), n, d, _ptr(codebook), _ptr(entry_importance), k,
                           float(decay), float(eps), int(bool(scale_normalize)), _tc_mode(tensor_cores),
                           _ptr(dist), _ptr(idx), _ptr(ws), ws.numel(), _stream(x))
    if rc:
        raise _err(rc)
    return dist, idx


def assign_accumulate(x, w, codebook, acc: torch.Tensor | None = None, rows=None, tensor_cores: bool = True,
                      return_dist: bool = False, return_idx: bool = True, rows_checked: bool = False,
                      out_idx: torch.Tensor | None = None):
    """Assignment + weighted sums added into acc, no EMA (multi-GPU building block).  c3vq_assign_accumulate."""
    _check(x, torch.float32, "x", 2)
    _check(w, torch.float32, "importance", 1)
    _check(codebook, torch.float32, "codebook", 2)
    if codebook.size(1) != x.size(1):
        raise RuntimeError("coefs and codebook must have same number of channels")
    x, w, codebook = x.contiguous(), w.contiguous(), codebook.contiguous()
    if w.numel() != x.size(0):
        raise RuntimeError("importance must have one entry per row of x")
    if rows is not None:
        rows = _check_rows(rows, x.size(0), rows_checked)
    n = rows.numel() if rows is not None else x.size(0)
    d, k = x.size(1), codebook.size(0)
    L = load()
    stride = L.c3vq_acc_stride(d)
    with torch.cuda.device(x.device):
        if acc is None:
            acc = torch.zeros(k, stride, dtype=torch.float32, device=x.device)
        elif acc.shape != (k, stride) or not acc.is_contiguous():
            raise RuntimeError(f"acc must be a contiguous [{k}, {stride}] float32 tensor")
        dist = torch.empty(n, dtype=torch.float32, device=x.device) if return_dist else None
        if out_idx is not None:
            _check(out_idx, torch.int64, "out_idx", 1)
            if out_idx.numel() != n or not out_idx.is_contiguous() or out_idx.device != x.device:
                raise RuntimeError("out_idx must be a contiguous int64 tensor with one entry per row on x's device")
            idx = out_idx
        else:
            idx = torch.empty(n, dtype=torch.int64, device=x.device) if return_idx else None
        nbytes = L.c3vq_update_workspace_bytes(n, d, k)
        ws = _workspace.get(x.device, nbytes)
        rc = L.c3vq_assign_accumulate(_ptr(x), _ptr(w), _ptr(rows), n, d, _ptr(codebook), k, _tc_mode(tensor_cores),
                                      _ptr(acc), _ptr(dist), _ptr(idx), _ptr(ws), ws.numel(), _stream(x))
    if rc:
        raise _err(rc)
    return acc, dist, idx


def launch_count() -> int:
    return int(load().c3vq_launch_count())


def acc_stride(d: int) -> int:
    return int(load().c3vq_acc_stride(int(d)))


def accumulate(x, w, idx, k: int, acc: torch.Tensor | None = None, rows=None, sorted_path: bool = False) -> torch.Tensor:
    """Weighted centroid sums into acc[k, acc_stride(d)] (vq.py:31,33).  c3vq_accumulate, or with
    sorted_path the atomic-free c3vq_accumulate_sorted (counting sort by codeword + segmented sum)."""
    _check(x, torch.float32, "x", 2)
    _check(idx, torch.int64, "idx", 1)
    x, idx = x.contiguous(), idx.contiguous()
    if w is not None:
        _check(w, torch.float32, "importance", 1)
        w = w.contiguous()
    if rows is not None:
        _check(rows, torch.int64, "rows", 1)
        rows = rows.contiguous()
    n, d = idx.numel(), x.size(1)
    L = load()
    stride = L.c3vq_acc_stride(d)
    with torch.cuda.device(x.device):
        if acc is None:
            acc = torch.zeros(k, stride, dtype=torch.float32, device=x.device)
        elif acc.shape != (k, stride) or not acc.is_contiguous():
            raise RuntimeError(f"acc must be a contiguous [{k}, {stride}] float32 tensor")
        if sorted_path:
            nbytes = L.c3vq_accumulate_sorted_workspace_bytes(n, k)
            ws = _workspace.get(x.device, nbytes, "sums")
            rc = L.c3vq_accumulate_sorted(_ptr(x), _ptr(w), _ptr(rows), _ptr(idx), n, d, k, _ptr(acc), _ptr(ws),
                                          nbytes, _stream(x))
        else:
            rc = L.c3vq_accumulate(_ptr(x), _ptr(w), _ptr(rows), _ptr(idx), n, d, k, _ptr(acc), _stream(x))
    if rc:
        raise _err(rc)
    return acc


def ema(codebook, entry_importance, acc, decay: float, eps: float, scale_normalize: bool = False) -> None:
    """EMA epilogue (vq.py:32,34,45-46,76-77) in place.  c3vq_ema."""
    _check(codebook, torch.float32, "codebook", 2)
    _check(entry_importance, torch.float32, "entry_importance", 1)
    _check(acc, torch.float32, "acc", 2)
    k, d = codebook.shape
    if acc.shape != (k, acc_stride(d)) or not (acc.is_contiguous() and codebook.is_contiguous()):
        raise RuntimeError("acc/codebook layout mismatch")
    with torch.cuda.device(codebook.device):
        rc = load().c3vq_ema(_ptr(codebook), _ptr(entry_importance), _ptr(acc), k, d, float(decay), float(eps),
                             int(bool(scale_normalize)), _stream(codebook))
    if rc:
        raise _err(rc)


def gather(table: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """out[i] = table[idx[i]] for a [R, ...] float32 table.  c3vq_gather / c3vq_gather_ws."""
    _check(table, torch.float32, "table")
    _check(idx, torch.int64, "idx", 1)
    if idx.device != table.device:
        raise RuntimeError("table and idx must be on the same device")
    table, idx = table.contiguous(), idx.contiguous()
    rows = table.size(0)
    d = table[0].numel() if rows > 0 else int(torch.Size(table.shape[1:]).numel())
    n = idx.numel()
    with torch.cuda.device(table.device):
        out = torch.empty((n,) + tuple(table.shape[1:]), dtype=torch.float32, device=table.device)
        L = load()
        nbytes = L.c3vq_gather_workspace_bytes(rows, d, n)     # > 0: 180-byte rows & co. decode from padded copies
        if nbytes:
            ws = _workspace.get(table.device, nbytes, "gather")
            rc = L.c3vq_gather_ws(_ptr(table), rows, d, _ptr(idx), n, _ptr(out), _ptr(ws), ws.numel(), _stream(table))
        else:
            rc = L.c3vq_gather(_ptr(table), rows, d, _ptr(idx), n, _ptr(out), _stream(table))
    if rc:
        raise _err(rc)
    return out


def join_features(all_features: torch.Tensor, keep_mask: torch.Tensor, codebook: torch.Tensor,
                  codebook_indices: torch.Tensor):
    """compression/vq.py:90-103 in two kernels: (cat(codebook, all_features[keep_mask]), remapped int64 indices).
    c3vq_join_count + c3vq_join_write."""
    _check(all_features, torch.float32, "all_features", 2)
    _check(codebook, torch.float32, "codebook", 2)
    _check(codebook_indices, torch.int64, "codebook_indices", 1)
    if keep_mask.dtype != torch.bool or keep_mask.dim() != 1 or not keep_mask.is_cuda:
        raise RuntimeError("keep_mask must be a 1-D CUDA bool tensor")
    n, d = all_features.shape
    if keep_mask.numel() != n:
        raise RuntimeError("keep_mask must have one entry per row of all_features")
    if codebook.numel() and codebook.size(1) != d:
        raise RuntimeError("codebook and all_features must have the same number of channels")
    all_features, keep_mask = all_features.contiguous(), keep_mask.contiguous()
    codebook, codebook_indices = codebook.contiguous(), codebook_indices.contiguous()
    k = codebook.size(0)
    L = load()
    with torch.cuda.device(all_features.device):
        nbytes = L.c3vq_join_workspace_bytes(n)
        ws = torch.empty(max(int(nbytes), 4), dtype=torch.uint8, device=all_features.device)
        kept = _c_i64(0)
        rc = L.c3vq_join_count(_ptr(keep_mask), n, ctypes.byref(kept), _ptr(ws), ws.numel(), _stream(all_features))
        if rc:
            raise _err(rc)
        kept = int(kept.value)
        if codebook_indices.numel() != n - kept:
            raise RuntimeError(f"codebook_indices has {codebook_indices.numel()} entries for {n - kept} rows "
                               "outside keep_mask")
        out_f = torch.empty(k + kept, d, dtype=torch.float32, device=all_features.device)
        out_i = torch.empty(n, dtype=torch.int64, device=all_features.device)
        rc = L.c3vq_join_write(_ptr(all_features), _ptr(keep_mask), n, d, _ptr(codebook), k, _ptr(codebook_indices),
                               _ptr(out_f), _ptr(out_i), _ptr(ws), _stream(all_features))
    if rc:
        raise _err(rc)
    return out_f, out_i


def extract_rot_scale(cov: torch.Tensor):
    """utils/splats.py:28-36 (with to_full_cov, :7-25, fused away when `cov` is [n,6]): cov [n,6] stored entries or
    [n,3,3] -> (rot [n,4] unit quaternions, scaling [n,3] ascending).  c3vq_extract_rot_scale."""
    _check(cov, torch.float32, "cov")
    if cov.dim() == 2 and cov.size(1) == 6:
        full = 0
    elif cov.dim() == 3 and tuple(cov.shape[1:]) == (3, 3):
        full = 1
    else:
        raise RuntimeError(f"cov must be [n,6] or [n,3,3], got {tuple(cov.shape)}")
    cov = cov.contiguous()
    n = cov.size(0)
    with torch.cuda.device(cov.device):
        rot = torch.empty(n, 4, dtype=torch.float32, device=cov.device)
        scale = torch.empty(n, 3, dtype=torch.float32, device=cov.device)
        rc = load().c3vq_extract_rot_scale(_ptr(cov), n, full, _ptr(rot), _ptr(scale), _stream(cov))
    if rc:
        raise _err(rc)
    return rot, scale


def morton_order(xyz: torch.Tensor, return_keys: bool = False):
    """scene/gaussian_model.py:997-1004: permutation that sorts the rows of xyz [n,3] by their 63-bit Morton key
    (stable).  Returns int64 [n] (and the sorted int64 keys).  c3vq_morton_order."""
    _check(xyz, torch.float32, "xyz", 2)
    if xyz.size(1) != 3:
        raise RuntimeError(f"xyz must be [n,3], got {tuple(xyz.shape)}")
    xyz = xyz.contiguous()
    n = xyz.size(0)
    L = load()
    with torch.cuda.device(xyz.device):
        order = torch.empty(n, dtype=torch.int64, device=xyz.device)
        keys = torch.empty(n, dtype=torch.int64, device=xyz.device) if return_keys else None
        nbytes = L.c3vq_morton_workspace_bytes(n)
        ws = torch.empty(max(int(nbytes), 4), dtype=torch.uint8, device=xyz.device)
        rc = L.c3vq_morton_order(_ptr(xyz), n, _ptr(order), _ptr(keys), _ptr(ws), ws.numel(), _stream(xyz))
    if rc:
        raise _err(rc)
    return (order, keys) if return_keys else order


def _aligned(t: torch.Tensor) -> torch.Tensor:
    """Contiguous and 16-byte aligned (a view into the middle of a storage may not be)."""
    t = t.contiguous()
    return t.clone() if t.data_ptr() % 16 else t


def quantize_i8(x: torch.Tensor, scale, zero_point) -> torch.Tensor:
    """torch.quantize_per_tensor(x, scale, zero_point, torch.qint8).int_repr() (scene/gaussian_model.py:528-620)
    -> int8 tensor of x's shape.  scale / zero_point: python numbers or one-element tensors.  c3vq_quantize_i8."""
    _check(x, torch.float32, "x")
    x = _aligned(x)
    scale = float(scale.item() if isinstance(scale, torch.Tensor) else scale)
    zero_point = int(zero_point.item() if isinstance(zero_point, torch.Tensor) else zero_point)
    with torch.cuda.device(x.device):
        out = torch.empty(x.shape, dtype=torch.int8, device=x.device)
        rc = load().c3vq_quantize_i8(_ptr(x), x.numel(), scale, zero_point, _ptr(out), _stream(x))
    if rc:
        raise _err(rc)
    return out


def dequantize_i8(q: torch.Tensor, scale, zero_point) -> torch.Tensor:
    """(q.int() - zero_point) * scale (scene/gaussian_model.py:644-716) -> float32.  c3vq_dequantize_i8."""
    _check(q, torch.int8, "q")
    q = _aligned(q)
    scale = float(scale.item() if isinstance(scale, torch.Tensor) else scale)
    zero_point = int(zero_point.item() if isinstance(zero_point, torch.Tensor) else zero_point)
    with torch.cuda.device(q.device):
        out = torch.empty(q.shape, dtype=torch.float32, device=q.device)
        rc = load().c3vq_dequantize_i8(_ptr(q), q.numel(), scale, zero_point, _ptr(out), _stream(q))
    if rc:
        raise _err(rc)
    return out


_CONVERSIONS = {  # (src dtype, dst dtype) -> C3VQ_* kind of include/c3vq.h
    (torch.float32, torch.float16): 0, (torch.float16, torch.float32): 1,
    (torch.int64, torch.int32): 2, (torch.int32, torch.int64): 3,
}


def convert(t: torch.Tensor, dtype: torch.dtype) -> torch.Tensor:
    """The element-type conversions of the scene file format (float32 <-> float16 positions, int64 <-> int32
    index arrays; scene/gaussian_model.py:522, :566-569, :639, :723-735).  c3vq_convert."""
    kind = _CONVERSIONS.get((t.dtype, dtype))
    if kind is None:
        raise RuntimeError(f"no wire-format conversion from {t.dtype} to {dtype}")
    _check(t, t.dtype, "t")
    t = _aligned(t)
    with torch.cuda.device(t.device):
        out = torch.empty(t.shape, dtype=dtype, device=t.device)
        rc = load().c3vq_convert(kind, _ptr(t), t.numel(), _ptr(out), _stream(t))
    if rc:
        raise _err(rc)
    return out


def build_covariance(scaling: torch.Tensor, rotation: torch.Tensor, scaling_modifier: float = 1.0,
                     activate: bool = False) -> torch.Tensor:
    """GaussianModel.get_normalized_covariance(strip_sym=True) (scene/gaussian_model.py:274-278): scaling [n,3],
    rotation [n,4] -> cov [n,6].  activate=True takes the raw parameters (:66, :77).  c3vq_build_covariance."""
    _check(scaling, torch.float32, "scaling", 2)
    _check(rotation, torch.float32, "rotation", 2)
    n = scaling.size(0)
    if scaling.size(1) != 3 or tuple(rotation.shape) != (n, 4):
        raise RuntimeError(f"scaling must be [n,3] and rotation [n,4], got {tuple(scaling.shape)} and "
                           f"{tuple(rotation.shape)}")
    scaling, rotation = scaling.contiguous(), rotation.contiguous()
    with torch.cuda.device(scaling.device):
        cov = torch.empty(n, 6, dtype=torch.float32, device=scaling.device)
        rc = load().c3vq_build_covariance(_ptr(scaling), _ptr(rotation), n, float(scaling_modifier),
                                          1 if activate else 0, _ptr(cov), _stream(scaling))
    if rc:
        raise _err(rc)
    return cov


def compact(tensors, keep_mask: torch.Tensor):
    """[t[keep_mask] for t in tensors] (GaussianModel.mask_splats, scene/gaussian_model.py:1027-1046): one scan of
    the mask (c3vq_join_count) shared by all tensors, one row-move kernel per tensor (c3vq_compact_rows).  Tensors
    are float32 / int32 / int64 with the mask's length as first dimension; None entries pass through."""
    if keep_mask.dtype != torch.bool or keep_mask.dim() != 1 or not keep_mask.is_cuda:
        raise RuntimeError("keep_mask must be a 1-D CUDA bool tensor")
    keep_mask = keep_mask.contiguous()
    n = keep_mask.numel()
    L = load()
    dev = keep_mask.device
    with torch.cuda.device(dev):
        ws = torch.empty(max(int(L.c3vq_join_workspace_bytes(n)), 4), dtype=torch.uint8, device=dev)
        kept = _c_i64(0)
        rc = L.c3vq_join_count(_ptr(keep_mask), n, ctypes.byref(kept), _ptr(ws), ws.numel(), _stream(keep_mask))
        if rc:
            raise _err(rc)
        kept = int(kept.value)
        outs = []
        for t in tensors:
            if t is None:
                outs.append(None)
                continue
            if not t.is_cuda or t.dtype not in (torch.float32, torch.int32, torch.int64):
                raise RuntimeError("compact() takes float32 / int32 / int64 CUDA tensors")
            if t.dim() < 1 or t.size(0) != n:
                raise RuntimeError(f"tensor of shape {tuple(t.shape)} does not match the mask ({n} rows)")
            t = t.detach().contiguous()
            out = torch.empty((kept,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
            row_bytes = (t.numel() // n) * t.element_size() if n else 0
            if kept == 0 or row_bytes == 0:
                outs.append(out)
                continue
            rc = L.c3vq_compact_rows(_ptr(t), _ptr(keep_mask), n, row_bytes, _ptr(out), _ptr(ws), _stream(keep_mask))
            if rc:
                raise _err(rc)
            outs.append(out)
    return outs


def preprocess_indexed(means3d, scales, scale_factors, rotations, opacities, shs, sh_indices, g_indices, viewmatrix,
                       projmatrix, cam_pos, width: int, height: int, tan_fovx: float, tan_fovy: float,
                       sh_degree: int = 3, scale_modifier: float = 1.0, prefiltered: bool = False,
                       clamp_color: bool = True, out=None):
    """FORWARD::preprocess_indexed of the reference's indexed rasteriser (cuda_rasterizer/forward_indexed.cu:166-272):
    returns a dict with radii [P] int32, xy [P,2], depths [P], cov3d [P,6], rgb [P,3], conic_opacity [P,4],
    tiles_touched [P] int32 (bit pattern of the reference's uint32) and clamped [P,3] bool.  The kernel writes every
    row: culled Gaussians read as zeros (the reference leaves them untouched).  `out`: the dict of an earlier call
    with the same number of Gaussians, to write into its buffers again.  viewmatrix / projmatrix: 4x4 CUDA tensors whose flat storage is the
    column-major matrix, exactly what the reference's Python side passes.  c3vq_preprocess_indexed."""
    p = means3d.size(0)
    for t, name, shape in ((means3d, "means3d", (p, 3)), (scale_factors, "scale_factors", None),
                           (opacities, "opacities", None), (scales, "scales", (scales.size(0), 3)),
                           (rotations, "rotations", (rotations.size(0), 4)), (viewmatrix, "viewmatrix", None),
                           (projmatrix, "projmatrix", None), (cam_pos, "cam_pos", None), (shs, "shs", None)):
        _check(t, torch.float32, name)
        if shape is not None and tuple(t.shape) != shape:
            raise RuntimeError(f"{name} must be {shape}, got {tuple(t.shape)}")
    _check(sh_indices, torch.int64, "sh_indices", 1)
    _check(g_indices, torch.int64, "g_indices", 1)
    if scale_factors.numel() != p or opacities.numel() != p or sh_indices.numel() != p or g_indices.numel() != p:
        raise RuntimeError("scale_factors, opacities, sh_indices and g_indices need one entry per Gaussian")
    if viewmatrix.numel() != 16 or projmatrix.numel() != 16 or cam_pos.numel() != 3:
        raise RuntimeError("viewmatrix / projmatrix must hold 16 values and cam_pos 3")
    if shs.dim() != 3 or shs.size(2) != 3:
        raise RuntimeError(f"shs must be [F, coefficients, 3], got {tuple(shs.shape)}")
    means3d, scales, scale_factors, rotations = (means3d.contiguous(), scales.contiguous(), scale_factors.contiguous(),
                                                 _aligned(rotations))
    opacities, shs, sh_indices, g_indices = opacities.contiguous(), shs.contiguous(), sh_indices.contiguous(), g_indices.contiguous()
    viewmatrix, projmatrix, cam_pos = viewmatrix.contiguous(), projmatrix.contiguous(), cam_pos.contiguous()
    dev = means3d.device
    import numpy as np
    # rasterizer_impl.cu:468-469: focal lengths from the field of view, in float32
    focal_y = float(np.float32(height) / (np.float32(2.0) * np.float32(tan_fovy)))
    focal_x = float(np.float32(width) / (np.float32(2.0) * np.float32(tan_fovx)))
    with torch.cuda.device(dev):
        if out is None or out["radii"].numel() != p or out["radii"].device != dev:
            # one allocation carved into the eight outputs (like the reference's geometry-state chunk), segments
            # aligned to 256 bytes
            spec = (("conic_opacity", (p, 4), torch.float32), ("cov3d", (p, 6), torch.float32),
                    ("rgb", (p, 3), torch.float32), ("xy", (p, 2), torch.float32), ("depths", (p,), torch.float32),
                    ("radii", (p,), torch.int32), ("tiles_touched", (p,), torch.int32), ("clamped", (p, 3), torch.bool))
            sizes = [((int(torch.Size(shape).numel()) * torch.empty(0, dtype=dt).element_size() + 255) // 256) * 256
                     for _, shape, dt in spec]
            chunk = torch.empty(sum(sizes), dtype=torch.uint8, device=dev)
            out, off = {}, 0
            for (name, shape, dt), size in zip(spec, sizes):
                nbytes = int(torch.Size(shape).numel()) * torch.empty(0, dtype=dt).element_size()
                out[name] = chunk[off:off + nbytes].view(dt).reshape(shape)
                off += size
        rc = load().c3vq_preprocess_indexed(
            p, int(sh_degree), shs.size(1), _ptr(means3d), _ptr(scales), _ptr(scale_factors), float(scale_modifier),
            _ptr(rotations), _ptr(opacities), _ptr(shs), _ptr(viewmatrix), _ptr(projmatrix), _ptr(cam_pos),
            int(width), int(height), focal_x, focal_y, float(tan_fovx), float(tan_fovy), 1 if prefiltered else 0,
            _ptr(sh_indices), _ptr(g_indices), 1 if clamp_color else 0, _ptr(out["radii"]), _ptr(out["xy"]),
            _ptr(out["depths"]), _ptr(out["cov3d"]), _ptr(out["rgb"]), _ptr(out["conic_opacity"]),
            _ptr(out["tiles_touched"]), _ptr(out["clamped"]), _stream(means3d))
    if rc:
        raise _err(rc)
    return out


def preprocess_indexed_backward(means3d, radii, shs, clamped, scales, scale_factors, rotations, cov3d, viewmatrix,
                                projmatrix, cam_pos, width: int, height: int, tan_fovx: float, tan_fovy: float,
                                dl_dmean2d, dl_dconic, dl_dcolor, sh_indices, g_indices, sh_degree: int = 3,
                                scale_modifier: float = 1.0):
    """BACKWARD::preprocess_indexed of the reference's indexed rasteriser (cuda_rasterizer/backward_indexed.cu:
    469-550): gradients of the per-Gaussian preprocess.  `radii`, `clamped`, `cov3d` are the forward's outputs
    (preprocess_indexed); dl_dmean2d [P,3], dl_dconic [P,2,2] (or [P,4]) and dl_dcolor [P,3] come from the render
    backward.  Returns a dict: means3d [P,3], cov3d [P,6], scale_factors [P,1] per Gaussian, and the TABLE gradients
    shs [F,M,3], scales [G,3], rotations [G,4] summed over the Gaussians that reference each row (allocated as zeros
    here, like rasterize_points_indexed.cu:166-175).  c3vq_preprocess_indexed_backward."""
    p = means3d.size(0)
    for t, name in ((means3d, "means3d"), (shs, "shs"), (scales, "scales"), (scale_factors, "scale_factors"),
                    (rotations, "rotations"), (cov3d, "cov3d"), (viewmatrix, "viewmatrix"), (projmatrix, "projmatrix"),
                    (cam_pos, "cam_pos"), (dl_dmean2d, "dl_dmean2d"), (dl_dconic, "dl_dconic"), (dl_dcolor, "dl_dcolor")):
        _check(t, torch.float32, name)
    _check(radii, torch.int32, "radii", 1)
    _check(sh_indices, torch.int64, "sh_indices", 1)
    _check(g_indices, torch.int64, "g_indices", 1)
    if clamped.dtype not in (torch.bool, torch.uint8) or not clamped.is_cuda:
        raise RuntimeError("clamped must be a CUDA bool tensor")
    if shs.dim() != 3 or shs.size(2) != 3:
        raise RuntimeError(f"shs must be [F, coefficients, 3], got {tuple(shs.shape)}")
    if tuple(scales.shape[1:]) != (3,) or tuple(rotations.shape) != (scales.size(0), 4):
        raise RuntimeError("scales must be [G,3] and rotations [G,4]")
    for t, name, numel in ((radii, "radii", p), (clamped, "clamped", 3 * p), (scale_factors, "scale_factors", p),
                           (cov3d, "cov3d", 6 * p), (dl_dmean2d, "dl_dmean2d", 3 * p), (dl_dconic, "dl_dconic", 4 * p),
                           (dl_dcolor, "dl_dcolor", 3 * p), (sh_indices, "sh_indices", p), (g_indices, "g_indices", p),
                           (means3d, "means3d", 3 * p)):
        if t.numel() != numel:
            raise RuntimeError(f"{name} has {t.numel()} elements for {p} Gaussians, expected {numel}")
    if viewmatrix.numel() != 16 or projmatrix.numel() != 16 or cam_pos.numel() != 3:
        raise RuntimeError("viewmatrix / projmatrix must hold 16 values and cam_pos 3")
    means3d, radii, shs, clamped = means3d.contiguous(), radii.contiguous(), shs.contiguous(), clamped.contiguous()
    scales, scale_factors, rotations, cov3d = scales.contiguous(), scale_factors.contiguous(), _aligned(rotations), cov3d.contiguous()
    viewmatrix, projmatrix, cam_pos = viewmatrix.contiguous(), projmatrix.contiguous(), cam_pos.contiguous()
    dl_dmean2d, dl_dconic, dl_dcolor = dl_dmean2d.contiguous(), _aligned(dl_dconic), dl_dcolor.contiguous()
    sh_indices, g_indices = sh_indices.contiguous(), g_indices.contiguous()
    dev = means3d.device
    import numpy as np
    focal_y = float(np.float32(height) / (np.float32(2.0) * np.float32(tan_fovy)))       # rasterizer_impl.cu:468-469
    focal_x = float(np.float32(width) / (np.float32(2.0) * np.float32(tan_fovx)))
    with torch.cuda.device(dev):
        out = {"means3d": torch.empty(p, 3, dtype=torch.float32, device=dev),
               "cov3d": torch.empty(p, 6, dtype=torch.float32, device=dev),
               "scale_factors": torch.empty(p, 1, dtype=torch.float32, device=dev),
               "shs": torch.zeros(shs.shape, dtype=torch.float32, device=dev),
               "scales": torch.zeros(scales.shape, dtype=torch.float32, device=dev),
               "rotations": torch.zeros(rotations.shape, dtype=torch.float32, device=dev)}
        rc = load().c3vq_preprocess_indexed_backward(
            p, int(sh_degree), shs.size(1), _ptr(means3d), _ptr(radii), _ptr(shs), _ptr(clamped), _ptr(scales),
            _ptr(scale_factors), _ptr(rotations), float(scale_modifier), _ptr(cov3d), _ptr(viewmatrix),
            _ptr(projmatrix), focal_x, focal_y, float(tan_fovx), float(tan_fovy), _ptr(cam_pos), _ptr(dl_dmean2d),
            _ptr(dl_dconic), _ptr(dl_dcolor), _ptr(out["means3d"]), _ptr(out["cov3d"]), _ptr(out["shs"]),
            _ptr(out["scales"]), _ptr(out["scale_factors"]), _ptr(out["rotations"]), _ptr(sh_indices),
            _ptr(g_indices), _stream(means3d))
    if rc:
        raise _err(rc)
    return out


def scatter_add(grad_out: torch.Tensor, idx: torch.Tensor, table_rows: int) -> torch.Tensor:
    """grad_table[idx[i]] += grad_out[i]  (backward of gather).  c3vq_scatter_add."""
    _check(grad_out, torch.float32, "grad_out")
    _check(idx, torch.int64, "idx", 1)
    grad_out, idx = grad_out.contiguous(), idx.contiguous()
    n = idx.numel()
    d = int(torch.Size(grad_out.shape[1:]).numel())
    with torch.cuda.device(grad_out.device):
        gt = torch.zeros((table_rows,) + tuple(grad_out.shape[1:]), dtype=torch.float32, device=grad_out.device)
        rc = load().c3vq_scatter_add(_ptr(grad_out), _ptr(idx), n, d, _ptr(gt), table_rows, _stream(grad_out))
    if rc:
        raise _err(rc)
    return gt


def gather_rows(x: torch.Tensor, w: torch.Tensor | None, rows: torch.Tensor):
    """(x[rows], w[rows]) — the mini-batch gather of vq.py:70-71.  c3vq_gather_rows."""
    _check(x, torch.float32, "x", 2)
    _check(rows, torch.int64, "rows", 1)
    x, rows = x.contiguous(), rows.contiguous()
    n, d = rows.numel(), x.size(1)
    with torch.cuda.device(x.device):
        ox = torch.empty(n, d, dtype=torch.float32, device=x.device)
        ow = None
        if w is not None:
            _check(w, torch.float32, "importance", 1)
            w = w.contiguous()
            ow = torch.empty(n, dtype=torch.float32, device=x.device)
        rc = load().c3vq_gather_rows(_ptr(x), _ptr(w), _ptr(rows), n, d, _ptr(ox), _ptr(ow), _stream(x))
    if rc:
        raise _err(rc)
    return ox, ow


def debug_scores(x: torch.Tensor, codebook: torch.Tensor):
    """Screening scores of the tensor-core path -> (scores [n,k], row_norm2 [n], row_band [n])."""
    _check(x, torch.float32, "x", 2)
    _check(codebook, torch.float32, "codebook", 2)
    x, codebook = x.contiguous(), codebook.contiguous()
    n, d, k = x.size(0), x.size(1), codebook.size(0)
    L = load()
    with torch.cuda.device(x.device):
        scores = torch.empty(n, k, dtype=torch.float32, device=x.device)
        norm2 = torch.empty(n, dtype=torch.float32, device=x.device)
        band = torch.empty(n, dtype=torch.float32, device=x.device)
        nbytes = L.c3vq_assign_workspace_bytes(n, d, k)
        ws = _workspace.get(x.device, nbytes)
        rc = L.c3vq_debug_scores(_ptr(x), n, d, _ptr(codebook), k, _ptr(scores), _ptr(norm2), _ptr(band),
                                 _ptr(ws), ws.numel(), _stream(x))
    if rc:
        raise _err(rc)
    return scores, norm2, band


def assign_stats(device=None):
    """(rows, candidates, fallback_rows) of the last tensor-core assignment on `device`."""
    device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    ws = _workspace.get(device, 256)
    out = (_c_i64 * 3)()
    with torch.cuda.device(device):
        rc = load().c3vq_assign_stats(_ptr(ws), out, torch.cuda.current_stream(device).cuda_stream)
    if rc:
        raise _err(rc)
    return int(out[0]), int(out[1]), int(out[2])
