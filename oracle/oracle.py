"""ctypes front-end of the CPU oracle (oracle/c3vq_oracle.c).

TEST INFRASTRUCTURE ONLY — see the header of c3vq_oracle.c.  Imported by tests/,
__graft_entry__.smoke() and bench.py's CPU-baseline legs; never by c3dgs_b200/.

Works on numpy arrays (float32 / int64, C-contiguous).  Each function names the
reference lines its C counterpart restates.
"""
from __future__ import annotations

import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libc3vq_oracle.so")
_lib = None


def build(force: bool = False) -> str:
    """Compile the oracle with the recipe in oracle/Makefile."""
    src = os.path.join(_HERE, "c3vq_oracle.c")
    if force or not os.path.exists(_LIB_PATH) or os.path.getmtime(_LIB_PATH) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "libc3vq_oracle.so"], stdout=subprocess.DEVNULL)
    return _LIB_PATH


def lib() -> ctypes.CDLL:
    global _lib
    if _lib is None:
        build()
        L = ctypes.CDLL(_LIB_PATH)
        f32p = ctypes.POINTER(ctypes.c_float)
        f64p = ctypes.POINTER(ctypes.c_double)
        i64p = ctypes.POINTER(ctypes.c_int64)
        i64, i32, f64 = ctypes.c_int64, ctypes.c_int, ctypes.c_double
        L.c3vq_oracle_num_threads.restype = i32
        L.c3vq_oracle_set_threads.argtypes = [i32]
        L.c3vq_oracle_distance.restype = ctypes.c_float
        L.c3vq_oracle_distance.argtypes = [i32, f32p, f32p]
        L.c3vq_oracle_assign.argtypes = [f32p, i64, i32, f32p, i32, f32p, i64p]
        L.c3vq_oracle_scatter_sum.argtypes = [f32p, f32p, i64p, i64, i32, i32, f32p, f32p]
        L.c3vq_oracle_scatter_sum_f64.argtypes = [f32p, f32p, i64p, i64, i32, i32, f64p, f64p]
        L.c3vq_oracle_ema.argtypes = [f32p, f32p, f32p, f32p, i32, i32, f64, f64]
        L.c3vq_oracle_trace_normalize.argtypes = [f32p, i32, i32]
        L.c3vq_oracle_update.argtypes = [f32p, f32p, i64, i32, f32p, f32p, i32, f64, f64, f32p, i64p, f32p]
        L.c3vq_oracle_gather.argtypes = [f32p, i64p, i64, i32, f32p]
        _lib = L
    return _lib


def _f32(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.float32)


def _i64(a) -> np.ndarray:
    return np.ascontiguousarray(a, dtype=np.int64)


def _p(a: np.ndarray, ty):
    return a.ctypes.data_as(ctypes.POINTER(ty))


def num_threads() -> int:
    return int(lib().c3vq_oracle_num_threads())


def set_threads(n: int) -> None:
    lib().c3vq_oracle_set_threads(int(n))


def distance(coefs, codeword) -> float:
    """weighted_distance.cu:9-18."""
    a, b = _f32(coefs), _f32(codeword)
    return float(lib().c3vq_oracle_distance(a.shape[0], _p(a, ctypes.c_float), _p(b, ctypes.c_float)))


def assign(x, codebook):
    """weightedDistance(x, codebook) -> (min_dist f32[N], idx i64[N]).  weighted_distance.cu:20-93."""
    x, cb = _f32(x), _f32(codebook)
    if x.ndim != 2 or cb.ndim != 2:
        raise RuntimeError("ceofs and codebook must have dimension 2")  # weighted_distance.cu:64-67
    if x.shape[1] != cb.shape[1]:
        raise RuntimeError("coefs and codebook must have same number of channels")  # :69-72
    n, d = x.shape
    k = cb.shape[0]
    dist = np.zeros(n, np.float32)
    idx = np.zeros(n, np.int64)
    lib().c3vq_oracle_assign(_p(x, ctypes.c_float), n, d, _p(cb, ctypes.c_float), k,
                             _p(dist, ctypes.c_float), _p(idx, ctypes.c_int64))
    return dist, idx


def scatter_sum(x, w, idx, k):
    """vq.py:31,33 -> (sum_x f32[K,D], sum_w f32[K]); fp32, row order."""
    x, idx = _f32(x), _i64(idx)
    n, d = x.shape
    wp = None
    if w is not None:
        w = _f32(w)
        wp = _p(w, ctypes.c_float)
    sx = np.zeros((k, d), np.float32)
    sw = np.zeros(k, np.float32)
    lib().c3vq_oracle_scatter_sum(_p(x, ctypes.c_float), wp, _p(idx, ctypes.c_int64), n, d, k,
                                  _p(sx, ctypes.c_float), _p(sw, ctypes.c_float))
    return sx, sw


def scatter_sum_f64(x, w, idx, k):
    x, idx = _f32(x), _i64(idx)
    n, d = x.shape
    wp = None
    if w is not None:
        w = _f32(w)
        wp = _p(w, ctypes.c_float)
    sx = np.zeros((k, d), np.float64)
    sw = np.zeros(k, np.float64)
    lib().c3vq_oracle_scatter_sum_f64(_p(x, ctypes.c_float), wp, _p(idx, ctypes.c_int64), n, d, k,
                                      _p(sx, ctypes.c_double), _p(sw, ctypes.c_double))
    return sx, sw


def ema(codebook, entry_importance, sum_x, sum_w, decay, eps=1e-5):
    """vq.py:32,34,45-46 — in place on float32 arrays."""
    assert codebook.dtype == np.float32 and entry_importance.dtype == np.float32
    k, d = codebook.shape
    sx, sw = _f32(sum_x), _f32(sum_w)
    lib().c3vq_oracle_ema(_p(codebook, ctypes.c_float), _p(entry_importance, ctypes.c_float),
                          _p(sx, ctypes.c_float), _p(sw, ctypes.c_float), k, d, float(decay), float(eps))


def trace_normalize(codebook):
    """vq.py:76-77 — in place."""
    assert codebook.dtype == np.float32 and codebook.flags.c_contiguous
    k, d = codebook.shape
    lib().c3vq_oracle_trace_normalize(_p(codebook, ctypes.c_float), k, d)


def update(x, w, codebook, entry_importance, decay, eps=1e-5):
    """VectorQuantize.update (vq.py:28-35): mutates codebook/entry_importance, returns (min_dists, idx)."""
    x, w = _f32(x), _f32(w)
    assert codebook.dtype == np.float32 and codebook.flags.c_contiguous
    assert entry_importance.dtype == np.float32
    n, d = x.shape
    k = codebook.shape[0]
    dist = np.zeros(n, np.float32)
    idx = np.zeros(n, np.int64)
    scratch = np.zeros(k * d + k, np.float32)
    lib().c3vq_oracle_update(_p(x, ctypes.c_float), _p(w, ctypes.c_float), n, d,
                             _p(codebook, ctypes.c_float), _p(entry_importance, ctypes.c_float), k,
                             float(decay), float(eps),
                             _p(dist, ctypes.c_float), _p(idx, ctypes.c_int64), _p(scratch, ctypes.c_float))
    return dist, idx


def gather(table, idx):
    """codebook[idx] decode (vq.py:40-42, gaussian_model.py:220/240/257)."""
    table, idx = _f32(table), _i64(idx)
    d = int(np.prod(table.shape[1:])) if table.ndim > 1 else 1
    out = np.zeros((idx.shape[0],) + table.shape[1:], np.float32)
    lib().c3vq_oracle_gather(_p(table, ctypes.c_float), _p(idx, ctypes.c_int64), idx.shape[0], d,
                             _p(out, ctypes.c_float))
    return out


def vq_features(features, importance, codebook0, batches, decay=0.8, scale_normalize=False, eps=1e-5):
    """The vq_features loop (vq.py:49-87) with the random draws made explicit.

    `codebook0` replaces uniform_init's device RNG (vq.py:24-26); `batches` is the list of
    index arrays `torch.randint` would have produced (vq.py:69).  Returns
    (codebook, entry_importance, final_idx, per-step (min_dists, idx))."""
    features = _f32(features)
    imp = _f32(importance)
    imp_n = imp / imp.max()                                   # vq.py:59
    cb = _f32(codebook0).copy()
    ei = np.zeros(cb.shape[0], np.float32)                    # vq.py:21
    trace = []
    for b in batches:                                         # vq.py:68
        b = _i64(b)
        dist, idx = update(features[b], imp_n[b], cb, ei, decay, eps)   # vq.py:70-71
        trace.append((dist, idx))
        if scale_normalize:                                   # vq.py:73-77
            trace_normalize(cb)
    _, final_idx = assign(features, cb)                       # vq.py:82
    return cb, ei, final_idx, trace


def join_features(all_features, keep_mask, codebook, codebook_indices):
    """compression/vq.py:90-103 restated in numpy (boolean-mask compaction is stable, like torch's):
    returns (cat(codebook, all_features[keep_mask]), int64 indices)."""
    all_features = np.asarray(all_features, np.float32)
    keep_mask = np.asarray(keep_mask, bool)
    codebook = np.asarray(codebook, np.float32).reshape(-1, all_features.shape[1])
    keep_features = all_features[keep_mask]                                        # vq.py:96
    compressed = np.concatenate([codebook, keep_features], 0)                      # vq.py:97
    indices = np.zeros(len(all_features), np.int64)                                # vq.py:99
    indices[~keep_mask] = np.asarray(codebook_indices, np.int64)                   # vq.py:100
    indices[keep_mask] = np.arange(len(keep_features), dtype=np.int64) + len(codebook)   # vq.py:101
    return compressed, indices


# ---- §8(f) rank 2: covariance codebook -> (rotation quaternion, scaling), utils/splats.py ----------------------------
def to_full_cov(cov6):
    """utils/splats.py:7-25: the 6 stored entries (xx, xy, xz, yy, yz, zz) to a symmetric 3x3, float32."""
    c = np.asarray(cov6, np.float32).reshape(-1, 6)
    full = np.zeros((len(c), 3, 3), np.float32)
    for (i, j), k in (((0, 0), 0), ((0, 1), 1), ((0, 2), 2), ((1, 1), 3), ((1, 2), 4), ((2, 2), 5)):
        full[:, i, j] = c[:, k]
        full[:, j, i] = c[:, k]
    return full


def _sqrt_positive_part(x):
    """utils/splats.py:103-110."""
    return np.sqrt(np.where(x > 0, x, 0.0))


def matrix_to_quaternion(m):
    """utils/splats.py:45-100 in float64: the four candidate quaternions (each scaled by one of r, i, j, k), pick the
    one with the largest denominator (first maximum, like argmax)."""
    m = np.asarray(m, np.float64)
    m00, m01, m02, m10, m11, m12, m20, m21, m22 = [m[:, i, j] for i in range(3) for j in range(3)]
    q_abs = _sqrt_positive_part(np.stack([1.0 + m00 + m11 + m22, 1.0 + m00 - m11 - m22,
                                          1.0 - m00 + m11 - m22, 1.0 - m00 - m11 + m22], -1))     # :59-69
    cand = np.stack([
        np.stack([q_abs[:, 0] ** 2, m21 - m12, m02 - m20, m10 - m01], -1),
        np.stack([m21 - m12, q_abs[:, 1] ** 2, m10 + m01, m02 + m20], -1),
        np.stack([m02 - m20, m10 + m01, q_abs[:, 2] ** 2, m12 + m21], -1),
        np.stack([m10 - m01, m20 + m02, m21 + m12, q_abs[:, 3] ** 2], -1)], -2)                   # :72-89
    cand = cand / (2.0 * np.maximum(q_abs[:, :, None], 0.1))                                      # :93-94
    return cand[np.arange(len(m)), np.argmax(q_abs, -1)]                                          # :99-101


def quaternion_to_matrix(q):
    """utils/splats.py:113-144 in float64 (real part first)."""
    q = np.asarray(q, np.float64)
    r, i, j, k = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    two_s = 2.0 / (q * q).sum(-1)
    o = np.stack([1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                  two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                  two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)], -1)
    return o.reshape(-1, 3, 3)


def build_covariance(rotation, scaling):
    """utils/splats.py:39-42: R diag(s^2) R^T, float64."""
    R = quaternion_to_matrix(rotation)
    s2 = np.asarray(scaling, np.float64) ** 2
    return np.einsum("nij,nj,nkj->nik", R, s2, R)


def extract_rot_scale(cov):
    """utils/splats.py:28-36.  The 1e-8 is added to the diagonal in float32 as the reference does; the
    eigendecomposition itself is LAPACK in float64 here (the reference's is float32 LAPACK / cuSOLVER), upper
    triangle only, eigenvalues ascending.  Eigenvector signs are whatever LAPACK returns: the outputs are defined
    up to the sign choices that leave R diag(s^2) R^T unchanged, and the tests compare through that product.
    Returns (rot [n,4] f32 unit quaternions, real part first; scaling [n,3] f32 ascending)."""
    a32 = np.asarray(cov, np.float32).reshape(-1, 3, 3) + (np.eye(3, dtype=np.float32) * np.float32(1e-8))
    S, R = np.linalg.eigh(a32.astype(np.float64), UPLO="U")                                       # :30
    with np.errstate(invalid="ignore"):
        scaling = np.sqrt(S.astype(np.float32))
    scaling = np.where(np.isnan(scaling), np.float32(1e-6), scaling).astype(np.float32)           # :31
    q = matrix_to_quaternion(R * np.linalg.det(R)[:, None, None])                                 # :35
    q = q / np.maximum(np.linalg.norm(q, axis=-1, keepdims=True), 1e-12)                          # F.normalize
    return q.astype(np.float32), scaling


# ---- §8(f) rank 3: Morton order of the Gaussians, scene/gaussian_model.py:997-1004, :1415-1432 -----------------------
def split_by_3(a):
    """scene/gaussian_model.py:1415-1423 on uint64 arrays."""
    x = np.asarray(a).astype(np.uint64) & np.uint64(0x1FFFFF)
    for sh, mask in ((32, 0x1F00000000FFFF), (16, 0x1F0000FF0000FF), (8, 0x100F00F00F00F00F),
                     (4, 0x10C30C30C30C30C3), (2, 0x1249249249249249)):
        x = (x | (x << np.uint64(sh))) & np.uint64(mask)
    return x


def morton_keys(xyz):
    """scene/gaussian_model.py:998-1003 + mortonEncode (:1426-1432): float32 quantisation to 21 bits per axis,
    interleave with the axis of smallest extent in bit 0 (stable argsort of the extents).  Returns int64 keys."""
    xyz = np.asarray(xyz, np.float32)
    pp_min = xyz.min(0)
    pp_diap = xyz.max(0) - pp_min
    with np.errstate(invalid="ignore", divide="ignore"):
        f = (np.float32(2 ** 21 - 1) * (xyz - pp_min)) / pp_diap
    q = np.where(np.isnan(f), np.float32(0), f).astype(np.int64)       # NaN.long() & 0x1FFFFF is 0 on CPU and GPU alike
    ordering = np.argsort(pp_diap, kind="stable")
    key = split_by_3(q[:, ordering[0]]) | (split_by_3(q[:, ordering[1]]) << np.uint64(1)) | \
        (split_by_3(q[:, ordering[2]]) << np.uint64(2))
    return key.astype(np.int64)


def morton_order(xyz):
    """`mortonEncode(...).sort().indices` (:1003) with ties in row order (stable)."""
    return np.argsort(morton_keys(xyz), kind="stable").astype(np.int64)


# ---- §8(f) ranks 1 and 3: the scene-file wire format and the glue around the path ------------------------------------
def quantize_i8(x, scale, zero_point):
    """torch.quantize_per_tensor(x, scale, zero_point, torch.qint8).int_repr() as save_npz calls it on the model's
    device (scene/gaussian_model.py:528-620).  torch's CUDA quantiser evaluates
    clamp(int64(nearbyint(double(x) / scale) + zero_point), -128, 127) with round-half-to-even; its CPU quantiser
    multiplies by float32(1/scale) instead, so the two differ on a few elements per million that sit on a rounding
    boundary (tests/test_oracle_golden.py bounds exactly that).  NaN quantises like 0 (the integer conversion)."""
    xd = np.asarray(x, np.float32).astype(np.float64)
    with np.errstate(invalid="ignore", over="ignore"):
        t = np.rint(xd / float(scale)) + float(int(zero_point))
    t = np.where(np.isnan(t), 0.0, t)
    return np.clip(t, -128, 127).astype(np.int8)


def dequantize_i8(q, scale, zero_point):
    """load_npz (scene/gaussian_model.py:644-716): (q.int() - zero_point) * scale, int32 difference times a float32."""
    d = np.asarray(q, np.int8).astype(np.int32) - np.int32(zero_point)
    return d.astype(np.float32) * np.float32(scale)


def build_rotation(r):
    """utils/general_utils.py:84-107: quaternion (real part first, normalised here) -> 3x3, float32."""
    r = np.asarray(r, np.float32)
    norm = np.sqrt(r[:, 0] * r[:, 0] + r[:, 1] * r[:, 1] + r[:, 2] * r[:, 2] + r[:, 3] * r[:, 3])
    q = r / norm[:, None]
    w, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
    one, two = np.float32(1), np.float32(2)
    R = np.zeros((len(q), 3, 3), np.float32)
    R[:, 0, 0] = one - two * (y * y + z * z)
    R[:, 0, 1] = two * (x * y - w * z)
    R[:, 0, 2] = two * (x * z + w * y)
    R[:, 1, 0] = two * (x * y + w * z)
    R[:, 1, 1] = one - two * (x * x + z * z)
    R[:, 1, 2] = two * (y * z - w * x)
    R[:, 2, 0] = two * (x * z - w * y)
    R[:, 2, 1] = two * (y * z + w * x)
    R[:, 2, 2] = one - two * (x * x + y * y)
    return R


def normalized_covariance(scaling, rotation, scaling_modifier=1.0, activate=False):
    """GaussianModel.get_normalized_covariance(strip_sym=True) (scene/gaussian_model.py:274-278, :55-63;
    utils/general_utils.py:68-119): L = R diag(modifier * s), upper triangle of L L^T as [n,6], float32.
    activate: scaling <- normalize(relu(scaling)), rotation <- normalize(rotation) first (:66, :77)."""
    s = np.asarray(scaling, np.float32)
    r = np.asarray(rotation, np.float32)
    if activate:
        s = np.maximum(s, np.float32(0))
        s = s / np.maximum(np.sqrt((s * s).sum(1, dtype=np.float32)), np.float32(1e-12))[:, None]
        r = r / np.maximum(np.sqrt((r * r).sum(1, dtype=np.float32)), np.float32(1e-12))[:, None]
    L = build_rotation(r) * (np.float32(scaling_modifier) * s)[:, None, :]
    cov = np.einsum("nij,nkj->nik", L, L).astype(np.float32)
    return np.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], 1)


def mask_rows(t, mask):
    """t[mask] of GaussianModel.mask_splats (scene/gaussian_model.py:1027-1046): stable boolean-mask compaction."""
    return np.asarray(t)[np.asarray(mask, bool)]


# ---- §8(f) rank 4: per-Gaussian preprocess of the indexed rasteriser ---------------------------------------------------
_SH_C0 = np.float32(0.28209479177387814)
_SH_C1 = np.float32(0.4886025119029199)
_SH_C2 = np.array([1.0925484305920792, -1.0925484305920792, 0.31539156525252005, -1.0925484305920792,
                   0.5462742152960396], np.float32)
_SH_C3 = np.array([-0.5900435899266435, 2.890611442640554, -0.4570457994644658, 0.3731763325901154,
                   -0.4570457994644658, 1.445305721320277, -0.5900435899266435], np.float32)


def preprocess_indexed(means3d, scales, scale_factors, rotations, opacities, shs, sh_indices, g_indices, viewmatrix,
                       projmatrix, cam_pos, width, height, tan_fovx, tan_fovy, sh_degree=3, scale_modifier=1.0,
                       clamp_color=True, dtype=np.float32):
    """preprocess_indexedCUDA restated in numpy float32 (submodules/diff-gaussian-rasterization/cuda_rasterizer/
    forward_indexed.cu:166-272; computeColorFromSH_indexed :20-77, computeCov2D_indexed :79-121, computeCov3D_indexed
    :123-159; in_frustum / ndc2Pix / getRect / transformPoint auxiliary.h:42-152; focal lengths rasterizer_impl.cu:
    468-469).  viewmatrix / projmatrix: 16 floats, column-major.  Returns the kernel's outputs (zeros where the
    kernel returns early).  dtype=np.float64 evaluates the same expressions in double (used to check the backward
    restatement below by finite differences)."""
    f = dtype
    P = len(means3d)
    m = np.asarray(means3d, f)
    v = np.asarray(viewmatrix, f).reshape(16)
    pj = np.asarray(projmatrix, f).reshape(16)
    cam = np.asarray(cam_pos, f).reshape(3)
    focal_y = f(height) / (f(2.0) * f(tan_fovy))
    focal_x = f(width) / (f(2.0) * f(tan_fovx))
    px, py, pz = m[:, 0], m[:, 1], m[:, 2]
    out = {"radii": np.zeros(P, np.int32), "xy": np.zeros((P, 2), f), "depths": np.zeros(P, f),
           "cov3d": np.zeros((P, 6), f), "rgb": np.zeros((P, 3), f), "conic_opacity": np.zeros((P, 4), f),
           "tiles_touched": np.zeros(P, np.uint32), "clamped": np.zeros((P, 3), bool)}
    with np.errstate(all="ignore"):
        vx = v[0] * px + v[4] * py + v[8] * pz + v[12]
        vy = v[1] * px + v[5] * py + v[9] * pz + v[13]
        vz = v[2] * px + v[6] * py + v[10] * pz + v[14]
        alive = ~(vz <= f(0.01))                                                    # auxiliary.h:139-152
        hx = pj[0] * px + pj[4] * py + pj[8] * pz + pj[12]
        hy = pj[1] * px + pj[5] * py + pj[9] * pz + pj[13]
        hw = pj[3] * px + pj[7] * py + pj[11] * pz + pj[15]
        p_w = f(1.0) / (hw + f(0.0000001))
        projx, projy = hx * p_w, hy * p_w
        # computeCov3D_indexed: quaternion used as stored, scale = factor * modifier * scaling row
        g = np.asarray(g_indices, np.int64)
        mod = np.asarray(scale_factors, f).reshape(P) * f(scale_modifier)
        sc = np.asarray(scales, f)[g] * mod[:, None]
        q = np.asarray(rotations, f)[g]
        r, x, y, z = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        one, two = f(1), f(2)
        R = np.empty((P, 3, 3), f)
        R[:, 0, 0] = one - two * (y * y + z * z); R[:, 0, 1] = two * (x * y - r * z); R[:, 0, 2] = two * (x * z + r * y)
        R[:, 1, 0] = two * (x * y + r * z); R[:, 1, 1] = one - two * (x * x + z * z); R[:, 1, 2] = two * (y * z - r * x)
        R[:, 2, 0] = two * (x * z - r * y); R[:, 2, 1] = two * (y * z + r * x); R[:, 2, 2] = one - two * (x * x + y * y)
        Mm = sc[:, :, None] * np.transpose(R, (0, 2, 1))                             # M[k][j] = s_k R[j][k]
        Sig = np.zeros((P, 3, 3), f)
        for k in range(3):
            Sig = Sig + Mm[:, k, :, None] * Mm[:, k, None, :]
        cov3 = np.stack([Sig[:, 0, 0], Sig[:, 0, 1], Sig[:, 0, 2], Sig[:, 1, 1], Sig[:, 1, 2], Sig[:, 2, 2]], 1)
        # computeCov2D_indexed
        limx, limy = f(1.3) * f(tan_fovx), f(1.3) * f(tan_fovy)
        tz = vz
        tx = np.minimum(limx, np.maximum(-limx, vx / tz)) * tz
        ty = np.minimum(limy, np.maximum(-limy, vy / tz)) * tz
        j00, j02 = focal_x / tz, -(focal_x * tx) / (tz * tz)
        j11, j12 = focal_y / tz, -(focal_y * ty) / (tz * tz)
        a0 = np.stack([v[0] * j00 + v[2] * j02, v[4] * j00 + v[6] * j02, v[8] * j00 + v[10] * j02], 1)
        a1 = np.stack([v[1] * j11 + v[2] * j12, v[5] * j11 + v[6] * j12, v[9] * j11 + v[10] * j12], 1)
        b0 = np.einsum("nij,nj->ni", Sig, a0).astype(f)
        b1 = np.einsum("nij,nj->ni", Sig, a1).astype(f)
        covx = (a0 * b0).sum(1, dtype=f) + f(0.3)
        covy = (a0 * b1).sum(1, dtype=f)
        covz = (a1 * b1).sum(1, dtype=f) + f(0.3)
        det = covx * covz - covy * covy
        alive2 = alive & ~(det == 0)
        det_inv = one / det
        mid = f(0.5) * (covx + covz)
        disc = np.sqrt(np.maximum(f(0.1), mid * mid - det))
        radius = np.ceil(f(3.0) * np.sqrt(np.maximum(mid + disc, mid - disc)))
        pix_x = (((projx.astype(np.float64) + 1.0) * width - 1.0) * 0.5).astype(f)   # ndc2Pix, in double
        pix_y = (((projy.astype(np.float64) + 1.0) * height - 1.0) * 0.5).astype(f)
        rad = np.nan_to_num(radius, nan=0.0, posinf=2.0 ** 31 - 1, neginf=0.0).astype(np.int64)
        gx, gy = (width + 15) // 16, (height + 15) // 16
        radf = rad.astype(f)

        def tile(val):                                                             # (int) truncation, clamp at 0
            t = np.nan_to_num(val / f(16), nan=0.0, posinf=2.0 ** 31 - 1, neginf=-2.0 ** 31)
            return np.maximum(0, np.trunc(t).astype(np.int64))

        rminx, rminy = np.minimum(gx, tile(pix_x - radf)), np.minimum(gy, tile(pix_y - radf))
        rmaxx, rmaxy = np.minimum(gx, tile(pix_x + radf + f(15))), np.minimum(gy, tile(pix_y + radf + f(15)))
        tiles = (rmaxx - rminx) * (rmaxy - rminy)
        vis = alive2 & (tiles != 0)
        # computeColorFromSH_indexed
        d = m - cam[None, :]
        d = d / np.sqrt((d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]))[:, None]
        dx, dy, dz = d[:, 0], d[:, 1], d[:, 2]
        sh = np.asarray(shs, f)[np.asarray(sh_indices, np.int64)]                    # [P, M, 3]
        res = _SH_C0 * sh[:, 0]
        if sh_degree > 0:
            res = res - (_SH_C1 * dy)[:, None] * sh[:, 1] + (_SH_C1 * dz)[:, None] * sh[:, 2] - (_SH_C1 * dx)[:, None] * sh[:, 3]
            if sh_degree > 1:
                xx, yy, zz, xy, yz, xz = dx * dx, dy * dy, dz * dz, dx * dy, dy * dz, dx * dz
                res = (res + (_SH_C2[0] * xy)[:, None] * sh[:, 4] + (_SH_C2[1] * yz)[:, None] * sh[:, 5]
                       + (_SH_C2[2] * (two * zz - xx - yy))[:, None] * sh[:, 6] + (_SH_C2[3] * xz)[:, None] * sh[:, 7]
                       + (_SH_C2[4] * (xx - yy))[:, None] * sh[:, 8])
                if sh_degree > 2:
                    res = (res + (_SH_C3[0] * dy * (f(3) * xx - yy))[:, None] * sh[:, 9]
                           + (_SH_C3[1] * xy * dz)[:, None] * sh[:, 10]
                           + (_SH_C3[2] * dy * (f(4) * zz - xx - yy))[:, None] * sh[:, 11]
                           + (_SH_C3[3] * dz * (two * zz - f(3) * xx - f(3) * yy))[:, None] * sh[:, 12]
                           + (_SH_C3[4] * dx * (f(4) * zz - xx - yy))[:, None] * sh[:, 13]
                           + (_SH_C3[5] * dz * (xx - yy))[:, None] * sh[:, 14]
                           + (_SH_C3[6] * dx * (xx - f(3) * yy))[:, None] * sh[:, 15])
        res = (res + f(0.5)).astype(f)
        clamped = (res < 0) if clamp_color else np.zeros_like(res, bool)
        if clamp_color:
            res = np.maximum(res, f(0))
    out["cov3d"][alive] = cov3[alive]
    out["radii"][vis] = rad[vis].astype(np.int32)
    out["tiles_touched"][vis] = tiles[vis].astype(np.uint32)
    out["xy"][vis] = np.stack([pix_x, pix_y], 1)[vis]
    out["depths"][vis] = vz[vis]
    out["rgb"][vis] = res[vis]
    out["clamped"][vis] = clamped[vis]
    out["conic_opacity"][vis] = np.stack([covz * det_inv, -covy * det_inv, covx * det_inv,
                                          np.asarray(opacities, f).reshape(P)], 1)[vis]
    return out


def preprocess_indexed_backward(means3d, radii, shs, clamped, scales, scale_factors, rotations, cov3d, viewmatrix,
                                projmatrix, cam_pos, width, height, tan_fovx, tan_fovy, dl_dmean2d, dl_dconic,
                                dl_dcolor, sh_indices, g_indices, sh_degree=3, scale_modifier=1.0, dtype=np.float32):
    """BACKWARD::preprocess_indexed restated in numpy (submodules/diff-gaussian-rasterization/cuda_rasterizer/
    backward_indexed.cu:469-550): computeCov2D_indexedCUDA :346-465, preprocess_indexedCUDA :288-341,
    computeColorFromSH_indexed :20-205, computeCov3D_indexed :210-283; dnormvdv / transformVec4x3Transpose
    auxiliary.h:83-110).  glm matrices are column-major: T[c][r] below is the reference's T[c][r].
    Returns dict(means3d [P,3], cov3d [P,6], scale_factors [P], shs [F,M,3], scales [G,3], rotations [G,4]);
    the table gradients are summed in float64 and rounded once (the reference's atomics are order dependent)."""
    f = dtype
    P = len(means3d)
    m = np.asarray(means3d, f)
    v = np.asarray(viewmatrix, f).reshape(16)
    pj = np.asarray(projmatrix, f).reshape(16)
    cam = np.asarray(cam_pos, f).reshape(3)
    h_y = f(height) / (f(2.0) * f(tan_fovy))
    h_x = f(width) / (f(2.0) * f(tan_fovx))
    vis = np.asarray(radii).reshape(P) > 0
    shs = np.asarray(shs, f)
    F_, M = shs.shape[0], shs.shape[1]
    G = len(scales)
    two, one, half = f(2), f(1), f(0.5)
    mx, my, mz = m[:, 0], m[:, 1], m[:, 2]
    out = {"means3d": np.zeros((P, 3), f), "cov3d": np.zeros((P, 6), f), "scale_factors": np.zeros(P, f)}
    with np.errstate(all="ignore"):
        # ---- computeCov2D_indexedCUDA ----
        c3 = np.asarray(cov3d, f).reshape(P, 6)
        c0, c1, c2, c3_, c4, c5 = (c3[:, i] for i in range(6))
        dcon = np.asarray(dl_dconic, f).reshape(P, 4)
        dcx, dcy, dcz = dcon[:, 0], dcon[:, 1], dcon[:, 3]
        tx = v[0] * mx + v[4] * my + v[8] * mz + v[12]
        ty = v[1] * mx + v[5] * my + v[9] * mz + v[13]
        tz = v[2] * mx + v[6] * my + v[10] * mz + v[14]
        limx, limy = f(1.3) * f(tan_fovx), f(1.3) * f(tan_fovy)
        txtz, tytz = tx / tz, ty / tz
        tx = np.minimum(limx, np.maximum(-limx, txtz)) * tz
        ty = np.minimum(limy, np.maximum(-limy, tytz)) * tz
        xg = np.where((txtz < -limx) | (txtz > limx), f(0), f(1))
        yg = np.where((tytz < -limy) | (tytz > limy), f(0), f(1))
        j00, j02 = h_x / tz, -(h_x * tx) / (tz * tz)
        j11, j12 = h_y / tz, -(h_y * ty) / (tz * tz)
        T00, T01, T02 = v[0] * j00 + v[2] * j02, v[4] * j00 + v[6] * j02, v[8] * j00 + v[10] * j02
        T10, T11, T12 = v[1] * j11 + v[2] * j12, v[5] * j11 + v[6] * j12, v[9] * j11 + v[10] * j12
        u0, u1, u2 = T00 * c0 + T01 * c1 + T02 * c2, T00 * c1 + T01 * c3_ + T02 * c4, T00 * c2 + T01 * c4 + T02 * c5
        w0, w1, w2 = T10 * c0 + T11 * c1 + T12 * c2, T10 * c1 + T11 * c3_ + T12 * c4, T10 * c2 + T11 * c4 + T12 * c5
        a = T00 * u0 + T01 * u1 + T02 * u2 + f(0.3)
        b = T00 * w0 + T01 * w1 + T02 * w2
        c = T10 * w0 + T11 * w1 + T12 * w2 + f(0.3)
        denom = a * c - b * b
        d2 = one / ((denom * denom) + f(0.0000001))
        ok = d2 != 0
        dL_da = np.where(ok, d2 * (-c * c * dcx + two * b * c * dcy + (denom - a * c) * dcz), f(0))
        dL_dc = np.where(ok, d2 * (-a * a * dcz + two * a * b * dcy + (denom - a * c) * dcx), f(0))
        dL_db = np.where(ok, d2 * two * (b * c * dcx - (denom + two * b * b) * dcy + a * b * dcz), f(0))
        dc3 = np.stack([
            T00 * T00 * dL_da + T00 * T10 * dL_db + T10 * T10 * dL_dc,
            two * T00 * T01 * dL_da + (T00 * T11 + T01 * T10) * dL_db + two * T10 * T11 * dL_dc,
            two * T00 * T02 * dL_da + (T00 * T12 + T02 * T10) * dL_db + two * T10 * T12 * dL_dc,
            T01 * T01 * dL_da + T01 * T11 * dL_db + T11 * T11 * dL_dc,
            two * T02 * T01 * dL_da + (T01 * T12 + T02 * T11) * dL_db + two * T11 * T12 * dL_dc,
            T02 * T02 * dL_da + T02 * T12 * dL_db + T12 * T12 * dL_dc], 1)
        dc3 = np.where(ok[:, None], dc3, f(0)).astype(f)
        dT00, dT01, dT02 = two * u0 * dL_da + w0 * dL_db, two * u1 * dL_da + w1 * dL_db, two * u2 * dL_da + w2 * dL_db
        dT10, dT11, dT12 = two * w0 * dL_dc + u0 * dL_db, two * w1 * dL_dc + u1 * dL_db, two * w2 * dL_dc + u2 * dL_db
        dJ00 = v[0] * dT00 + v[4] * dT01 + v[8] * dT02
        dJ02 = v[2] * dT00 + v[6] * dT01 + v[10] * dT02
        dJ11 = v[1] * dT10 + v[5] * dT11 + v[9] * dT12
        dJ12 = v[2] * dT10 + v[6] * dT11 + v[10] * dT12
        itz = one / tz
        itz2 = itz * itz
        itz3 = itz2 * itz
        dtx = xg * -h_x * itz2 * dJ02
        dty = yg * -h_y * itz2 * dJ12
        dtz = -h_x * itz2 * dJ00 - h_y * itz2 * dJ11 + (two * h_x * tx) * itz3 * dJ02 + (two * h_y * ty) * itz3 * dJ12
        dm = np.stack([v[0] * dtx + v[1] * dty + v[2] * dtz, v[4] * dtx + v[5] * dty + v[6] * dtz,
                       v[8] * dtx + v[9] * dty + v[10] * dtz], 1)
        # ---- preprocess_indexedCUDA: screen position ----
        hw = pj[3] * mx + pj[7] * my + pj[11] * mz + pj[15]
        m_w = one / (hw + f(0.0000001))
        mul1 = (pj[0] * mx + pj[4] * my + pj[8] * mz + pj[12]) * m_w * m_w
        mul2 = (pj[1] * mx + pj[5] * my + pj[9] * mz + pj[13]) * m_w * m_w
        g2 = np.asarray(dl_dmean2d, f).reshape(P, 3)
        g2x, g2y = g2[:, 0], g2[:, 1]
        dm = dm + np.stack([(pj[0] * m_w - pj[3] * mul1) * g2x + (pj[1] * m_w - pj[3] * mul2) * g2y,
                            (pj[4] * m_w - pj[7] * mul1) * g2x + (pj[5] * m_w - pj[7] * mul2) * g2y,
                            (pj[8] * m_w - pj[11] * mul1) * g2x + (pj[9] * m_w - pj[11] * mul2) * g2y], 1)
        # ---- computeColorFromSH_indexed ----
        si = np.asarray(sh_indices, np.int64)
        o = m - cam[None, :]
        ln = np.sqrt(o[:, 0] * o[:, 0] + o[:, 1] * o[:, 1] + o[:, 2] * o[:, 2])
        x, y, z = o[:, 0] / ln, o[:, 1] / ln, o[:, 2] / ln
        sh = shs[si]
        gRGB = np.asarray(dl_dcolor, f).reshape(P, 3) * np.where(np.asarray(clamped).reshape(P, 3), f(0), f(1))
        bw = np.zeros((P, M), f)
        bw[:, 0] = _SH_C0
        dRx, dRy, dRz = np.zeros((P, 3), f), np.zeros((P, 3), f), np.zeros((P, 3), f)
        C1, C2, C3 = f(_SH_C1), _SH_C2.astype(f), _SH_C3.astype(f)
        col = lambda t: t[:, None]                                                   # noqa: E731
        if sh_degree > 0:
            bw[:, 1], bw[:, 2], bw[:, 3] = -C1 * y, C1 * z, -C1 * x
            dRx, dRy, dRz = -C1 * sh[:, 3], -C1 * sh[:, 1], C1 * sh[:, 2]
            if sh_degree > 1:
                xx, yy, zz, xy, yz, xz = x * x, y * y, z * z, x * y, y * z, x * z
                bw[:, 4], bw[:, 5], bw[:, 6] = C2[0] * xy, C2[1] * yz, C2[2] * (two * zz - xx - yy)
                bw[:, 7], bw[:, 8] = C2[3] * xz, C2[4] * (xx - yy)
                dRx = dRx + col(C2[0] * y) * sh[:, 4] + col(C2[2] * two * -x) * sh[:, 6] + col(C2[3] * z) * sh[:, 7] + col(C2[4] * two * x) * sh[:, 8]
                dRy = dRy + col(C2[0] * x) * sh[:, 4] + col(C2[1] * z) * sh[:, 5] + col(C2[2] * two * -y) * sh[:, 6] + col(C2[4] * two * -y) * sh[:, 8]
                dRz = dRz + col(C2[1] * y) * sh[:, 5] + col(C2[2] * two * two * z) * sh[:, 6] + col(C2[3] * x) * sh[:, 7]
                if sh_degree > 2:
                    t3, t4 = f(3), f(4)
                    bw[:, 9] = C3[0] * y * (t3 * xx - yy)
                    bw[:, 10] = C3[1] * xy * z
                    bw[:, 11] = C3[2] * y * (t4 * zz - xx - yy)
                    bw[:, 12] = C3[3] * z * (two * zz - t3 * xx - t3 * yy)
                    bw[:, 13] = C3[4] * x * (t4 * zz - xx - yy)
                    bw[:, 14] = C3[5] * z * (xx - yy)
                    bw[:, 15] = C3[6] * x * (xx - t3 * yy)
                    dRx = dRx + (col(C3[0] * t3 * two * xy) * sh[:, 9] + col(C3[1] * yz) * sh[:, 10] + col(C3[2] * -two * xy) * sh[:, 11]
                                 + col(C3[3] * -t3 * two * xz) * sh[:, 12] + col(C3[4] * (-t3 * xx + t4 * zz - yy)) * sh[:, 13]
                                 + col(C3[5] * two * xz) * sh[:, 14] + col(C3[6] * t3 * (xx - yy)) * sh[:, 15])
                    dRy = dRy + (col(C3[0] * t3 * (xx - yy)) * sh[:, 9] + col(C3[1] * xz) * sh[:, 10]
                                 + col(C3[2] * (-t3 * yy + t4 * zz - xx)) * sh[:, 11] + col(C3[3] * -t3 * two * yz) * sh[:, 12]
                                 + col(C3[4] * -two * xy) * sh[:, 13] + col(C3[5] * -two * yz) * sh[:, 14]
                                 + col(C3[6] * -t3 * two * xy) * sh[:, 15])
                    dRz = dRz + (col(C3[1] * xy) * sh[:, 10] + col(C3[2] * t4 * two * yz) * sh[:, 11]
                                 + col(C3[3] * t3 * (two * zz - xx - yy)) * sh[:, 12] + col(C3[4] * t4 * two * xz) * sh[:, 13]
                                 + col(C3[5] * (xx - yy)) * sh[:, 14])
        contrib_sh = bw[:, :, None] * gRGB[:, None, :]                                # [P, M, 3]
        if sh_degree > 0:
            ddir = np.stack([(dRx * gRGB).sum(1), (dRy * gRGB).sum(1), (dRz * gRGB).sum(1)], 1).astype(f)
        else:
            ddir = np.zeros((P, 3), f)
        ox, oy, oz = o[:, 0], o[:, 1], o[:, 2]
        sum2 = ox * ox + oy * oy + oz * oz
        inv32 = one / np.sqrt(sum2 * sum2 * sum2)
        dm = dm + np.stack([((sum2 - ox * ox) * ddir[:, 0] - oy * ox * ddir[:, 1] - oz * ox * ddir[:, 2]) * inv32,
                            (-ox * oy * ddir[:, 0] + (sum2 - oy * oy) * ddir[:, 1] - oz * oy * ddir[:, 2]) * inv32,
                            (-ox * oz * ddir[:, 0] - oy * oz * ddir[:, 1] + (sum2 - oz * oz) * ddir[:, 2]) * inv32], 1)
        # ---- computeCov3D_indexed ----
        gi = np.asarray(g_indices, np.int64)
        fac = np.asarray(scale_factors, f).reshape(P)
        scl = np.asarray(scales, f)[gi]
        q = np.asarray(rotations, f)[gi]
        r, qx, qy, qz = q[:, 0], q[:, 1], q[:, 2], q[:, 3]
        s = (fac * f(scale_modifier))[:, None] * scl
        Rs = np.empty((P, 3, 3), f)                                                  # standard rotation matrix
        Rs[:, 0, 0] = one - two * (qy * qy + qz * qz); Rs[:, 0, 1] = two * (qx * qy - r * qz); Rs[:, 0, 2] = two * (qx * qz + r * qy)
        Rs[:, 1, 0] = two * (qx * qy + r * qz); Rs[:, 1, 1] = one - two * (qx * qx + qz * qz); Rs[:, 1, 2] = two * (qy * qz - r * qx)
        Rs[:, 2, 0] = two * (qx * qz - r * qy); Rs[:, 2, 1] = two * (qy * qz + r * qx); Rs[:, 2, 2] = one - two * (qx * qx + qy * qy)
        dS = np.empty((P, 3, 3), f)
        dS[:, 0, 0], dS[:, 1, 1], dS[:, 2, 2] = dc3[:, 0], dc3[:, 3], dc3[:, 5]
        dS[:, 0, 1] = dS[:, 1, 0] = half * dc3[:, 1]
        dS[:, 0, 2] = dS[:, 2, 0] = half * dc3[:, 2]
        dS[:, 1, 2] = dS[:, 2, 1] = half * dc3[:, 4]
        Mm = two * (s[:, :, None] * np.transpose(Rs, (0, 2, 1)))                      # 2 M[i][k] = 2 s_i Rs[k][i]
        N = np.einsum("nik,nkc->nic", Mm, dS).astype(f)                               # dL_dM[i][c]
        ds = np.einsum("nci,nic->ni", Rs, N).astype(f)                                # sum_c Rs[c][i] N[i][c]
        contrib_scale = ds * fac[:, None]
        dfac = (ds * scl).sum(1).astype(f)
        N = N * s[:, :, None]
        dq = np.stack([
            two * qz * (N[:, 0, 1] - N[:, 1, 0]) + two * qy * (N[:, 2, 0] - N[:, 0, 2]) + two * qx * (N[:, 1, 2] - N[:, 2, 1]),
            two * qy * (N[:, 1, 0] + N[:, 0, 1]) + two * qz * (N[:, 2, 0] + N[:, 0, 2]) + two * r * (N[:, 1, 2] - N[:, 2, 1]) - f(4) * qx * (N[:, 2, 2] + N[:, 1, 1]),
            two * qx * (N[:, 1, 0] + N[:, 0, 1]) + two * r * (N[:, 2, 0] - N[:, 0, 2]) + two * qz * (N[:, 1, 2] + N[:, 2, 1]) - f(4) * qy * (N[:, 2, 2] + N[:, 0, 0]),
            two * r * (N[:, 0, 1] - N[:, 1, 0]) + two * qx * (N[:, 2, 0] + N[:, 0, 2]) + two * qy * (N[:, 1, 2] + N[:, 2, 1]) - f(4) * qz * (N[:, 1, 1] + N[:, 0, 0])], 1)
    out["means3d"][vis] = dm[vis]
    out["cov3d"][vis] = dc3[vis]
    out["scale_factors"][vis] = dfac[vis]
    acc_sh = np.zeros((F_, M, 3), np.float64)
    np.add.at(acc_sh, si[vis], contrib_sh[vis].astype(np.float64))
    acc_s = np.zeros((G, 3), np.float64)
    np.add.at(acc_s, gi[vis], contrib_scale[vis].astype(np.float64))
    acc_q = np.zeros((G, 4), np.float64)
    np.add.at(acc_q, gi[vis], dq[vis].astype(np.float64))
    out["shs"], out["scales"], out["rotations"] = acc_sh.astype(f), acc_s.astype(f), acc_q.astype(f)
    return out
