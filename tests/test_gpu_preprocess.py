"""GPU (-m gpu): the indexed rasteriser's per-Gaussian preprocess through the C ABI against
(1) the numpy oracle and (2) the reference's own kernel — cuda_rasterizer/forward_indexed.cu compiled unmodified into
oracle/_ref/libpreprocess_ref.so — on the same seeded scene.

Tolerances (float32 arithmetic, the three implementations differ only in the order of sums inside 3x3 products and in
fma contraction): world covariance 1e-6 of its largest entry; screen position 2e-3 px + 3e-5 relative; depth 4e-6 absolute; colour
2e-6; conic 2e-4 relative to the row's largest entry (it is an inverse: the determinant amplifies rounding);
culling decisions identical except where the determinant or the tile count sits on a boundary; the integer radius is
ceil() of a float and may differ by one on at most 0.1 % of the Gaussians, the tile count only where the radius does
or the rectangle edge falls on a tile boundary (also <= 0.1 %)."""
import ctypes
import os

import numpy as np
import pytest
import torch

from conftest import ROOT
from preprocess_scene import camera, scene

pytestmark = pytest.mark.gpu
DEV = "cuda:0"
REF_SO = os.path.join(ROOT, "oracle", "_ref", "libpreprocess_ref.so")


@pytest.fixture(scope="module")
def L(lib):
    from c3dgs_b200 import _lib
    return _lib


def _run(L, sc, cam, deg=3, mod=1.0, clamp=True):
    t = {k: torch.from_numpy(v).to(DEV) for k, v in sc.items()}
    c = {k: torch.from_numpy(v).to(DEV) for k, v in cam.items() if isinstance(v, np.ndarray)}
    out = L.preprocess_indexed(t["means3d"], t["scales"], t["scale_factors"], t["rotations"], t["opacities"], t["shs"],
                               t["sh_indices"], t["g_indices"], c["viewmatrix"], c["projmatrix"], c["cam_pos"],
                               cam["width"], cam["height"], cam["tan_fovx"], cam["tan_fovy"], sh_degree=deg,
                               scale_modifier=mod, clamp_color=clamp)
    res = {k: v.cpu().numpy() for k, v in out.items()}
    res["tiles_touched"] = res["tiles_touched"].view(np.uint32)
    return res


def _run_reference(sc, cam, deg=3, mod=1.0, clamp=True):
    lib = ctypes.CDLL(REF_SO)
    p = len(sc["means3d"])
    t = {k: torch.from_numpy(v).to(DEV) for k, v in sc.items()}
    c = {k: torch.from_numpy(v).to(DEV) for k, v in cam.items() if isinstance(v, np.ndarray)}
    out = {"radii": torch.zeros(p, dtype=torch.int32, device=DEV), "xy": torch.zeros(p, 2, device=DEV),
           "depths": torch.zeros(p, device=DEV), "cov3d": torch.zeros(p, 6, device=DEV),
           "rgb": torch.zeros(p, 3, device=DEV), "conic_opacity": torch.zeros(p, 4, device=DEV),
           "tiles_touched": torch.zeros(p, dtype=torch.int32, device=DEV),
           "clamped": torch.zeros(p, 3, dtype=torch.bool, device=DEV)}
    vp, fl, it = ctypes.c_void_p, ctypes.c_float, ctypes.c_int
    lib.ref_preprocess_indexed.argtypes = [it, it, it, vp, vp, vp, fl, vp, vp, vp, vp, vp, vp, vp, it, it, fl, fl, fl,
                                           fl, vp, vp, vp, vp, vp, vp, vp, vp, vp, it]
    f32 = np.float32
    focal_y = float(f32(cam["height"]) / (f32(2.0) * f32(cam["tan_fovy"])))
    focal_x = float(f32(cam["width"]) / (f32(2.0) * f32(cam["tan_fovx"])))
    torch.cuda.synchronize()
    rc = lib.ref_preprocess_indexed(
        p, deg, sc["shs"].shape[1], t["means3d"].data_ptr(), t["scales"].data_ptr(), t["scale_factors"].data_ptr(), mod,
        t["rotations"].data_ptr(), t["opacities"].data_ptr(), t["shs"].data_ptr(), out["clamped"].data_ptr(),
        c["viewmatrix"].data_ptr(), c["projmatrix"].data_ptr(), c["cam_pos"].data_ptr(), cam["width"], cam["height"],
        focal_x, focal_y, cam["tan_fovx"], cam["tan_fovy"], out["radii"].data_ptr(), out["xy"].data_ptr(),
        out["depths"].data_ptr(), out["cov3d"].data_ptr(), out["rgb"].data_ptr(), out["conic_opacity"].data_ptr(),
        out["tiles_touched"].data_ptr(), t["sh_indices"].data_ptr(), t["g_indices"].data_ptr(), 1 if clamp else 0)
    assert rc == 0
    res = {k: v.cpu().numpy() for k, v in out.items()}
    res["tiles_touched"] = res["tiles_touched"].view(np.uint32)
    return res


def _compare(a, b, what):
    p = len(a["radii"])
    va, vb = a["radii"] > 0, b["radii"] > 0
    assert (va != vb).mean() <= 1e-4, f"{what}: visibility differs on {(va != vb).sum()} of {p}"
    both = va & vb
    assert both.sum() > p // 10
    scale3 = np.abs(b["cov3d"]).max(1, keepdims=True) + 1e-30
    near = np.abs(b["cov3d"]).sum(1) > 0                     # written for everything past the near plane
    assert np.array_equal(near, np.abs(a["cov3d"]).sum(1) > 0)
    assert (np.abs(a["cov3d"] - b["cov3d"]) / scale3).max() <= 1e-6
    # 2e-3 px on screen; far off screen (a Gaussian a few cm from the camera lands thousands of pixels out) the
    # reciprocal of its tiny w carries the float32 rounding: 3e-5 relative
    assert (np.abs(a["xy"][both] - b["xy"][both]) <= 2e-3 + 3e-5 * np.abs(b["xy"][both])).all()
    assert np.abs(a["depths"][both] - b["depths"][both]).max() <= 4e-6     # a sum of four terms of size <= ~10
    assert np.abs(a["rgb"][both] - b["rgb"][both]).max() <= 2e-6
    assert np.array_equal(a["clamped"][both] & (np.abs(b["rgb"][both]) > 1e-5), b["clamped"][both] & (np.abs(b["rgb"][both]) > 1e-5)) \
        or (a["clamped"][both] != b["clamped"][both]).mean() <= 1e-4
    cs = np.abs(b["conic_opacity"][both, :3]).max(1, keepdims=True)
    assert (np.abs(a["conic_opacity"][both, :3] - b["conic_opacity"][both, :3]) / cs).max() <= 2e-4
    assert np.array_equal(a["conic_opacity"][both, 3], b["conic_opacity"][both, 3])
    dr = a["radii"][both].astype(np.int64) - b["radii"][both]
    assert np.abs(dr).max() <= 1 and (dr != 0).mean() <= 1e-3, f"{what}: radii differ on {(dr != 0).sum()}"
    dt = a["tiles_touched"][both] != b["tiles_touched"][both]
    assert dt.mean() <= 1e-3, f"{what}: tile counts differ on {dt.sum()}"


@pytest.mark.parametrize("deg,mod,clamp", [(3, 1.0, True), (0, 1.0, True), (2, 0.6, False), (1, 1.7, True)])
def test_preprocess_vs_oracle(L, oracle, deg, mod, clamp):
    sc, cam = scene(60_000, seed=deg), camera()
    got = _run(L, sc, cam, deg, mod, clamp)
    want = oracle.preprocess_indexed(**sc, **cam, sh_degree=deg, scale_modifier=mod, clamp_color=clamp)
    assert (got["radii"] > 0).mean() > 0.2 and (got["radii"] == 0).mean() > 0.05       # both branches exercised
    _compare(got, want, "oracle")


@pytest.mark.skipif(not os.path.exists(REF_SO), reason="oracle/_ref/libpreprocess_ref.so not built (make -C oracle ref_preprocess)")
@pytest.mark.parametrize("deg,mod,clamp", [(3, 1.0, True), (2, 0.6, False)])
def test_reference_kernel_pins_oracle_and_product(L, oracle, deg, mod, clamp):
    sc, cam = scene(300_000, g=20_000, f=50_000, seed=10 + deg), camera(width=1297, height=840, fovx=0.9)
    ref = _run_reference(sc, cam, deg, mod, clamp)
    got = _run(L, sc, cam, deg, mod, clamp)
    want = oracle.preprocess_indexed(**sc, **cam, sh_degree=deg, scale_modifier=mod, clamp_color=clamp)
    _compare(got, ref, "product vs reference kernel")
    _compare(want, ref, "oracle vs reference kernel")


def test_preprocess_edge_cases_and_errors(L, oracle):
    cam = camera(width=64, height=48)
    sc = scene(1, g=1, f=1, seed=3)
    for p in (0, 1, 255, 257):
        s2 = scene(max(p, 1), g=3, f=2, seed=p)
        if p == 0:
            s2 = {k: (v[:0] if k in ("means3d", "scale_factors", "opacities", "sh_indices", "g_indices") else v) for k, v in s2.items()}
        got = _run(L, s2, cam)
        assert got["radii"].shape == (p,)
        if p:
            _compare_small = oracle.preprocess_indexed(**s2, **cam)
            assert np.array_equal(got["radii"] > 0, _compare_small["radii"] > 0)
    t = {k: torch.from_numpy(v).to(DEV) for k, v in sc.items()}
    c = {k: torch.from_numpy(v).to(DEV) for k, v in cam.items() if isinstance(v, np.ndarray)}
    # buffer reuse: a second call into the first call's outputs (dirtied in between) gives the same result in place
    s3 = scene(5000, g=50, f=40, seed=8)
    t3 = {k: torch.from_numpy(v).to(DEV) for k, v in s3.items()}
    args3 = (t3["means3d"], t3["scales"], t3["scale_factors"], t3["rotations"], t3["opacities"], t3["shs"],
             t3["sh_indices"], t3["g_indices"], c["viewmatrix"], c["projmatrix"], c["cam_pos"], 64, 48,
             cam["tan_fovx"], cam["tan_fovy"])
    first = L.preprocess_indexed(*args3)
    keep = {k: v.clone() for k, v in first.items()}
    for v in first.values():
        v.fill_(1)
    again = L.preprocess_indexed(*args3, out=first)
    assert all(again[k].data_ptr() == first[k].data_ptr() and torch.equal(again[k], keep[k]) for k in keep)
    with pytest.raises(RuntimeError):                          # degree 3 needs 16 coefficients per row
        L.preprocess_indexed(t["means3d"], t["scales"], t["scale_factors"], t["rotations"], t["opacities"],
                             t["shs"][:, :9].contiguous(), t["sh_indices"], t["g_indices"], c["viewmatrix"],
                             c["projmatrix"], c["cam_pos"], 64, 48, 0.5, 0.4, sh_degree=3)
    with pytest.raises(RuntimeError, match="CUDA"):
        L.preprocess_indexed(t["means3d"].cpu(), t["scales"], t["scale_factors"], t["rotations"], t["opacities"], t["shs"],
                             t["sh_indices"], t["g_indices"], c["viewmatrix"], c["projmatrix"], c["cam_pos"], 64, 48, 0.5, 0.4)


# ------------------------------------------------------------------------------------------------
# Backward (second half): c3vq_preprocess_indexed_backward against the numpy oracle and against
# the reference's own backward_indexed.cu compiled unmodified (oracle/_ref/libpreprocess_bwd_ref.so).
#
# Tolerances.  float32 throughout; the three implementations differ in fma contraction and in the order of sums,
# and the table gradients are sums over Gaussians in atomic order (reference and product alike).  Errors are measured
# per row, relative to the row's largest entry.  The float32 and float64 evaluations of the oracle itself differ by
# 2e-7 (median), 1e-4 (99.9th percentile) and 1e-2 (worst row: the size-factor gradient is a sum of three cancelling
# terms), so the test asks for median <= 2e-6, 99.9th percentile <= 5e-4, worst row <= 5e-2.
BWD_SO = os.path.join(ROOT, "oracle", "_ref", "libpreprocess_bwd_ref.so")
BWD_KEYS = ("means3d", "cov3d", "scale_factors", "shs", "scales", "rotations")


def _grads(p, seed):
    rng = np.random.default_rng(seed)
    return {"dl_dmean2d": rng.standard_normal((p, 3)).astype(np.float32),
            "dl_dconic": rng.standard_normal((p, 2, 2)).astype(np.float32),
            "dl_dcolor": rng.standard_normal((p, 3)).astype(np.float32)}


def _run_bwd(L, sc, cam, fwd, gr, deg=3, mod=1.0):
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(DEV) for k, v in {**sc, **gr}.items()}
    c = {k: torch.from_numpy(v).to(DEV) for k, v in cam.items() if isinstance(v, np.ndarray)}
    out = L.preprocess_indexed_backward(
        t["means3d"], torch.from_numpy(fwd["radii"]).to(DEV), t["shs"], torch.from_numpy(fwd["clamped"]).to(DEV),
        t["scales"], t["scale_factors"], t["rotations"], torch.from_numpy(fwd["cov3d"]).to(DEV), c["viewmatrix"],
        c["projmatrix"], c["cam_pos"], cam["width"], cam["height"], cam["tan_fovx"], cam["tan_fovy"], t["dl_dmean2d"],
        t["dl_dconic"], t["dl_dcolor"], t["sh_indices"], t["g_indices"], sh_degree=deg, scale_modifier=mod)
    return {k: v.cpu().numpy().reshape(v.shape[0], int(np.prod(v.shape[1:]))) for k, v in out.items()}


def _run_bwd_reference(sc, cam, fwd, gr, deg=3, mod=1.0):
    lib = ctypes.CDLL(BWD_SO)
    p = len(sc["means3d"])
    t = {k: torch.from_numpy(np.ascontiguousarray(v)).to(DEV) for k, v in {**sc, **gr}.items()}
    c = {k: torch.from_numpy(v).to(DEV) for k, v in cam.items() if isinstance(v, np.ndarray)}
    radii, clamped, cov3d = (torch.from_numpy(fwd[k]).to(DEV) for k in ("radii", "clamped", "cov3d"))
    out = {"means3d": torch.zeros(p, 3, device=DEV), "cov3d": torch.zeros(p, 6, device=DEV),
           "scale_factors": torch.zeros(p, 1, device=DEV), "shs": torch.zeros_like(t["shs"]),
           "scales": torch.zeros_like(t["scales"]), "rotations": torch.zeros_like(t["rotations"])}
    vp, fl, it = ctypes.c_void_p, ctypes.c_float, ctypes.c_int
    lib.ref_preprocess_indexed_backward.argtypes = [it, it, it, vp, vp, vp, vp, vp, vp, vp, fl, vp, vp, vp, fl, fl, fl,
                                                    fl, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    f32 = np.float32
    focal_y = float(f32(cam["height"]) / (f32(2.0) * f32(cam["tan_fovy"])))
    focal_x = float(f32(cam["width"]) / (f32(2.0) * f32(cam["tan_fovx"])))
    torch.cuda.synchronize()
    rc = lib.ref_preprocess_indexed_backward(
        p, deg, sc["shs"].shape[1], t["means3d"].data_ptr(), radii.data_ptr(), t["shs"].data_ptr(), clamped.data_ptr(),
        t["scales"].data_ptr(), t["scale_factors"].data_ptr(), t["rotations"].data_ptr(), mod, cov3d.data_ptr(),
        c["viewmatrix"].data_ptr(), c["projmatrix"].data_ptr(), focal_x, focal_y, cam["tan_fovx"], cam["tan_fovy"],
        c["cam_pos"].data_ptr(), t["dl_dmean2d"].data_ptr(), t["dl_dconic"].data_ptr(), out["means3d"].data_ptr(),
        t["dl_dcolor"].data_ptr(), out["cov3d"].data_ptr(), out["shs"].data_ptr(), out["scales"].data_ptr(),
        out["scale_factors"].data_ptr(), out["rotations"].data_ptr(), t["sh_indices"].data_ptr(),
        t["g_indices"].data_ptr())
    assert rc == 0
    return {k: v.cpu().numpy().reshape(v.shape[0], int(np.prod(v.shape[1:]))) for k, v in out.items()}


def _oracle_bwd(oracle, sc, cam, fwd, gr, deg=3, mod=1.0, dtype=np.float32):
    out = oracle.preprocess_indexed_backward(
        sc["means3d"], fwd["radii"], sc["shs"], fwd["clamped"], sc["scales"], sc["scale_factors"], sc["rotations"],
        fwd["cov3d"], cam["viewmatrix"], cam["projmatrix"], cam["cam_pos"], cam["width"], cam["height"],
        cam["tan_fovx"], cam["tan_fovy"], gr["dl_dmean2d"], gr["dl_dconic"], gr["dl_dcolor"], sc["sh_indices"],
        sc["g_indices"], sh_degree=deg, scale_modifier=mod, dtype=dtype)
    return {k: np.asarray(v).reshape(len(v), int(np.prod(np.asarray(v).shape[1:]))) for k, v in out.items()}


def _compare_bwd(a, b, what):
    for k in BWD_KEYS:
        x, y = a[k].astype(np.float64), b[k].astype(np.float64)
        assert x.shape == y.shape, f"{what}: {k} shape {x.shape} vs {y.shape}"
        assert np.array_equal(np.abs(x).sum(1) > 0, np.abs(y).sum(1) > 0) or \
            ((np.abs(x).sum(1) > 0) != (np.abs(y).sum(1) > 0)).mean() <= 1e-4, f"{what}: {k} support differs"
        live = np.abs(y).max(1) > 0
        assert live.sum() > 0, f"{what}: {k} is all zero"
        e = (np.abs(x - y)[live] / np.abs(y)[live].max(1, keepdims=True)).max(1)
        assert np.isfinite(x).all(), f"{what}: {k} holds non-finite values"
        assert np.median(e) <= 2e-6 and np.quantile(e, 0.999) <= 5e-4 and e.max() <= 5e-2, \
            f"{what}: {k} row-relative error median {np.median(e):.2e}, p99.9 {np.quantile(e, 0.999):.2e}, max {e.max():.2e}"


@pytest.mark.parametrize("deg,mod", [(3, 1.0), (0, 1.0), (2, 0.6), (1, 1.7)])
def test_preprocess_backward_vs_oracle(L, oracle, deg, mod):
    sc, cam = scene(60_000, seed=20 + deg), camera()
    fwd = _run(L, sc, cam, deg, mod, True)
    gr = _grads(60_000, deg)
    got = _run_bwd(L, sc, cam, fwd, gr, deg, mod)
    want = _oracle_bwd(oracle, sc, cam, fwd, gr, deg, mod)
    culled = fwd["radii"] <= 0
    assert culled.any() and (~culled).any()
    for k in ("means3d", "cov3d", "scale_factors"):                 # culled Gaussians: exact zeros
        assert not got[k][culled].any()
    if deg < 3:                                                     # coefficients beyond the active degree: untouched
        assert not got["shs"].reshape(len(sc["shs"]), 16, 3)[:, (deg + 1) ** 2:].any()
    _compare_bwd(got, want, "oracle")


@pytest.mark.skipif(not os.path.exists(BWD_SO), reason="oracle/_ref/libpreprocess_bwd_ref.so not built (make -C oracle ref_preprocess_bwd)")
@pytest.mark.parametrize("deg,mod", [(3, 1.0), (2, 0.6)])
def test_reference_backward_kernels_pin_oracle_and_product(L, oracle, deg, mod):
    """The reference's own computeCov2D_indexedCUDA + preprocess_indexedCUDA (backward_indexed.cu, unmodified) on the
    B200 next to the product kernel and the numpy restatement, on a 300 k-Gaussian scene with 20 k shape rows and
    50 k colour rows (so table rows collect several Gaussians each)."""
    sc, cam = scene(300_000, g=20_000, f=50_000, seed=30 + deg), camera(width=1297, height=840, fovx=0.9)
    fwd = _run(L, sc, cam, deg, mod, True)
    gr = _grads(300_000, 7 + deg)
    ref = _run_bwd_reference(sc, cam, fwd, gr, deg, mod)
    got = _run_bwd(L, sc, cam, fwd, gr, deg, mod)
    want = _oracle_bwd(oracle, sc, cam, fwd, gr, deg, mod)
    _compare_bwd(got, ref, "product vs reference kernels")
    _compare_bwd(want, ref, "oracle vs reference kernels")


def test_preprocess_backward_edge_cases_and_errors(L, oracle):
    cam = camera(width=64, height=48)
    c = {k: torch.from_numpy(v).to(DEV) for k, v in cam.items() if isinstance(v, np.ndarray)}
    for p in (0, 1, 255, 257):
        s2 = scene(max(p, 1), g=3, f=2, seed=40 + p)
        if p == 0:
            s2 = {k: (v[:0] if k in ("means3d", "scale_factors", "opacities", "sh_indices", "g_indices") else v) for k, v in s2.items()}
        fwd = _run(L, s2, cam)
        got = _run_bwd(L, s2, cam, fwd, _grads(p, p))
        assert got["means3d"].shape == (p, 3) and got["shs"].shape == (2, 48) and got["rotations"].shape == (3, 4)
        if p:
            _compare_bwd_small = _oracle_bwd(oracle, s2, cam, fwd, _grads(p, p))
            for k in BWD_KEYS:
                assert np.allclose(got[k], _compare_bwd_small[k], rtol=2e-3, atol=1e-5 * (np.abs(_compare_bwd_small[k]).max() + 1e-30))
    # a colour table with 9 coefficients per row (degree 2 scenes): the scalar-atomic path
    s9 = scene(5000, g=50, f=40, seed=9, coeffs=9)
    fwd = _run(L, s9, cam, deg=2)
    gr = _grads(5000, 9)
    got = _run_bwd(L, s9, cam, fwd, gr, deg=2)
    want = _oracle_bwd(oracle, s9, cam, fwd, gr, deg=2)
    _compare_bwd(got, want, "9-coefficient table")
    t = {k: torch.from_numpy(v).to(DEV) for k, v in s9.items()}
    g = {k: torch.from_numpy(v).to(DEV) for k, v in gr.items()}
    f = {k: torch.from_numpy(fwd[k]).to(DEV) for k in ("radii", "clamped", "cov3d")}
    with pytest.raises(RuntimeError):                          # degree 3 needs 16 coefficients per row
        L.preprocess_indexed_backward(t["means3d"], f["radii"], t["shs"], f["clamped"], t["scales"], t["scale_factors"],
                                      t["rotations"], f["cov3d"], c["viewmatrix"], c["projmatrix"], c["cam_pos"], 64, 48,
                                      0.5, 0.4, g["dl_dmean2d"], g["dl_dconic"], g["dl_dcolor"], t["sh_indices"],
                                      t["g_indices"], sh_degree=3)
    with pytest.raises(RuntimeError, match="elements"):        # gradient array of the wrong length
        L.preprocess_indexed_backward(t["means3d"], f["radii"], t["shs"], f["clamped"], t["scales"], t["scale_factors"],
                                      t["rotations"], f["cov3d"], c["viewmatrix"], c["projmatrix"], c["cam_pos"], 64, 48,
                                      0.5, 0.4, g["dl_dmean2d"][:10], g["dl_dconic"], g["dl_dcolor"], t["sh_indices"],
                                      t["g_indices"], sh_degree=2)
