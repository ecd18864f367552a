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
