"""Indexed preprocess on a 6M-Gaussian indexed scene: product kernel vs the reference's own kernel
(oracle/_ref/libpreprocess_ref.so, compiled unmodified) on the same inputs, CUDA-event timed.
Algorithmic bytes per Gaussian: 36 in (position, factor, opacity, two int64 indices) + 28 + 192 gathered table bytes
+ 75 out for a visible one."""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from c3dgs_b200 import _lib  # noqa: E402
from preprocess_scene import camera  # noqa: E402

dev = torch.device("cuda:0")
p, g, f = 6_000_000, 1_500_000, 720_000
gen = torch.Generator(device=dev).manual_seed(0)
cam = camera(width=1297, height=840, fovx=0.9)
c = {k: torch.from_numpy(v).to(dev) for k, v in cam.items() if isinstance(v, np.ndarray)}
means = torch.randn(p, 3, device=dev, generator=gen) * torch.tensor([2.5, 1.5, 3.0], device=dev)
scales = torch.nn.functional.normalize(torch.rand(g, 3, device=dev, generator=gen) + 0.05)
rots = torch.nn.functional.normalize(torch.randn(g, 4, device=dev, generator=gen))
factors = torch.exp(torch.randn(p, device=dev, generator=gen) * 0.7 - 4.5)
opac = torch.rand(p, device=dev, generator=gen)
shs = torch.randn(f, 16, 3, device=dev, generator=gen) * 0.3
shi = torch.randint(0, f, (p,), device=dev, generator=gen)
gi = torch.randint(0, g, (p,), device=dev, generator=gen)


def timed(fn, reps=10):
    fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def ours():
    return _lib.preprocess_indexed(means, scales, factors, rots, opac, shs, shi, gi, c["viewmatrix"], c["projmatrix"],
                                   c["cam_pos"], cam["width"], cam["height"], cam["tan_fovx"], cam["tan_fovy"])


out = ours()
ms = timed(ours)
vis = int((out["radii"] > 0).sum())
nbytes = p * 36 + int((out["cov3d"].abs().sum(1) != 0).sum()) * 52 + vis * (192 + 51)
print(f"product: {ms:.3f} ms per call for {p} Gaussians ({vis} visible), through the Python wrapper (allocates the "
      f"eight outputs every call); {nbytes / ms * 1e-6:.0f} GB/s algorithmic; kernel-only times: ncu")
so = os.path.join(ROOT, "oracle", "_ref", "libpreprocess_ref.so")
if os.path.exists(so):
    lib = ctypes.CDLL(so)
    vp, fl, it = ctypes.c_void_p, ctypes.c_float, ctypes.c_int
    lib.ref_preprocess_indexed.argtypes = [it, it, it, vp, vp, vp, fl, vp, vp, vp, vp, vp, vp, vp, it, it, fl, fl, fl, fl,
                                           vp, vp, vp, vp, vp, vp, vp, vp, vp, it]
    o = {k: torch.zeros_like(v) for k, v in out.items()}
    fy = float(np.float32(cam["height"]) / (np.float32(2.0) * np.float32(cam["tan_fovy"])))
    fx = float(np.float32(cam["width"]) / (np.float32(2.0) * np.float32(cam["tan_fovx"])))

    def ref():
        lib.ref_preprocess_indexed(p, 3, 16, means.data_ptr(), scales.data_ptr(), factors.data_ptr(), 1.0, rots.data_ptr(),
                                   opac.data_ptr(), shs.data_ptr(), o["clamped"].data_ptr(), c["viewmatrix"].data_ptr(),
                                   c["projmatrix"].data_ptr(), c["cam_pos"].data_ptr(), cam["width"], cam["height"], fx, fy,
                                   cam["tan_fovx"], cam["tan_fovy"], o["radii"].data_ptr(), o["xy"].data_ptr(),
                                   o["depths"].data_ptr(), o["cov3d"].data_ptr(), o["rgb"].data_ptr(),
                                   o["conic_opacity"].data_ptr(), o["tiles_touched"].data_ptr(), shi.data_ptr(),
                                   gi.data_ptr(), 1)

    torch.cuda.synchronize()
    print(f"reference kernel (forward_indexed.cu, unmodified, legacy stream, host-synchronised per call): {timed(ref):.3f} ms")

# ---- backward (backward_indexed.cu: computeCov2D_indexedCUDA + preprocess_indexedCUDA vs k_preprocess_indexed_bwd) ----
g2d = torch.randn(p, 3, device=dev, generator=gen)
gcon = torch.randn(p, 2, 2, device=dev, generator=gen)
gcol = torch.randn(p, 3, device=dev, generator=gen)


def ours_bwd():
    return _lib.preprocess_indexed_backward(means, out["radii"], shs, out["clamped"], scales, factors, rots, out["cov3d"],
                                            c["viewmatrix"], c["projmatrix"], c["cam_pos"], cam["width"], cam["height"],
                                            cam["tan_fovx"], cam["tan_fovy"], g2d, gcon, gcol, shi, gi)


gb = ours_bwd()
ms = timed(ours_bwd)
print(f"backward, product: {ms:.3f} ms per call ({vis} visible), through the Python wrapper (allocates the three "
      f"per-Gaussian outputs and zero-fills the three table gradients, {sum(v.numel() for v in gb.values()) * 4 / 1e6:.0f} MB, every call)")
so = os.path.join(ROOT, "oracle", "_ref", "libpreprocess_bwd_ref.so")
if os.path.exists(so):
    lib = ctypes.CDLL(so)
    vp, fl, it = ctypes.c_void_p, ctypes.c_float, ctypes.c_int
    lib.ref_preprocess_indexed_backward.argtypes = [it, it, it, vp, vp, vp, vp, vp, vp, vp, fl, vp, vp, vp, fl, fl, fl,
                                                    fl, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, vp]
    ob = {k: torch.zeros_like(v) for k, v in gb.items()}
    fy = float(np.float32(cam["height"]) / (np.float32(2.0) * np.float32(cam["tan_fovy"])))
    fx = float(np.float32(cam["width"]) / (np.float32(2.0) * np.float32(cam["tan_fovx"])))

    def ref_bwd():
        for k_ in ("shs", "scales", "rotations", "means3d", "cov3d", "scale_factors"):
            ob[k_].zero_()                     # rasterize_points_indexed.cu:166-175 allocates zeros every call
        lib.ref_preprocess_indexed_backward(
            p, 3, 16, means.data_ptr(), out["radii"].data_ptr(), shs.data_ptr(), out["clamped"].data_ptr(),
            scales.data_ptr(), factors.data_ptr(), rots.data_ptr(), 1.0, out["cov3d"].data_ptr(),
            c["viewmatrix"].data_ptr(), c["projmatrix"].data_ptr(), fx, fy, cam["tan_fovx"], cam["tan_fovy"],
            c["cam_pos"].data_ptr(), g2d.data_ptr(), gcon.data_ptr(), ob["means3d"].data_ptr(), gcol.data_ptr(),
            ob["cov3d"].data_ptr(), ob["shs"].data_ptr(), ob["scales"].data_ptr(), ob["scale_factors"].data_ptr(),
            ob["rotations"].data_ptr(), shi.data_ptr(), gi.data_ptr())

    torch.cuda.synchronize()
    print(f"backward, reference kernels (backward_indexed.cu, unmodified, two launches, legacy stream, host-synchronised "
          f"per call, incl. zero-filling the six outputs): {timed(ref_bwd, reps=5):.3f} ms")
    worst = 0.0
    for k_ in ("shs", "scales", "rotations"):
        a_, b_ = gb[k_].reshape(gb[k_].shape[0], -1), ob[k_].reshape(ob[k_].shape[0], -1)
        worst = max(worst, float(((a_ - b_).abs().max(1).values / (b_.abs().max(1).values + 1e-30)).max()))
    print(f"table gradients, product vs reference kernels: largest row-relative difference {worst:.2e}")
