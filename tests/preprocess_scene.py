"""Synthetic camera + indexed scene for the preprocess tests (numpy, seeded): Gaussians in a box in front of and
behind a look-at camera, a shape table (normalised scalings, unit quaternions), a colour table of SH rows."""
import numpy as np


def camera(width=979, height=546, fovx=1.1, znear=0.01, zfar=100.0, eye=(0.3, -0.2, -4.0)):
    eye = np.asarray(eye, np.float64)
    fwd = -eye / np.linalg.norm(eye)                      # looks at the origin, +z forward (COLMAP convention)
    right = np.cross(np.array([0.0, 1.0, 0.0]), fwd)
    right /= np.linalg.norm(right)
    up = np.cross(fwd, right)
    Rm = np.stack([right, up, fwd], 0)                    # world -> camera
    V = np.eye(4)
    V[:3, :3] = Rm
    V[:3, 3] = -Rm @ eye
    tan_fovx = np.tan(fovx * 0.5)
    tan_fovy = tan_fovx * height / width
    Pm = np.zeros((4, 4))
    Pm[0, 0], Pm[1, 1] = 1.0 / tan_fovx, 1.0 / tan_fovy
    Pm[3, 2] = 1.0
    Pm[2, 2], Pm[2, 3] = zfar / (zfar - znear), -(zfar * znear) / (zfar - znear)
    full = Pm @ V
    # the reference's Python side hands over the transposed matrices, i.e. flat storage = column-major
    return {"viewmatrix": V.T.astype(np.float32).reshape(16).copy(), "projmatrix": full.T.astype(np.float32).reshape(16).copy(),
            "cam_pos": eye.astype(np.float32), "width": width, "height": height,
            "tan_fovx": float(np.float32(tan_fovx)), "tan_fovy": float(np.float32(tan_fovy))}


def scene(p, g=5000, f=3000, seed=0, coeffs=16):
    rng = np.random.default_rng(seed)
    means = (rng.standard_normal((p, 3)) * np.array([2.5, 1.5, 3.0])).astype(np.float32)
    sc = np.abs(rng.standard_normal((g, 3))).astype(np.float32) + 0.05
    sc /= np.linalg.norm(sc, axis=1, keepdims=True)
    rot = rng.standard_normal((g, 4)).astype(np.float32)
    rot /= np.linalg.norm(rot, axis=1, keepdims=True)
    factors = np.exp(rng.standard_normal(p) * 0.7 - 3.5).astype(np.float32)
    opac = rng.random(p).astype(np.float32)
    shs = (rng.standard_normal((f, coeffs, 3)) * 0.3).astype(np.float32)
    return {"means3d": means, "scales": sc.astype(np.float32), "scale_factors": factors, "rotations": rot.astype(np.float32),
            "opacities": opac, "shs": shs, "sh_indices": rng.integers(0, f, p).astype(np.int64),
            "g_indices": rng.integers(0, g, p).astype(np.int64)}
