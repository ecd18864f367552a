"""Generate tests/golden/vq_*.npz by running the reference's own compression/vq.py.

Run in the build container (where /root/reference exists):

    python -m oracle.make_golden

The reference module is imported unchanged through oracle/ref_loader.py.  Its
`vq_features` (compression/vq.py:49-87) is executed on CPU tensors with
  * torch.cuda.synchronize / empty_cache patched to no-ops (no driver here),
  * torch.randint, VectorQuantize.uniform_init and VectorQuantize.update wrapped by
    recorders, so the fixture holds every random draw and every intermediate state.
The fixtures pin the Python half of the path (weight normalisation, batch draw,
update algebra, EMA order, eps, trace normalisation, final assignment).  The CUDA
half (weighted_distance.cu) is pinned on the GPU box by oracle/Makefile (target `ref`).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)

from oracle import ref_loader  # noqa: E402

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")

# name, N, D, K, vq_chunk, steps, scale_normalize, seed
CASES = [
    ("sh45", 1000, 45, 64, 256, 4, False, 0),       # BASELINE config 1/2 width
    ("sh48", 1000, 48, 64, 256, 3, False, 1),       # fork default colour width (arguments/__init__.py:97)
    ("shape7", 1500, 7, 96, 512, 4, False, 2),      # BASELINE config 3 width
    ("cov6", 1500, 6, 96, 512, 4, True, 3),         # fork reality: 6-D covariance + trace normalise (vq.py:159-172)
]


def synth(n, d, seed, cov=False):
    """Seeded CPU inputs: clustered features, heavy-tailed sensitivities."""
    g = torch.Generator().manual_seed(seed)
    centres = torch.randn(37, d, generator=g)
    x = centres[torch.randint(0, 37, (n,), generator=g)] + 0.25 * torch.randn(n, d, generator=g)
    if cov:
        # symmetric positive-definite 3x3, upper triangle (xx,xy,xz,yy,yz,zz), trace-normalised like
        # get_normalized_covariance(strip_sym=True) would give
        a = torch.randn(n, 3, 3, generator=g)
        m = a @ a.transpose(1, 2) + 0.05 * torch.eye(3)
        m = m / m.diagonal(dim1=1, dim2=2).sum(-1)[:, None, None]
        x = torch.stack([m[:, 0, 0], m[:, 0, 1], m[:, 0, 2], m[:, 1, 1], m[:, 1, 2], m[:, 2, 2]], 1)
    w = torch.rand(n, generator=g) ** 4
    return x.contiguous().float(), w.contiguous().float()


def run_case(name, n, d, k, chunk, steps, scale_normalize, seed):
    mod = ref_loader.load()
    x, w = synth(n, d, seed, cov=scale_normalize)
    rec = {"batches": [], "codebook0": None, "upd_x": [], "upd_w": [], "upd_dist": [], "upd_idx": [],
           "cb_after": [], "ei_after": []}

    VQ = mod.VectorQuantize
    orig_init, orig_update = VQ.uniform_init, VQ.update
    orig_randint = torch.randint
    orig_sync, orig_empty = torch.cuda.synchronize, torch.cuda.empty_cache
    orig_wd = mod.weightedDistance
    last = {}

    def wd(coefs, codebook):
        dist, idx = orig_wd(coefs, codebook)
        last["idx"] = idx.clone()
        return dist, idx

    def uniform_init(self, xx):
        orig_init(self, xx)
        rec["codebook0"] = self.codebook.data.clone()

    def update(self, xx, importance):
        out = orig_update(self, xx, importance)
        rec["upd_dist"].append(out.clone())
        rec["upd_idx"].append(last["idx"])
        rec["cb_after"].append(self.codebook.data.clone())
        rec["ei_after"].append(self.entry_importance.data.clone())
        return out

    def randint(*a, **kw):
        r = orig_randint(*a, **kw)
        rec["batches"].append(r.clone())
        return r

    torch.manual_seed(1000 + seed)
    try:
        VQ.uniform_init, VQ.update = uniform_init, update
        mod.weightedDistance = wd
        torch.randint = randint
        torch.cuda.synchronize = lambda *a, **kw: None
        torch.cuda.empty_cache = lambda *a, **kw: None
        codebook, indices = mod.vq_features(x, w, k, chunk, steps, scale_normalize=scale_normalize, silent=True)
    finally:
        VQ.uniform_init, VQ.update = orig_init, orig_update
        torch.randint = orig_randint
        torch.cuda.synchronize, torch.cuda.empty_cache = orig_sync, orig_empty

    # after cb_after was recorded the reference may have trace-normalised (vq.py:73-77); store both views
    out = dict(
        features=x.numpy(), importance=w.numpy(), codebook0=rec["codebook0"].numpy(),
        batches=torch.stack(rec["batches"]).numpy(),
        upd_dist=torch.stack(rec["upd_dist"]).numpy(), upd_idx=torch.stack(rec["upd_idx"]).numpy(),
        cb_after_update=torch.stack(rec["cb_after"]).numpy(), ei_after_update=torch.stack(rec["ei_after"]).numpy(),
        codebook=codebook.numpy(), indices=indices.numpy(),
        meta=np.array([n, d, k, chunk, steps, int(scale_normalize), seed], np.int64),
        decay=np.array(0.8), eps=np.array(1e-5),
    )
    os.makedirs(GOLDEN_DIR, exist_ok=True)
    path = os.path.join(GOLDEN_DIR, f"vq_{name}.npz")
    np.savez_compressed(path, **out)
    return path


def run_join_cases():
    """join_features (compression/vq.py:90-103) of the reference, run unchanged on CPU tensors: inputs and outputs
    of a few shapes incl. nothing kept, everything kept, an empty codebook and a ragged last block."""
    mod = ref_loader.load()
    out = {}
    cases = [("a", 5000, 45, 64, 0.07), ("b", 2049, 6, 17, 0.5), ("none", 300, 7, 5, 0.0), ("all", 300, 7, 5, 1.0),
             ("emptycb", 1500, 48, 0, 0.3), ("one", 1, 3, 2, 1.0)]
    for name, n, d, k, frac in cases:
        g = torch.Generator().manual_seed(n + d)
        feats = torch.randn(n, d, generator=g)
        keep = torch.rand(n, generator=g) < frac
        cb = torch.randn(k, d, generator=g)
        cb_idx = torch.randint(0, max(k, 1), (int((~keep).sum()),), generator=g)
        comp, idx = mod.join_features(feats, keep, cb, cb_idx)
        out.update({f"{name}_feats": feats.numpy(), f"{name}_keep": keep.numpy(), f"{name}_cb": cb.numpy(),
                    f"{name}_cbidx": cb_idx.numpy(), f"{name}_out": comp.numpy(), f"{name}_idx": idx.numpy()})
    out["names"] = np.array([c[0] for c in cases])
    path = os.path.join(GOLDEN_DIR, "join_features.npz")
    np.savez_compressed(path, **out)
    return path


def synth_covariances(seed=0):
    """Inputs for extract_rot_scale: trace-normalised covariances like get_normalized_covariance(strip_sym=True)
    (scene/gaussian_model.py:274-278) plus the degenerate shapes the decomposition has to survive."""
    g = torch.Generator().manual_seed(seed)
    n = 4000
    q = torch.nn.functional.normalize(torch.randn(n, 4, generator=g))
    r, x, y, z = q.unbind(-1)
    R = torch.stack([1 - 2 * (y * y + z * z), 2 * (x * y - r * z), 2 * (x * z + r * y),
                     2 * (x * y + r * z), 1 - 2 * (x * x + z * z), 2 * (y * z - r * x),
                     2 * (x * z - r * y), 2 * (y * z + r * x), 1 - 2 * (x * x + y * y)], -1).reshape(n, 3, 3)
    s = torch.exp(torch.randn(n, 3, generator=g) * 1.5 - 3.0)
    s[:500, 1] = s[:500, 0]                                  # a repeated eigenvalue
    s[500:700] = s[500:700, :1]                              # isotropic
    s[700:900, 0] = 0.0                                      # flat splats (rank 2)
    s[900:1000, :2] = 0.0                                    # needles (rank 1)
    s[1000:1100] *= 1e-4                                     # tiny
    s[1100:1200] *= 1e3                                      # large
    cov = R @ torch.diag_embed(s * s) @ R.transpose(-1, -2)
    cov[1200:3000] /= cov[1200:3000].diagonal(dim1=-1, dim2=-2).sum(-1)[:, None, None]   # trace-normalised
    cov6 = torch.stack([cov[:, 0, 0], cov[:, 0, 1], cov[:, 0, 2], cov[:, 1, 1], cov[:, 1, 2], cov[:, 2, 2]], -1)
    special = torch.tensor([[0, 0, 0, 0, 0, 0], [1, 0, 0, 1, 0, 1], [3, 0, 0, 2, 0, 1], [1, 0, 0, 2, 0, 3],
                            [1, 1, 1, 1, 1, 1], [-1, 0, 0, -2, 0, -3], [1, 2, 0, 1, 0, -5], [0, 1, 0, 0, 0, 0],
                            [1e-12, 0, 0, 1e-12, 0, 1e-12], [2, -1, 0, 2, -1, 2]], dtype=torch.float32)
    return torch.cat([special, cov6.float()], 0).contiguous()


def run_rot_scale_case():
    """extract_rot_scale(to_full_cov(.)) of the reference (utils/splats.py:7-36, called at compression/vq.py:186),
    run unchanged on CPU tensors (float32 LAPACK)."""
    mod = ref_loader.load()
    cov6 = synth_covariances()
    rot, scale = mod.extract_rot_scale(mod.to_full_cov(cov6))
    path = os.path.join(GOLDEN_DIR, "rot_scale.npz")
    np.savez_compressed(path, cov6=cov6.numpy(), rot=rot.numpy(), scale=scale.numpy())
    return path


def _reference_functions(path, names):
    """Compile the named top-level functions of a reference file in isolation (the module itself imports the
    rasterisers and simple_knn, which are not built here); nothing is copied into this repository."""
    import ast
    src = open(path).read()
    tree = ast.parse(src)
    ns = {"torch": torch}
    for node in tree.body:
        if isinstance(node, ast.FunctionDef) and node.name in names:
            exec(compile(ast.Module([node], []), path, "exec"), ns)
    return [ns[n] for n in names]


def run_morton_cases():
    """GaussianModel._sort_morton's permutation (scene/gaussian_model.py:997-1004) with the reference's own
    mortonEncode / splitBy3 (:1415-1432) executed from the checkout, on CPU tensors."""
    (morton_encode, _) = _reference_functions(os.path.join(ref_loader.REFERENCE_ROOT, "scene", "gaussian_model.py"),
                                              ["mortonEncode", "splitBy3"])
    out = {}
    g = torch.Generator().manual_seed(77)
    cases = {
        "normal": torch.randn(20000, 3, generator=g) * torch.tensor([3.0, 1.0, 7.0]),
        "ties": torch.randn(500, 3, generator=g)[torch.randint(0, 500, (6000,), generator=g)],
        "flat": torch.cat([torch.rand(3000, 2, generator=g), torch.full((3000, 1), 0.25)], 1),
        "five": torch.tensor([[0., 0, 0], [1, 1, 1], [0.5, 0.25, 0.75], [1, 0, 0], [0.5, 0.25, 0.75]]),
        "one": torch.tensor([[1.5, -2.0, 3.0]]),
    }
    for name, xyz in cases.items():
        xyz = xyz.float().contiguous()
        pp_min = xyz.min(0).values                                                     # :999
        pp_diap = xyz.max(0).values - pp_min                                           # :1000
        xyz_q = ((2 ** 21 - 1) * (xyz - pp_min) / pp_diap).long()                      # :1001
        keys = morton_encode(xyz_q, pp_diap.argsort())                                 # :1003
        order = keys.sort().indices
        out.update({f"{name}_xyz": xyz.numpy(), f"{name}_keys": keys.numpy(), f"{name}_order": order.numpy()})
    out["names"] = np.array(list(cases))
    path = os.path.join(GOLDEN_DIR, "morton.npz")
    np.savez_compressed(path, **out)
    return path


def run_wire_cases():
    """The scene-file arithmetic and the covariance construction, produced by torch / the reference's own functions
    on CPU tensors: torch.quantize_per_tensor(...).int_repr() and (q.int() - zp) * scale as save_npz / load_npz
    write them (scene/gaussian_model.py:528-620, :644-716), Tensor.half(), and
    strip_symmetric(L @ L^T) with L = build_scaling_rotation(s, q) from utils/general_utils.py (imported unchanged)
    as GaussianModel.get_normalized_covariance evaluates it (:55-63, :274-278)."""
    sys.path.insert(0, ref_loader.REFERENCE_ROOT)
    import utils.general_utils as gu   # the reference's module
    g = torch.Generator().manual_seed(11)
    out = {}
    x = torch.randn(20011, generator=g) * 0.8
    x[:64] = (torch.arange(-32, 32).float() + 0.5) * 0.0123          # rounding boundaries
    x[64:70] = torch.tensor([1e9, -1e9, 3.3e38, -3.3e38, 0.0, -0.0])
    for i, (scale, zp) in enumerate([(0.0123, -7), (0.0041, 12), (1.7, 0)]):
        scale = float(np.float32(scale))
        q = torch.quantize_per_tensor(x, scale, zp, torch.qint8).int_repr()
        out[f"q{i}_scale"] = np.float32(scale)
        out[f"q{i}_zp"] = np.int32(zp)
        out[f"q{i}_q"] = q.numpy()
        out[f"q{i}_deq"] = ((q.int() - torch.tensor([zp], dtype=torch.int32)) * torch.tensor([scale])).numpy()
    out["q_x"] = x.numpy()
    out["half_x"] = (torch.randn(4099, generator=g) * 37.0).numpy()
    out["half_y"] = torch.from_numpy(out["half_x"]).half().numpy()
    n = 3001
    s_raw = torch.randn(n, 3, generator=g)
    r_raw = torch.randn(n, 4, generator=g) * 2.0
    s_act = torch.nn.functional.normalize(torch.nn.functional.relu(s_raw))
    r_act = torch.nn.functional.normalize(r_raw)
    for mod in (1.0, 0.7):
        L = gu.build_scaling_rotation(mod * s_act, r_act)
        out[f"cov_m{mod}"] = gu.strip_symmetric(L @ L.transpose(1, 2)).numpy()
    out["cov_s_raw"], out["cov_r_raw"] = s_raw.numpy(), r_raw.numpy()
    out["cov_s_act"], out["cov_r_act"] = s_act.numpy(), r_act.numpy()
    np.savez_compressed(os.path.join(GOLDEN_DIR, "wire.npz"), **out)
    print("wire.npz written")


def main():
    if not ref_loader.available():
        raise SystemExit("reference checkout not present; fixtures can only be generated in the build container")
    if "--wire-only" in sys.argv:
        run_wire_cases()
        return
    if "--morton-only" in sys.argv:
        p = run_morton_cases()
        print("wrote", p, os.path.getsize(p) // 1024, "KiB")
        return
    if "--rot-scale-only" in sys.argv:
        p = run_rot_scale_case()
        print("wrote", p, os.path.getsize(p) // 1024, "KiB")
        return
    if "--join-only" not in sys.argv:
        for case in CASES:
            p = run_case(*case)
            print("wrote", p, os.path.getsize(p) // 1024, "KiB")
    p = run_join_cases()
    print("wrote", p, os.path.getsize(p) // 1024, "KiB")
    p = run_rot_scale_case()
    print("wrote", p, os.path.getsize(p) // 1024, "KiB")
    p = run_morton_cases()
    print("wrote", p, os.path.getsize(p) // 1024, "KiB")
    run_wire_cases()


if __name__ == "__main__":
    main()
