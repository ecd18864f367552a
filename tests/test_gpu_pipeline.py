"""GPU (-m gpu): the callers and data formats either side of the clustering path through
the C ABI — int8 scene-file quantisation, float16 / int32 conversions, normalised covariance, mask compaction, the
SplatScene file round trip and compress_gaussians end to end.

Bar: bit-exact for the integer / byte / copy work (quantised bytes against torch's own CUDA quantiser — the call the
reference's save_npz makes — and against the oracle); 1e-6 absolute on covariance entries (|entries| <= 1, float32
sum order); the clustering inside compress_gaussians is checked through its fixed point (final indices == exact
argmin of the returned codebook) because uniform_init draws from the device generator."""
import io
import os

import numpy as np
import pytest
import torch

from conftest import ROOT

pytestmark = pytest.mark.gpu

DEV = "cuda:0"


@pytest.fixture(scope="module")
def L(lib):
    from c3dgs_b200 import _lib
    assert torch.cuda.is_available()
    return _lib


def _wire():
    return np.load(os.path.join(ROOT, "tests", "golden", "wire.npz"))


def _t(a):
    return torch.from_numpy(np.ascontiguousarray(a)).to(DEV)


# ---------------------------------------------------------------- int8 quantisation ----------------
@pytest.mark.parametrize("scale,zp", [(0.0123, -7), (0.0041, 12), (1.7, 0), (3.0e-4, -128), (0.25, 127)])
@pytest.mark.parametrize("n", [1, 3, 4, 1001, 2_000_003])
def test_quantize_bit_exact_vs_torch_cuda_and_oracle(L, oracle, scale, zp, n):
    scale = float(np.float32(scale))
    g = torch.Generator().manual_seed(n)
    x = torch.randn(n, generator=g) * (60 * scale)
    m = min(n, 300)
    x[:m] = ((torch.arange(m) - m // 2).float() + 0.5) * scale           # rounding boundaries
    if n > 400:
        x[300:308] = torch.tensor([1e9, -1e9, 3.3e38, -3.3e38, float("inf"), float("-inf"), 0.0, -0.0])
    xd = x.to(DEV)
    q = L.quantize_i8(xd, scale, zp)
    assert q.dtype == torch.int8 and q.shape == xd.shape
    ref = torch.quantize_per_tensor(xd, scale, zp, torch.qint8).int_repr()          # the reference's call, on the GPU
    assert torch.equal(q, ref), f"{(q != ref).sum().item()} of {n} bytes differ from torch's CUDA quantiser"
    assert np.array_equal(q.cpu().numpy(), oracle.quantize_i8(x.numpy(), scale, zp))


def test_quantize_non_float32_scale_tensor_qparams_and_shapes(L, oracle):
    g = torch.Generator().manual_seed(5)
    x = torch.randn(777, 15, 3, generator=g)
    xd = x.to(DEV)
    # a scale that is not a float32 value takes the double expression for every element
    q = L.quantize_i8(xd, 0.1, 3)
    assert torch.equal(q, torch.quantize_per_tensor(xd, 0.1, 3, torch.qint8).int_repr())
    assert np.array_equal(q.cpu().numpy(), oracle.quantize_i8(x.numpy(), 0.1, 3).reshape(x.shape))
    # FakeQuantize keeps scale / zero_point as one-element tensors (gaussian_model.py:531-533)
    s_t, z_t = torch.tensor([0.02], device=DEV), torch.tensor([-5], dtype=torch.int32, device=DEV)
    assert torch.equal(L.quantize_i8(xd, s_t, z_t), torch.quantize_per_tensor(xd, s_t, z_t, torch.qint8).int_repr())
    # a view that starts in the middle of a storage
    v = xd.reshape(-1)[1:]
    assert torch.equal(L.quantize_i8(v, 0.02, -5), torch.quantize_per_tensor(v, 0.02, -5, torch.qint8).int_repr())
    assert L.quantize_i8(torch.empty(0, 3, device=DEV), 0.1, 0).shape == (0, 3)
    with pytest.raises(RuntimeError):
        L.quantize_i8(xd, 0.0, 0)
    with pytest.raises(RuntimeError, match="CUDA"):
        L.quantize_i8(x, 0.1, 0)


def test_quantize_golden_and_dequantize_bit_exact(L, oracle):
    g = _wire()
    x = _t(g["q_x"])
    for i in range(3):
        scale, zp = float(g[f"q{i}_scale"]), int(g[f"q{i}_zp"])
        q = L.quantize_i8(x, scale, zp)
        assert np.array_equal(q.cpu().numpy(), oracle.quantize_i8(g["q_x"], scale, zp))
        deq = L.dequantize_i8(_t(g[f"q{i}_q"]), scale, zp)
        assert np.array_equal(deq.cpu().numpy(), g[f"q{i}_deq"])                   # torch's (q.int() - zp) * scale
    q = torch.randint(-128, 128, (1_000_003,), dtype=torch.int8, device=DEV)
    s_t, z_t = torch.tensor([0.0173], device=DEV), torch.tensor([9], dtype=torch.int32, device=DEV)
    assert torch.equal(L.dequantize_i8(q, s_t, z_t), (q.int() - z_t) * s_t)        # load_npz's expression


def test_quantize_round_trip_error_bound(L):
    x = torch.rand(3_000_000, device=DEV) * 2 - 1
    scale, zp = 2.0 / 255.0, 0
    scale = float(np.float32(scale))
    back = L.dequantize_i8(L.quantize_i8(x, scale, zp), scale, zp)
    inside = (x / scale).abs() < 127
    assert ((back - x).abs()[inside] <= scale * 0.5 * (1 + 1e-6)).all()


# ---------------------------------------------------------------- conversions ----------------
@pytest.mark.parametrize("n", [0, 1, 5, 4099, 1_000_001])
def test_conversions_bit_exact(L, n):
    g = _wire()
    x = torch.randn(n, 3, device=DEV) * 100
    h = L.convert(x, torch.float16)
    assert h.dtype == torch.float16 and torch.equal(h, x.half())
    assert torch.equal(L.convert(h, torch.float32), h.float())
    idx = torch.randint(0, 2 ** 31 - 1, (n,), device=DEV)
    i32 = L.convert(idx, torch.int32)
    assert i32.dtype == torch.int32 and torch.equal(i32, idx.int())
    assert torch.equal(L.convert(i32, torch.int64), idx)
    if n == 4099:
        assert np.array_equal(L.convert(_t(g["half_x"]), torch.float16).cpu().numpy(), g["half_y"])
    with pytest.raises(RuntimeError):
        L.convert(x, torch.int32)


# ---------------------------------------------------------------- covariance ----------------
def test_build_covariance_vs_reference_golden(L, oracle):
    g = _wire()
    for mod in (1.0, 0.7):
        ref = g[f"cov_m{mod}"]
        a = L.build_covariance(_t(g["cov_s_act"]), _t(g["cov_r_act"]), mod).cpu().numpy()
        b = L.build_covariance(_t(g["cov_s_raw"]), _t(g["cov_r_raw"]), mod, activate=True).cpu().numpy()
        assert a.shape == ref.shape
        assert np.abs(a - ref).max() <= 1e-6 and np.abs(b - ref).max() <= 1e-6
    with pytest.raises(RuntimeError):
        L.build_covariance(_t(g["cov_s_act"]), _t(g["cov_r_act"][:-1]))
    assert L.build_covariance(torch.empty(0, 3, device=DEV), torch.empty(0, 4, device=DEV)).shape == (0, 6)


def test_build_covariance_large_properties(L, oracle):
    n = 1_000_003
    s = torch.randn(n, 3, device=DEV)
    r = torch.randn(n, 4, device=DEV)
    cov = L.build_covariance(s, r, activate=True)
    # trace of the normalised covariance = |normalize(relu(s))|^2 = 1 unless relu zeroed the whole row
    live = (s > 0).any(1)
    tr = cov[:, 0] + cov[:, 3] + cov[:, 5]
    assert (tr[live] - 1).abs().max() < 1e-5 and (tr[~live] == 0).all()
    sl = slice(0, 50_000)
    ref = oracle.normalized_covariance(s[sl].cpu().numpy(), r[sl].cpu().numpy(), 1.0, activate=True)
    assert np.abs(cov[sl].cpu().numpy() - ref).max() <= 2e-6      # |entries| <= 1; ~10 float32 roundings per entry
    # rotation + scaling recovered from it rebuild it (extract_rot_scale is the consumer, vq.py:186)
    rot, sc = L.extract_rot_scale(cov[:200_000].contiguous())
    again = L.build_covariance(sc, rot)
    assert (again - cov[:200_000]).abs().max() < 1e-5


# ---------------------------------------------------------------- mask compaction ----------------
@pytest.mark.parametrize("n,frac", [(0, 0.5), (1, 1.0), (1023, 0.3), (1024, 0.0), (1025, 1.0), (300_007, 0.6)])
def test_compact_bit_exact_vs_boolean_indexing(L, n, frac):
    g = torch.Generator().manual_seed(n)
    mask = (torch.rand(n, generator=g) < frac).to(DEV)
    tensors = [torch.randn(n, 3, device=DEV), torch.randn(n, 15, 3, device=DEV), torch.randn(n, 1, device=DEV),
               torch.randint(0, 10 ** 12, (n,), device=DEV), None, torch.randint(0, 99, (n,), device=DEV).int()]
    outs = L.compact(tensors, mask)
    for t, o in zip(tensors, outs):
        if t is None:
            assert o is None
        else:
            assert o.dtype == t.dtype and torch.equal(o, t[mask])
    with pytest.raises(RuntimeError):
        L.compact([torch.randn(n + 1, 3, device=DEV)], mask)


# ---------------------------------------------------------------- scene + pipeline ----------------
def _scene(n, seed=0, factor=True):
    from c3dgs_b200.scene import SplatScene
    g = torch.Generator().manual_seed(seed)
    centres = torch.randn(64, 48, generator=g)
    feats = centres[torch.randint(0, 64, (n,), generator=g)] + 0.05 * torch.randn(n, 48, generator=g)
    feats = feats.reshape(n, 16, 3)
    sc = SplatScene(
        xyz=(torch.randn(n, 3, generator=g) * 3).to(DEV),
        features_dc=feats[:, :1].contiguous().to(DEV), features_rest=feats[:, 1:].contiguous().to(DEV),
        opacity=torch.randn(n, 1, generator=g).to(DEV),
        scaling=(torch.rand(n, 3, generator=g) + 0.05).to(DEV), rotation=torch.randn(n, 4, generator=g).to(DEV),
        scaling_factor=(torch.randn(n, 1, generator=g) - 3).to(DEV) if factor else None)
    color_imp = (torch.rand(n, generator=g) ** 4).to(DEV)
    gauss_imp = (torch.rand(n, generator=g) ** 4).to(DEV)
    return sc, color_imp, gauss_imp


def test_compress_gaussians_end_to_end(L, oracle, monkeypatch):
    from c3dgs_b200 import CompressionSettings, compress_gaussians, splats
    seen = []
    real_extract = splats.extract_rot_scale
    monkeypatch.setattr(splats, "extract_rot_scale", lambda cov: (seen.append(cov.clone()), real_extract(cov))[1])
    n, k = 30_000, 256
    # compress_gaussians draws from the global generators like the reference (uniform_init on the device, the batch
    # indices on the CPU): seeded here so that the outcome does not depend on which tests ran before
    torch.manual_seed(20250)
    scene, cimp, gimp = _scene(n)
    orig_feats = scene.get_features.clone()
    orig_cov = scene.get_normalized_covariance().clone()
    prune = 1e-7
    alive = cimp > prune
    cc = CompressionSettings(k, 0.0, None, 0.9, steps=30, decay=0.8, batch_size=4096)
    gc = CompressionSettings(k, 0.0, 0.3, 0.9, steps=30, decay=0.8, batch_size=4096)
    compress_gaussians(scene, cimp, gimp, cc, gc, color_compress_non_dir=True, prune_threshold=prune, silent=True)
    m = int(alive.sum())
    assert scene._xyz.shape[0] == m and scene.is_color_indexed and scene.is_gaussian_indexed
    cimp_a, gimp_a, feats_a, cov_a = cimp[alive], gimp[alive], orig_feats[alive], orig_cov[alive]
    # thresholds: the colour one was derived by torch.quantile (vq.py:212)
    assert cc.importance_include == pytest.approx(torch.quantile(cimp_a, 0.9).item())
    # colour table = [codebook ; kept rows in order], indices as join_features defines them (vq.py:90-103)
    keep = cimp_a > cc.importance_include
    fi = scene._feature_indices
    table = torch.cat((scene._features_dc, scene._features_rest), 1).flatten(1)
    assert fi.dtype == torch.int64 and table.shape[0] == k + int(keep.sum())
    assert torch.equal(fi[keep], torch.arange(int(keep.sum()), device=DEV) + k)
    assert torch.equal(table[k:], feats_a[keep].flatten(1))
    assert torch.equal(scene.get_features.flatten(1)[keep], feats_a[keep].flatten(1))
    # the clustered rows carry the exact fp32 arg-min of the returned codebook (vq.py:82)
    _, idx_o = oracle.assign(feats_a[~keep].flatten(1).cpu().numpy(), table[:k].cpu().numpy())
    assert np.array_equal(fi[~keep].cpu().numpy(), idx_o)
    # shape table: trace-normalised codewords (vq.py:76-77) behind them the kept rows, then rotation + scaling of
    # every table row (vq.py:186) — `seen` holds the joined table extract_rot_scale was given
    keep_g = gimp_a > 0.3
    gi = scene._gaussian_indices
    (table_g,) = seen
    assert table_g.shape == (k + int(keep_g.sum()), 6)
    assert torch.equal(table_g[k:], cov_a[keep_g]) and torch.equal(table_g[gi][keep_g], cov_a[keep_g])
    tr = table_g[:k, 0] + table_g[:k, 3] + table_g[:k, 5]
    # x / tr(x) has trace 1 up to the rounding of its three terms: an unused codeword keeps its uniform-box draw, whose
    # diagonal can nearly cancel (seen once: 2e-4 off with |diag| sums 2000 times the trace)
    mag = table_g[:k, 0].abs() + table_g[:k, 3].abs() + table_g[:k, 5].abs()
    assert ((tr - 1).abs() <= 4e-7 * mag + 1e-6).all()
    _, idx_g = oracle.assign(cov_a[~keep_g].cpu().numpy(), table_g[:k].contiguous().cpu().numpy())
    assert np.array_equal(gi[~keep_g].cpu().numpy(), idx_g)
    rot, sc = L.extract_rot_scale(table_g)
    assert torch.equal(scene._rotation, rot) and torch.equal(scene._scaling, sc)
    rebuilt = L.build_covariance(scene._scaling, scene._rotation)            # stored scaling is already activated
    assert (rebuilt[gi][keep_g] - cov_a[keep_g]).abs().max() < 1e-5


@pytest.mark.parametrize("quantization", [True, False])
def test_scene_file_round_trip_and_format(L, oracle, quantization, tmp_path):
    from c3dgs_b200 import CompressionSettings, compress_gaussians
    from c3dgs_b200.scene import SplatScene
    scene, cimp, gimp = _scene(12_000, seed=3)
    scene.quantization = quantization
    cc = CompressionSettings(128, 0.0, 0.6, 0.9, steps=10, decay=0.8, batch_size=2048)
    gc = CompressionSettings(128, 0.0, 0.3, 0.9, steps=10, decay=0.8, batch_size=2048)
    compress_gaussians(scene, cimp, gimp, cc, gc, color_compress_non_dir=True, prune_threshold=0.0, silent=True)
    path = str(tmp_path / "scene" / "point_cloud.npz")
    xyz_before = scene._xyz.clone()
    scene.save_npz(path, sort_morton=True, half_precision=not quantization)
    # Morton order was applied to every per-Gaussian tensor (gaussian_model.py:1005-1023)
    order = L.morton_order(xyz_before)
    assert torch.equal(scene._xyz, xyz_before[order])
    sd = np.load(path)
    n, kc, kg = scene._xyz.shape[0], scene._features_dc.shape[0], scene._rotation.shape[0]
    want = {"xyz": (n, 3), "features_dc": (kc, 1, 3), "features_rest": (kc, 15, 3), "opacity": (n, 1),
            "scaling": (kg, 3), "scaling_factor": (n, 1), "rotation": (kg, 4), "feature_indices": (n,),
            "gaussian_indices": (n,)}
    for key, shape in want.items():
        assert sd[key].shape == shape, key
    assert sd["feature_indices"].dtype == np.int32 and sd["gaussian_indices"].dtype == np.int32
    assert bool(sd["quantization"]) == quantization
    if quantization:
        assert sd["xyz"].dtype == np.float16
        for key in ("features_dc", "features_rest", "opacity", "scaling", "scaling_factor", "rotation"):
            assert sd[key].dtype == np.int8
            assert sd[key + "_scale"].dtype == np.float32 and sd[key + "_zero_point"].dtype == np.int32
            stored = scene._stored(key)
            assert np.array_equal(sd[key], oracle.quantize_i8(stored.cpu().numpy(), sd[key + "_scale"][0],
                                                              sd[key + "_zero_point"][0]).reshape(sd[key].shape))
    else:
        assert sd["xyz"].dtype == np.float32 and sd["features_rest"].dtype == np.float32
        assert sd["opacity"].dtype == np.float16 and sd["rotation"].dtype == np.float16
    back = SplatScene.load_npz(path, device=DEV)
    assert torch.equal(back._feature_indices, scene._feature_indices)
    assert torch.equal(back._gaussian_indices, scene._gaussian_indices)
    if quantization:
        assert torch.equal(back._xyz, scene._xyz.half().float())
        for key in ("features_dc", "features_rest", "scaling_factor", "rotation"):
            scale = float(sd[key + "_scale"][0])
            err = (getattr(back, "_" + key) - scene._stored(key)).abs().max().item()
            assert err <= scale * 0.5 * (1 + 1e-5), (key, err, scale)
        assert (torch.sigmoid(back._opacity) - torch.sigmoid(scene._opacity)).abs().max() <= \
            float(sd["opacity_scale"][0]) * 0.5 * (1 + 1e-4) + 1e-6
        # a second save of the loaded scene reproduces the bytes (quantise o dequantise is idempotent)
        buf = io.BytesIO()
        back.save_npz(buf)
        buf.seek(0)
        sd2 = np.load(buf)
        for key in ("features_dc", "features_rest", "scaling_factor", "xyz"):        # the others are re-normalised
            assert np.array_equal(sd2[key], sd[key]), key
    else:
        assert torch.equal(back._xyz, scene._xyz)
        assert torch.equal(back._features_rest, scene._features_rest)
        assert torch.equal(back._rotation, scene._rotation.half().float())


# ---------------------------------------------------------------- reference-side interop of the scene file -----------
# tests/golden/scene_interop.npz was made by oracle/make_scene_golden.py, which imports the reference's own
# scene/gaussian_model.py on CPU and runs ITS save_npz (:505-623), load_npz (:625-740) and decode properties
# (:213-257) on a small indexed, quantisation-aware model.
def _interop():
    return np.load(os.path.join(ROOT, "tests", "golden", "scene_interop.npz"))


def _subfile(g, prefix):
    return {k[len(prefix):]: g[k] for k in g.files if k.startswith(prefix)}


def test_reference_written_scene_file_loads_into_the_same_tensors(L, tmp_path):
    """A file written by the reference's save_npz, read by SplatScene.load_npz: every tensor equals what the
    reference's own load_npz produced from it; then the indexed decode sites (get_features / get_rotation /
    get_scaling, gaussian_model.py:213-257) equal the reference's on that model."""
    from c3dgs_b200 import SplatScene
    g = _interop()
    path = str(tmp_path / "ref_written.npz")
    np.savez_compressed(path, **_subfile(g, "fileA/"))
    sc = SplatScene.load_npz(path, device=DEV)
    for name in ("xyz", "features_dc", "features_rest", "scaling", "scaling_factor", "rotation"):
        got, want = getattr(sc, "_" + name).cpu().numpy(), g["loaded/" + name]
        assert got.shape == want.shape and got.dtype == want.dtype
        assert np.array_equal(got, want), name                        # (q - zero_point) * scale in float32: exact
    assert np.allclose(sc._opacity.cpu().numpy(), g["loaded/opacity"], rtol=3e-6, atol=1e-6)   # logit: libm vs CUDA log
    for name in ("feature_indices", "gaussian_indices"):
        got = getattr(sc, "_" + name)
        assert got.dtype == torch.int64 and np.array_equal(got.cpu().numpy(), g["loaded/" + name])
    assert sc.quantization and sc.is_color_indexed and sc.is_gaussian_indexed
    assert np.array_equal(sc.get_features.cpu().numpy(), g["decoded/features"])                 # pure gather
    assert np.allclose(sc.get_rotation.cpu().numpy(), g["decoded/rotation"], rtol=2e-6, atol=1e-7)
    scaling = torch.exp(sc._scaling_factor) * sc.get_scaling_normalized                         # :213-222
    assert np.allclose(scaling.cpu().numpy(), g["decoded/scaling"], rtol=4e-6, atol=1e-9)


@pytest.mark.parametrize("sort_morton", [False, True])
def test_scene_file_written_here_holds_the_reference_files_arrays(L, tmp_path, sort_morton):
    """The other direction: SplatScene.save_npz from the same tensors and (scale, zero_point) pairs writes the same
    keys, dtypes, shapes and values as the reference's save_npz did (so the reference's load_npz, pinned above on
    its own file, reads it the same way).  The int8 values go through sigmoid / normalize evaluated on the GPU here
    and on the CPU there, so a value on a rounding boundary may land one step apart (bounded at 0.5 % of a tensor)."""
    from c3dgs_b200 import SplatScene
    g = _interop()
    want = _subfile(g, "fileB/" if sort_morton else "fileA/")
    t = {k[3:]: torch.from_numpy(g[k]).to(DEV) for k in g.files if k.startswith("in/") and not k.endswith(("_scale", "_zero_point"))}
    qp = {n: (float(g[f"in/{n}_scale"][0]), int(g[f"in/{n}_zero_point"][0]))
          for n in ("features_dc", "features_rest", "opacity", "scaling", "scaling_factor", "rotation")}
    sc = SplatScene(t["xyz"], t["features_dc"], t["features_rest"], t["opacity"], t["scaling"], t["rotation"],
                    scaling_factor=t["scaling_factor"], quantization=True, qparams=qp)
    sc._feature_indices, sc._gaussian_indices = t["feature_indices"], t["gaussian_indices"]
    path = str(tmp_path / "ours.npz")
    sc.save_npz(path, sort_morton=sort_morton)
    got = dict(np.load(path))
    assert sorted(got) == sorted(want)
    for k in sorted(want):
        assert got[k].dtype == want[k].dtype and got[k].shape == want[k].shape, (k, got[k].dtype, got[k].shape)
        if want[k].dtype == np.int8:
            diff = np.abs(got[k].astype(np.int32) - want[k].astype(np.int32))
            assert diff.max() <= 1 and (diff != 0).mean() <= 5e-3, (k, int(diff.max()), float((diff != 0).mean()))
        else:
            assert np.array_equal(got[k], want[k]), k


def test_out_of_range_index_arrays_are_rejected_where_they_enter_the_scene(L, tmp_path):
    """The decode gathers do not check indices (an out-of-range one would decode to zeros where torch indexing
    raises): SplatScene validates index arrays once, in set_*_indexed and load_npz."""
    from c3dgs_b200 import SplatScene
    g = _interop()
    t = {k[3:]: torch.from_numpy(g[k]).to(DEV) for k in g.files if k.startswith("in/") and not k.endswith(("_scale", "_zero_point"))}
    sc = SplatScene(t["xyz"], t["features_dc"], t["features_rest"], t["opacity"], t["scaling"], t["rotation"],
                    scaling_factor=t["scaling_factor"])
    bad = t["feature_indices"].clone()
    bad[3] = t["features_dc"].shape[0]
    with pytest.raises(IndexError):
        sc.set_color_indexed(torch.cat((t["features_dc"], t["features_rest"]), 1), bad)
    bad = t["gaussian_indices"].clone()
    bad[0] = -1
    with pytest.raises(IndexError):
        sc.set_gaussian_indexed(t["rotation"], t["scaling"], bad)
    corrupt = _subfile(g, "fileA/")
    corrupt["gaussian_indices"] = corrupt["gaussian_indices"].copy()
    corrupt["gaussian_indices"][5] = 10_000
    path = str(tmp_path / "corrupt.npz")
    np.savez_compressed(path, **corrupt)
    with pytest.raises(IndexError, match="gaussian_indices"):
        SplatScene.load_npz(path, device=DEV)
