"""CPU: the C-ABI library loads and exports every symbol include/c3vq.h declares; host-only entry
points behave; the Python host fails loudly without CUDA (no fallback); the product never touches
oracle/."""
import ctypes
import os
import re

import pytest
import torch

from conftest import ROOT


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "c3vq.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(c3vq_[a-z0-9_]+)\s*\(", text)))


def test_header_symbols_all_exported(lib):
    syms = _declared_symbols()
    assert len(syms) >= 14
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/c3vq.h but not exported by libc3vq.so"


def test_python_signature_table_matches_header(lib):
    from c3dgs_b200 import _lib
    assert sorted(_lib.SIGNATURES) == _declared_symbols()


def test_host_only_entry_points(lib):
    assert lib.c3vq_version() >= 100
    for d, want in [(6, 8), (7, 8), (45, 48), (48, 52), (3, 4), (63, 64)]:
        assert lib.c3vq_acc_stride(d) == want
    small = lib.c3vq_update_workspace_bytes(1000, 45, 4096)
    big = lib.c3vq_update_workspace_bytes(1000000, 45, 4096)
    assert 0 < small <= big
    assert lib.c3vq_assign_workspace_bytes(0, 45, 4096) >= 0


def test_argument_errors_are_reported_not_thrown(lib):
    rc = lib.c3vq_assign_exact(None, None, 10, 45, None, 16, None, None, None)
    assert rc == -1 and b"null" in lib.c3vq_last_error()
    rc = lib.c3vq_assign_exact(None, None, 0, 1000, None, 16, None, None, None)
    assert rc == -1 and b"d=1000" in lib.c3vq_last_error()
    rc = lib.c3vq_ema(None, None, None, 4, 0, 0.8, 1e-5, 0, None)
    assert rc == -1


def test_cpu_tensors_raise(lib):
    from c3dgs_b200 import _lib
    from c3dgs_b200.weighted_distance._C import weightedDistance
    x = torch.zeros(4, 6)
    cb = torch.zeros(3, 6)
    with pytest.raises(RuntimeError, match="CUDA"):
        weightedDistance(x, cb)
    with pytest.raises(RuntimeError, match="CUDA"):
        _lib.gather(cb, torch.zeros(2, dtype=torch.int64))
    from c3dgs_b200 import vq_features
    with pytest.raises(RuntimeError, match="CUDA"):
        vq_features(x, torch.ones(4), 4, steps=1, silent=True)


def test_missing_library_is_loud(monkeypatch, tmp_path):
    from c3dgs_b200 import _lib
    monkeypatch.setattr(_lib, "_lib", None)
    monkeypatch.setattr(_lib, "LIB_PATH", str(tmp_path / "nope.so"))
    with pytest.raises(RuntimeError, match="no fallback"):
        _lib.load()


def test_install_registers_reference_module_names(lib):
    import sys
    import c3dgs_b200
    saved = {k: sys.modules.get(k) for k in ("weighted_distance", "weighted_distance._C")}
    try:
        c3dgs_b200.install()
        from weighted_distance._C import weightedDistance  # the import line of compression/vq.py:12
        assert weightedDistance is c3dgs_b200.weighted_distance._C.weightedDistance
    finally:
        for k, v in saved.items():
            if v is None:
                sys.modules.pop(k, None)
            else:
                sys.modules[k] = v


def test_reference_signatures_are_mirrored():
    import inspect
    from c3dgs_b200 import VectorQuantize, vq_features
    sig = inspect.signature(vq_features)
    assert list(sig.parameters) == ["features", "importance", "codebook_size", "vq_chunk", "steps", "decay",
                                    "scale_normalize", "silent"]                      # vq.py:49-58
    assert sig.parameters["vq_chunk"].default == 2 ** 16 and sig.parameters["steps"].default == 1000
    assert sig.parameters["decay"].default == 0.8 and sig.parameters["scale_normalize"].default is False
    init = inspect.signature(VectorQuantize.__init__)
    assert list(init.parameters) == ["self", "channels", "codebook_size", "decay"]   # vq.py:16
    assert init.parameters["codebook_size"].default == 2 ** 12 and init.parameters["decay"].default == 0.5
    vq = VectorQuantize(7, 16)
    assert vq.eps == 1e-5 and vq.codebook.shape == (16, 7) and vq.entry_importance.shape == (16,)
    assert not vq.codebook.requires_grad and not vq.entry_importance.requires_grad
    assert set(vq.state_dict()) == {"codebook", "entry_importance"}


def test_product_never_imports_the_oracle():
    """oracle/ is test infrastructure: nothing under c3dgs_b200/ may reference it."""
    pkg = os.path.join(ROOT, "c3dgs_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp")):
                text = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", text, flags=re.M), f
                assert "libc3vq_oracle" not in text and "c3vq_oracle" not in text, f


def test_sass_targets_sm100a_only(lib):
    import subprocess
    from c3dgs_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs
