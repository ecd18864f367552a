"""Build libc3vq.so in-tree with nvcc for sm_100a (the only target).

    python -m c3dgs_b200.build [--force]

The library is plain CUDA C++ behind an extern "C" ABI (include/c3vq.h); it links no
torch symbols, so the build is one nvcc command and takes seconds.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
LIB = os.path.join(HERE, "libc3vq.so")
SOURCES = ["api.cu", "assign_exact.cu", "assign_tc.cu", "update.cu", "gather.cu", "rot_scale.cu", "morton.cu"]
NVCC_FLAGS = [
    "-gencode", "arch=compute_100a,code=sm_100a",
    "-O3", "-std=c++17", "-lineinfo",
    "-Xcompiler", "-fPIC", "-Xcompiler", "-fvisibility=hidden",
    "--expt-relaxed-constexpr",
    "-shared",
]


def nvcc() -> str:
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    raise RuntimeError("nvcc not found; libc3vq.so cannot be built")


def stale() -> bool:
    if not os.path.exists(LIB):
        return True
    t = os.path.getmtime(LIB)
    deps = [os.path.join(CSRC, f) for f in os.listdir(CSRC)] + [os.path.join(HERE, "..", "include", "c3vq.h")]
    return any(os.path.getmtime(p) > t for p in deps)


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    cmd = [nvcc()] + NVCC_FLAGS + (["-Xptxas", "-v"] if verbose else []) + \
          [os.path.join(CSRC, s) for s in SOURCES] + ["-o", LIB]
    env = dict(os.environ)
    # the image exports CC/CXX wrappers that lack some spec files; let nvcc pick the distro g++
    ccbin = "/usr/bin/g++" if os.path.exists("/usr/bin/g++") else None
    if ccbin:
        cmd[1:1] = ["-ccbin", ccbin]
    res = subprocess.run(cmd, env=env, capture_output=True, text=True)
    if verbose:
        sys.stderr.write(res.stderr)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))


def build_dev(defines, out_name="libc3vq_dev.so") -> str:
    """Developer build with extra -D flags (e.g. C3VQ_TC_SINGLE, C3VQ_TC_PROF); load it with C3VQ_LIB=<path>."""
    out = os.path.join(HERE, out_name)
    cmd = [nvcc(), "-ccbin", "/usr/bin/g++"] + NVCC_FLAGS + ["-D" + d for d in defines] + \
          [os.path.join(CSRC, s) for s in SOURCES] + ["-o", out]
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + res.stdout + res.stderr)
    return out
