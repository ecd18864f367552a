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
SOURCES = ["api.cu", "assign_exact.cu", "assign_tc.cu", "update.cu", "gather.cu", "rot_scale.cu", "morton.cu",
           "wire.cu", "preprocess.cu", "preprocess_bwd.cu"]
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


def _run(cmd):
    res = subprocess.run(cmd, capture_output=True, text=True)
    if res.returncode != 0:
        raise RuntimeError("nvcc failed:\n" + " ".join(cmd) + "\n" + res.stdout + res.stderr)
    return res


def _compile_objects(sources, objdir, extra, verbose=False):
    """One object per source, rebuilt only when the source, a header or the flags are newer; sources compile
    in parallel (assign_tc.cu alone takes minutes, the rest seconds)."""
    from concurrent.futures import ThreadPoolExecutor

    os.makedirs(objdir, exist_ok=True)
    # the image exports CC/CXX wrappers that lack some spec files; let nvcc pick the distro g++
    ccbin = ["-ccbin", "/usr/bin/g++"] if os.path.exists("/usr/bin/g++") else []
    headers = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))] + \
              [os.path.join(HERE, "..", "include", "c3vq.h"), os.path.abspath(__file__)]
    hdr_t = max(os.path.getmtime(h) for h in headers)
    flags = [f for f in NVCC_FLAGS if f != "-shared"] + extra + (["-Xptxas", "-v"] if verbose else [])
    jobs, objs = [], []
    for src in sources:
        obj = os.path.join(objdir, src.replace(".cu", ".o"))
        objs.append(obj)
        srcp = os.path.join(CSRC, src)
        if not os.path.exists(obj) or os.path.getmtime(obj) < max(hdr_t, os.path.getmtime(srcp)):
            jobs.append([nvcc()] + ccbin + flags + ["-c", srcp, "-o", obj])
    with ThreadPoolExecutor(max_workers=max(1, min(len(jobs), os.cpu_count() or 1))) as ex:
        for res in ex.map(_run, jobs):
            if verbose:
                sys.stderr.write(res.stderr)
    return objs, ccbin


def build(force: bool = False, verbose: bool = False) -> str:
    if not force and not stale():
        return LIB
    objdir = os.path.join(HERE, "_obj")
    if force and os.path.isdir(objdir):
        shutil.rmtree(objdir)
    objs, ccbin = _compile_objects(SOURCES, objdir, [], verbose)
    _run([nvcc()] + ccbin + NVCC_FLAGS + objs + ["-o", LIB])
    return LIB


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))


def build_dev(defines, out_name="libc3vq_dev.so") -> str:
    """Developer build with extra -D flags (e.g. C3VQ_TC_SINGLE, C3VQ_TC_PROF); load it with C3VQ_LIB=<path>."""
    out = os.path.join(HERE, out_name)
    objdir = os.path.join(HERE, "_obj_" + out_name.replace(".so", ""))
    if os.path.isdir(objdir):
        shutil.rmtree(objdir)
    objs, ccbin = _compile_objects(SOURCES, objdir, ["-D" + d for d in defines])
    _run([nvcc()] + ccbin + NVCC_FLAGS + objs + ["-o", out])
    return out
