"""Experiment: does cudaLimitMaxL2FetchGranularity = 32 B reduce the over-fetch of the row gather in k_seg_sum
(180-byte rows fetched at 64-byte granularity: 1.55 GB of DRAM reads for 1.09 GB of rows)?"""
import ctypes
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from c3dgs_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
x, w = bench.synth(5_800_000, 45, 0, device=dev)
cb = bench.init_codebook(x, 4096, 1)
ei = torch.zeros(4096, device=dev)
for _ in range(3):
    _lib.update(x, w, cb, ei, 0.8, 1e-5)
_, idx = _lib.assign(x, cb)
rt = ctypes.CDLL("libcudart.so.12")
LIMIT = 0x05          # cudaLimitMaxL2FetchGranularity


def timed(fn, reps=20):
    fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for gran in (None, 32, 64, 128, 32):
    if gran is not None:
        rc = rt.cudaDeviceSetLimit(LIMIT, ctypes.c_size_t(gran))
        got = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(got), LIMIT)
        label = f"limit {gran} (rc {rc}, now {got.value})"
    else:
        got = ctypes.c_size_t(0)
        rt.cudaDeviceGetLimit(ctypes.byref(got), LIMIT)
        label = f"default ({got.value})"
    t_sum = timed(lambda: _lib.accumulate(x, w, idx, 4096, sorted_path=True))
    t_asg = timed(lambda: _lib.assign(x, cb, return_dist=False), reps=5)
    t_gather = timed(lambda: _lib.gather(cb, idx))
    print(f"{label}: accumulate_sorted {t_sum:.3f} ms, assign {t_asg:.3f} ms, gather {t_gather:.3f} ms")
