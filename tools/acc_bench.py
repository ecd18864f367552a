"""Weighted centroid sums: atomic (red.global) vs sorted (counting sort + segmented sum) path.
python tools/acc_bench.py [rows] [dim]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c3dgs_b200 import _lib  # noqa: E402
from bench import synth, init_codebook  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_800_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 45
K = 4096
dev = torch.device("cuda", 0)
x, w = synth(n, d, 1234, device=dev)
wn = (w / w.max()).contiguous()
cb = init_codebook(x, K, 99)
_, idx = _lib.assign(x, cb)
acc_a = torch.zeros(K, _lib.acc_stride(d), device=dev)
acc_s = torch.zeros_like(acc_a)
for name, sp, acc in (("atomic", False, acc_a), ("sorted", True, acc_s)):
    for it in range(4):
        acc.zero_()
        a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
        a.record()
        _lib.accumulate(x, wn, idx, K, acc=acc, sorted_path=sp)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b)
    gbs = n * (4 * d + 4 + 8) / ms / 1e6
    print(f"{name}: {ms:.3f} ms for {n} rows x {d}  -> {gbs:.0f} GB/s algorithmic (4d+4+8 B/row)")
ref = torch.zeros(K, d + 1, device=dev, dtype=torch.float64)
ref.index_add_(0, idx, torch.cat([x * wn[:, None], wn[:, None]], 1).double())
for name, acc in (("atomic", acc_a), ("sorted", acc_s)):
    err = ((acc[:, :d + 1].double() - ref).abs().max(1).values / (ref.abs().max(1).values + 1e-30)).max().item()
    print(f"{name}: max row-normalised error vs fp64 = {err:.2e}")
