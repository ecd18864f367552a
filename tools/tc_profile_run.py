"""Minimal driver for ncu: a few fused assign+sums launches on synthetic rows.
python tools/tc_profile_run.py [rows] [dim] [reps]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c3dgs_b200 import _lib  # noqa: E402
from bench import synth, init_codebook  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 45
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
dev = torch.device("cuda", 0)
x, w = synth(n, d, 1234, device=dev)
wn = (w / w.max()).contiguous()
cb = init_codebook(x, 4096, 99)
ei = torch.zeros(4096, device=dev)
acc = torch.zeros(4096, _lib.acc_stride(d), device=dev)
for it in range(reps):
    acc.zero_()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    _lib.assign_accumulate(x, wn, cb, acc=acc)
    b.record()
    _lib.ema(cb, ei, acc, 0.8, 1e-5)
    torch.cuda.synchronize()
    print(f"iter {it}: assign_accumulate {a.elapsed_time(b):.3f} ms for {n} rows x {d}", _lib.assign_stats())
