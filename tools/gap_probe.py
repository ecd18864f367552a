"""Where does the time between the end of a fused assign+sums launch and the next stream operation go at 12-24 M rows?
(bench.py's strong-scaling block sees ~2 ms between two back-to-back event records there, nothing at 5.8 M rows.)"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from c3dgs_b200 import _lib  # noqa: E402
from bench import synth, init_codebook  # noqa: E402

dev = torch.device("cuda:0")
K = 4096
for n in (5_800_000, 12_000_000, 24_000_000):
    x, w = synth(n, 45, 1, device=dev)
    wn = (w / w.max()).contiguous()
    cb = init_codebook(x, K, 99)
    acc = torch.zeros(K, _lib.acc_stride(45), device=dev)
    idx = torch.empty(n, dtype=torch.int64, device=dev)
    for mode, kw in (("fused", dict(tensor_cores="fused")), ("sorted", dict(tensor_cores="sorted")),
                     ("fused, no out_idx", dict(tensor_cores="fused"))):
        for rep in range(3):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
            acc.zero_()
            ev[0].record()
            if "no out_idx" in mode:
                _lib.assign_accumulate(x, wn, cb, acc=acc, return_idx=False, **kw)
            else:
                _lib.assign_accumulate(x, wn, cb, acc=acc, out_idx=idx, **kw)
            ev[1].record()
            ev[2].record()
            acc.add_(0.0)                      # any small kernel
            ev[3].record()
            torch.cuda.synchronize()
        print(f"n={n} {mode:18s}: launch {ev[0].elapsed_time(ev[1]):8.3f} ms, event->event {ev[1].elapsed_time(ev[2]):7.3f} ms, "
              f"small kernel {ev[2].elapsed_time(ev[3]):7.3f} ms")
    del x, w, wn, idx
