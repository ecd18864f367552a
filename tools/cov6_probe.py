"""The trace-normalised 6-D shape codebook (BASELINE configs[2], fork reality) iteration by iteration: time and
screening statistics of every fused assignment launch.  python tools/cov6_probe.py [rows] [iters]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c3dgs_b200 import _lib  # noqa: E402
from bench import init_codebook  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_800_000
iters = int(sys.argv[2]) if len(sys.argv) > 2 else 10
K = 4096
dev = torch.device("cuda", 0)
gen = torch.Generator(device=dev).manual_seed(5)
s3 = torch.nn.functional.normalize(torch.rand(n, 3, device=dev, generator=gen) + 0.05)
r4 = torch.randn(n, 4, device=dev, generator=gen)
cov = _lib.build_covariance(s3, r4)
w = torch.rand(n, device=dev, generator=gen) ** 4
wn = (w / w.max()).contiguous()
cb = init_codebook(cov, K, 99)
ei = torch.zeros(K, device=dev)
acc = torch.zeros(K, _lib.acc_stride(6), device=dev)
for it in range(iters):
    acc.zero_()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    _lib.assign_accumulate(cov, wn, cb, acc=acc)
    b.record()
    _lib.ema(cb, ei, acc, 0.8, 1e-5, scale_normalize=True)
    torch.cuda.synchronize()
    tr = cb[:, 0] + cb[:, 3] + cb[:, 5]
    n2 = (cb * cb).sum(1)
    print(f"iter {it}: assign_accumulate {a.elapsed_time(b):.3f} ms", _lib.assign_stats(),
          f"codeword |c|^2 median {n2.median().item():.3g} max {n2.max().item():.3g}, live {(ei > 0).sum().item()}")
