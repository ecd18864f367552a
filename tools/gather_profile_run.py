"""Minimal driver for ncu: the decode gather codebook[idx] for 5.8M rows x 45-D (k_gather_span) and x 48-D (float4)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from c3dgs_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
n = 5_800_000
idx = torch.randint(0, 4096, (n,), device=dev, generator=g)
for d in (45, 48, 27):
    cb = torch.randn(4096, d, device=dev, generator=g)
    for _ in range(3):
        out = _lib.gather(cb, idx)
torch.cuda.synchronize()
print("ok")
