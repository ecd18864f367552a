"""Minimal driver for ncu: one launch each of the wire-format / glue kernels on bench-sized inputs."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from c3dgs_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
n = 5_800_000
g = torch.Generator(device=dev).manual_seed(0)
x = torch.randn(n, 45, device=dev, generator=g)
for _ in range(2):
    q = _lib.quantize_i8(x, 0.03125, 3)
    back = _lib.dequantize_i8(q, 0.03125, 3)
    cov = _lib.build_covariance(torch.rand(n, 3, device=dev, generator=g) + 0.05, torch.randn(n, 4, device=dev, generator=g),
                                activate=True)
    keep = torch.rand(n, device=dev, generator=g) < 0.6
    out = _lib.compact([x], keep)
    half = _lib.convert(x, torch.float16)
torch.cuda.synchronize()
print("ok")
