"""Per-step device time and re-rank statistics of mini-batch VectorQuantize.update_indexed on two input
distributions (A = iid N(0,1), B = clustered), colour (48-D) and shape (6-D, trace-normalised) widths."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from c3dgs_b200 import _lib  # noqa: E402
from c3dgs_b200.vq import VectorQuantize  # noqa: E402

dev = torch.device("cuda:0")
K, B = 4096, 2 ** 18
gen = torch.Generator(device=dev).manual_seed(0)


def data(kind, n, d):
    if kind == "iid":
        return torch.randn(n, d, device=dev, generator=gen)
    if kind == "clustered":
        c = torch.randn(8192, d, device=dev, generator=gen)
        return c[torch.randint(0, 8192, (n,), device=dev, generator=gen)] + 0.05 * torch.randn(n, d, device=dev, generator=gen)
    s = torch.nn.functional.normalize(torch.rand(n, 3, device=dev, generator=gen) + 0.05)
    r = torch.randn(n, 4, device=dev, generator=gen)
    return _lib.build_covariance(s, r)


for kind, d, norm in [("iid", 48, False), ("clustered", 48, False), ("iid", 45, False), ("cov", 6, True)]:
    n = 3_600_000
    x = data(kind, n, d)
    w = torch.rand(n, device=dev, generator=gen) ** 4
    m = VectorQuantize(d, K, 0.8).to(dev)
    m.uniform_init(x)
    print(f"--- {kind} {d}-D, {n} rows, batch {B}")
    for step in range(12):
        batch = torch.randint(0, n, (B,), device=dev, generator=gen)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        m.update_indexed(x, w, batch, scale_normalize=norm)
        b.record()
        torch.cuda.synchronize()
        rows, cand, fb = _lib.assign_stats(dev)
        if step in (0, 1, 2, 5, 11):
            print(f"step {step:2d}: {a.elapsed_time(b):7.3f} ms   rows {rows} candidates/row {cand / max(rows, 1):6.2f} fallback rows {fb}")
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    m.assign(x)
    b.record()
    torch.cuda.synchronize()
    rows, cand, fb = _lib.assign_stats(dev)
    print(f"final assign: {a.elapsed_time(b):7.3f} ms   rows {rows} candidates/row {cand / max(rows, 1):6.2f} fallback rows {fb}")
