"""Decode gather timings (codebook[idx], 5.8 M rows): 45-D (k_gather_span), 48-D from the codebook and from a large
table (k_gather_span4), with torch indexing beside them.  python tools/gather_probe.py"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from c3dgs_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(0)
n = 5_800_000


def timed(fn, reps=10):
    fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


for d, rows in ((45, 4096), (45, 294096), (48, 4096), (48, 294096), (7, 4096), (6, 4096), (13, 4096), (27, 512)):
    tab = torch.randn(rows, d, device=dev, generator=g)
    idx = torch.randint(0, rows, (n,), device=dev, generator=g)
    out = _lib.gather(tab, idx)
    assert torch.equal(out, tab[idx])
    t = timed(lambda: _lib.gather(tab, idx))
    t_ref = timed(lambda: tab[idx])
    gb = n * (8 + 4 * d) / 1e9
    lib = _lib.load()
    t_plain = timed(lambda: lib.c3vq_gather(tab.data_ptr(), rows, d, idx.data_ptr(), n, out.data_ptr(), None))
    ws = lib.c3vq_gather_workspace_bytes(rows, d, n)
    print(f"d={d} table rows {rows}: {t:.3f} ms = {gb / t:.2f} TB/s (torch indexing {t_ref:.3f} ms; "
          f"c3vq_gather without scratch {t_plain:.3f} ms; scratch {ws} B)")
