"""Morton order timing and a stability / sortedness check at full size: python tools/morton_probe.py [rows]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c3dgs_b200 import _lib  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 5_800_000
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev).manual_seed(5)
xyz = torch.randn(n, 3, device=dev, generator=g) * torch.tensor([3.0, 1.0, 7.0], device=dev)
xyz[: n // 50] = xyz[n // 2: n // 2 + n // 50]          # duplicated points: equal keys, stability matters
order, keys = _lib.morton_order(xyz, return_keys=True)
torch.cuda.synchronize()
assert bool((keys[1:] >= keys[:-1]).all()), "keys not sorted"
assert bool(torch.equal(torch.sort(order).values, torch.arange(n, device=dev))), "not a permutation"
same = keys[1:] == keys[:-1]
assert bool((order[1:][same] > order[:-1][same]).all()), "not stable"
# against torch's stable sort of the same keys (keys in row order = keys of the inverse permutation)
krow = torch.empty_like(keys)
krow[order] = keys
ref = torch.sort(krow, stable=True).indices
assert bool(torch.equal(ref, order)), "differs from torch's stable sort"


def timed(fn, reps=10):
    fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


t = timed(lambda: _lib.morton_order(xyz))
t_sort = timed(lambda: torch.sort(krow))
print(f"morton_order {n} rows: {t:.3f} ms (workspace allocation included); torch.sort of the same int64 keys: {t_sort:.3f} ms; checks ok")
