"""Centroid sums under different codeword-usage distributions: assignment alone, assignment + sorted sums,
assignment with the sums fused (one bulk reduce-add per row), and the stand-alone sum kernels on the same indices.
The bench workload starts from a uniform-box codebook, so its first iterations put most rows on a handful of
codewords; a codebook sampled from the rows uses every codeword about equally.
python tools/sums_probe.py [rows] [dim]"""
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


def timed(fn, reps=4):
    fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


books = {}
cb = init_codebook(x, K, 99)
ei = torch.zeros(K, device=dev)
acc = torch.zeros(K, _lib.acc_stride(d), device=dev)
for it in range(10):
    if it in (0, 2, 4, 6, 9):
        books[f"bench iteration {it}"] = cb.clone()
    acc.zero_()
    _lib.assign_accumulate(x, wn, cb, acc=acc)
    _lib.ema(cb, ei, acc, 0.8, 1e-5)
g = torch.Generator(device=dev).manual_seed(7)
books["sampled from the rows"] = x[torch.randint(0, n, (K,), device=dev, generator=g)].clone()

for name, book in books.items():
    _, idx = _lib.assign(x, book, return_dist=False)
    rows, cands, fb = _lib.assign_stats(dev)
    h = torch.bincount(idx, minlength=K).sort(descending=True).values.double() / n
    print(f"== {name}: live {int((h > 0).sum())}, top-1 {h[0]:.3f}, top-16 {h[:16].sum():.3f}, "
          f"top-148 {h[:148].sum():.3f}; candidates/row {cands / max(rows, 1):.2f}, fallback rows {fb}")
    t_assign = timed(lambda: _lib.assign(x, book, return_dist=False))
    res = {}
    for label, mode in (("sorted", True), ("fused", "fused")):
        def run():
            acc.zero_()
            _lib.assign_accumulate(x, wn, book, acc=acc, tensor_cores=mode)
        res[label] = timed(run)
    # accuracy of the sums (as centroids sum_x / (sum_w + eps), row-normalised) against float64 accumulation of the
    # same float32 products: fused bulk reduce-adds with warp combining, sorted segmented sum, plain atomics
    prod = (x * wn[:, None]).double()
    ref_x = torch.zeros(K, d, device=dev, dtype=torch.float64).index_add_(0, idx, prod)
    ref_w = torch.zeros(K, device=dev, dtype=torch.float64).index_add_(0, idx, wn.double())
    del prod
    ref_c = ref_x / (ref_w[:, None] + 1e-5)
    errs = {}
    for label, fn in (("fused", lambda: _lib.assign_accumulate(x, wn, book, acc=acc, tensor_cores="fused")),
                      ("sorted", lambda: _lib.accumulate(x, wn, idx, K, acc=acc, sorted_path=True)),
                      ("atomic", lambda: _lib.accumulate(x, wn, idx, K, acc=acc, sorted_path=False))):
        acc.zero_()
        fn()
        c = acc[:, :d].double() / (acc[:, d].double()[:, None] + 1e-5)
        e = (c - ref_c).abs().max(1).values / (ref_c.abs().max(1).values + 1e-30)
        live = ref_w > 0
        errs[label] = float(e[live].max())
        worst = int(torch.argmax(torch.where(live, e, torch.zeros_like(e))))
        errs[label + "_rows_of_worst"] = int((idx == worst).sum())
    print("   centroid error vs float64 (row-normalised max): " + ", ".join(
        f"{k} {errs[k]:.2e} (worst codeword owns {errs[k + '_rows_of_worst']} rows)" for k in ("fused", "sorted", "atomic")))
    t_atomic = timed(lambda: _lib.accumulate(x, wn, idx, K, acc=acc, sorted_path=False), reps=2)
    t_sorted = timed(lambda: _lib.accumulate(x, wn, idx, K, acc=acc, sorted_path=True))
    print(f"   assign {t_assign:.3f} ms | assign+sorted sums {res['sorted']:.3f} ms (+{res['sorted'] - t_assign:.3f}) | "
          f"assign with fused sums {res['fused']:.3f} ms (+{res['fused'] - t_assign:.3f}) | "
          f"stand-alone sums: atomic {t_atomic:.3f} ms, sorted {t_sorted:.3f} ms")
