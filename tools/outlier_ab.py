"""A/B of the outlier-codeword exclusion on the bench workload: per-iteration device time of c3vq_update and the
re-rank statistics.  Run once with the product library and once with a -DC3VQ_TC_NO_OUTLIERS developer build
(C3VQ_LIB=...)."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench  # noqa: E402
from c3dgs_b200 import _lib  # noqa: E402

dev = torch.device("cuda:0")
d = int(sys.argv[1]) if len(sys.argv) > 1 else 45
x, w = bench.synth(5_800_000, d, 0, device=dev)
for rep in range(2):
    cb = bench.init_codebook(x, 4096, 1)
    ei = torch.zeros(4096, device=dev)
    line = []
    for it in range(10):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        _lib.update(x, w, cb, ei, 0.8, 1e-5, tensor_cores=True, return_dist=False)
        b.record()
        torch.cuda.synchronize()
        rows, cand, fb = _lib.assign_stats(dev)
        line.append(f"{a.elapsed_time(b):.3f}ms c/r {cand / rows:.2f} fb {fb}")
    print(os.environ.get("C3VQ_LIB", "product"), f"D={d} rep {rep}:", " | ".join(line))
