"""Read the cycle counters of the C3VQ_TC_PROF developer build (c3dgs_b200/libc3vq_prof.so).
C3VQ_LIB=c3dgs_b200/libc3vq_prof.so python tools/tc_prof.py [rows] [dim]"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c3dgs_b200 import _lib  # noqa: E402
from bench import synth, init_codebook  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
d = int(sys.argv[2]) if len(sys.argv) > 2 else 45
dev = torch.device("cuda", 0)
x, w = synth(n, d, 1234, device=dev)
wn = (w / w.max()).contiguous()
cb = init_codebook(x, 4096, 99)
ei = torch.zeros(4096, device=dev)
acc = torch.zeros(4096, _lib.acc_stride(d), device=dev)
names = {0: "tma other", 1: "tma wait B_EMPTY",
         8: "mma issue/other", 9: "mma wait A_FULL", 10: "mma wait B_FULL", 11: "mma wait ACC_EMPTY", 12: "mma issue 6 MMAs", 13: "mma commit",
         16: "prod other", 17: "prod wait A_EMPTY", 18: "prod convert+transpose+band", 19: "prod rerank(prev pass) work",
         20: "prod prefetch issue", 21: "prod last rerank", 22: "prod rerank wait Q_FULL", 23: "prod A -> TMEM",
         24: "drain other", 25: "drain wait A_FULL", 26: "drain band", 27: "drain wait ACC_FULL",
         28: "drain LDTM+min+lists", 29: "drain wait Q_EMPTY", 30: "drain queue push"}
for it in range(3):
    acc.zero_()
    a = torch.cuda.Event(enable_timing=True); b = torch.cuda.Event(enable_timing=True)
    a.record()
    _lib.assign_accumulate(x, wn, cb, acc=acc)
    b.record()
    torch.cuda.synchronize()
    ws = _lib._workspace.get(dev, 256)
    prof = ws[1368:1368 + 256].view(torch.int64).cpu().tolist()
    ms = a.elapsed_time(b)
    npairs = (n + 255) // 256
    ctas = min(148, npairs)
    print(f"iter {it}: {ms:.3f} ms; cycles per CTA (sum over CTAs / {ctas}), per pass ({npairs / ctas:.1f} passes/CTA):")
    for i, nm in names.items():
        if prof[i]:
            print(f"   {nm:28s} {prof[i] / ctas:12.0f} cyc/CTA  {prof[i] / npairs:10.0f} cyc/pass")
    _lib.ema(cb, ei, acc, 0.8, 1e-5)
