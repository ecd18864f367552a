"""Time the reference's own CUDA kernel (oracle/_ref, weighted_distance.cu compiled unmodified for
sm_100a) next to libc3vq's kernels on the same inputs — the "GPU reference baseline" of BASELINE.md §4.
Run on the GPU box:  python tools/gpu_ref_kernel_bench.py [rows]"""
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle", "_ref"))


def timed(fn, reps=3):
    fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def main():
    from c3dgs_b200 import _lib
    try:
        from weighted_distance_ref import _C as ref
    except Exception as e:  # noqa: BLE001
        ref = None
        print("reference kernel unavailable:", e)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 262144
    out = {}
    for d in (45, 7):
        g = torch.Generator(device="cuda").manual_seed(d)
        x = torch.randn(n, d, device="cuda", generator=g)
        cb = torch.randn(4096, d, device="cuda", generator=g)
        r = {}
        if ref is not None:
            r["reference_ms"] = timed(lambda: ref.weightedDistance(x, cb), reps=2)
            dr, ir = ref.weightedDistance(x, cb)
        r["simt_exact_ms"] = timed(lambda: _lib.assign(x, cb, exact=True))
        de, ie = _lib.assign(x, cb, exact=True)
        if ref is not None:
            r["simt_equals_reference"] = bool(torch.equal(ie, ir) and torch.equal(de, dr))
        try:
            r["tc_ms"] = timed(lambda: _lib.assign(x, cb, exact=False))
            dt, it = _lib.assign(x, cb, exact=False)
            r["tc_equals_simt"] = bool(torch.equal(it, ie) and torch.equal(dt, de))
        except RuntimeError as e:
            r["tc_error"] = str(e)
        for key in ("reference_ms", "simt_exact_ms", "tc_ms"):
            if key in r:
                r[key.replace("_ms", "_rows_per_s")] = n / (r[key] * 1e-3)
        out[f"d{d}"] = r
    out["rows"] = n
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
