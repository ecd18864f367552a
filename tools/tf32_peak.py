"""Measured dense TF32 matmul peak of this GPU, the way MEASURED_PEAKS.json measures bf16: torch.matmul of fp32
8192^3 with allow_tf32 (cuBLAS TF32 tensor-core kernels), best of 10 (burst) and back to back for 4 s (sustained).
python tools/tf32_peak.py -> one JSON line"""
import json
import time

import torch


def measure(seconds: float = 4.0, n: int = 8192, device=None):
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(n, n, device=dev)
        b = torch.randn(n, n, device=dev)
        c = torch.empty(n, n, device=dev)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize(dev)
        flops = 2.0 * n ** 3
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b, out=c)
            e1.record()
            torch.cuda.synchronize(dev)
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(seconds * 1e3 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.time()
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize(dev)
        sustained_ms = e0.elapsed_time(e1) / reps
        return {"tf32_tflops": flops / best * 1e-9, "tf32_tflops_sustained": flops / sustained_ms * 1e-9,
                "how": f"torch.matmul fp32 {n}^3 with allow_tf32: best of 10 (burst) and {reps} back to back "
                       f"({time.time() - t0:.1f} s, sustained), CUDA events"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


if __name__ == "__main__":
    print(json.dumps(measure()))
