#!/usr/bin/env python
"""bench.py — Gaussian*iterations/s of the sensitivity-aware VQ clustering hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one 10-iteration clustering job (10 x VectorQuantize.update over ALL rows, full batch,
K = 4096, decay 0.8) on synthetic inputs shaped like BASELINE.json configs[1]: 5.8 M Gaussians x 45-D
SH features with heavy-tailed sensitivities.  With N GPUs every rank holds its own 5.8 M rows (weak
scaling), the codebook is replicated and each iteration all-reduces the K x (D+1) partial sums (NCCL).
`--scaling strong --total-rows 24000000` splits a fixed number of rows over the ranks instead (configs[3]);
the default run reports that configuration under `extras.strong_24M` at every N.

The JSON line carries, besides the base contract: `roofline` (assignment kernels, tensor roofline),
`roofline_update` (centroid sums + EMA, HBM roofline), `roofline_gather` (decode gather, HBM roofline) — all
from CUDA events inside this run —, `cpu_baseline` (the CPU oracle port of the reference path on the box's host
cores, rank 0, N=1), `e2e` (same job through the public API from pinned HOST buffers, H2D/D2H inside the timed
region), `gpu_launches`, `assign_stats` (candidates per row / rows in the exact scan), at N > 1 `parity`
(replica agreement + agreement with a one-rank run of the same rows) and `h2d_probe`, and `extras` with the other
BASELINE configs (shape 7-D, trace-normalised 6-D, 48-D, iid rows, the reference's own CUDA kernel on the same
inputs, the compress pipeline).

`--impl reference` times the reference path's CPU port (oracle/, all host threads) on a bounded sample
(100 k rows, all 10 iterations per step) of the same workload.  oracle/ is used here only as that baseline.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "Gaussian*iter/sec for VQ k-means (K=4096)"
UNIT = "Gaussian*iter/s"
K = 4096
ITERS = 10
DECAY, EPS = 0.8, 1e-5
TF32_SPEC_TFLOPS = 1125.0          # dense TF32 of the B200 data sheet (half the 2.25 PFLOP/s bf16 figure)


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return {"hbm_gbs": j["hbm_gbs"], "bf16_tflops": j["bf16_tflops"],
                "bf16_tflops_sustained": j.get("bf16_tflops_sustained", j["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


def measure_tf32_peak(device, seconds: float = 2.0, n: int = 8192):
    """Dense TF32 matmul rate of this GPU, measured the way MEASURED_PEAKS.json measures bf16: torch.matmul of fp32
    n^3 with allow_tf32 (cuBLAS), best of 10 (burst) and back to back for `seconds` (sustained)."""
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = True
    try:
        a = torch.randn(n, n, device=device)
        b = torch.randn(n, n, device=device)
        c = torch.empty(n, n, device=device)
        for _ in range(3):
            torch.matmul(a, b, out=c)
        torch.cuda.synchronize(device)
        flops = 2.0 * n ** 3
        best = 1e9
        for _ in range(10):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            torch.matmul(a, b, out=c)
            e1.record()
            torch.cuda.synchronize(device)
            best = min(best, e0.elapsed_time(e1))
        reps = max(10, int(seconds * 1e3 / best))
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            torch.matmul(a, b, out=c)
        e1.record()
        torch.cuda.synchronize(device)
        sustained_ms = e0.elapsed_time(e1) / reps
        return {"tf32_tflops": flops / best * 1e-9, "tf32_tflops_sustained": flops / sustained_ms * 1e-9,
                "how": f"torch.matmul fp32 {n}^3 with allow_tf32 (cuBLAS): best of 10 (burst) and {reps} back to back "
                       f"({sustained_ms * reps * 1e-3:.1f} s, sustained), CUDA events, this run"}
    finally:
        torch.backends.cuda.matmul.allow_tf32 = old


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except OSError:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def stop(self, t0: float, t1: float):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [l.split(", ") for t, l in self.lines if t0 <= t <= t1 + 0.2] or [l.split(", ") for _, l in self.lines]
        sm, smax, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in rows:
            try:
                sm.append(float(r[1])); smax.append(float(r[2]))
                for nm, v in zip(names, r[5:9]):
                    if v.strip().lower() == "active":
                        reasons.add(nm)
            except (ValueError, IndexError):
                pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(smax) if smax else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def synth(n: int, d: int, seed: int, device=None, kind: str = "clustered"):
    """Seeded synthetic rows.  `clustered`: 8192 centres ~N(0,1) with the per-band decay of SH
    coefficients, points = centre + 0.05 N(0,1); `iid`: N(0,1).  w = rand**4 (heavy-tailed sensitivities).
    The same expressions on either generator: the CUDA generator on the GPU arm (CPU generation of 5.8 M x 45 takes
    seconds per rank, and BASELINE.md §3's "identical bits on every box" needs a 1 GB upload per run for no benefit
    to a throughput number), the CPU generator for the CPU legs; the two streams differ, the distribution does not."""
    g = torch.Generator(device=device if device is not None else "cpu").manual_seed(seed)
    kw = {"device": device, "generator": g} if device is not None else {"generator": g}
    if kind == "iid":
        x = torch.randn(n, d, **kw)
    else:
        centres = torch.randn(8192, d, **kw)
        band = torch.arange(d, device=device) // 3
        centres = centres / (1.0 + band.float().sqrt())
        x = centres[torch.randint(0, 8192, (n,), **kw)]
        x = x + 0.05 * torch.randn(n, d, **kw)
    w = torch.rand(n, **kw) ** 4
    return x.contiguous(), w.contiguous()


def init_codebook(x: torch.Tensor, k: int, seed: int):
    """uniform_init (vq.py:24-26) with a fixed seed."""
    amin, amax = x.aminmax()
    g = torch.Generator(device=x.device).manual_seed(seed)
    return (torch.rand(k, x.shape[1], device=x.device, generator=g) * (amax - amin) + amin).contiguous()


# ------------------------------------------------------------------------------------------------
def workload_config(n, d, world, scaling="weak"):
    """The workload only (identical for both arms); how each arm computes it is in `implementation`."""
    which = "configs[2]: shape codebook clustering" if d <= 8 else "configs[1]: SH codebook clustering"
    if scaling == "strong":
        which = "configs[3]: sharded SH clustering, fixed total rows"
    return {"workload": f"{which}, {n / 1e6:.1f}M Gaussians x {d}-D per GPU, K=4096, "
                        "10 full-batch k-means iterations per step",
            "rows_per_gpu": n, "dim": d, "codebook_size": K, "iterations_per_step": ITERS, "decay": DECAY,
            "distribution": "clustered (8192 centres, SH band decay, sigma 0.05), w = rand**4, uniform-box initial codebook",
            "parallelism": f"rows sharded x{world}, codebook replicated, 1 all-reduce of K x (D+1) sums per iteration",
            "l2": f"inputs ({n * (d + 1) * 4 / 1e9:.2f} GB per GPU) exceed the 126 MB L2; no explicit flush"}


SAMPLE_ROWS = 100_000     # CPU legs: a bounded sample of the workload's rows (the metric is per row and iteration)


def cpu_sample(d):
    """The CPU legs' inputs: SAMPLE_ROWS rows of the workload's distribution, normalised weights, uniform-box codebook
    (same expressions and seeds as the GPU arm; the CPU generator's stream)."""
    x, w = synth(SAMPLE_ROWS, d, 1234)
    cb0 = init_codebook(x, K, 99)
    return x.numpy(), (w / w.max()).numpy().astype(np.float32), cb0.numpy()


def run_reference_arm(args, rank):
    """The reference path's CPU implementation (oracle port: OpenMP C restatement of weighted_distance.cu +
    vq.py's update) on the box's host cores, on a bounded sample of OUR arm's workload; rank 0 only."""
    if rank != 0:
        return
    from oracle import oracle
    oracle.build()
    n, d = SAMPLE_ROWS, args.dim
    x, wn, cb0 = cpu_sample(d)
    # torchrun exports OMP_NUM_THREADS=1 to every rank; this leg is the only CPU work of the job and takes every core
    oracle.set_threads(len(os.sched_getaffinity(0)))
    cores = oracle.num_threads()

    def job():
        cb, ei = cb0.copy(), np.zeros(K, np.float32)
        for _ in range(ITERS):
            oracle.update(x, wn, cb, ei, DECAY, EPS)

    # each step is a bounded sample (100k rows x 10 iterations, ~seconds); cap the run at a few minutes
    t_probe = time.time(); job(); t_probe = time.time() - t_probe
    warm = max(0, min(args.warmup, int(20 / max(t_probe, 1e-3))) - 1)
    steps = max(1, min(args.steps, int(120 / max(t_probe, 1e-3))))
    for _ in range(warm):
        job()
    t0 = time.time()
    for _ in range(steps):
        job()
    dt = (time.time() - t0) / steps
    value = n * ITERS / dt
    sample = (f"{n} of the workload's {args.rows} rows per GPU (x {d}-D, K=4096), all {ITERS} iterations per step; "
              "oracle/c3vq_oracle.c (OpenMP restatement of weighted_distance.cu + vq.py update), "
              f"{cores} host threads; ONE CPU box whatever --gpus says (the metric is per row and iteration)")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm + 1, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.rows, d, max(1, args.gpus)),
        "implementation": {"assignment": "cpu-exact (reference arithmetic)", "sums": "cpu scatter",
                           "sampled_rows": n},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)


# ------------------------------------------------------------------------------------------------
def cpu_baseline_leg(d=45):
    from oracle import oracle
    oracle.build()
    n = SAMPLE_ROWS
    x, wn, cb = cpu_sample(d)
    ei = np.zeros(K, np.float32)
    oracle.set_threads(len(os.sched_getaffinity(0)))    # all host cores, whatever OMP_NUM_THREADS says
    oracle.update(x, wn, cb, ei, DECAY, EPS)            # warm-up iteration
    t0 = time.time()
    done = 0
    while done < 40 * ITERS and (done < ITERS or time.time() - t0 < 12):
        oracle.update(x, wn, cb, ei, DECAY, EPS)
        done += 1
    dt = time.time() - t0
    return {"value": n * done / dt, "unit": UNIT, "cores": oracle.num_threads(), "kind": "port",
            "sample": f"{n} rows of the workload (x {d}-D, K=4096), {done} update iterations, {dt:.1f} s of CPU work; "
                      "oracle/c3vq_oracle.c (OpenMP restatement of weighted_distance.cu + vq.py update)"}


class Clustering:
    """10 full-batch iterations over this rank's rows through the same kernels c3vq_update launches, as separate
    calls so that the assignment (prep + k_assign_tc + fallback scan) and the update (sums + EMA) can be bracketed
    by CUDA events on their own; the all-reduce of the sums sits between the two halves of the update."""

    def __init__(self, x, wn, cb0, dev, simt=False, sums="fused", scale_normalize=False):
        from c3dgs_b200 import _lib
        from c3dgs_b200.sharded import CudaOps, ShardedVectorQuantize
        self.L, self.x, self.wn, self.cb0 = _lib, x, wn, cb0
        self.n, self.d = x.shape
        self.scale_normalize = scale_normalize
        self.ops = CudaOps()
        self.ops.tensor_cores = False if simt else sums          # "sorted" / "fused": explicit, whatever the default
        self.split = (not simt) and sums == "sorted"
        self.vq = ShardedVectorQuantize(self.d, K, DECAY, device=dev, ops=self.ops)
        # the assignments land in one buffer for the whole job: a fresh 8 n-byte tensor per iteration makes the caching
        # allocator return and re-request blocks at 12-24 M rows, and cudaFree waits for the device
        self.idx = torch.empty(self.n, dtype=torch.int64, device=dev)
        self.events = []           # per iteration (e0, e1, e2, e3, e4)

    def step(self, record: bool):
        vq, L = self.vq, self.L
        vq.codebook.copy_(self.cb0)
        vq.entry_importance.zero_()
        for _ in range(ITERS):
            vq._acc.zero_()
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(5)] if record else None
            if record:
                ev[0].record()
            if self.split:
                L.assign(self.x, vq.codebook, return_dist=False, out_idx=self.idx)
                if record:
                    ev[1].record()
                L.accumulate(self.x, self.wn, self.idx, K, acc=vq._acc, sorted_path=True)
            else:
                L.assign_accumulate(self.x, self.wn, vq.codebook, acc=vq._acc, tensor_cores=self.ops.tensor_cores,
                                    out_idx=self.idx)
                if record:
                    ev[1].record()
            if record:
                ev[2].record()
            vq._all_reduce(vq._acc)
            if record:
                ev[3].record()
            self.ops.ema(vq.codebook, vq.entry_importance, vq._acc, DECAY, EPS, self.scale_normalize)
            if record:
                ev[4].record()
                self.events.append(ev)

    def kernel_ms(self):
        """(assignment ms, update ms) per iteration, averaged over the recorded iterations.  With the sums fused into
        the assignment kernel the first figure holds both and the second only the EMA; `fused_split` separates them."""
        a = float(np.mean([e[0].elapsed_time(e[1]) for e in self.events]))
        u = float(np.mean([e[1].elapsed_time(e[2]) + e[3].elapsed_time(e[4]) for e in self.events]))
        return a, u

    def fused_split(self):
        """Fused mode only: one more 10-iteration job in which every iteration's codebook is first assigned WITHOUT
        the sums (timed), then with them (timed): (assignment alone ms, increment for the sums ms) per iteration."""
        vq, L = self.vq, self.L
        vq.codebook.copy_(self.cb0)
        vq.entry_importance.zero_()
        alone, fused = [], []
        for _ in range(ITERS):
            ev = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
            vq._acc.zero_()
            ev[0].record()
            L.assign(self.x, vq.codebook, return_dist=False, out_idx=self.idx)
            ev[1].record()
            L.assign_accumulate(self.x, self.wn, vq.codebook, acc=vq._acc, tensor_cores=self.ops.tensor_cores,
                                out_idx=self.idx)
            ev[2].record()
            vq._all_reduce(vq._acc)
            self.ops.ema(vq.codebook, vq.entry_importance, vq._acc, DECAY, EPS, self.scale_normalize)
            alone.append(ev[:2])
            fused.append(ev[1:])
        torch.cuda.synchronize()
        a = float(np.mean([e[0].elapsed_time(e[1]) for e in alone]))
        f = float(np.mean([e[0].elapsed_time(e[1]) for e in fused]))
        return a, f - a


def timed(fn, reps=5):
    fn()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps


def screening_pipe():
    """Which tensor pipe the screening GEMM runs on (libc3vq reads C3VQ_SCREEN per call; default fp16 operands)."""
    return "tf32" if os.environ.get("C3VQ_SCREEN", "f16").lower().startswith("t") else "f16"


def assignment_roofline(n, d, kern_ms, share, pk, tf32, simt, split):
    """Tensor roofline for the wide rows, the score-drain (TMEM read) roofline for D <= 8 (that width
    is compare-rate bound, the MMA is a single k-step).  `peak` is the measured sustained rate of the tensor pipe
    the kernel actually issues to: cuBLAS TF32 measured in this run for kind::tf32, MEASURED_PEAKS.json's bf16 figure
    for kind::f16 (fp16 and bf16 share the pipe and the rate).  The north star words its target against the TF32
    peak whatever the pipe, so that fraction is printed beside it."""
    flops = 2.0 * K * d * n                               # assign flops = 2*K*D per Gaussian*iteration
    pipe = screening_pipe()
    kernel = ("k_assign_exact (SIMT)" if simt else
              (f"k_tc_prep_* + k_assign_tc (kind::{pipe}) + k_assign_fallback_* (assignment alone)"
               if split else f"k_assign_tc (kind::{pipe}) with the centroid sums fused in + prep + fallback"))
    drain_peak = 148 * 4 * 64 * 1.965e9 / 4 / 1e12        # 1e12 scores/s: 64 B/clk per SM sub-partition, 4 B per score
    drain = {"achieved_tscore_s": n * K / (kern_ms * 1e-3) / 1e12, "peak_tscore_s": drain_peak,
             "frac": n * K / (kern_ms * 1e-3) / 1e12 / drain_peak,
             "note": "every one of the n*K fp32 scores leaves TMEM once: 64 B/clk per SM sub-partition x 592 x 1.965 GHz"}
    if d > 8:
        achieved = flops / (kern_ms * 1e-3) / 1e12
        if pipe == "tf32" or simt:
            peak, src = tf32["tf32_tflops_sustained"], ("measured in this run: cuBLAS TF32 matmul 8192^3, sustained "
                                                         "(kernel timed inside a long step)")
        else:
            peak, src = pk["bf16_tflops_sustained"], (f"{pk['source']}: MEASURED_PEAKS.json bf16_tflops_sustained "
                                                      "(the kind::f16 pipe; kernel timed inside a long step)")
        return {"bound": "tensor", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "traffic": None,
                "traffic_source": "not measured in this run; ncu dram__bytes of one k_assign_tc launch at 5.8M x 45 under "
                                  "profiles/ (1.09 GB = the algorithmic bytes)",
                "kernel": kernel, "kernel_ms": kern_ms, "kernel_share_of_step": share, "pipe": "kind::" + pipe,
                "peak_source": src,
                "frac_vs_tf32_sustained_measured": achieved / tf32["tf32_tflops_sustained"],
                "frac_vs_tf32_burst_measured": achieved / tf32["tf32_tflops"],
                "frac_vs_half_bf16_sustained": achieved / (pk["bf16_tflops_sustained"] / 2.0),
                "frac_vs_tf32_spec_1125": achieved / TF32_SPEC_TFLOPS,
                "tf32_peak_measured": tf32, "score_drain": drain, "algorithmic_flops_per_launch": flops,
                "note": "the kernel is bound by the score drain (TMEM read + min reduction), not by the MMA: see score_drain"}
    return {"bound": "tensor", "achieved": drain["achieved_tscore_s"], "peak": drain_peak, "unit": "Tscore/s",
            "frac": drain["frac"], "traffic": None, "kernel": kernel, "kernel_ms": kern_ms,
            "kernel_share_of_step": share, "pipe": "kind::" + pipe,
            "peak_source": "score-drain roofline for D <= 8 (SURVEY §8d: compare-rate bound): " + drain["note"],
            "tensor_tflops": flops / (kern_ms * 1e-3) / 1e12,
            "hbm_floor_ms": n * (4 * d + 12) / pk["hbm_gbs"] * 1e-6, "algorithmic_scores_per_launch": float(n) * K}


def update_roofline(n, d, upd_ms, share, pk, fused=False):
    nbytes = float(n) * (4 * d + 4 + 8)                   # update bytes = 4*D + 4 (w) + 8 (idx) per row
    achieved = nbytes / (upd_ms * 1e-3) / 1e9
    kernel = ("what the centroid sums add to k_assign_tc when fused in (rows read from HBM once per iteration; timed as "
              "fused launch minus assignment alone on the same codebooks) + k_ema" if fused else
              "k_seg_hist + k_seg_scan + k_seg_scatter + k_seg_sum + k_ema (sorted centroid sums + EMA)")
    note = ("algorithmic bytes (4 D + 12 per row) over the time the update ADDS; with the sums fused into the assignment "
            "kernel the rows are not read a second time, so this can exceed the HBM peak: the update then costs less "
            "than one pass over x would" if fused else
            "algorithmic bytes (4 D + 12 per row) over the time of the sort + sum + EMA kernels")
    return {"bound": "hbm", "achieved": achieved, "peak": pk["hbm_gbs"], "unit": "GB/s", "frac": achieved / pk["hbm_gbs"],
            "traffic": None, "kernel": kernel, "note": note,
            "kernel_ms": upd_ms, "kernel_share_of_step": share, "algorithmic_bytes_per_launch": nbytes,
            "peak_source": f"{pk['source']}: MEASURED_PEAKS.json hbm_gbs (copy bandwidth)"}


def run_config(name, x, w, dev, steps=3, scale_normalize=False, cb0=None, pk=None, tf32=None):
    """One of the other BASELINE configurations on this GPU: ms per 10-iteration step, kernel times, screening stats."""
    from c3dgs_b200 import _lib
    n, d = x.shape
    wn = (w / w.max()).contiguous()
    cb0 = init_codebook(x, K, 99) if cb0 is None else cb0
    job = Clustering(x, wn, cb0, dev, sums=_lib.default_sums(), scale_normalize=scale_normalize)
    for _ in range(2):
        job.step(False)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        job.step(True)
    b.record()
    torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    rows, cands, fb = _lib.assign_stats(dev)              # last iteration's assignment
    am, um = job.kernel_ms()
    fused = not job.split
    if fused:
        alone, incr = job.fused_split()
        um = max(incr, 0.0) + um
    out = {"workload": name, "rows": n, "dim": d, "ms_per_step": ms, "value": n * ITERS / (ms * 1e-3),
           "assign_ms": am, "update_ms": um, "candidates_per_row": cands / max(rows, 1), "fallback_rows": fb,
           "roofline": assignment_roofline(n, d, am, am * ITERS / ms, pk, tf32, False, job.split),
           "roofline_update": update_roofline(n, d, um, um * ITERS / ms, pk, fused=fused)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=5_800_000, help="rows per GPU (default: configs[1])")
    ap.add_argument("--dim", type=int, default=45)
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"])
    ap.add_argument("--total-rows", type=int, default=24_000_000, help="--scaling strong: rows split over the ranks")
    ap.add_argument("--simt", action="store_true", help="use the SIMT exact assignment instead of tensor cores")
    ap.add_argument("--sums", default=None, choices=["sorted", "fused"],
                    help="centroid sums: fused into the assignment kernel (one bulk reduce-add per distinct codeword of "
                         "a warp; rows read once) or the sorted segmented sum after it; default: the library's default")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true")
    args = ap.parse_args()

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference_arm(args, rank)
        return

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU port)")
    import torch.distributed as dist
    from c3dgs_b200 import _lib

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    _lib.load()
    pk = peaks()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(ms):
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    if args.sums is None:
        args.sums = _lib.default_sums()
    d = args.dim
    n = args.rows if args.scaling == "weak" else args.total_rows // world
    warm = max(3, args.warmup)
    x, w = synth(n, d, 1234 + rank, device=dev)
    wmax = w.max().clone()
    if world > 1:
        dist.all_reduce(wmax, op=dist.ReduceOp.MAX)
    wn = (w / wmax).contiguous()
    cb0 = init_codebook(x, K, 99)                       # same seed on every rank -> identical replicas
    if world > 1:
        dist.broadcast(cb0, src=0)                      # the box bounds differ per shard: rank 0's codebook everywhere
    job = Clustering(x, wn, cb0, dev, simt=args.simt, sums=args.sums)
    vq, ops = job.vq, job.ops

    # the TF32 tensor peak the assignment is quoted against, measured on this GPU before the timed region
    # (C3VQ_BENCH_TF32_SECONDS shortens it for runs under ncu, where every GEMM launch counts against the capture)
    tf32 = measure_tf32_peak(dev, seconds=float(os.environ.get("C3VQ_BENCH_TF32_SECONDS", "2.0")))

    for _ in range(warm):
        job.step(False)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = _lib.launch_count()
    marks = [torch.cuda.Event(enable_timing=True) for _ in range(args.steps + 1)]
    barrier()
    t_wall0 = time.time()
    marks[0].record()
    for s in range(args.steps):
        job.step(True)
        marks[s + 1].record()
    barrier()
    t_wall1 = time.time()
    launches = _lib.launch_count() - launches0
    ms_per_step = max_over_ranks(marks[0].elapsed_time(marks[-1])) / args.steps
    per_step = [marks[i].elapsed_time(marks[i + 1]) for i in range(args.steps)]
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    value = world * n * ITERS / (ms_per_step * 1e-3)
    stats_rows, stats_cands, stats_fb = (_lib.assign_stats(dev) if not args.simt else (n, 0, 0))
    replica_hash = int(vq.codebook.view(torch.int32).to(torch.int64).sum().item()) & 0xFFFFFFFFFFFF

    assign_ms, update_ms = job.kernel_ms()
    if job.split or args.simt:
        roofline = assignment_roofline(n, d, assign_ms, assign_ms * ITERS / ms_per_step, pk, tf32, args.simt, job.split)
        roofline_update = update_roofline(n, d, update_ms, update_ms * ITERS / ms_per_step, pk) if job.split else None
    else:
        # sums fused into the assignment kernel: the launch is timed as one; a second job splits it per iteration
        alone_ms, incr_ms = job.fused_split()
        roofline = assignment_roofline(n, d, assign_ms, assign_ms * ITERS / ms_per_step, pk, tf32, False, False)
        roofline["assignment_alone_ms"] = alone_ms
        upd = max(incr_ms, 0.0) + update_ms                     # update_ms holds only the EMA here
        roofline_update = update_roofline(n, d, upd, upd * ITERS / ms_per_step, pk, fused=True)
        roofline_update["sums_increment_ms"], roofline_update["ema_ms"] = incr_ms, update_ms

    # write-only DRAM rate (cudaMemset-class fill of 1 GiB): the decode gather reads its rows out of the L2 and is a pure
    # write stream to DRAM, so the copy peak (read + write bytes) is not what bounds it
    fill = torch.empty(1 << 28, dtype=torch.float32, device=dev)
    ms_fill = timed(lambda: fill.fill_(1.0))
    write_peak_gbs = fill.numel() * 4 / ms_fill * 1e-6
    del fill
    # decode gather: codebook[idx] -> [n, d], 8 B index in + 4 d B out per row
    gen = torch.Generator(device=dev).manual_seed(5 + rank)
    idx_dec = torch.randint(0, K, (n,), device=dev, generator=gen)
    ms_g = timed(lambda: _lib.gather(vq.codebook, idx_dec), reps=20)
    gbytes = float(n) * (8 + 4 * d)
    roofline_gather = {"bound": "hbm", "achieved": gbytes / ms_g * 1e-6, "peak": pk["hbm_gbs"], "unit": "GB/s",
                       "frac": gbytes / ms_g * 1e-6 / pk["hbm_gbs"], "traffic": None,
                       "kernel": "k_gather_pad_prep + k_gather_padded (c3vq_gather_ws: four padded copies of the "
                                 "table, aligned 16-byte loads)" if d % 4 and d >= 12 else "k_gather_span4 / k_gather_span",
                       "kernel_ms": ms_g, "algorithmic_bytes_per_launch": gbytes,
                       "write_only_peak_measured_gbs": write_peak_gbs,
                       "frac_vs_write_only_peak": gbytes / ms_g * 1e-6 / write_peak_gbs,
                       "note": "the table rows come out of the L2: DRAM sees the indices and the output rows, i.e. a "
                               "write stream; write_only_peak = torch fill_ of 1 GiB timed in this run",
                       "peak_source": f"{pk['source']}: MEASURED_PEAKS.json hbm_gbs (copy bandwidth)"}
    del idx_dec

    # ---- N > 1: replica agreement, and agreement with a one-rank run of the same rows at a reduced size ----
    parity = None
    if world > 1:
        hashes = [None] * world
        dist.all_gather_object(hashes, replica_hash)
        sub = 262_144                                     # rows per rank in the reduced problem
        xs, ws = x[:sub].contiguous(), wn[:sub].contiguous()
        # sharded: all ranks, NCCL all-reduce per iteration; every iteration's starting codebook and all-reduced sums
        # are kept for the teacher-forced comparison (k-means feeds rounding back into later assignments,
        # so only per-iteration comparisons from a COMMON codebook have a tolerance that means something)
        pj = Clustering(xs, ws, cb0, dev, sums=args.sums)
        pv = pj.vq
        pv.codebook.copy_(cb0)
        pv.entry_importance.zero_()
        books, sums = [], []
        for _ in range(ITERS):
            books.append(pv.codebook.clone())
            pv._acc.zero_()
            _lib.assign_accumulate(xs, ws, pv.codebook, acc=pv._acc, tensor_cores=args.sums, out_idx=pj.idx)
            pv._all_reduce(pv._acc)
            sums.append(pv._acc.clone())
            pj.ops.ema(pv.codebook, pv.entry_importance, pv._acc, DECAY, EPS, False)
        cb_sharded = pv.codebook.clone()
        gx = [torch.empty_like(xs) for _ in range(world)] if rank == 0 else None
        gw = [torch.empty_like(ws) for _ in range(world)] if rank == 0 else None
        dist.gather(xs, gx, dst=0)
        dist.gather(ws, gw, dst=0)
        if rank == 0:
            xa, wa = torch.cat(gx), torch.cat(gw)
            one = Clustering(xa, wa, cb0, dev, sums=args.sums)
            one.vq._world = lambda: 1                     # no collectives: the same rows on one GPU
            # (1) teacher-forced: each iteration of the one-rank run starts from the sharded run's codebook
            tf_err, acc1 = 0.0, torch.zeros_like(pv._acc)
            for cb_i, acc_s in zip(books, sums):
                acc1.zero_()
                _lib.assign_accumulate(xa, wa, cb_i, acc=acc1, tensor_cores=args.sums, out_idx=one.idx)
                c1 = acc1[:, :d] / (acc1[:, d:d + 1] + EPS)
                cs = acc_s[:, :d] / (acc_s[:, d:d + 1] + EPS)
                live = acc1[:, d] > 0
                e = ((c1 - cs).abs().max(1).values / (c1.abs().max(1).values + 1e-30))[live]
                tf_err = max(tf_err, float(e.max()) if e.numel() else 0.0)
            # (2) free-running: ten iterations of feedback on each side (a single near-tie that flips moves a small
            # codeword by ~1/members; reported, not held to the tolerance)
            one.step(False)
            scale = one.vq.codebook.abs().max(1, keepdim=True).values + 1e-30
            fin = torch.isfinite(one.vq.codebook).all(1)
            err = float(((one.vq.codebook - cb_sharded).abs() / scale)[fin].max())
            agree = float((_lib.assign(xa, one.vq.codebook)[1] == _lib.assign(xa, cb_sharded)[1]).float().mean())
            parity = {"replicas_identical": len(set(hashes)) == 1, "codebook_hashes": hashes,
                      "one_rank_rows": int(xa.shape[0]), "teacher_forced_centroid_max_rel_err": tf_err,
                      "tolerance": 1e-5, "ok": bool(len(set(hashes)) == 1 and tf_err < 1e-5),
                      "free_running_codebook_max_rel_err_after_10_iterations": err,
                      "free_running_assignments_equal_frac": agree,
                      "note": "headline run: 48-bit checksum of the final codebook bits of every rank; reduced problem: "
                              f"{sub} rows of every rank, 10 iterations sharded over {world} ranks (NCCL all-reduce of the "
                              "sums); rank 0 then forms every iteration's sums alone from the same codebook and the "
                              "centroids must agree within 1e-5 (row-normalised); the free-running comparison lets the "
                              "two runs diverge through their own rounding"}
            del xa, wa, one
        del pj, xs, ws
        barrier()

    # ---- e2e: the same job from pinned HOST buffers through the public API, copies inside the timed region ----
    from c3dgs_b200.sharded import gpu_local_host_memory
    with gpu_local_host_memory(local_rank):          # pinned staging buffers on this GPU's NUMA node
        xh = torch.empty(x.shape, dtype=torch.float32).pin_memory()
        wh = torch.empty(w.shape, dtype=torch.float32).pin_memory()
        cbh = torch.empty(K, d, dtype=torch.float32).pin_memory()
        idxh = torch.empty(n, dtype=torch.int64).pin_memory()
    xh.copy_(x); wh.copy_(wn)
    xd, wd = torch.empty_like(x), torch.empty_like(wn)

    def e2e_job():
        vq.codebook.copy_(cb0)
        vq.entry_importance.zero_()
        # iteration 1 straight from the pinned host rows: chunked upload, each chunk assigned as it lands
        idx = vq.update_from_host(xh, wh, xd, wd)
        for it in range(1, ITERS):
            vq._acc.zero_()
            _, idx = ops.assign_accumulate(xd, wd, vq.codebook, vq._acc)
            vq._all_reduce(vq._acc)
            ops.ema(vq.codebook, vq.entry_importance, vq._acc, DECAY, EPS, False)
        cbh.copy_(vq.codebook, non_blocking=True)
        idxh.copy_(idx, non_blocking=True)

    e2e_job()
    barrier()
    s2, t2 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2e_steps = max(2, min(args.steps, 5))
    s2.record()
    for _ in range(e2e_steps):
        e2e_job()
    t2.record()
    barrier()
    ms_e2e = max_over_ranks(s2.elapsed_time(t2)) / e2e_steps
    e2e = {"value": world * n * ITERS / (ms_e2e * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": int(xh.numel() * 4 + wh.numel() * 4),
           "d2h_bytes_per_step": int(cbh.numel() * 4 + idxh.numel() * 8), "ms_per_step": ms_e2e}

    # pinned host -> device rate of this rank alone and of all ranks at once: what bounds e2e at N > 1
    h2d_probe = None
    if world > 1:
        def h2d_once():
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            xd.copy_(xh, non_blocking=True)
            b.record()
            torch.cuda.synchronize()
            return xh.numel() * 4 / a.elapsed_time(b) * 1e-6
        h2d_once()
        solo = torch.zeros(world, device=dev, dtype=torch.float64)
        for r in range(world):
            barrier()
            if r == rank:
                solo[r] = h2d_once()
        barrier()
        conc = torch.zeros(world, device=dev, dtype=torch.float64)
        conc[rank] = h2d_once()
        dist.all_reduce(solo)
        dist.all_reduce(conc)
        h2d_probe = {"bytes": int(xh.numel() * 4), "solo_gbs_per_rank": [round(v, 1) for v in solo.tolist()],
                     "concurrent_gbs_per_rank": [round(v, 1) for v in conc.tolist()],
                     "concurrent_aggregate_gbs": float(conc.sum()),
                     "e2e_upload_floor_ms": float(xh.numel() * 4 / (conc.min().item() * 1e6)),
                     "note": "pinned host -> device copy of this rank's rows: one rank at a time, then all ranks at once"}
    del xh, wh, xd, wd, idxh

    # ---- BASELINE configs[3]: 24 M rows split over the ranks (strong scaling), and the same rows on rank 0 alone ----
    strong = None
    if not args.no_extras and args.scaling == "weak" and not args.simt and world in (1, 2, 4, 8):
        total, blocks = args.total_rows, 8
        per = total // blocks
        mine = range(rank * blocks // world, (rank + 1) * blocks // world)
        parts = [synth(per, d, 5000 + b, device=dev) for b in mine]
        xs, ws = torch.cat([p_[0] for p_ in parts]), torch.cat([p_[1] for p_ in parts])
        del parts
        cbs = init_codebook(xs, K, 99)
        if world > 1:
            dist.broadcast(cbs, src=0)
        sj = Clustering(xs, ws, cbs, dev, sums=args.sums)
        for _ in range(2):
            sj.step(False)
        barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(3):
            sj.step(True)
        b.record()
        barrier()
        ms_n = max_over_ranks(a.elapsed_time(b)) / 3
        am, um = sj.kernel_ms()
        ar_ms = float(np.mean([e[2].elapsed_time(e[3]) for e in sj.events]))
        intervals = [float(np.mean([e[i].elapsed_time(e[i + 1]) for e in sj.events])) for i in range(4)]
        del sj, xs, ws
        ms_1 = ms_n
        if world > 1:
            if rank == 0:
                parts = [synth(per, d, 5000 + b, device=dev) for b in range(blocks)]
                xa, wa = torch.cat([p_[0] for p_ in parts]), torch.cat([p_[1] for p_ in parts])
                del parts
                oj = Clustering(xa, wa, cbs, dev, sums=args.sums)
                oj.vq._world = lambda: 1
                for _ in range(2):
                    oj.step(False)
                torch.cuda.synchronize()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                for _ in range(3):
                    oj.step(False)
                b.record()
                torch.cuda.synchronize()
                ms_1 = a.elapsed_time(b) / 3
                del oj, xa, wa
            t = torch.tensor([ms_1], device=dev, dtype=torch.float64)
            dist.broadcast(t, src=0)
            ms_1 = float(t.item())
        strong = {"workload": f"configs[3]: {total / 1e6:.0f}M Gaussians x {d}-D split over {world} GPU(s), K=4096, "
                              "10 full-batch iterations per step, NCCL all-reduce of the sums per iteration",
                  "total_rows": total, "rows_per_gpu": total // world, "n_gpus": world, "scaling": "strong",
                  "ms_per_step": ms_n, "value": total * ITERS / (ms_n * 1e-3),
                  "one_gpu_ms_per_step": ms_1, "speedup_vs_one_gpu": ms_1 / ms_n,
                  "assign_ms_per_iteration": am, "update_ms_per_iteration": um, "allreduce_ms_per_iteration": ar_ms,
                  "intervals_ms": dict(zip(("assign (+ fused sums)", "separate sums kernels", "allreduce", "ema"), intervals)),
                  "note": "the one-GPU time is measured in this run: rank 0 clusters all rows alone while the other ranks wait"}
        torch.cuda.empty_cache()

    extras = None
    if world == 1 and not args.simt and not args.no_extras:
        pk["write_only_gbs"] = write_peak_gbs
        extras = single_gpu_extras(x, w, wn, cb0, n, d, dev, pk, tf32)
    if strong is not None:
        extras = extras or {}
        extras["strong_24M"] = strong

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_leg(d)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_per_step, "ms_per_step_median": float(np.median(per_step)), "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None,
            "dtype": (f"f32 ({screening_pipe()} tensor-core screening + exact fp32 re-rank: results bit-identical to the "
                      "fp32 reference)") if not args.simt else "f32",
            "data": "synthetic",
            "config": workload_config(n, d, world, args.scaling),
            "implementation": {"assignment": "simt-exact" if args.simt else f"tcgen05-{screening_pipe()}+rerank",
                               "sums": "atomic" if args.simt else args.sums},
            "roofline": roofline, "roofline_update": roofline_update, "roofline_gather": roofline_gather,
            "assign_stats": {"rows": stats_rows, "candidates_per_row": stats_cands / max(stats_rows, 1),
                             "fallback_rows": stats_fb, "of": "the last assignment of the timed region"},
            "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches), "clocks": clocks,
            "parity": parity, "h2d_probe": h2d_probe, "extras": extras,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def single_gpu_extras(x, w, wn, cb0, n, d, dev, pk, tf32):
    """The other BASELINE configurations and the stages either side of the path, on one GPU (rank 0, N = 1)."""
    from c3dgs_b200 import _lib as ops2
    extras = {}
    gen = torch.Generator(device=dev).manual_seed(5)

    # ---- the other north-star configurations, each a 10-iteration full-batch clustering of 5.8 M rows ----
    x7, w7 = synth(n, 7, 77, device=dev)
    extras["shape7"] = run_config("configs[2]: shape codebook clustering, 5.8M x 7-D (scale + rotation)", x7, w7, dev,
                                  pk=pk, tf32=tf32)
    del x7, w7
    s3 = torch.nn.functional.normalize(torch.rand(n, 3, device=dev, generator=gen) + 0.05)
    r4 = torch.randn(n, 4, device=dev, generator=gen)
    cov = ops2.build_covariance(s3, r4)
    extras["cov6_trace_normalised"] = run_config(
        "configs[2], fork reality: 5.8M normalised covariances x 6-D with the trace normalisation after every step "
        "(vq.py:76-77)", cov, w, dev, scale_normalize=True, pk=pk, tf32=tf32)
    extras["cov6_trace_normalised"]["fallback_rows_limit"] = n // 50
    del s3, r4, cov
    x48, w48 = synth(n, 48, 48, device=dev)
    extras["sh48"] = run_config("configs[1], fork reality: 5.8M x 48-D (DC + 15 SH coefficients x 3)", x48, w48, dev,
                                pk=pk, tf32=tf32)
    del x48, w48
    xi, wi = synth(n, d, 45, device=dev, kind="iid")
    cbi = xi[torch.randint(0, n, (K,), device=dev, generator=gen)].clone()
    extras["iid45"] = run_config("5.8M iid N(0,1) rows x 45-D, initial codebook sampled from the rows (every codeword in use)",
                                 xi, wi, dev, cb0=cbi, pk=pk, tf32=tf32)
    del xi, wi, cbi

    # ---- the "before" number (SURVEY §8d last row): the reference's own weighted_distance.cu
    # compiled unmodified (oracle/_ref) + index_add_ for torch_scatter, one iteration on the bench's own rows ----
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    if os.path.exists(os.path.join(ref_dir, "weighted_distance_ref", "_C.so")):
        sys.path.insert(0, ref_dir)
        try:
            from weighted_distance_ref import _C as ref
            cb, ei = cb0.clone(), torch.zeros(K, device=dev)

            def ref_update():
                # compression/vq.py:28-35 with the reference kernel and index_add_ standing in for torch_scatter
                _, idx = ref.weightedDistance(x, cb)
                acc_w = torch.zeros(K, device=dev).index_add_(0, idx, wn)
                ei.mul_(DECAY).add_(acc_w, alpha=1 - DECAY)
                acc_x = torch.zeros(K, d, device=dev).index_add_(0, idx, x * wn[:, None])
                cb.mul_(DECAY).add_(acc_x / (acc_w[:, None] + EPS), alpha=1 - DECAY)

            ref.weightedDistance(x[:4096], cb)                       # module load / first launch
            torch.cuda.synchronize()
            a, m, b = (torch.cuda.Event(enable_timing=True) for _ in range(3))
            a.record(torch.cuda.default_stream(dev))
            _, idx_r = ref.weightedDistance(x, cb0)                    # legacy default stream (weighted_distance.cu:52)
            m.record(torch.cuda.default_stream(dev))
            torch.cuda.synchronize()
            ms_assign = a.elapsed_time(m)
            ms_update = timed(ref_update, reps=1)
            _, idx_o = ops2.assign(x, cb0, return_dist=False)
            extras["gpu_reference_kernel"] = {
                "workload": f"one update iteration on the headline rows ({n} x {d}, K=4096, iteration-0 codebook)",
                "assign_ms": ms_assign, "update_iteration_ms": ms_update,
                "value": n / (ms_update * 1e-3), "unit": UNIT,
                "assignments_equal_ours": bool(torch.equal(idx_r, idx_o)),
                "note": "submodules/weighted_distance/weighted_distance.cu compiled unmodified for sm_100a "
                        "(make -C oracle ref) + torch index_add_ for the two torch_scatter calls; same box, same rows"}
            del idx_r, idx_o
        except Exception as exc:  # noqa: BLE001
            extras["gpu_reference_kernel"] = {"unavailable": repr(exc)}
        finally:
            sys.path.remove(ref_dir)
    else:
        extras["gpu_reference_kernel"] = {"unavailable": "oracle/_ref/weighted_distance_ref/_C.so not built"}

    # the reference's real default, mini-batch mode, through the drop-in vq_features
    # (CPU randint batches with replacement, 100 x 2^18 rows, then the final full assignment); reported, not the metric
    from c3dgs_b200.vq import vq_features
    mb_steps, mb_chunk = 100, 2 ** 18
    vq_features(x, w, K, vq_chunk=mb_chunk, steps=3, decay=DECAY, silent=True)      # warm-up
    torch.cuda.synchronize()
    t0 = time.time()
    vq_features(x, w, K, vq_chunk=mb_chunk, steps=mb_steps, decay=DECAY, silent=True)
    torch.cuda.synchronize()
    dt = time.time() - t0
    extras["vq_features_minibatch"] = {"batch_rows": mb_chunk, "steps": mb_steps, "final_assign_rows": n,
                                       "wall_ms": dt * 1e3, "rows_per_s": (mb_steps * mb_chunk + n) / dt,
                                       "note": "host wall clock incl. the reference's per-step CPU torch.randint"}

    # SURVEY §8f ranks 1-2, the steps right after the clustering: join_features (5 % of the rows kept) and
    # extract_rot_scale on a joined covariance table of n/4 rows; HBM-bound, reported against the copy peak
    keep = torch.rand(n, device=dev, generator=gen) < 0.05
    cbi = torch.randint(0, K, (int((~keep).sum()),), device=dev, generator=gen)
    cbj = torch.randn(K, d, device=dev, generator=gen)
    ms = timed(lambda: ops2.join_features(x, keep, cbj, cbi))
    jbytes = n * (1 + 8) + int((~keep).sum()) * 8 + (K + int(keep.sum())) * d * 4 * 2
    extras["join_features"] = {"rows": n, "kept_frac": 0.05, "ms": ms, "gbs": jbytes / ms * 1e-6,
                               "frac_of_hbm_peak": jbytes / ms * 1e-6 / pk["hbm_gbs"],
                               "note": "two kernels + one host sync for the kept count (output size is data dependent)"}
    nr = n // 4
    a6 = torch.randn(nr, 3, 3, device=dev, generator=gen)
    a6 = a6 @ a6.transpose(-1, -2)
    cov6 = torch.stack([a6[:, 0, 0], a6[:, 0, 1], a6[:, 0, 2], a6[:, 1, 1], a6[:, 1, 2], a6[:, 2, 2]], -1).contiguous()
    ms = timed(lambda: ops2.extract_rot_scale(cov6))
    rbytes = nr * (24 + 28)
    ms_torch, n_torch = None, 0
    for n_torch in (100_000, 32_768, 4_096):             # cuSOLVER's batched syev rejects large batches on some builds
        sub = a6[:n_torch].contiguous()
        try:
            ms_torch = timed(lambda: torch.linalg.eigh(sub, UPLO="U"), reps=1)
            break
        except RuntimeError:
            continue
    extras["extract_rot_scale"] = {"rows": nr, "ms": ms, "gbs": rbytes / ms * 1e-6,
                                   "frac_of_hbm_peak": rbytes / ms * 1e-6 / pk["hbm_gbs"],
                                   "torch_eigh_ms": ms_torch, "torch_eigh_rows": n_torch,
                                   "note": "torch.linalg.eigh (the reference's call, cuSOLVER batched) timed on "
                                           "a slice for scale; it is not on the product path"}
    xyz = torch.randn(n, 3, device=dev, generator=gen)
    ms = timed(lambda: ops2.morton_order(xyz))
    mbytes = n * 12 * 2 + 8 * n * (8 + 8 + 12 + 12)       # min/max + keys, then 8 radix passes of (hist read, scatter read+write)
    ms_torch = timed(lambda: torch.sort(torch.randint(0, 2 ** 62, (n,), device=dev, generator=gen)), reps=2)
    extras["morton_order"] = {"rows": n, "ms": ms, "gbs": mbytes / ms * 1e-6,
                              "frac_of_hbm_peak": mbytes / ms * 1e-6 / pk["hbm_gbs"],
                              "torch_sort_int64_ms": ms_torch,
                              "note": "keys + stable LSD radix sort of (u64 key, u32 row), tiles digit-sorted in shared "
                                      "memory before the write-out; torch.sort (cub) "
                                      "of n random int64 incl. their generation, timed beside it for scale"}
    del keep, cbi, cbj, a6, cov6, sub, xyz
    tab48 = torch.randn(K + n // 20, 16, 3, device=dev, generator=gen)
    idx48 = torch.randint(0, tab48.shape[0], (n,), device=dev, generator=gen)
    ms = timed(lambda: ops2.gather(tab48, idx48))
    gbytes = n * (8 + 4 * 48)
    extras["decode_gather_sh48"] = {"rows": n, "dim": 48, "table_rows": int(tab48.shape[0]), "ms": ms,
                                    "gbs": gbytes / ms * 1e-6, "frac_of_hbm_peak": gbytes / ms * 1e-6 / pk["hbm_gbs"],
                                    "frac_vs_write_only_peak": gbytes / ms * 1e-6 / pk["write_only_gbs"]
                                    if pk.get("write_only_gbs") else None}
    del tab48, idx48

    # SURVEY §8f ranks 1 and 3, the formats either side of the path: normalised covariance, mask compaction,
    # int8 scene-file quantisation (algorithmic bytes / time against the measured copy peak)
    s3 = torch.rand(n, 3, device=dev, generator=gen) + 0.05
    r4 = torch.randn(n, 4, device=dev, generator=gen)
    ms = timed(lambda: ops2.build_covariance(s3, r4, activate=True))
    extras["build_covariance"] = {"rows": n, "ms": ms, "gbs": n * 52 / ms * 1e-6,
                                  "frac_of_hbm_peak": n * 52 / ms * 1e-6 / pk["hbm_gbs"]}
    keep2 = torch.rand(n, device=dev, generator=gen) < 0.6
    ms = timed(lambda: ops2.compact([x], keep2))
    cbytes = n + int(keep2.sum()) * d * 4 * 2
    extras["mask_splats_compact"] = {"rows": n, "kept_frac": 0.6, "row_bytes": d * 4, "ms": ms,
                                     "gbs": cbytes / ms * 1e-6, "frac_of_hbm_peak": cbytes / ms * 1e-6 / pk["hbm_gbs"],
                                     "note": "mask scan + host sync for the kept count + row move"}
    ms = timed(lambda: ops2.quantize_i8(x, 0.03125, 3))
    ms_t = timed(lambda: torch.quantize_per_tensor(x, 0.03125, 3, torch.qint8).int_repr())
    extras["quantize_i8"] = {"elements": x.numel(), "ms": ms, "gbs": x.numel() * 5 / ms * 1e-6,
                             "frac_of_hbm_peak": x.numel() * 5 / ms * 1e-6 / pk["hbm_gbs"],
                             "torch_quantize_per_tensor_ms": ms_t,
                             "note": "bit-exact with torch's CUDA quantiser (fp64 division there; fp32 fast path "
                                     "with an exact fp64 slow path here), timed beside it"}
    del s3, r4, keep2

    # BASELINE configs[4]: the full compress pipeline on a synthetic 6M-Gaussian scene — prune, thresholds, both
    # codebooks with the reference's default schedule shape (mini-batches), join, rotation/scaling extraction,
    # decode of every attribute, Morton order + int8 scene tensors (file IO itself excluded: host numpy/zlib)
    from c3dgs_b200 import CompressionSettings, SplatScene, compress_gaussians
    npipe = 6_000_000
    feats = torch.randn(npipe, 16, 3, device=dev, generator=gen)
    scene = SplatScene(torch.randn(npipe, 3, device=dev, generator=gen), feats[:, :1].contiguous(),
                       feats[:, 1:].contiguous(), torch.randn(npipe, 1, device=dev, generator=gen),
                       torch.rand(npipe, 3, device=dev, generator=gen) + 0.05,
                       torch.randn(npipe, 4, device=dev, generator=gen),
                       torch.randn(npipe, 1, device=dev, generator=gen) - 3)
    del feats
    cimp = torch.rand(npipe, device=dev, generator=gen) ** 4
    gimp = torch.rand(npipe, device=dev, generator=gen) ** 4
    p_steps, p_batch = 100, 2 ** 18
    cc = CompressionSettings(K, 0.0, 0.6, 0.9, p_steps, DECAY, p_batch)
    gc_ = CompressionSettings(K, 0.0, 0.3, 0.9, p_steps, DECAY, p_batch)
    torch.cuda.synchronize()
    t0 = time.time()
    compress_gaussians(scene, cimp, gimp, cc, gc_, color_compress_non_dir=True, prune_threshold=1e-8, silent=True)
    torch.cuda.synchronize()
    t1 = time.time()
    from c3dgs_b200.decode import decode_gaussians
    sdec, rdec, fdec = decode_gaussians(scene._scaling, scene._rotation,
                                        torch.cat((scene._features_dc, scene._features_rest), 1),
                                        scene._gaussian_indices, scene._feature_indices)
    torch.cuda.synchronize()
    t2 = time.time()
    scene._sort_morton()
    qbytes = 0
    for name in ("features_dc", "features_rest", "opacity", "scaling", "scaling_factor", "rotation"):
        t = scene._stored(name)
        qbytes += ops2.quantize_i8(t, 0.01, 0).numel()
    half = ops2.convert(scene._xyz, torch.float16)
    i32 = ops2.convert(scene._feature_indices, torch.int32)
    torch.cuda.synchronize()
    t3 = time.time()
    extras["compress_pipeline"] = {
        "workload": "configs[4]: synthetic 6M-Gaussian scene, SH 48-D + shape 6-D codebooks (K=4096, "
                    f"{p_steps} mini-batch steps of {p_batch} rows each + final full assignment), join, "
                    "rot/scale, decode gather, Morton order, int8/fp16/int32 scene tensors",
        "gaussians": npipe, "after_prune": int(scene._xyz.shape[0]),
        "color_table_rows": int(scene._features_dc.shape[0]), "shape_table_rows": int(scene._rotation.shape[0]),
        "compress_gaussians_ms": (t1 - t0) * 1e3, "decode_ms": (t2 - t1) * 1e3,
        "morton_and_wire_ms": (t3 - t2) * 1e3, "total_ms": (t3 - t0) * 1e3,
        "decoded_bytes": int(sdec.numel() + rdec.numel() + fdec.numel()) * 4, "int8_bytes": int(qbytes),
        "note": "host wall clock, single run, includes workspace allocation and the host syncs the data-dependent "
                "output sizes need"}
    extras["compress_pipeline_reference"] = reference_pipeline(dev, gen, npipe, cimp, gimp, p_steps, p_batch)
    # the indexed rasteriser's per-Gaussian preprocess on the compressed scene just built,
    # forward and backward (the backward is where finetuning writes gradients into the two codebooks)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from preprocess_scene import camera
    cam = camera(width=1297, height=840, fovx=0.9)
    camt = {k_: torch.from_numpy(v_).to(dev) for k_, v_ in cam.items() if hasattr(v_, "dtype")}
    table = torch.cat((scene._features_dc, scene._features_rest), 1).contiguous()
    npre = scene._xyz.shape[0]
    pre = {}
    factors, opac = torch.exp(scene._scaling_factor), torch.sigmoid(scene._opacity)

    def run_pre():
        pre.update(ops2.preprocess_indexed(scene._xyz, scene._scaling, factors, scene._rotation, opac, table,
                                           scene._feature_indices, scene._gaussian_indices, camt["viewmatrix"],
                                           camt["projmatrix"], camt["cam_pos"], cam["width"], cam["height"],
                                           cam["tan_fovx"], cam["tan_fovy"]))
    ms = timed(run_pre)
    nvis = int((pre["radii"] > 0).sum())
    pbytes = npre * 36 + nvis * (52 + 192 + 51)
    extras["preprocess_indexed"] = {"gaussians": npre, "visible": nvis, "ms": ms, "gbs": pbytes / ms * 1e-6,
                                    "frac_of_hbm_peak": pbytes / ms * 1e-6 / pk["hbm_gbs"],
                                    "note": "forward; zero-filling of the culled rows included"}
    g2d = torch.randn(npre, 3, device=dev, generator=gen)
    gcon = torch.randn(npre, 2, 2, device=dev, generator=gen)
    gcol = torch.randn(npre, 3, device=dev, generator=gen)
    ms = timed(lambda: ops2.preprocess_indexed_backward(
        scene._xyz, pre["radii"], table, pre["clamped"], scene._scaling, factors, scene._rotation, pre["cov3d"],
        camt["viewmatrix"], camt["projmatrix"], camt["cam_pos"], cam["width"], cam["height"], cam["tan_fovx"],
        cam["tan_fovy"], g2d, gcon, gcol, scene._feature_indices, scene._gaussian_indices))
    bbytes = npre * 4 + nvis * (12 + 24 + 16 + 12 + 12 + 3 + 4 + 16 + 28 + 192) + npre * (12 + 24 + 4) + nvis * (192 + 28)
    extras["preprocess_indexed_backward"] = {
        "gaussians": npre, "visible": nvis, "ms": ms, "gbs": bbytes / ms * 1e-6,
        "frac_of_hbm_peak": bbytes / ms * 1e-6 / pk["hbm_gbs"],
        "note": "one kernel (the reference launches two) + zero-filling of the three table gradients; bytes = inputs "
                "read + per-Gaussian gradients written + table rows read and reduced into"}
    return extras


def reference_pipeline(dev, gen, npipe, cimp, gimp, p_steps, p_batch):
    """BASELINE configs[4] "end-to-end vs reference": the reference's OWN compress_gaussians (compression/vq.py:194-223,
    staged unmodified under oracle/_ref/c3dgs_ref by `make -C oracle ref_py`) with its own vq_features loop, its own
    CUDA kernel (oracle/_ref/weighted_distance_ref, compiled unmodified) and index_add_ standing in for the un-vendored
    torch_scatter, on a scene with the same shapes, settings and schedule as extras.compress_pipeline.  Baseline leg
    only; nothing here is on the product path."""
    ref_dir = os.path.join(ROOT, "oracle", "_ref")
    py_dir = os.path.join(ref_dir, "c3dgs_ref")
    if not (os.path.exists(os.path.join(py_dir, "compression", "vq.py")) and
            os.path.exists(os.path.join(ref_dir, "weighted_distance_ref", "_C.so"))):
        return {"unavailable": "oracle/_ref/c3dgs_ref or oracle/_ref/weighted_distance_ref not built (make -C oracle ref ref_py)"}
    import importlib
    import types
    saved = {k_: sys.modules.get(k_) for k_ in ("torch_scatter", "weighted_distance", "weighted_distance._C",
                                                 "compression", "compression.vq", "utils", "utils.splats")}
    sys.path.insert(0, ref_dir)
    sys.path.insert(0, py_dir)
    try:
        from weighted_distance_ref import _C as refk

        def scatter(src, index, dim=0, reduce="sum", dim_size=None):         # torch_scatter.scatter(reduce="sum"), dim 0
            out = torch.zeros((dim_size,) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
            return out.index_add_(0, index, src)

        ts = types.ModuleType("torch_scatter")
        ts.scatter = scatter
        wd, wdc = types.ModuleType("weighted_distance"), types.ModuleType("weighted_distance._C")
        wdc.weightedDistance = refk.weightedDistance
        wd._C = wdc
        sys.modules.update({"torch_scatter": ts, "weighted_distance": wd, "weighted_distance._C": wdc})
        for k_ in ("compression", "compression.vq", "utils", "utils.splats"):
            sys.modules.pop(k_, None)
        rvq = importlib.import_module("compression.vq")
        from c3dgs_b200 import SplatScene
        feats = torch.randn(npipe, 16, 3, device=dev, generator=gen)
        scene = SplatScene(torch.randn(npipe, 3, device=dev, generator=gen), feats[:, :1].contiguous(),
                           feats[:, 1:].contiguous(), torch.randn(npipe, 1, device=dev, generator=gen),
                           torch.rand(npipe, 3, device=dev, generator=gen) + 0.05,
                           torch.randn(npipe, 4, device=dev, generator=gen),
                           torch.randn(npipe, 1, device=dev, generator=gen) - 3)
        del feats
        cc = rvq.CompressionSettings(K, 0.0, 0.6, 0.9, p_steps, DECAY, p_batch)
        gc_ = rvq.CompressionSettings(K, 0.0, 0.3, 0.9, p_steps, DECAY, p_batch)
        out = {"workload": "the reference's compress_gaussians on a scene of the same shapes, settings and schedule",
               "gaussians": npipe}
        torch.cuda.synchronize()
        t0 = time.time()
        with torch.no_grad():
            # compress_gaussians, :194-223, stage by stage so that a failure of the last one (torch.linalg.eigh on a
            # table of 10^6 3x3 matrices: cuSOLVER's batched syev rejects large batches) still leaves the rest timed
            mask = cimp > 1e-8
            scene.mask_splats(mask)
            ci, gi = cimp[mask], gimp[mask]
            rvq.compress_color(scene, ci, cc, True, silent=True)
            torch.cuda.synchronize()
            t1 = time.time()
            out["compress_color_ms"] = (t1 - t0) * 1e3
            try:
                rvq.compress_covariance(scene, gi, gc_, silent=True)
                torch.cuda.synchronize()
                out["compress_covariance_ms"] = (time.time() - t1) * 1e3
            except Exception as exc:  # noqa: BLE001
                torch.cuda.synchronize()
                out["compress_covariance_ms"] = None
                out["compress_covariance_error"] = repr(exc)[:300]
                # the clustering half of it, which is what the path replaces (vq.py:158-171)
                cov = scene.get_normalized_covariance(strip_sym=True).detach()
                keep = gi > gc_.importance_include
                t2 = time.time()
                rvq.vq_features(cov[~keep], gi[~keep], gc_.codebook_size, gc_.batch_size, gc_.steps,
                                scale_normalize=True, silent=True)
                torch.cuda.synchronize()
                out["covariance_vq_features_ms"] = (time.time() - t2) * 1e3
        out["total_ms"] = (time.time() - t0) * 1e3
        out["note"] = ("host wall clock, single run; reference Python + reference CUDA kernel (both unmodified) + torch "
                       "index_add_ for torch_scatter; mask_splats / get_normalized_covariance are this package's "
                       "(SplatScene), as in extras.compress_pipeline")
        return out
    except Exception as exc:  # noqa: BLE001
        return {"unavailable": repr(exc)[:300]}
    finally:
        sys.path.remove(ref_dir)
        sys.path.remove(py_dir)
        for k_, v_ in saved.items():
            if v_ is None:
                sys.modules.pop(k_, None)
            else:
                sys.modules[k_] = v_
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
