#!/usr/bin/env python
"""bench.py — Gaussian*iterations/s of the sensitivity-aware VQ clustering hot path.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

One "step" = one 10-iteration clustering job (10 x VectorQuantize.update over ALL rows, full batch,
K = 4096, decay 0.8) on synthetic inputs shaped like BASELINE.json configs[1]: 5.8 M Gaussians x 45-D
SH features with heavy-tailed sensitivities.  With N GPUs every rank holds its own 5.8 M rows (weak
scaling), the codebook is replicated and each iteration all-reduces the K x (D+1) partial sums (NCCL).

The JSON line carries, besides the base contract: `roofline` (dominant kernel, measured live with
CUDA events around every launch of it inside the timed region), `cpu_baseline` (the CPU oracle port
of the reference path on the box's host cores, rank 0, N=1), `e2e` (same job through the public API
from pinned HOST buffers, H2D/D2H inside the timed region) and `gpu_launches`.

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


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return {"hbm_gbs": j["hbm_gbs"], "bf16_tflops": j["bf16_tflops"],
                "bf16_tflops_sustained": j.get("bf16_tflops_sustained", j["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


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


def synth(n: int, d: int, seed: int, device=None, pin=False):
    """Seeded synthetic scene: clustered SH-like features with band decay, w = rand**4.
    Generated on the device generator for speed (CPU generation of 5.8M x 45 takes seconds per rank)."""
    if device is None:
        g = torch.Generator().manual_seed(seed)
        centres = torch.randn(8192, d, generator=g)
        x = centres[torch.randint(0, 8192, (n,), generator=g)] + 0.05 * torch.randn(n, d, generator=g)
        w = torch.rand(n, generator=g) ** 4
    else:
        g = torch.Generator(device=device).manual_seed(seed)
        centres = torch.randn(8192, d, device=device, generator=g)
        band = torch.arange(d, device=device) // 3
        centres = centres / (1.0 + band.float().sqrt())
        x = centres[torch.randint(0, 8192, (n,), device=device, generator=g)]
        x = x + 0.05 * torch.randn(n, d, device=device, generator=g)
        w = torch.rand(n, device=device, generator=g) ** 4
    return x.contiguous(), w.contiguous()


def init_codebook(x: torch.Tensor, k: int, seed: int):
    """uniform_init (vq.py:24-26) with a fixed seed."""
    amin, amax = x.aminmax()
    g = torch.Generator(device=x.device).manual_seed(seed)
    return (torch.rand(k, x.shape[1], device=x.device, generator=g) * (amax - amin) + amin).contiguous()


# ------------------------------------------------------------------------------------------------
def workload_config(n, d, world, assignment, sums):
    which = "configs[2]: shape codebook clustering" if d <= 8 else "configs[1]: SH codebook clustering"
    return {"workload": f"{which}, {n / 1e6:.1f}M Gaussians x {d}-D per GPU, K=4096, "
                        "10 full-batch k-means iterations per step",
            "rows_per_gpu": n, "dim": d, "codebook_size": K, "iterations_per_step": ITERS, "decay": DECAY,
            "parallelism": f"rows sharded x{world}, codebook replicated, 1 all-reduce of K x (D+1) sums per iteration",
            "l2": f"inputs ({n * (d + 1) * 4 / 1e9:.2f} GB per GPU) exceed the 126 MB L2; no explicit flush",
            "assignment": assignment, "sums": sums}


# (rows, dim, simt) -> DRAM bytes of one launch of the dominant kernel, measured with ncu (see profiles/)
NCU_DRAM_BYTES_PER_LAUNCH = {(5_800_000, 45, False): 1_092_362_240}

SAMPLE_ROWS = 100_000     # CPU legs: a bounded sample of the workload's rows (the metric is per row and iteration)


def run_reference_arm(args, rank):
    """The reference path's CPU implementation (oracle port: OpenMP C restatement of weighted_distance.cu +
    vq.py's update) on the box's host cores, on a bounded sample of OUR arm's workload; rank 0 only."""
    if rank != 0:
        return
    from oracle import oracle
    oracle.build()
    n, d = SAMPLE_ROWS, args.dim
    x, w = synth(n, d, 0)
    x, w = x.numpy(), w.numpy()
    wn = (w / w.max()).astype(np.float32)
    rng = np.random.default_rng(0)
    cb0 = (rng.random((K, d), dtype=np.float32) * (x.max() - x.min()) + x.min()).astype(np.float32)
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
              f"{cores} host threads")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": steps, "warmup": warm + 1, "ms_per_step": dt * 1e3, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.rows, d, max(1, args.gpus), "cpu-exact (reference arithmetic)", "cpu scatter"),
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
    x, w = synth(n, d, 0)
    x, w = x.numpy(), w.numpy()
    wn = (w / w.max()).astype(np.float32)
    rng = np.random.default_rng(0)
    cb = (rng.random((K, d), dtype=np.float32) * (x.max() - x.min()) + x.min()).astype(np.float32)
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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--rows", type=int, default=5_800_000, help="rows per GPU (default: configs[1])")
    ap.add_argument("--dim", type=int, default=45)
    ap.add_argument("--simt", action="store_true", help="use the SIMT exact assignment instead of tensor cores")
    ap.add_argument("--sums", default="sorted", choices=["sorted", "fused"],
                    help="centroid sums: sorted segmented sum after the assignment (default) or fused bulk reduce-adds")
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    from c3dgs_b200.sharded import ShardedVectorQuantize, CudaOps

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus or world == 1, f"--gpus {args.gpus} but WORLD_SIZE={world}"
    _lib.load()

    n, d = args.rows, args.dim
    warm = max(3, args.warmup)
    x, w = synth(n, d, 1234 + rank, device=dev)
    wmax = w.max().clone()
    if world > 1:
        dist.all_reduce(wmax, op=dist.ReduceOp.MAX)
    wn = (w / wmax).contiguous()
    cb0 = init_codebook(x, K, 99)                       # same seed on every rank -> identical replicas
    ops = CudaOps()
    ops.tensor_cores = False if args.simt else ("fused" if args.sums == "fused" else True)
    vq = ShardedVectorQuantize(d, K, DECAY, device=dev, ops=ops)

    ev_pairs = []
    split = (not args.simt) and args.sums == "sorted"

    def job(record: bool):
        """10 iterations of update over all local rows; the dominant kernel is bracketed by events."""
        vq.codebook.copy_(cb0)
        vq.entry_importance.zero_()
        for _ in range(ITERS):
            vq._acc.zero_()
            if record:
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
            if split:
                # the same kernels c3vq_assign_accumulate launches, as two calls so that the assignment
                # (prep + k_assign_tc + fallback scan) can be bracketed on its own
                idx = _lib.assign(x, vq.codebook, return_dist=False)[1]
                if record:
                    e1.record()
                _lib.accumulate(x, wn, idx, K, acc=vq._acc, sorted_path=True)
            else:
                ops.assign_accumulate(x, wn, vq.codebook, vq._acc)
                if record:
                    e1.record()
            if record:
                ev_pairs.append((e0, e1))
            vq._all_reduce(vq._acc)
            ops.ema(vq.codebook, vq.entry_importance, vq._acc, DECAY, EPS, False)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    for _ in range(warm):
        job(False)
    barrier()
    sampler = ClockSampler(local_rank)
    if rank == 0:
        sampler.start()
        time.sleep(0.3)
    launches0 = _lib.launch_count()
    start, stop = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    t_wall0 = time.time()
    start.record()
    for _ in range(args.steps):
        job(True)
    stop.record()
    barrier()
    t_wall1 = time.time()
    launches = _lib.launch_count() - launches0
    ms_total = torch.tensor([start.elapsed_time(stop)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_total, op=dist.ReduceOp.MAX)
    ms_per_step = float(ms_total.item()) / args.steps
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    value = world * n * ITERS / (ms_per_step * 1e-3)

    # dominant kernel: fused assignment(+sums) launch, average duration over the timed region
    kern_ms = float(np.mean([a.elapsed_time(b) for a, b in ev_pairs]))
    pk = peaks()
    flops = 2.0 * K * d * n                               # assign flops = 2*K*D per Gaussian*iteration
    tf32_peak = pk["bf16_tflops_sustained"] / 2.0         # dense TF32 = half the bf16 rate; kernel timed inside a long step
    achieved = flops / (kern_ms * 1e-3) / 1e12
    # dram__bytes_read.sum + dram__bytes_write.sum of ONE k_assign_tc launch from the committed `ncu --set full`
    # capture (profiles/r1n_ncu_assign_tc_summary.txt); only valid for the configuration that capture ran
    traffic = NCU_DRAM_BYTES_PER_LAUNCH.get((n, d, bool(args.simt)))
    roofline = {"bound": "tensor", "achieved": achieved, "peak": tf32_peak, "unit": "TFLOP/s",
                "frac": achieved / tf32_peak, "traffic": traffic,
                "kernel": "k_assign_exact+k_accumulate (SIMT)" if args.simt else
                          ("k_assign_tc (assignment; sums by the sorted segmented-sum kernels)" if split
                           else "k_assign_tc (fused assign+sums)"),
                "kernel_ms": kern_ms, "kernel_share_of_step": kern_ms * ITERS / ms_per_step,
                "peak_source": f"{pk['source']}: bf16_tflops_sustained/2 (TF32 dense = half bf16)",
                "algorithmic_flops_per_launch": flops}

    # e2e: the same job from pinned HOST buffers through the public API, copies inside the timed region
    e2e = None
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
    ms_e2e = torch.tensor([s2.elapsed_time(t2)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms_e2e, op=dist.ReduceOp.MAX)
    ms_e2e = float(ms_e2e.item()) / e2e_steps
    e2e = {"value": world * n * ITERS / (ms_e2e * 1e-3), "unit": UNIT,
           "h2d_bytes_per_step": int(xh.numel() * 4 + wh.numel() * 4),
           "d2h_bytes_per_step": int(cbh.numel() * 4 + idxh.numel() * 8), "ms_per_step": ms_e2e}

    # the reference's real default, mini-batch mode, through the drop-in vq_features
    # (CPU randint batches with replacement, 100 x 2^18 rows, then the final full assignment); reported, not the metric
    extras = None
    if world == 1 and not args.simt:
        from c3dgs_b200.vq import vq_features
        mb_steps, mb_chunk = 100, 2 ** 18
        vq_features(x, w, K, vq_chunk=mb_chunk, steps=3, decay=DECAY, silent=True)      # warm-up
        torch.cuda.synchronize()
        t0 = time.time()
        vq_features(x, w, K, vq_chunk=mb_chunk, steps=mb_steps, decay=DECAY, silent=True)
        torch.cuda.synchronize()
        dt = time.time() - t0
        extras = {"vq_features_minibatch": {"batch_rows": mb_chunk, "steps": mb_steps, "final_assign_rows": n,
                                            "wall_ms": dt * 1e3,
                                            "rows_per_s": (mb_steps * mb_chunk + n) / dt,
                                            "note": "host wall clock incl. the reference's per-step CPU torch.randint"}}

        # SURVEY §8f ranks 1-2, the steps right after the clustering: join_features (5 % of the rows kept) and
        # extract_rot_scale on a joined covariance table of n/4 rows; HBM-bound, reported against the copy peak
        from c3dgs_b200 import _lib as ops2
        pk = peaks()

        def timed(fn, reps=5):
            fn()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            for _ in range(reps):
                fn()
            b.record()
            torch.cuda.synchronize()
            return a.elapsed_time(b) / reps

        gen = torch.Generator(device=dev).manual_seed(5)
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
                                  "note": "keys + stable 8x8-bit LSD radix sort of (u64 key, u32 row); torch.sort (cub) "
                                          "of n random int64 incl. their generation, timed beside it for scale"}
        del keep, cbi, cbj, a6, cov6, sub, xyz
        # the indexed-codebook decode, codebook[idx] -> [n, d] (8 B index in + 4 d B out per row)
        idx_dec = torch.randint(0, K, (n,), device=dev, generator=gen)
        cb_dec = torch.randn(K, d, device=dev, generator=gen)
        ms = timed(lambda: ops2.gather(cb_dec, idx_dec))
        gbytes = n * (8 + 4 * d)
        extras["decode_gather"] = {"rows": n, "dim": d, "ms": ms, "gbs": gbytes / ms * 1e-6,
                                   "frac_of_hbm_peak": gbytes / ms * 1e-6 / pk["hbm_gbs"]}
        tab48 = torch.randn(K + n // 20, 16, 3, device=dev, generator=gen)
        idx48 = torch.randint(0, tab48.shape[0], (n,), device=dev, generator=gen)
        ms = timed(lambda: ops2.gather(tab48, idx48))
        gbytes = n * (8 + 4 * 48)
        extras["decode_gather_sh48"] = {"rows": n, "dim": 48, "table_rows": int(tab48.shape[0]), "ms": ms,
                                        "gbs": gbytes / ms * 1e-6, "frac_of_hbm_peak": gbytes / ms * 1e-6 / pk["hbm_gbs"]}
        del idx_dec, cb_dec, tab48, idx48

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
        # the indexed rasteriser's per-Gaussian preprocess on the compressed scene just built
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from preprocess_scene import camera
        cam = camera(width=1297, height=840, fovx=0.9)
        camt = {k_: torch.from_numpy(v_).to(dev) for k_, v_ in cam.items() if hasattr(v_, "dtype")}
        table = torch.cat((scene._features_dc, scene._features_rest), 1).contiguous()
        npre = scene._xyz.shape[0]
        pre = {}

        def run_pre():
            pre.update(ops2.preprocess_indexed(scene._xyz, scene._scaling, torch.exp(scene._scaling_factor), scene._rotation,
                                               torch.sigmoid(scene._opacity), table, scene._feature_indices,
                                               scene._gaussian_indices, camt["viewmatrix"], camt["projmatrix"],
                                               camt["cam_pos"], cam["width"], cam["height"], cam["tan_fovx"],
                                               cam["tan_fovy"]))
        ms = timed(run_pre)
        nvis = int((pre["radii"] > 0).sum())
        pbytes = npre * 36 + nvis * (52 + 192 + 51)
        extras["preprocess_indexed"] = {"gaussians": npre, "visible": nvis, "ms": ms, "gbs": pbytes / ms * 1e-6,
                                        "frac_of_hbm_peak": pbytes / ms * 1e-6 / pk["hbm_gbs"],
                                        "note": "incl. the activations (exp, sigmoid: torch) and zero-filling the outputs"}
        del table, pre
        del scene, cimp, gimp, sdec, rdec, fdec, half, i32

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        cpu = cpu_baseline_leg(d)

    if rank == 0:
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": warm,
            "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f32 (tf32 tensor-core screening + exact fp32 re-rank)" if not args.simt else "f32",
            "data": "synthetic",
            "config": workload_config(n, d, world, "simt-exact" if args.simt else "tcgen05-tf32+rerank",
                                      "atomic" if args.simt else args.sums),
            "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": int(launches),
            "clocks": clocks, "extras": extras,
        }
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
