"""Row-sharded clustering across GPUs (the reference has no distributed code).

One process per GPU.  Gaussians (rows of features / importance) are split into contiguous
blocks, the codebook and entry_importance are replicated.  Each iteration every rank runs the
fused assignment + weighted-sum kernel on its rows, the K x (D+1) partial sums are combined by
ONE all-reduce (NCCL over NVLink/NVSwitch; 768 KB at K=4096, D=45), and every rank applies the
identical EMA epilogue to bitwise-identical inputs, so replicas stay in sync without a broadcast.
Two more scalar collectives happen once per clustering: max(importance) (vq.py:59) and
min/max(features) for uniform_init (vq.py:25).

`ops` is the set of local kernels; the default is the CUDA library.  Tests inject a CPU
implementation to exercise the collective plumbing under gloo — the product itself never runs
on CPU tensors (the default ops raise for them).
"""
from __future__ import annotations

import contextlib
import os
from typing import Optional, Tuple

import torch
import torch.distributed as dist

from . import _lib


@contextlib.contextmanager
def gpu_local_host_memory(device_index: int):
    """Run the enclosed host allocations on the CPUs next to GPU `device_index` (NVML's ideal CPU set), so that pinned
    staging buffers land on that GPU's NUMA node: with one process per GPU and eight uploads at once, buffers on the
    far socket share one inter-socket link.  The previous affinity is restored on exit (the pages stay where they
    were allocated).  Does nothing when NVML or the topology is unavailable."""
    old = None
    try:
        import pynvml
        pynvml.nvmlInit()
        props = torch.cuda.get_device_properties(device_index)
        bus_id = f"{props.pci_domain_id:08x}:{props.pci_bus_id:02x}:{props.pci_device_id:02x}.0"
        handle = pynvml.nvmlDeviceGetHandleByPciBusId(bus_id.encode())
        words = (os.cpu_count() + 63) // 64
        mask = pynvml.nvmlDeviceGetCpuAffinity(handle, words)
        cpus = {64 * i + b for i, word in enumerate(mask) for b in range(64) if (int(word) >> b) & 1}
        allowed = os.sched_getaffinity(0)
        target = cpus & allowed
        if target and target != allowed:
            old = allowed
            os.sched_setaffinity(0, target)
    except Exception:                            # no NVML / no topology in this container: leave the affinity alone
        old = None
    try:
        yield
    finally:
        if old is not None:
            os.sched_setaffinity(0, old)


def shard_bounds(n: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [lo, hi) of rank `rank`; the first n % world ranks get one extra row."""
    base, rem = divmod(n, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


class CudaOps:
    """Local kernels backed by libc3vq.so."""

    def __init__(self):
        self.tensor_cores = _lib.default_tensor_cores()

    def new_acc(self, k: int, d: int, device) -> torch.Tensor:
        return torch.zeros(k, _lib.acc_stride(d), dtype=torch.float32, device=device)

    def assign_accumulate(self, x, w, codebook, acc, rows=None, return_dist=False, rows_checked=False):
        _, dist_, idx = _lib.assign_accumulate(x, w, codebook, acc=acc, rows=rows, tensor_cores=self.tensor_cores,
                                               return_dist=return_dist, return_idx=True, rows_checked=rows_checked)
        return dist_, idx

    def ema(self, codebook, entry_importance, acc, decay, eps, scale_normalize):
        _lib.ema(codebook, entry_importance, acc, decay, eps, scale_normalize)

    def assign(self, x, codebook, out_idx=None):
        return _lib.assign(x, codebook, exact=not self.tensor_cores, return_dist=False, out_idx=out_idx)[1]

    def accumulate(self, x, w, idx, acc):
        _lib.accumulate(x, w, idx, acc.size(0), acc=acc, sorted_path=bool(self.tensor_cores))

    def assign_accumulate_chunk(self, x, w, codebook, acc, out_idx):
        """Assignment of a chunk of rows into `out_idx` (a slice of the caller's index array) with the chunk's
        weighted sums added into acc (update_from_host)."""
        _lib.assign_accumulate(x, w, codebook, acc=acc, tensor_cores=self.tensor_cores, return_idx=True, out_idx=out_idx)


class ShardedVectorQuantize:
    """VectorQuantize (compression/vq.py:15-42) with rows sharded over a process group."""

    def __init__(self, channels: int, codebook_size: int = 2 ** 12, decay: float = 0.5, device=None,
                 group: Optional[dist.ProcessGroup] = None, ops=None) -> None:
        self.decay = decay
        self.eps = 1e-5
        self.group = group
        self.ops = ops or CudaOps()
        self.codebook = torch.zeros(codebook_size, channels, dtype=torch.float32, device=device)
        self.entry_importance = torch.zeros(codebook_size, dtype=torch.float32, device=device)
        self._acc = self.ops.new_acc(codebook_size, channels, device)

    # -- collectives -------------------------------------------------------------------------
    def _world(self) -> int:
        return dist.get_world_size(self.group) if dist.is_initialized() else 1

    def _all_reduce(self, t: torch.Tensor, op=dist.ReduceOp.SUM) -> None:
        if self._world() > 1:
            dist.all_reduce(t, op=op, group=self.group)

    # -- reference-shaped API ------------------------------------------------------------------
    def uniform_init(self, x_local: torch.Tensor, generator: Optional[torch.Generator] = None) -> None:
        """vq.py:24-26 with a global min/max.  The random codebook is drawn on rank 0 and broadcast
        (device RNG streams differ per process), so all replicas start identical."""
        if x_local.numel():
            amin, amax = x_local.aminmax()
        else:
            amin = torch.tensor(float("inf"), device=x_local.device)
            amax = torch.tensor(float("-inf"), device=x_local.device)
        amin, amax = amin.clone(), amax.clone()
        self._all_reduce(amin, dist.ReduceOp.MIN)
        self._all_reduce(amax, dist.ReduceOp.MAX)
        cb = torch.rand(self.codebook.shape, dtype=torch.float32, device=self.codebook.device,
                        generator=generator) * (amax - amin) + amin
        if self._world() > 1:
            dist.broadcast(cb, src=dist.get_global_rank(self.group, 0) if self.group is not None else 0,
                           group=self.group)
        self.codebook.copy_(cb)

    def update(self, x_local: torch.Tensor, importance_local: torch.Tensor, rows: Optional[torch.Tensor] = None,
               scale_normalize: bool = False, return_dist: bool = False):
        """One clustering iteration over all ranks' rows (vq.py:28-35 + :73-77)."""
        self._acc.zero_()
        dist_, idx = self.ops.assign_accumulate(x_local, importance_local, self.codebook, self._acc, rows=rows,
                                                return_dist=return_dist)
        self._all_reduce(self._acc)                                   # the one exchange step per iteration
        self.ops.ema(self.codebook, self.entry_importance, self._acc, self.decay, self.eps, scale_normalize)
        return dist_, idx

    def update_from_host(self, x_host: torch.Tensor, importance_host: torch.Tensor, x_dev: torch.Tensor,
                         importance_dev: torch.Tensor, chunk_rows: int = 1 << 19, scale_normalize: bool = False):
        """The first iteration when the rows still sit in (pinned) host memory: the upload runs in chunks on a copy
        stream and every chunk is assigned as soon as it has landed, so the assignment hides under the PCIe
        transfer; the sums, the all-reduce and the EMA follow once all rows are resident.  Same kernels and the
        same result as `update(x_dev, importance_dev)` after a plain copy; later iterations use `update`.
        `importance_host` is already normalised (vq.py:59).  Returns the assignments (int64 [n])."""
        dev = x_dev.device
        n = x_host.shape[0]
        main = torch.cuda.current_stream(dev)
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
        copy = self._copy_stream
        copy.wait_stream(main)                                   # earlier work may still read x_dev
        idx = torch.empty(n, dtype=torch.int64, device=dev)
        self._acc.zero_()
        fused = getattr(self.ops, "assign_accumulate_chunk", None) is not None
        if fused:
            with torch.cuda.stream(copy):                        # 4 bytes per row: ahead of the rows themselves
                importance_dev.copy_(importance_host, non_blocking=True)
        for lo in range(0, n, chunk_rows):
            hi = min(lo + chunk_rows, n)
            with torch.cuda.stream(copy):
                x_dev[lo:hi].copy_(x_host[lo:hi], non_blocking=True)
                landed = copy.record_event()
            main.wait_event(landed)
            if fused:       # assignment with the chunk's weighted sums added into acc by the same kernel
                self.ops.assign_accumulate_chunk(x_dev[lo:hi], importance_dev[lo:hi], self.codebook, self._acc,
                                                 idx[lo:hi])
            else:
                self.ops.assign(x_dev[lo:hi], self.codebook, out_idx=idx[lo:hi])
        if not fused:
            with torch.cuda.stream(copy):
                importance_dev.copy_(importance_host, non_blocking=True)
                landed = copy.record_event()
            main.wait_event(landed)
            self.ops.accumulate(x_dev, importance_dev, idx, self._acc)
        self._all_reduce(self._acc)
        self.ops.ema(self.codebook, self.entry_importance, self._acc, self.decay, self.eps, scale_normalize)
        return idx

    def assign(self, x_local: torch.Tensor) -> torch.Tensor:
        return self.ops.assign(x_local, self.codebook)


def vq_features_sharded(
        features_local: torch.Tensor,
        importance_local: torch.Tensor,
        codebook_size: int,
        vq_chunk: int = 2 ** 16,
        steps: int = 1000,
        decay: float = 0.8,
        scale_normalize: bool = False,
        n_global: Optional[int] = None,
        row_offset: Optional[int] = None,
        full_batch: bool = False,
        group: Optional[dist.ProcessGroup] = None,
        ops=None,
        codebook0: Optional[torch.Tensor] = None,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """vq_features (compression/vq.py:49-87) on row shards; returns (codebook, local indices).

    Mini-batch mode reproduces the reference's draw: every rank draws the SAME global batch
    `torch.randint(0, N, [vq_chunk])` from its CPU generator (callers seed all ranks alike) and keeps
    the indices that fall in its shard, so the union over ranks is exactly the reference's batch.
    `full_batch=True` uses every local row each iteration (the benchmark's "k-means iteration").
    """
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    n_local = features_local.shape[0]
    dev = features_local.device
    if n_global is None or row_offset is None:
        counts = torch.zeros(world, dtype=torch.int64, device=dev)
        counts[rank] = n_local
        if world > 1:
            dist.all_reduce(counts, group=group)
        n_global = int(counts.sum().item())
        row_offset = int(counts[:rank].sum().item())

    vq = ShardedVectorQuantize(features_local.shape[-1], codebook_size, decay, device=dev, group=group, ops=ops)
    wmax = importance_local.max().clone() if n_local else torch.tensor(float("-inf"), device=dev)
    vq._all_reduce(wmax, dist.ReduceOp.MAX)
    importance_n = (importance_local / wmax).contiguous()                                  # vq.py:59
    if codebook0 is not None:
        vq.codebook.copy_(codebook0)
    else:
        vq.uniform_init(features_local)                                                    # vq.py:62

    for _ in range(steps):
        if full_batch:
            vq.update(features_local, importance_n, scale_normalize=scale_normalize)
        else:
            batch = torch.randint(low=0, high=n_global, size=[vq_chunk])                   # vq.py:69, same on all ranks
            mine = batch[(batch >= row_offset) & (batch < row_offset + n_local)] - row_offset
            vq.update(features_local, importance_n, rows=mine.to(dev), scale_normalize=scale_normalize)
    return vq.codebook, vq.assign(features_local)                                          # vq.py:82,87
