"""Host-side mirror of the reference's clustering interface (compression/vq.py:15-103).

Same names, positional order, defaults, return arity and dtypes as the reference, so
`compress_color` / `compress_covariance` / `compress_gaussians` (vq.py:117-223) and their
callers (compress.py:254, scene/gaussian_model.py:984) run unchanged on top of it.  All
arithmetic happens in libc3vq.so (CUDA, sm_100a) through c3dgs_b200._lib; there is no
CPU or torch fallback for it.
"""
from __future__ import annotations

import queue
import threading
import time
from typing import Optional, Tuple

import torch
from torch import nn

from . import _lib


class VectorQuantize(nn.Module):
    """Drop-in for compression/vq.py:15-42 (codebook [K, D] + entry_importance [K], eps = 1e-5)."""

    def __init__(self, channels: int, codebook_size: int = 2 ** 12, decay: float = 0.5) -> None:
        super().__init__()
        self.decay = decay
        self.codebook = nn.Parameter(torch.empty(codebook_size, channels), requires_grad=False)
        nn.init.kaiming_uniform_(self.codebook)
        self.entry_importance = nn.Parameter(torch.zeros(codebook_size), requires_grad=False)
        self.eps = 1e-5
        # product knobs (not in the reference): tensor-core screening on/off
        self.tensor_cores = _lib.default_tensor_cores()

    def uniform_init(self, x: torch.Tensor):
        # vq.py:24-26 — global scalar min/max, device RNG
        amin, amax = x.aminmax()
        self.codebook.data = torch.rand_like(self.codebook) * (amax - amin) + amin

    def update(self, x: torch.Tensor, importance: torch.Tensor) -> torch.Tensor:
        """vq.py:28-35: assignment, weighted scatter sums, two EMAs; returns min_dists."""
        with torch.no_grad():
            min_dists, _ = _lib.update(x.detach(), importance.detach(), self.codebook.data,
                                       self.entry_importance.data, self.decay, self.eps,
                                       tensor_cores=self.tensor_cores, return_dist=True)
            return min_dists

    def update_indexed(self, features: torch.Tensor, importance: torch.Tensor, batch: torch.Tensor,
                       scale_normalize: bool = False, return_dist: bool = False, batch_checked: bool = False):
        """`update(features[batch], importance[batch])` (+ optional vq.py:73-77 trace normalise) without
        materialising the gathered rows: the kernels read rows through `batch`.  `batch_checked`: the caller
        guarantees 0 <= batch < len(features) (skips the range check and its host sync)."""
        with torch.no_grad():
            min_dists, _ = _lib.update(features.detach(), importance.detach(), self.codebook.data,
                                       self.entry_importance.data, self.decay, self.eps, rows=batch,
                                       scale_normalize=scale_normalize, tensor_cores=self.tensor_cores,
                                       return_dist=return_dist, rows_checked=batch_checked)
            return min_dists

    def forward(self, x: torch.Tensor, return_dists: bool = False
                ) -> Tuple[torch.Tensor, torch.Tensor, Optional[torch.Tensor]]:
        # vq.py:37-42
        min_dists, idx = _lib.assign(x.detach(), self.codebook.detach(), exact=not self.tensor_cores,
                                     return_dist=return_dists)
        quantized = _lib.gather(self.codebook.detach(), idx)
        if return_dists:
            return quantized, idx, min_dists
        else:
            return quantized, idx

    def assign(self, x: torch.Tensor) -> torch.Tensor:
        """Indices only (what vq_features keeps of forward(), vq.py:82)."""
        _, idx = _lib.assign(x.detach(), self.codebook.detach(), exact=not self.tensor_cores, return_dist=False)
        return idx


def ema_inplace(moving_avg: torch.Tensor, new: torch.Tensor, decay: float):
    # vq.py:45-46 (kept for callers that import it; the clustering loop fuses it into c3vq_ema)
    moving_avg.data.mul_(decay).add_(new, alpha=(1 - decay))


def vq_features(
        features: torch.Tensor,
        importance: torch.Tensor,
        codebook_size: int,
        vq_chunk: int = 2 ** 16,
        steps: int = 1000,
        decay: float = 0.8,
        scale_normalize: bool = False,
        silent: bool = False,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Drop-in for compression/vq.py:49-87.

    Same random stream as the reference: uniform_init draws from the device generator,
    every step draws `torch.randint(low=0, high=N, size=[vq_chunk])` from the CPU generator
    (with replacement; the draws are made in step order on one helper thread, ahead of the launches).  The per-step `.mean().item()` of the reference only feeds a list
    that is never read (vq.py:64,71-72), so no host sync is made here.
    """
    if not features.is_cuda:
        raise RuntimeError("vq_features needs CUDA tensors (libc3vq has no CPU path)")
    features = features.detach().contiguous()
    importance_n = (importance / importance.max()).contiguous()                      # vq.py:59

    vq_model = VectorQuantize(channels=features.shape[-1], codebook_size=codebook_size,
                              decay=decay).to(device=features.device)                # vq.py:61
    vq_model.uniform_init(features)                                                  # vq.py:62

    diap = range(steps)
    if not silent:
        from tqdm import tqdm
        diap = tqdm(diap)
    # vq.py:69 draws every batch with the CPU generator (0.8 ms per 2^18 indices, more than the GPU needs for the
    # step), so the draws run ahead on a helper thread — one thread, in step order, hence the same random stream —
    # into a small ring of pinned buffers while this thread uploads and launches.
    ring = 4
    pinned = [torch.empty(vq_chunk, dtype=torch.int64).pin_memory() for _ in range(ring)]
    uploaded = [None] * ring                    # event: the H2D copy that last read pinned[slot] has finished
    drawn: "queue.Queue" = queue.Queue(maxsize=ring - 1)
    free: "queue.Queue" = queue.Queue()
    for slot in range(ring):
        free.put(slot)
    n_rows = features.shape[0]

    def draw_batches():
        try:
            for _ in range(steps):
                slot = free.get()
                if slot is None:
                    return
                if uploaded[slot] is not None:
                    uploaded[slot].synchronize()
                torch.randint(low=0, high=n_rows, size=[vq_chunk], out=pinned[slot])       # vq.py:69
                drawn.put(slot)
        except BaseException as exc:            # surface the failure in the consumer instead of hanging it
            drawn.put(exc)

    drawer = threading.Thread(target=draw_batches, name="c3vq-batch-draw", daemon=True)
    drawer.start()
    try:
        for step in diap:
            slot = drawn.get()
            if isinstance(slot, BaseException):
                raise slot
            batch = pinned[slot].to(features.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(features.device))
            uploaded[slot] = ev
            free.put(slot)
            # vq.py:70-77: gather + update (+ trace normalise), one fused call
            vq_model.update_indexed(features, importance_n, batch, scale_normalize=scale_normalize,
                                    batch_checked=True)              # drawn by randint(0, n_rows) above
    finally:
        free.put(None)                          # normal end: the drawer has left already; after an error: stop it
        while drawer.is_alive():
            try:
                drawn.get_nowait()              # it may be blocked handing over a batch nobody will take
            except queue.Empty:
                pass
            drawer.join(timeout=0.01)
    # vq.py:79-80 runs gc.collect() + torch.cuda.empty_cache() here to make room for the three [N, D] temporaries
    # of its update; this path has none (the kernels read rows in place), and returning the allocator's blocks to the
    # driver costs more than the final assignment, so the cache is left alone.

    start = time.time()
    vq_indices = vq_model.assign(features)                                           # vq.py:82
    torch.cuda.synchronize(device=vq_indices.device)
    end = time.time()
    if not silent:
        print(f"calculating indices took {end - start} seconds ")
    return vq_model.codebook.data.detach(), vq_indices.detach()


def join_features(
        all_features: torch.Tensor,
        keep_mask: torch.Tensor,
        codebook: torch.Tensor,
        codebook_indices: torch.Tensor,
) -> Tuple[torch.Tensor, torch.Tensor]:
    """Drop-in for compression/vq.py:90-103: stable compaction of the kept rows behind the
    codebook and the index remap, in two kernels of libc3vq (c3vq_join_count / c3vq_join_write) instead of two
    boolean-mask gathers, a cat, a zeros, two masked scatters and an arange."""
    if codebook.dim() != 2:                      # the reference's empty covariance codebook is [0, C, 1] (vq.py:172)
        codebook = codebook.reshape(codebook.shape[0], -1)
    return _lib.join_features(all_features.detach(), keep_mask, codebook.detach(), codebook_indices)
