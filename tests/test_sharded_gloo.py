"""CPU, world_size 2, gloo: the row-sharded clustering plumbing (c3dgs_b200/sharded.py).

The local kernels are replaced by an oracle-backed CPU implementation (test-only injection), so
what is checked here is the host logic: shard bounds, the global batch draw split by shard, the
one all-reduce per iteration, replicated EMA — against the single-process oracle."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, golden_path


class OracleOps:
    """CPU stand-in for CudaOps built on the oracle (checker side only)."""

    def __init__(self):
        from oracle import oracle
        self.o = oracle

    def new_acc(self, k, d, device):
        return torch.zeros(k, (d + 1 + 3) & ~3, dtype=torch.float32)

    def assign_accumulate(self, x, w, codebook, acc, rows=None, return_dist=False):
        xn, wn = x.numpy(), w.numpy()
        if rows is not None:
            xn, wn = xn[rows.numpy()], wn[rows.numpy()]
        k, d = codebook.shape
        dist_, idx = self.o.assign(xn, codebook.numpy())
        sx, sw = self.o.scatter_sum(xn, wn, idx, k)
        acc[:, :d] += torch.from_numpy(sx)
        acc[:, d] += torch.from_numpy(sw)
        return torch.from_numpy(dist_), torch.from_numpy(idx)

    def ema(self, codebook, ei, acc, decay, eps, scale_normalize):
        k, d = codebook.shape
        cb, e = codebook.numpy(), ei.numpy()
        self.o.ema(cb, e, acc[:, :d].contiguous().numpy(), acc[:, d].contiguous().numpy(), decay, eps)
        if scale_normalize:
            self.o.trace_normalize(cb)

    def assign(self, x, codebook):
        return torch.from_numpy(self.o.assign(x.numpy(), codebook.numpy())[1])


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _worker(rank, world, port, case, full_batch, out_dir):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from c3dgs_b200.sharded import shard_bounds, vq_features_sharded
        g = np.load(golden_path(case))
        n, d, k, chunk, steps, sn, seed = [int(v) for v in g["meta"]]
        lo, hi = shard_bounds(n, rank, world)
        f = torch.from_numpy(g["features"][lo:hi].copy())
        w = torch.from_numpy(g["importance"][lo:hi].copy())
        torch.manual_seed(4242)                                    # same CPU stream on every rank
        cb, idx = vq_features_sharded(f, w, k, chunk, steps, decay=0.8, scale_normalize=bool(sn),
                                      full_batch=full_batch, ops=OracleOps(),
                                      codebook0=torch.from_numpy(g["codebook0"].copy()))
        np.savez(os.path.join(out_dir, f"r{rank}.npz"), cb=cb.numpy(), idx=idx.numpy(), lo=lo, hi=hi)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("case,full_batch", [("sh45", False), ("cov6", False), ("shape7", True)])
def test_two_rank_matches_single_process_oracle(oracle, tmp_path, case, full_batch):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), case, full_batch, str(tmp_path)), nprocs=world, join=True)
    g = np.load(golden_path(case))
    n, d, k, chunk, steps, sn, seed = [int(v) for v in g["meta"]]
    # single-process oracle with the same draws
    torch.manual_seed(4242)
    if full_batch:
        batches = [np.arange(n)] * steps
    else:
        batches = [torch.randint(low=0, high=n, size=[chunk]).numpy() for _ in range(steps)]
    cb1, _, idx1, _ = oracle.vq_features(g["features"], g["importance"], g["codebook0"], batches, 0.8, bool(sn))
    parts = [np.load(tmp_path / f"r{r}.npz") for r in range(world)]
    # replicas agree bit for bit with each other
    assert np.array_equal(parts[0]["cb"], parts[1]["cb"])
    # shards tile the row range
    assert parts[0]["lo"] == 0 and parts[0]["hi"] == parts[1]["lo"] and parts[1]["hi"] == n
    # codebook within tolerance of the 1-process result (partial sums are added in a different order)
    cb2 = parts[0]["cb"]
    scale = np.abs(cb1).max(1, keepdims=True) + 1e-12
    fin = np.isfinite(cb1).all(1)
    assert np.max(np.abs(cb2[fin] - cb1[fin]) / scale[fin]) < 1e-5
    idx2 = np.concatenate([p["idx"] for p in parts])
    assert (idx2 == idx1).mean() > 0.995                       # near-ties may flip with 1-ulp codebook changes


def test_shard_bounds_cover_and_balance():
    from c3dgs_b200.sharded import shard_bounds
    for n in [0, 1, 7, 8, 9, 24_000_000]:
        for world in [1, 2, 3, 8]:
            spans = [shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(spans[i][1] == spans[i + 1][0] for i in range(world - 1))
            sizes = [b - a for a, b in spans]
            assert max(sizes) - min(sizes) <= 1
