"""Multi-GPU parity: python -m torch.distributed.run --nproc-per-node N --master-addr 127.0.0.1 tools/nccl_check.py
Every rank clusters its row shard (NCCL all-reduce of the centroid sums per iteration); rank 0 also clusters the
whole data set alone.  Codebooks must agree within 1e-5 (row-normalised) and the shard's assignments must equal
the single-GPU assignments (given the sharded codebook they are recomputed exactly, so equality is bit-exact)."""
import os
import sys

import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c3dgs_b200 import _lib  # noqa: E402
from c3dgs_b200.sharded import vq_features_sharded, shard_bounds  # noqa: E402
from bench import synth  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
ok = True
for d, k, n in ((45, 4096, 1_000_003), (7, 512, 300_001)):
    x, w = synth(n, d, 7)                        # CPU generator: identical on every rank
    g = torch.Generator().manual_seed(1)
    cb0 = (torch.rand(k, d, generator=g) * (x.max() - x.min()) + x.min())
    lo, hi = shard_bounds(n, rank, world)
    cb_s, idx_s = vq_features_sharded(x[lo:hi].to(dev), w[lo:hi].to(dev), k, steps=6, decay=0.8, full_batch=True,
                                      codebook0=cb0.to(dev))
    if rank == 0:
        saved = dist.group.WORLD
        # single-GPU run of the same job (no collectives: a private ShardedVectorQuantize with world 1 semantics)
        from c3dgs_b200.sharded import ShardedVectorQuantize
        vq = ShardedVectorQuantize(d, k, 0.8, device=dev)
        vq._world = lambda: 1
        vq.codebook.copy_(cb0.to(dev))
        xa, wa = x.to(dev), (w / w.max()).to(dev)
        for _ in range(6):
            vq.update(xa, wa)
        scale = vq.codebook.abs().max(1, keepdim=True).values + 1e-30
        err = float(((vq.codebook - cb_s).abs() / scale).max())
        _, idx_1 = _lib.assign(xa[lo:hi], cb_s)
        same = bool(torch.equal(idx_1, idx_s))
        agree = float((_lib.assign(xa[lo:hi], vq.codebook)[1] == idx_s).float().mean())
        print(f"[{world} ranks] d={d} k={k} n={n}: codebook max rel err {err:.2e}, shard idx == exact assign of the "
              f"sharded codebook: {same}, agreement with the single-GPU clustering: {agree:.6f}")
        ok = ok and err < 1e-5 and same
    dist.barrier()
if rank == 0:
    print("NCCL_CHECK", "OK" if ok else "FAILED")
dist.destroy_process_group()
sys.exit(0 if ok else 1)
