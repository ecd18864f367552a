"""Diagnostics for the tensor-core assignment on the GPU box: screening scores vs fp64, band soundness,
assignment parity vs the SIMT exact kernel.  python tools/tc_debug.py"""
import os
import sys
import traceback

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from c3dgs_b200 import _lib  # noqa: E402


def scores_check(n, d, k, seed, scale=1.0, offset=0.0):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(n, d, generator=g) * scale + offset).contiguous()
    cb = (torch.randn(k, d, generator=g) * scale + offset).contiguous()
    s, norm2, band = _lib.debug_scores(x.cuda(), cb.cuda())
    torch.cuda.synchronize()
    s, norm2, band = s.cpu().double().numpy(), norm2.cpu().double().numpy(), band.cpu().double().numpy()
    xd, cd = x.double().numpy(), cb.double().numpy()
    mu = cd.mean(0)
    xc, cc = xd - mu, cd - mu
    true_d = ((xd[:, None, :] - cd[None, :, :]) ** 2).sum(-1) if n * k * d < 5e7 else None
    exact_s = (cc ** 2).sum(1)[None, :] - 2 * xc @ cc.T
    err = np.abs(s - exact_s)
    print(f"[scores n={n} d={d} k={k} scale={scale} off={offset}] max|s-s_exact|={err.max():.3e} "
          f"mean={err.mean():.3e}  |s|max={np.abs(exact_s).max():.3e}  band[min,max]=({band.min():.3e},{band.max():.3e}) "
          f"X err={np.abs(norm2 - (xc ** 2).sum(1)).max():.3e}")
    if true_d is not None:
        e2 = np.abs(s + norm2[:, None] - true_d)            # what the band must cover (per row: 2E <= band)
        ratio = (2 * e2.max(1)) / band
        print(f"    max 2*|s+X-d| / band = {ratio.max():.4f}  (must be < 1)")
    if err.max() > 1e-2 * max(1.0, np.abs(exact_s).max()):
        bad = np.argwhere(err > 1e-2 * max(1.0, np.abs(exact_s).max()))
        print("    BAD entries:", len(bad), "first:", bad[:8].tolist())
        print("    rows with errors:", np.unique(bad[:, 0])[:20], " cols:", np.unique(bad[:, 1])[:20])
        print("    sample got/exp:", s[bad[0][0], bad[0][1]], exact_s[bad[0][0], bad[0][1]])
    return err.max()


def assign_check(n, d, k, seed, clustered=True):
    g = torch.Generator().manual_seed(seed)
    if clustered:
        c = torch.randn(max(k // 2, 1), d, generator=g)
        x = c[torch.randint(0, c.shape[0], (n,), generator=g)] + 0.1 * torch.randn(n, d, generator=g)
    else:
        x = torch.randn(n, d, generator=g)
    amin, amax = x.aminmax()
    cb = torch.rand(k, d, generator=g) * (amax - amin) + amin
    xg, cg = x.cuda(), cb.cuda()
    de, ie = _lib.assign(xg, cg, exact=True)
    dt, it = _lib.assign(xg, cg, exact=False)
    torch.cuda.synchronize()
    rows, cands, fb = _lib.assign_stats()
    ok_i = bool(torch.equal(ie, it))
    ok_d = bool(torch.equal(de, dt))
    print(f"[assign n={n} d={d} k={k} clustered={clustered}] idx_equal={ok_i} dist_equal={ok_d} "
          f"mismatch={(ie != it).sum().item()} stats rows={rows} cands/row={cands / max(rows, 1):.2f} fallback={fb}")
    return ok_i and ok_d


def main():
    print(torch.cuda.get_device_name(0))
    for args in [(256, 45, 256, 0), (300, 45, 700, 1), (256, 6, 256, 2), (256, 7, 512, 3), (512, 48, 512, 4),
                 (256, 45, 256, 5, 100.0, 0.0), (256, 45, 256, 6, 1.0, 50.0), (256, 20, 256, 7)]:
        try:
            scores_check(*args)
        except Exception:  # noqa: BLE001
            traceback.print_exc()
    for args in [(1000, 45, 256, 0), (5003, 45, 4096, 1), (5003, 6, 4096, 2), (5003, 7, 512, 3), (5003, 48, 4096, 4),
                 (100000, 45, 4096, 5), (100000, 45, 4096, 6, False), (100000, 6, 4096, 7), (1, 45, 1, 8),
                 (777, 33, 100, 9), (777, 64, 300, 10), (777, 3, 10, 11)]:
        try:
            assign_check(*args)
        except Exception:  # noqa: BLE001
            traceback.print_exc()


if __name__ == "__main__":
    main()
