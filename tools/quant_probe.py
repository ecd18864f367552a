"""Which arithmetic does torch's CUDA quantize_per_tensor use?  Prints mismatch counts of candidate formulas
(evaluated with torch ops on the GPU in the stated precision) against torch.quantize_per_tensor(...).int_repr()."""
import numpy as np
import torch

dev = "cuda:0"
g = torch.Generator().manual_seed(0)
for scale, zp in [(0.0123, -7), (0.0041, 12), (3.0e-4, -128)]:
    scale = float(np.float32(scale))
    x = torch.randn(4_000_000, generator=g) * (60 * scale)
    x[:300] = ((torch.arange(300) - 150).float() + 0.5) * scale
    xd = x.to(dev)
    ref = torch.quantize_per_tensor(xd, scale, zp, torch.qint8).int_repr().int()
    cands = {
        "double_div": torch.clamp(torch.round(xd.double() / scale) + zp, -128, 127).int(),
        "float_div": torch.clamp(torch.round(xd / torch.tensor(scale, device=dev)) + zp, -128, 127).int(),
        "float_inv_mul": torch.clamp(torch.round(xd * (torch.tensor(1.0, device=dev) / torch.tensor(scale, device=dev))) + zp, -128, 127).int(),
        "double_inv_mul": torch.clamp(torch.round(xd.double() * (1.0 / scale)) + zp, -128, 127).int(),
    }
    print(scale, zp, {k: int((v != ref).sum()) for k, v in cands.items()})
