"""Measured dense TF32 matmul peak of this GPU (bench.measure_tf32_peak: torch.matmul fp32 8192^3 with allow_tf32,
best of 10 = burst, back to back for 4 s = sustained).  python tools/tf32_peak.py -> one JSON line"""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from bench import measure_tf32_peak  # noqa: E402

if __name__ == "__main__":
    print(json.dumps(measure_tf32_peak(torch.device("cuda", 0), seconds=4.0)))
