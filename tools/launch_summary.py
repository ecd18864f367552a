"""Per-kernel totals of an ncu launch list (--metrics gpu__time_duration.sum --csv):
python tools/launch_summary.py launches.csv "<how it was captured>" """
import csv
import sys
from collections import defaultdict

rows = [r for r in csv.reader(open(sys.argv[1], errors="replace")) if len(r) > 5]
h = rows[0]
ki, mi, vi, ui = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("Metric Unit")
tot, cnt = defaultdict(float), defaultdict(int)
for r in rows[1:]:
    if r[mi] != "gpu__time_duration.sum":
        continue
    v = float(r[vi].replace(",", ""))
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(r[ui], 1e-6)
    tot[r[ki]] += v
    cnt[r[ki]] += 1
total = sum(tot.values())
print(sys.argv[2] if len(sys.argv) > 2 else sys.argv[1])
print(f"{'kernel':60s} {'launches':>8s} {'total ms':>10s} {'avg us':>10s} {'share':>7s}")
for k, v in sorted(tot.items(), key=lambda kv: -kv[1])[:24]:
    print(f"{k[:60]:60s} {cnt[k]:8d} {v:10.3f} {1e3 * v / cnt[k]:10.1f} {100 * v / total:6.1f}%")
