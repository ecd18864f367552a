"""Summarise the SASS page of an ncu report: python tools/ncu_sass_hot.py src.csv [top]
(src.csv = `ncu -i rep.ncu-rep --page source --csv`).  Prints total samples, the hottest instructions
and a running listing with sample counts so loops can be attributed to roles."""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {n: i for i, n in enumerate(hdr)}
stall_cols = [n for n in hdr if n.startswith("stall_") and "Not Issued" not in n]
ins = []
for r in rows[2:]:
    if len(r) < len(hdr):
        continue
    s = int(r[ix["# Samples"]] or 0)
    ex = int(r[ix["Instructions Executed"]] or 0)
    st = sorted(((int(r[ix[c]] or 0), c[6:]) for c in stall_cols), reverse=True)[:3]
    ins.append((s, ex, r[ix["Source"]].strip(), st))
tot = sum(i[0] for i in ins)
print("total samples", tot, "instructions", len(ins), "executed", sum(i[1] for i in ins))
mode = sys.argv[2] if len(sys.argv) > 2 else "list"
if mode == "list":
    for n, (s, ex, src, st) in enumerate(ins):
        print(f"{n:5d} {s:7d} {100.0 * s / tot:5.2f}% {ex:10d}  {src[:90]:90s} {' '.join(f'{c}:{v}' for v, c in st if v)}")
else:
    for n, (s, ex, src, st) in sorted(enumerate(ins), key=lambda t: -t[1][0])[:int(mode)]:
        print(f"{n:5d} {s:7d} {100.0 * s / tot:5.2f}% {ex:10d}  {src[:90]:90s} {' '.join(f'{c}:{v}' for v, c in st if v)}")
