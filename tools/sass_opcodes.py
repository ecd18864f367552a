"""Opcode histogram of libc3vq.so per kernel (cuobjdump -sass): the Blackwell-only instructions that prove the
tcgen05 / TMEM / bulk-copy path is compiled in, next to the arch of every cubin.
python tools/sass_opcodes.py [lib] > profiles/r2_sass_opcodes.txt"""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
lib = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "c3dgs_b200", "libc3vq.so")
WATCH = ("UTCHMMA", "UTCMMA", "LDTM", "STTM", "UTCBAR", "UTCATOM", "UBLKCP", "UBLKRED", "UBLKPF", "UTMALDG", "UTMASTG",
         "SYNCS", "REDG", "RED.", "ATOMG", "ATOMS", "MATCH", "CCTL", "FMNMX", "FFMA", "LDG", "STG", "LDS", "STS", "SHFL",
         "USETMAXREG", "ELECT")
out = subprocess.run(["cuobjdump", "-sass", lib], capture_output=True, text=True, check=True).stdout
arch = collections.Counter(re.findall(r"arch = (sm_\w+)", out))
print(f"cuobjdump -sass {os.path.relpath(lib, ROOT)}: cubin architectures {dict(arch)}")
kern, counts, total = None, {}, collections.Counter()
for line in out.splitlines():
    m = re.match(r"\s*Function : (\S+)", line)
    if m:
        kern = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0]
        counts[kern] = collections.Counter()
        continue
    m = re.match(r"\s*/\*[0-9a-f]+\*/\s+(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m and kern:
        op = m.group(1)
        counts[kern]["_all"] += 1
        for wname in WATCH:
            if op.startswith(wname):
                counts[kern][wname] += 1
                total[wname] += 1
print("totals over the library:", ", ".join(f"{k} {v}" for k, v in sorted(total.items())))
print()
print(f"{'kernel':78s} {'instr':>7s}  watched opcodes")
for kern, c in sorted(counts.items(), key=lambda kv: -kv[1]["_all"]):
    w = ", ".join(f"{k} {v}" for k, v in sorted(c.items()) if k != "_all" and k not in ("FFMA", "LDG", "STG", "LDS", "STS", "SHFL", "FMNMX"))
    base = ", ".join(f"{k} {c[k]}" for k in ("FFMA", "FMNMX", "LDG", "STG", "LDS", "STS", "SHFL") if c[k])
    print(f"{kern[:78]:78s} {c['_all']:7d}  {w}{' | ' if w and base else ''}{base}")
