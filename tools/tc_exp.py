"""Timing-only experiments with developer builds of the tensor-core kernel (results are garbage by design):
python tools/tc_exp.py  -> builds nothing; expects c3dgs_b200/libc3vq_exp_<name>.so built by tools/tc_exp_build.py"""
import glob
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for lib in [os.path.join(ROOT, "c3dgs_b200", "libc3vq.so")] + sorted(glob.glob(os.path.join(ROOT, "c3dgs_b200", "libc3vq_exp_*.so"))):
    env = dict(os.environ, C3VQ_LIB=lib)
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "tc_profile_run.py"), "2000000", "45", "4"],
                       env=env, capture_output=True, text=True, timeout=120)
    print(os.path.basename(lib), "|", " | ".join(l.split("assign_accumulate")[1].split("for")[0].strip()
                                                 for l in r.stdout.splitlines() if "assign_accumulate" in l), r.stderr[-300:])
