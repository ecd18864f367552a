#!/bin/bash
mkdir -p gpurun_out
T=${TAG:-r2m}
O=gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "assign_bit_exact" > $O/${T}_pytest_first.log 2>&1
echo "== first (BN=112 sanity): rc=$?"; tail -5 $O/${T}_pytest_first.log
if tail -1 $O/${T}_pytest_first.log | grep -q passed; then
  timeout 1800 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1
  echo "== pytest"; tail -6 $O/${T}_pytest_gpu.log
  for d in 45 48 7; do timeout 300 python tools/tc_profile_run.py 5800000 $d 8 > $O/${T}_tc_run_d$d.txt 2>&1; echo "== main (BN112, two chains) d$d"; tail -8 $O/${T}_tc_run_d$d.txt; done
  for d in 45 48 7; do C3VQ_LIB=c3dgs_b200/libc3vq_vA.so timeout 300 python tools/tc_profile_run.py 5800000 $d 8 > $O/${T}_tc_run_vA_d$d.txt 2>&1; echo "== vA (BN112, one chain) d$d"; tail -8 $O/${T}_tc_run_vA_d$d.txt; done
  timeout 900 python bench.py --steps 5 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err
  python - <<PY
import json
j = json.load(open("gpurun_out/${T}_bench.json"))
print("ms/step", round(j["ms_per_step"], 2), "assign", round(j["roofline"]["kernel_ms"], 3), "e2e", round(j["e2e"]["ms_per_step"], 2))
for k in ("shape7", "cov6_trace_normalised", "sh48", "iid45"):
    v = j["extras"][k]; print(k, round(v["ms_per_step"], 2), round(v["assign_ms"], 3))
PY
  tail -n 3 $O/${T}_bench.err
fi
for d in 45 48 7; do C3VQ_LIB=c3dgs_b200/libc3vq_vB.so timeout 300 python tools/tc_profile_run.py 5800000 $d 8 > $O/${T}_tc_run_vB_d$d.txt 2>&1; echo "== vB (BN96, two chains) d$d"; tail -8 $O/${T}_tc_run_vB_d$d.txt; done
