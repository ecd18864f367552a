#!/bin/bash
mkdir -p gpurun_out
T=r2j
O=gpurun_out
timeout 300 python tools/tc_profile_run.py 5800000 45 8 > $O/${T}_tc_run_probe.txt 2>&1
C3VQ_LIB=c3dgs_b200/libc3vq_lcap8.so timeout 300 python tools/tc_profile_run.py 5800000 45 8 > $O/${T}_tc_run_probe_lcap8.txt 2>&1
timeout 300 python tools/tc_profile_run.py 5800000 48 8 > $O/${T}_tc_run_probe_d48.txt 2>&1
timeout 300 python tools/tc_profile_run.py 5800000 7 8 > $O/${T}_tc_run_probe_d7.txt 2>&1
timeout 1500 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err
for f in tc_run_probe tc_run_probe_lcap8 tc_run_probe_d48 tc_run_probe_d7; do echo "== $f"; cat $O/${T}_$f.txt; done
echo "== pytest"; tail -6 $O/${T}_pytest_gpu.log
python - <<'PY'
import json
j = json.load(open("gpurun_out/r2j_bench.json"))
print("ms/step", round(j["ms_per_step"], 2), "assign", round(j["roofline"]["kernel_ms"], 3), "alone", round(j["roofline"]["assignment_alone_ms"], 3), "update", round(j["roofline_update"]["kernel_ms"], 3), "gather", round(j["roofline_gather"]["frac"], 3), "e2e", round(j["e2e"]["ms_per_step"], 2))
for k in ("shape7", "cov6_trace_normalised", "sh48", "iid45"):
    v = j["extras"][k]; print(k, round(v["ms_per_step"], 2), round(v["assign_ms"], 3), round(v["candidates_per_row"], 2), v["fallback_rows"])
for k in ("compress_pipeline", "compress_pipeline_reference", "strong_24M"):
    print(k, j["extras"][k])
PY
tail -n 3 $O/${T}_bench.err
