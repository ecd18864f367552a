#!/bin/bash
# The batch of GPU work for the next gpurun call (edited between calls; outputs under gpurun_out/).
mkdir -p gpurun_out
T=r2f
O=gpurun_out
timeout 900 python tools/sums_probe.py > $O/${T}_sums_probe.txt 2>&1
timeout 1500 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > $O/${T}_bench.json 2> $O/${T}_bench.err
cat $O/${T}_sums_probe.txt | tail -20
echo "== pytest"; tail -6 $O/${T}_pytest_gpu.log
python - <<'PY'
import json
j = json.load(open("gpurun_out/r2f_bench.json"))
print("ms/step", round(j["ms_per_step"], 2), "assign", round(j["roofline"]["kernel_ms"], 3), "update", round(j["roofline_update"]["kernel_ms"], 3), "gather", round(j["roofline_gather"]["frac"], 3), round(j["roofline_gather"]["kernel_ms"], 3), "e2e", round(j["e2e"]["ms_per_step"], 2))
for k in ("decode_gather_sh48", "strong_24M"):
    print(k, j["extras"][k])
PY
tail -n 3 $O/${T}_bench.err
