#!/bin/bash
# end-of-round evidence: GPU tests, smoke, bench, ncu launch list of the bench command
mkdir -p gpurun_out
T=${TAG:-r4j}
O=gpurun_out
timeout 1800 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.txt 2>&1
echo "== pytest rc=$?: $(tail -1 $O/${T}_pytest_gpu.txt)"
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > $O/${T}_pytest_no_alloc_alone.txt 2>&1
echo "== no_alloc alone rc=$?: $(tail -1 $O/${T}_pytest_no_alloc_alone.txt)"
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
timeout 1500 python bench.py > $O/${T}_bench.json 2> $O/${T}_bench.err
echo "== bench (default flags) rc=$?"
python - <<PY
import json
j = json.load(open("gpurun_out/${T}_bench.json"))
print("steps", j["steps"], "warmup", j["warmup"], "ms/step", round(j["ms_per_step"], 2), "assign", round(j["roofline"]["kernel_ms"], 3), "frac", round(j["roofline"]["frac"],3), "vs tf32", round(j["roofline"]["frac_vs_tf32_sustained_measured"],3), "e2e", round(j["e2e"]["ms_per_step"], 2), "launches", j["gpu_launches"])
for k in ("shape7", "cov6_trace_normalised", "sh48", "iid45"):
    v = j["extras"][k]; print(k, round(v["ms_per_step"], 2), round(v["assign_ms"], 3), v.get("candidates_per_row"), v.get("fallback_rows"))
e = j["extras"]
print("strong_24M", e["strong_24M"]["ms_per_step"], "pipeline", e["compress_pipeline"]["total_ms"], "minibatch", e["vq_features_minibatch"]["wall_ms"], "morton", e["morton_order"]["ms"], "gather", j["roofline_gather"]["kernel_ms"])
PY
tail -n 2 $O/${T}_bench.err
