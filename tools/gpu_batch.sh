#!/bin/bash
# The batch of GPU work for the next gpurun call (edited between calls; outputs under gpurun_out/).
mkdir -p gpurun_out
T=r2c
O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/${T}_smoke.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > $O/${T}_bench.json 2> $O/${T}_bench.err
C3VQ_SCREEN=tf32 timeout 600 python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > $O/${T}_bench_tf32.json 2> $O/${T}_bench_tf32.err
timeout 600 python bench.py --steps 5 --warmup 3 --sums sorted --no-extras --no-cpu-baseline > $O/${T}_bench_sorted.json 2> $O/${T}_bench_sorted.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > $O/${T}_bench_reference_arm.json 2> $O/${T}_bench_reference_arm.err
C3VQ_BENCH_TF32_SECONDS=0.02 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file $O/${T}_launches_bench.csv python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > $O/${T}_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_assign_tc -s 2 -c 1 -o $O/${T}_assign_tc_fused python tools/tc_profile_run.py 5800000 45 4 > $O/${T}_ncu2.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:k_preprocess_indexed_bwd -s 1 -c 1 -o $O/${T}_preprocess_bwd python tools/preprocess_bench.py > $O/${T}_ncu3.log 2>&1
timeout 600 ncu --set full --clock-control none -k regex:k_gather_span -s 1 -c 2 -o $O/${T}_gather python tools/gather_profile_run.py > $O/${T}_ncu4.log 2>&1
echo "== pytest"; tail -6 $O/${T}_pytest_gpu.log; echo "== smoke"; tail -3 $O/${T}_smoke.log
python - <<'PY'
import json
for n in ("", "_tf32", "_sorted"):
    try:
        j = json.load(open(f"gpurun_out/r2c_bench{n}.json"))
        print(n or "default", "ms/step", round(j["ms_per_step"], 2), "assign", round(j["roofline"]["kernel_ms"], 3), "frac", round(j["roofline"]["frac"], 3),
              "update", j["roofline_update"] and round(j["roofline_update"]["kernel_ms"], 3), "gather", round(j["roofline_gather"]["frac"], 3), "e2e", round(j["e2e"]["ms_per_step"], 2), j["assign_stats"])
        if j.get("extras"):
            for k, v in j["extras"].items():
                print("   ", k, {a: (round(b, 3) if isinstance(b, float) else b) for a, b in v.items() if a in ("ms_per_step", "assign_ms", "update_ms", "candidates_per_row", "fallback_rows", "ms", "frac_of_hbm_peak", "update_iteration_ms", "assign_ms", "total_ms", "wall_ms")})
    except Exception as e:
        print(n, "failed", e)
PY
tail -n 3 $O/${T}_bench.err
