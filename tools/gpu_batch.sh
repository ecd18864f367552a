#!/bin/bash
# The batch of GPU work for the next gpurun call (edited between calls; outputs under gpurun_out/).
mkdir -p gpurun_out
T=r2b
O=gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > $O/${T}_pytest_gpu.log 2>&1
C3VQ_SCREEN=tf32 timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "assign and not full_size" > $O/${T}_pytest_assign_tf32.log 2>&1
timeout 600 python tools/sums_probe.py > $O/${T}_sums_probe.txt 2>&1
C3VQ_SUMS_WINDOWED=1 timeout 600 python tools/sums_probe.py > $O/${T}_sums_probe_windowed.txt 2>&1
timeout 300 python tools/preprocess_bench.py > $O/${T}_preprocess_bench.txt 2>&1
C3VQ_LIB=c3dgs_b200/libc3vq_prof.so timeout 300 python tools/tc_prof.py 5800000 45 > $O/${T}_tc_prof_f16.txt 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > $O/${T}_bench_sorted.json 2> $O/${T}_bench_sorted.err
timeout 600 python bench.py --steps 5 --warmup 3 --sums fused --no-extras --no-cpu-baseline > $O/${T}_bench_fused.json 2> $O/${T}_bench_fused.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 700 --csv --log-file $O/${T}_launches_bench.csv python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > $O/${T}_ncu1.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_assign_tc -s 2 -c 1 -o $O/${T}_assign_tc python tools/tc_profile_run.py 5800000 45 4 > $O/${T}_ncu2.log 2>&1
for f in pytest_gpu pytest_assign_tf32; do echo "== $f"; tail -8 $O/${T}_$f.log; done
for f in sums_probe sums_probe_windowed preprocess_bench; do echo "== $f"; tail -12 $O/${T}_$f.txt; done
python - <<'PY'
import json
for n in ("sorted", "fused"):
    try:
        j = json.load(open(f"gpurun_out/r2b_bench_{n}.json"))
        print(n, "ms/step", round(j["ms_per_step"], 2), "assign", round(j["roofline"]["kernel_ms"], 3), "frac", round(j["roofline"]["frac"], 3),
              "update", j["roofline_update"] and round(j["roofline_update"]["kernel_ms"], 3), "e2e", round(j["e2e"]["ms_per_step"], 2), j["assign_stats"])
    except Exception as e:
        print(n, "failed", e)
PY
tail -3 $O/${T}_bench_sorted.err $O/${T}_bench_fused.err
