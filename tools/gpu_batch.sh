#!/bin/bash
# The batch of GPU work for the next gpurun call (edited between calls; outputs under gpurun_out/).
mkdir -p gpurun_out
T=r2a
O=gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,power.limit --format=csv > $O/${T}_gpu.txt 2>&1
timeout 300 python tools/tc_debug.py > $O/${T}_tc_debug_f16.txt 2>&1
C3VQ_SCREEN=tf32 timeout 300 python tools/tc_debug.py > $O/${T}_tc_debug_tf32.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "assign and not full_size" > $O/${T}_pytest_assign_f16.log 2>&1
timeout 600 python tools/sums_probe.py > $O/${T}_sums_probe_f16_windowed.txt 2>&1
C3VQ_SCREEN=tf32 C3VQ_SUMS_WINDOWED=0 timeout 600 python tools/sums_probe.py > $O/${T}_sums_probe_tf32_plain.txt 2>&1
C3VQ_SCREEN=tf32 C3VQ_LIB=c3dgs_b200/libc3vq_prof.so timeout 300 python tools/tc_prof.py 5800000 45 > $O/${T}_tc_prof_tf32.txt 2>&1
C3VQ_LIB=c3dgs_b200/libc3vq_prof.so timeout 300 python tools/tc_prof.py 5800000 45 > $O/${T}_tc_prof_f16.txt 2>&1
timeout 300 python tools/tf32_peak.py > $O/${T}_tf32_peak.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_preprocess.py -x -q -m gpu > $O/${T}_pytest_preprocess.log 2>&1
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "accumulate or full_size or modes_agree" > $O/${T}_pytest_sums.log 2>&1
timeout 900 python bench.py --steps 5 --warmup 3 > $O/${T}_bench_f16.json 2> $O/${T}_bench_f16.err
C3VQ_SCREEN=tf32 timeout 600 python bench.py --steps 5 --warmup 3 --no-extras --no-cpu-baseline > $O/${T}_bench_tf32.json 2> $O/${T}_bench_tf32.err
for f in tc_debug_f16 pytest_assign_f16 sums_probe_f16_windowed sums_probe_tf32_plain tf32_peak pytest_preprocess pytest_sums; do echo "== $f"; tail -8 $O/${T}_$f.txt $O/${T}_$f.log 2>/dev/null; done
head -c 1500 $O/${T}_bench_f16.json; tail -5 $O/${T}_bench_f16.err
