#!/bin/bash
# The batch of GPU work for the next gpurun call (edited between calls; outputs under gpurun_out/).
mkdir -p gpurun_out
T=r2a
python tools/sums_probe.py > gpurun_out/${T}_sums_probe.txt 2>&1
C3VQ_LIB=c3dgs_b200/libc3vq_prof.so python tools/tc_prof.py 5800000 45 > gpurun_out/${T}_tc_prof.txt 2>&1
python tools/tf32_peak.py > gpurun_out/${T}_tf32_peak.txt 2>&1
timeout 900 python -m pytest tests/test_gpu_preprocess.py -x -q -m gpu > gpurun_out/${T}_pytest_preprocess.log 2>&1
tail -30 gpurun_out/${T}_sums_probe.txt; cat gpurun_out/${T}_tf32_peak.txt; tail -15 gpurun_out/${T}_pytest_preprocess.log
