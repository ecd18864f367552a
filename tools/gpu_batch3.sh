#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out; T=${TAG:-r4i}
C3VQ_LIB=c3dgs_b200/libc3vq_vA.so timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/${T}_pytest_vA.log 2>&1
echo "== vA parity rc=$?: $(tail -1 $O/${T}_pytest_vA.log)"
for v in new vA; do
  lib=c3dgs_b200/libc3vq_$v.so; [ $v = new ] && lib=c3dgs_b200/libc3vq.so
  for d in 45 48 7; do
    C3VQ_LIB=$lib timeout 300 python tools/tc_profile_run.py 5800000 $d 10 > $O/${T}_tc_run_${v}_d$d.txt 2>&1
    echo "-- $v d$d: $(tail -9 $O/${T}_tc_run_${v}_d$d.txt | awk '{print $4}' | tr '\n' ' ')"
  done
done
