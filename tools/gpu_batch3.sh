#!/bin/bash
mkdir -p gpurun_out
O=gpurun_out; T=${TAG:-r4s}
for v in new ${EXTRA_LIBS}; do
  lib=c3dgs_b200/libc3vq_$v.so; [ $v = new ] && lib=c3dgs_b200/libc3vq.so
  C3VQ_LIB=$lib timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/${T}_pytest_$v.log 2>&1
  echo "== $v parity rc=$?: $(tail -1 $O/${T}_pytest_$v.log)"
  for d in ${DIMS:-45 48 7}; do
    C3VQ_LIB=$lib timeout 300 python tools/tc_profile_run.py 5800000 $d 10 > $O/${T}_tc_run_${v}_d$d.txt 2>&1
    echo "-- $v d$d: $(tail -9 $O/${T}_tc_run_${v}_d$d.txt | awk '{print $4}' | tr '\n' ' ')"
  done
done
