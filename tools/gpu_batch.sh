#!/bin/bash
# scratch driver for one gpurun call: A/B of the working library against the round's starting point (libc3vq_base.so)
mkdir -p gpurun_out
T=${TAG:-r2t}
O=gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu > $O/${T}_pytest_parity.log 2>&1
echo "== parity rc=$?: $(tail -1 $O/${T}_pytest_parity.log)"
for v in new base ${EXTRA_LIBS}; do
  lib=c3dgs_b200/libc3vq_$v.so; [ $v = new ] && lib=c3dgs_b200/libc3vq.so
  [ -f $lib ] || continue
  for d in ${DIMS:-45 48 7}; do
    C3VQ_LIB=$lib timeout 300 python tools/tc_profile_run.py 5800000 $d 10 > $O/${T}_tc_run_${v}_d$d.txt 2>&1
    echo "-- $v d$d: $(tail -9 $O/${T}_tc_run_${v}_d$d.txt | awk '{print $4}' | tr '\n' ' ')"
  done
done
if [ -n "$NCU" ]; then
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_assign_tc -s 5 -c 1 -o $O/${T}_assign_tc python tools/tc_profile_run.py 5800000 ${NCU_D:-45} 7 > $O/${T}_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i $O/${T}_assign_tc.ncu-rep --page source --csv > $O/${T}_assign_tc_source.csv 2>/dev/null
ncu -i $O/${T}_assign_tc.ncu-rep --page raw --csv > $O/${T}_assign_tc_raw.csv 2>/dev/null
fi
