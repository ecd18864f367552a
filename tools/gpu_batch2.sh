#!/bin/bash
mkdir -p gpurun_out
T=${TAG:-r3v}
O=gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_assign_tc -s ${IT:-9} -c 1 -o $O/${T}_assign_tc python tools/tc_profile_run.py 5800000 ${D:-45} 10 > $O/${T}_ncu.log 2>&1
echo "ncu rc=$?"
ncu -i $O/${T}_assign_tc.ncu-rep --page source --csv > $O/${T}_assign_tc_source.csv 2>/dev/null
ncu -i $O/${T}_assign_tc.ncu-rep --page raw --csv > $O/${T}_assign_tc_raw.csv 2>/dev/null
