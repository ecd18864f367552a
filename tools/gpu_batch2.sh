#!/bin/bash
# final evidence of a state of the tree: launch list of the bench command and full captures of the assignment kernel
mkdir -p gpurun_out
T=${TAG:-r3l}
O=gpurun_out
timeout 900 python bench.py --steps 2 --warmup 1 > $O/${T}_bench_short.json 2> $O/${T}_bench_short.err
echo "bench rc=$?"
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file $O/${T}_launches_bench.csv python bench.py --steps 2 --warmup 1 > $O/${T}_ncu_launch.log 2>&1
echo "launch list rc=$?"
for d in 45 7; do
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_assign_tc -s 5 -c 1 -o $O/${T}_assign_tc_d$d python tools/tc_profile_run.py 5800000 $d 7 > $O/${T}_ncu_d$d.log 2>&1
echo "ncu d$d rc=$?"
ncu -i $O/${T}_assign_tc_d$d.ncu-rep --page source --csv > $O/${T}_assign_tc_d${d}_source.csv 2>/dev/null
done
ls -la $O | tail -8
