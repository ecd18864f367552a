#!/bin/bash
# Multi-GPU batch: bash tools/gpu_batch_multi.sh N TAG   (run under gpurun --gpus N)
N=$1; T=$2; O=gpurun_out
mkdir -p $O
run() { python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 "$@"; }
run bench.py --gpus $N --steps 5 --warmup 3 > $O/${T}_bench_${N}gpu.json 2> $O/${T}_bench_${N}gpu.err
run tools/nccl_check.py > $O/${T}_nccl_check_${N}gpu.txt 2>&1
python - <<PY
import json
j = [json.loads(l) for l in open("$O/${T}_bench_${N}gpu.json") if l.startswith("{")][-1]
print("n_gpus", j["n_gpus"], "ms/step", round(j["ms_per_step"], 2), "value", round(j["value"] / 1e9, 3), "G; e2e ms", round(j["e2e"]["ms_per_step"], 2))
print("parity", j["parity"])
print("h2d", j["h2d_probe"])
print("strong", j["extras"]["strong_24M"] if j.get("extras") else None)
PY
tail -n 5 $O/${T}_bench_${N}gpu.err; tail -n 4 $O/${T}_nccl_check_${N}gpu.txt
