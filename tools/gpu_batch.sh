#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gap_probe.py > gpurun_out/r2h_gap_probe.txt 2>&1
cat gpurun_out/r2h_gap_probe.txt
