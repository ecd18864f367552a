#!/bin/bash
# usage: [GPUS=N] tools/gpurun_retry.sh <timeout-seconds> <log> '<command>'  — retries while the pod answers "busy" (exit 3)
t=$1; log=$2; shift 2
extra=""
if [ -n "$GPUS" ]; then extra="--gpus $GPUS"; fi
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun $extra --timeout "$t" -- "$@" > "$log" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then exit $rc; fi
  sleep 60
done
exit 3
