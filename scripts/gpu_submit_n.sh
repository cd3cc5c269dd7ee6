#!/bin/bash
# usage: scripts/gpu_submit_n.sh <gpus> <timeout_s> <log> <command...>
G=$1; T=$2; LOG=$3; shift 3
for i in $(seq 1 60); do
  /usr/local/graft/bin/gpurun --gpus "$G" --timeout "$T" -- "$@" > "$LOG" 2>&1
  rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc" >> "$LOG"; exit $rc; fi
  sleep 120
done
