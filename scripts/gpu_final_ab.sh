#!/bin/bash
# A/B of the blocked forward replay, then the round-end pass with the faster setting exported (recorded in gpurun_out/fwdc_choice.txt)
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
for v in 0 1; do
  R3DGW_FWDC_BLOCKED=$v timeout 600 python bench.py --steps 12 --warmup 4 --no-cpu-baseline > gpurun_out/abw_$v.json 2> gpurun_out/abw_$v.err
done
CHOICE=$(python - <<'PY'
import json
r = {}
for v in (0, 1):
    d = json.loads(open(f"gpurun_out/abw_{v}.json").read().strip().splitlines()[-1])
    r[v] = (d["kernels"]["blend_forward_cached"]["avg_launch_ms"], d["ms_per_step"])
import sys
print(f"blocked=0 {r[0]} blocked=1 {r[1]}", file=sys.stderr)
print(1 if r[1][0] < 0.98 * r[0][0] else 0)
PY
)
echo "R3DGW_FWDC_BLOCKED=$CHOICE" | tee gpurun_out/fwdc_choice.txt
export R3DGW_FWDC_BLOCKED=$CHOICE
bash scripts/gpu_final.sh
