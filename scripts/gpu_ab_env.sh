#!/bin/bash
# usage: gpu_ab_env.sh VAR v1 v2 ...  - drop-in bench per-kernel numbers with VAR set to each value, then the GPU suite with the last
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
VAR=$1; shift
for v in "$@"; do
  env $VAR=$v timeout 600 python bench.py --steps 12 --warmup 4 --no-cpu-baseline > gpurun_out/ab_$v.json 2> gpurun_out/ab_$v.err; echo "$VAR=$v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/ab_$v.json").read().strip().splitlines()[-1])
k=d["kernels"]
print("$VAR=$v step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), {n: round(k[n]["avg_launch_ms"],4) for n in ("blend_backward_color_cached","blend_forward_cached","blend_backward_geometry","blend_forward")}, "fused", d.get("fused_single_pass",{}).get("ms_per_step"))
PY
done
env $VAR=$v timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -15
