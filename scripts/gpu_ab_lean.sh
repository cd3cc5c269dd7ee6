#!/bin/bash
# A/B of the colour-only cached backward walk (R3DGW_BWDC_LEAN=0: pre-diet, 1: lean) + GPU tests on the lean build
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
for v in 0 1; do
  R3DGW_BWDC_LEAN=$v timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/ab_lean$v.json 2> gpurun_out/ab_lean$v.err; echo "lean=$v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/ab_lean$v.json").read().strip().splitlines()[-1])
k=d["kernels"]
print("lean=$v step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), "bwd_color", k["blend_backward_color_cached"], "fwd_cached", k["blend_forward_cached"], "fused", d.get("fused_single_pass",{}).get("ms_per_step"))
PY
done
