#!/bin/bash
# A/B of the colour-only cached backward of the fused 16-channel call (R3DGW_BWDC_WIDE_CFG), then the GPU suite
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
for v in "$@"; do
  R3DGW_BWDC_WIDE_CFG=$v timeout 600 python bench.py --mode fused --steps 12 --warmup 4 --no-cpu-baseline > gpurun_out/abf_$v.json 2> gpurun_out/abf_$v.err; echo "cfg=$v rc=$?"
  python - <<PY
import json
d=json.loads(open("gpurun_out/abf_$v.json").read().strip().splitlines()[-1])
k=d["kernels"]
print("cfg=$v step", round(d["ms_per_step"],3), "e2e", round(d["e2e"]["ms_per_step"],3), {n: round(v["ms_per_step"],4) for n,v in k.items() if n.startswith("blend")})
PY
done
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4
