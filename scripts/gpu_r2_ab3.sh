#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_gradient_f64.py -m gpu -q -s 2>&1 | grep "^config\|new-f64\|passed\|failed\|Error" | cut -c1-200
timeout 900 python -m pytest tests -m gpu -q 2>&1 | tail -4
for v in 0 1; do
  echo "== R3DGW_BWDG_CACHED=$v"
  R3DGW_BWDG_CACHED=$v python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3), round(d['fused_single_pass']['ms_per_step'],3), {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items() if 'blend' in k})"
done
R3DGW_GEOMETRY_CACHE=0 python -m pytest tests -m gpu -q -x 2>&1 | grep "FAILED\|Error\|assert" | head -5
