#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_shade_knn.py tests/test_gpu_gaussian_ops.py tests/test_gpu_stock_render.py -m gpu -q 2>&1 | tail -3
for v in 3 4; do
  echo "== R3DGW_SHADE_BWD_MINB=$v"
  R3DGW_SHADE_BWD_MINB=$v python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],3), round(d['fused_single_pass']['ms_per_step'],3), {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items() if 'shade' in k})"
done
