#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_image_ops.py -m gpu -q 2>&1 | tail -3
for v in 1 5; do
  echo "== R3DGW_BWDG_MINB=$v"
  R3DGW_BWDG_MINB=$v python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],3), round(d['fused_single_pass']['ms_per_step'],3), {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items() if 'backward' in k})"
done
R3DGW_BWDG_MINB=5 timeout 900 python -m pytest tests/test_gpu_gradient_f64.py tests/test_gpu_raster_vs_reference.py -m gpu -q -x 2>&1 | tail -3
