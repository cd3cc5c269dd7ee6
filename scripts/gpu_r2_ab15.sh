#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python -m pytest tests/test_gpu_gradient_f64.py -m gpu -q -s 2>&1 | grep "^config\|new-f64\|passed\|failed" | cut -c1-220 | head -20
python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3), 'fused', round(d['fused_single_pass']['ms_per_step'],3), 'u1', d['u1_raster_call'], {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items() if 'blend' in k or 'ssim' in k})"
