#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
for m in ballot; do
  echo "== R3DGW_OS_MATCH=$m"
  R3DGW_OS_MATCH=$m timeout 600 python -m pytest tests/test_gpu_raster_vs_reference.py tests/test_gpu_shade_knn.py -m gpu -q -x 2>&1 | tail -1
  R3DGW_OS_MATCH=$m python scripts/diag_E_kernels.py E 2>&1 | grep "^fwd" | cut -c1-400
  R3DGW_OS_MATCH=$m python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],3), 'u1', d['u1_raster_call'], {k:round(v['ms_per_step'],3) for k,v in d['kernels'].items() if k in ('sort_onesweep_pass','depth_sort','sort_histogram')})"
done
