#!/bin/bash
# GPU suite with the documented switches flipped (each must stay green), then the gradient L2 profile on the current kernels
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
for sw in ${SWITCHES:-R3DGW_SPLIT_FUSED_BACKWARD=0 R3DGW_BWDC_LEAN=0 R3DGW_FAST_PLUMBING=0 R3DGW_DEFERRED_BACKWARD=0 R3DGW_ALPHA_CACHE=0 R3DGW_GEOMETRY_CACHE=0 R3DGW_GEOM_BWD_F32=1 R3DGW_GRAD_PAD=0 R3DGW_BWDG_CACHED=0 R3DGW_BWDG_IN_PLACE=0 R3DGW_FWD_FLAT=0 R3DGW_CULL_V2=0 R3DGW_EARLY_R=0 R3DGW_OS_MATCH=any R3DGW_OS_ITEMS=8}; do
  echo "== $sw"; env $sw timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -1
done
timeout 600 python scripts/diag_l2.py > gpurun_out/grad_l2.txt 2>&1; echo "diag_l2 rc=$?"; head -12 gpurun_out/grad_l2.txt | cut -c1-250
timeout 600 python bench.py --steps 12 --warmup 4 --no-cpu-baseline 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('step', d['ms_per_step'], {k: round(v['avg_launch_ms'],4) for k,v in d['kernels'].items() if 'blend' in k})"
