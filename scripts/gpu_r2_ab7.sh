#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q 2>&1 | tail -2
for sw in R3DGW_CULL_V2=0 R3DGW_BWDG_IN_PLACE=0 R3DGW_FWD_FLAT=0; do
  echo "== $sw"; env $sw timeout 600 python -m pytest tests -m gpu -q -x 2>&1 | tail -1
done
run() {
  echo "== $*"
  env "$@" python bench.py --steps 20 --warmup 5 --no-cpu-baseline 2>/dev/null | python -c "
import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print(round(d['ms_per_step'],3), round(d['e2e']['ms_per_step'],3), 'fused', round(d['fused_single_pass']['ms_per_step'],3), 'u1', d['u1_raster_call'], 'fps', round(d['fps_fwd'],1), 'launches', d['gpu_launches'])"
}
run R3DGW_NOP=1
python scripts/diag_timeline.py 2>&1 | grep -v Warn | head -50 | cut -c1-150
