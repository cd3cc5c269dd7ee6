#!/bin/bash
# moment reduction in the full K5 (COLOR = true): U1 fwd+bwd (C = 3) and the fused single-pass step
mkdir -p gpurun_out
for q in 0 8 16; do
  echo "== R3DGW_BWDC_Q=$q"
  R3DGW_BWDG=3 R3DGW_BWDC_Q=$q DIAG_MODE=fused python scripts/diag_kernels.py 10 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('fused: blend_backward', d.get('blend_backward'), 'blend_forward', d.get('blend_forward'), 'step', d['_step_ms'], 'ksum', d['_kernel_sum_ms'])"
  R3DGW_BWDG=3 R3DGW_BWDC_Q=$q python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-profile 2>/dev/null | tail -1 | python -c "
import json,sys; d=json.loads(sys.stdin.read()); print('dropin: step', d['ms_per_step'], 'u1', d['u1_raster_call'], 'fused', d['fused_single_pass']['ms_per_step'])"
done
R3DGW_BWDG=3 R3DGW_BWDC_Q=16 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
