#!/bin/bash
# kernel-variant sweep for the alpha-cache colour backward
mkdir -p gpurun_out
: > gpurun_out/variants.jsonl
for pad in 0 1; do for w in 32 16; do
  R3DGW_GRAD_PAD=$pad R3DGW_BWDC_WROWS=$w python scripts/diag_kernels.py 10 2>/dev/null | tail -1 >> gpurun_out/variants.jsonl
done; done
python - <<'PY'
import json
for l in open('gpurun_out/variants.jsonl'):
    d=json.loads(l); print(d['variant'], 'bwd_color', d.get('blend_backward_color_cached'), 'step', d['_step_ms'], 'ksum', d['_kernel_sum_ms'])
PY
python -m pytest tests/test_gpu_raster_vs_reference.py -x -q -m gpu 2>&1 | tail -3
