#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
python bench.py --steps 40 --warmup 5 --no-cpu-baseline > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; grep "step ms" gpurun_out/bench_native.err | cut -c1-500
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_native.json').read().strip().splitlines()[-1])
print({k:d.get(k) for k in ('value','ms_per_step','gpu_launches','fps_fwd','fps_fwd_relit_same_view','u1_raster_call','fused_single_pass','stock_render_on_dropins','allocator')}, d['e2e']['value'])
print(d.get('roofline')); print(d.get('roofline_fwd_call'))
print({k:round(v['ms_per_step'],3) for k,v in d['kernels'].items()})
PY
python scripts/diag_timeline.py > gpurun_out/timeline.txt 2>&1; head -3 gpurun_out/timeline.txt | tail -1; grep "idle/step" gpurun_out/timeline.txt | cut -c1-200
