#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 5 --no-cpu-baseline > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; tail -4 gpurun_out/bench_native.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_native.json'))
print({k:d[k] for k in ('value','ms_per_step','gpu_launches','fps_fwd','u1_raster_call')}, d['e2e']['value'], d.get('fused_single_pass'))
print(d.get('kernels'))
PY
