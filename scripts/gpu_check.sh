#!/bin/bash
# parity tests, then a short bench of the native arm (and optionally more via $EXTRA)
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -6 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"
python - <<'PY'
import json
try:
    d=json.load(open('gpurun_out/bench_native.json'))
    print({k:d[k] for k in ('value','ms_per_step','fps_fwd','fps_fwd_relit_same_view','u1_raster_call','gpu_launches')})
    print('e2e',d['e2e']['value']); print(json.dumps(d.get('kernels'),indent=0)); print(d.get('roofline')); print(d.get('roofline_fwd_call'))
except Exception as e: print('parse fail',e); print(open('gpurun_out/bench_native.err').read()[-2000:])
PY
if [ -n "$EXTRA" ]; then eval "$EXTRA"; fi
