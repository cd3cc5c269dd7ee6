#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -s > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log; grep -n "^config\|new-f64" gpurun_out/pytest_gpu.log > gpurun_out/grad_f64.txt
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; tail -4 gpurun_out/bench_native.err | cut -c1-300
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -3 gpurun_out/bench_ref.err | cut -c1-300
timeout 600 python bench.py --views 8 --steps 10 --warmup 3 > gpurun_out/bench_C_n1.json 2> gpurun_out/bench_C_n1.err; echo "C rc=$?"; tail -2 gpurun_out/bench_C_n1.err | cut -c1-300
timeout 600 python scripts/bench_configs.py A > gpurun_out/config_A.json 2> gpurun_out/config_A.err; echo "A rc=$?"; tail -2 gpurun_out/config_A.err | cut -c1-300
timeout 900 python scripts/bench_configs.py E > gpurun_out/config_E.json 2> gpurun_out/config_E.err; echo "E rc=$?"; tail -2 gpurun_out/config_E.err | cut -c1-300
python - <<'PY'
import json
for f in ('bench_native','bench_ref'):
    try:
        d=json.load(open(f'gpurun_out/{f}.json'))
        print(f, {k:d.get(k) for k in ('value','ms_per_step','gpu_launches','fps_fwd','u1_raster_call','fused_single_pass','stock_render_on_dropins')}, d['e2e']['value'], d['config'].get('caller'))
    except Exception as e: print(f, 'ERR', e)
for f in ('bench_C_n1','config_A','config_E'):
    try: print(f, open(f'gpurun_out/{f}.json').read()[:1500])
    except Exception as e: print(f, 'ERR', e)
PY
