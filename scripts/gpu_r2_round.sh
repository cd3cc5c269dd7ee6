#!/bin/bash
# Round-2 single-GPU pass: parity tests, smoke, both bench arms, fused mode, config A / C / D / E lines.
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -1 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; grep "step ms" gpurun_out/bench_native.err | cut -c1-300
timeout 900 python bench.py --steps 20 --warmup 5 --mode fused --no-cpu-baseline > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err; echo "fused rc=$?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; grep "step ms" gpurun_out/bench_ref.err | cut -c1-300
timeout 600 python bench.py --views 8 --steps 10 --warmup 3 > gpurun_out/C_n1.json 2> gpurun_out/C_n1.err; echo "C rc=$?"; grep "step ms" gpurun_out/C_n1.err | cut -c1-300
timeout 600 python bench.py --impl reference --views 8 --steps 4 --warmup 2 > gpurun_out/C_ref_n1.json 2> gpurun_out/C_ref_n1.err; echo "C ref rc=$?"; grep "step ms" gpurun_out/C_ref_n1.err | cut -c1-300
timeout 600 python scripts/bench_configs.py A > gpurun_out/config_A.json 2> gpurun_out/config_A.err; echo "A rc=$?"
timeout 600 python scripts/bench_configs.py D > gpurun_out/D_n1.json 2> gpurun_out/D_n1.err; echo "D rc=$?"
timeout 600 python scripts/bench_configs.py E --arm native > gpurun_out/config_E_native.json 2> gpurun_out/config_E_native.err; echo "E rc=$?"
python - <<'PY'
import json
def last(f): return json.loads(open(f).read().strip().splitlines()[-1])
for f in ('bench_native','bench_fused','bench_ref'):
    try:
        d=last(f'gpurun_out/{f}.json')
        print(f, {k:d.get(k) for k in ('value','ms_per_step','gpu_launches','fps_fwd','fps_fwd_relit_same_view','u1_raster_call','fused_single_pass','stock_render_on_dropins','allocator')}, 'e2e', d['e2e']['value'], d['config'].get('caller'))
    except Exception as e: print(f, 'ERR', e)
for f in ('C_n1','C_ref_n1','config_A','D_n1','config_E_native'):
    try: print(f, open(f'gpurun_out/{f}.json').read().strip().splitlines()[-1][:700])
    except Exception as e: print(f, 'ERR', e)
PY
