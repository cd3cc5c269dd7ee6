#!/bin/bash
# last pass of the round: GPU suite, smoke, native bench line, config E
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -1 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" 2>&1 | tail -1
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; grep "step ms" gpurun_out/bench_native.err | cut -c1-200
timeout 600 python scripts/bench_configs.py E --arm native > gpurun_out/config_E_native.json 2> gpurun_out/config_E_native.err; cat gpurun_out/config_E_native.json
python -c "
import json; d=json.loads(open('gpurun_out/bench_native.json').read().strip().splitlines()[-1]); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['u1_raster_call'], d['fused_single_pass']['ms_per_step'], d['roofline_fwd_call']['frac'], d['roofline']['frac'])"
