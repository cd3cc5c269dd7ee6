#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
python scripts/diag_alpha_sectors.py B > gpurun_out/alpha_sectors.txt 2>&1; tail -5 gpurun_out/alpha_sectors.txt
python bench.py --steps 40 --warmup 5 --quick 2>&1 >/dev/null | grep "step ms (resident)" | cut -c1-400
R3DGW_BENCH_NO_SAMPLER=1 python bench.py --steps 40 --warmup 5 --quick 2>&1 >/dev/null | grep "step ms (resident)" | cut -c1-400
python bench.py --views 8 --steps 10 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C', d['ms_per_step'], d['step_ms'], d['allocator'])"
R3DGW_BENCH_NO_SAMPLER=1 python bench.py --views 8 --steps 10 --warmup 3 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('C nosampler', d['ms_per_step'], d['step_ms'], d['allocator'])"
python scripts/bench_configs.py E --arm native 2> gpurun_out/config_E_native.err | tee gpurun_out/config_E_native.json | cut -c1-500
python scripts/bench_configs.py E --arm reference 2> gpurun_out/config_E_ref.err | tee gpurun_out/config_E_ref.json | cut -c1-500; tail -2 gpurun_out/config_E_ref.err | cut -c1-200
