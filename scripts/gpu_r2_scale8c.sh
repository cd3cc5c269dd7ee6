#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) "$@"; }
run 8 bench.py --gpus 8 --steps 20 --warmup 5 --quick > gpurun_out/scale_n8.json 2> gpurun_out/scale_n8.err; echo "weak n=8 rc=$?"; grep "step ms" gpurun_out/scale_n8.err | cut -c1-500
run 8 bench.py --gpus 8 --views 8 --steps 10 --warmup 3 --quick > gpurun_out/C_n8.json 2> gpurun_out/C_n8.err; echo "C n=8 rc=$?"; grep "step ms" gpurun_out/C_n8.err | cut -c1-500
python - <<'PY'
import json
for f in ('gpurun_out/scale_n8.json','gpurun_out/C_n8.json'):
    d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, round(d['value'],1), round(d['ms_per_step'],3), d.get('e2e') and round(d['e2e']['value'],1))
PY
