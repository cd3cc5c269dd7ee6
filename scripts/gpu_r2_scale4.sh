#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
run() { n=$1; shift; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) "$@"; }
for rep in 1 2; do for n in 2 4; do
  run $n bench.py --gpus $n --steps 20 --warmup 5 --quick > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "weak n=$n rc=$?"; grep "step ms" gpurun_out/scale_n$n.err | cut -c1-400
  python -c "
import json; d=json.loads(open('gpurun_out/scale_n$n.json').read().strip().splitlines()[-1]); print($n, round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d['clocks'])"
done; done
