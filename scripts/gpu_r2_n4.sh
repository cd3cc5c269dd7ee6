#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
for i in 1 2 3; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port $((29600 + i)) bench.py --gpus 4 --steps 20 --warmup 5 --quick > gpurun_out/scale4_$i.json 2> gpurun_out/scale4_$i.err; echo "rc=$?"; grep "per-rank\|step ms (res" gpurun_out/scale4_$i.err | cut -c1-400; tail -1 gpurun_out/scale4_$i.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'])"
done
