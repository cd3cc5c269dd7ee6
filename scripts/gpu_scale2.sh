#!/bin/bash
# N = 1 and N = 2 of the weak-scaling run (cheap check of the gradient bucket path before the 8-GPU run)
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 400 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/s2_n1.json 2> gpurun_out/s2_n1.err; echo "N=1 rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/s2_n2.json 2> gpurun_out/s2_n2.err; echo "N=2 rc=$?"; tail -3 gpurun_out/s2_n2.err
python - <<'PY'
import json
for n in (1, 2):
    d = json.loads(open(f"gpurun_out/s2_n{n}.json").read().strip().splitlines()[-1])
    print(n, round(d["value"], 1), "it/s", round(d["ms_per_step"], 3), "ms  e2e", round(d["e2e"]["value"], 1))
PY
