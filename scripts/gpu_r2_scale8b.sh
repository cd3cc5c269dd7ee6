#!/bin/bash
# final-state scaling pass on one 8-GPU box: weak 1/2/4/8 (quick), config C and D at N = 8
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
run() { n=$1; shift; if [ "$n" = 1 ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) "$@"; fi; }
for n in 1 2 4 8; do
  run $n bench.py --gpus $n --steps 20 --warmup 5 --quick > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "weak n=$n rc=$?"; grep "step ms" gpurun_out/scale_n$n.err | cut -c1-300
done
run 8 bench.py --gpus 8 --views 8 --steps 10 --warmup 3 --quick > gpurun_out/C_n8.json 2> gpurun_out/C_n8.err; echo "C n=8 rc=$?"; grep "step ms" gpurun_out/C_n8.err | cut -c1-300
run 8 scripts/bench_configs.py D > gpurun_out/D_n8.json 2> gpurun_out/D_n8.err; echo "D n=8 rc=$?"; cut -c1-300 gpurun_out/D_n8.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/scale_n*.json'))+sorted(glob.glob('gpurun_out/C_n8.json')):
    try:
        d=json.loads(open(f).read().strip().splitlines()[-1]); print(f, d['n_gpus'], round(d['value'],2), round(d['ms_per_step'],3), d.get('e2e',{}) and round(d['e2e']['value'],2), d.get('per_rank_mean_max_step_ms'))
    except Exception as e: print(f,'ERR',e)
PY
