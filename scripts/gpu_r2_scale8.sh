#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
nvidia-smi topo -m > gpurun_out/topo.txt 2>&1
run() { n=$1; shift; if [ "$n" = 1 ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29500 + n)) "$@"; fi; }
for n in 1 2 4 8; do
  run $n bench.py --gpus $n --steps 20 --warmup 5 --quick > gpurun_out/scale_n$n.json 2> gpurun_out/scale_n$n.err; echo "weak n=$n rc=$?"; grep "step ms" gpurun_out/scale_n$n.err | cut -c1-400
done
for n in 2 4 8; do
  run $n bench.py --gpus $n --views 8 --steps 10 --warmup 3 > gpurun_out/C_n$n.json 2> gpurun_out/C_n$n.err; echo "C n=$n rc=$?"; grep "step ms" gpurun_out/C_n$n.err | cut -c1-300
done
run 8 scripts/bench_configs.py D > gpurun_out/D_n8.json 2> gpurun_out/D_n8.err; echo "D n=8 rc=$?"; cut -c1-300 gpurun_out/D_n8.json
python - <<'PY'
import json,glob
for f in sorted(glob.glob('gpurun_out/scale_n*.json'))+sorted(glob.glob('gpurun_out/C_n*.json')):
    try:
        d=json.load(open(f)); print(f, d['n_gpus'], round(d['value'],2), round(d['ms_per_step'],2), d.get('e2e',{}) and round(d['e2e']['value'],2))
    except Exception as e: print(f,'ERR',e)
PY
