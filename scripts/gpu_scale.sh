#!/bin/bash
# 1 -> 8 GPU weak-scaling run of bench.py (one view per rank per step + NCCL all-reduce of the gradient bucket)
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8 > gpurun_out/scale_gpus.txt
for N in 1 2 4 8; do
  if [ "$N" = 1 ]; then
    timeout 400 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/scale_n1.json 2> gpurun_out/scale_n1.err
  else
    timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps 10 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/scale_n$N.json 2> gpurun_out/scale_n$N.err
  fi
  echo "N=$N rc=$?"; tail -c 700 gpurun_out/scale_n$N.json | head -c 700; echo; tail -2 gpurun_out/scale_n$N.err
done
