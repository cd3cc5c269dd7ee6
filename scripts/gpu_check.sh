#!/bin/bash
# full GPU test suite + per-kernel step breakdown + the bench line (native arm)
mkdir -p gpurun_out
python -m pytest tests -x -q -m gpu 2>&1 | tail -4 | tee gpurun_out/pytest_gpu_tail.log
python scripts/diag_kernels.py 10 2>/dev/null | tail -1 > gpurun_out/diag_kernels.json
cat gpurun_out/diag_kernels.json
python bench.py --steps 20 --warmup 5 > gpurun_out/bench_g.json 2> gpurun_out/bench_g.err; tail -c 1500 gpurun_out/bench_g.json
