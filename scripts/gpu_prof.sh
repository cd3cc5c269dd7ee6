#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 300 python scripts/prof_u1.py B 3 > gpurun_out/plain_u1.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:blend_ -s 2 -c 4 -o gpurun_out/prof_blend -f python scripts/prof_u1.py B 3 > gpurun_out/ncu_blend.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu_blend.log
