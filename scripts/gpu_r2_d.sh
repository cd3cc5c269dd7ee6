#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
python scripts/diag_f64_split.py A > gpurun_out/f64_split.txt 2>&1; python scripts/diag_f64_split.py B >> gpurun_out/f64_split.txt 2>&1; cat gpurun_out/f64_split.txt | tail -12
timeout 900 python -m pytest tests/test_gpu_stock_render.py -m gpu -q -x > gpurun_out/pytest_stock.log 2>&1; echo "pytest rc=$?"; tail -30 gpurun_out/pytest_stock.log
