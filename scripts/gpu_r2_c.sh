#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_stock_render.py tests/test_gpu_gradient_f64.py tests/test_gpu_raster_vs_reference.py -m gpu -q -s > gpurun_out/pytest_new.log 2>&1; echo "pytest rc=$?"; grep -v "^$" gpurun_out/pytest_new.log | tail -120
