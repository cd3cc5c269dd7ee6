#!/bin/bash
# Round-2 opening pass: parity suite + smoke on the rebuilt tree, then `ncu --set full` of the binning chain,
# shading and geometry-backward kernels (no counter evidence for them yet).
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
nproc > gpurun_out/host.txt; free -g >> gpurun_out/host.txt; nvidia-smi -L >> gpurun_out/host.txt
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 300 python scripts/prof_u2.py 3 > gpurun_out/plain_u2.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on \
  -k regex:"preprocess_scan|duplicate_sorted|radix_histogram|radix_scan|onesweep_pass|ranges_gather|shade_|geometry_backward|depth_keys" \
  -s 26 -c 26 -o gpurun_out/prof_chain -f python scripts/prof_u2.py 3 > gpurun_out/ncu_chain.log 2>&1
echo "ncu chain rc=$?"; tail -2 gpurun_out/ncu_chain.log
