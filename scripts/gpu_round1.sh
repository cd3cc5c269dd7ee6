#!/bin/bash
# First GPU pass of the round: parity tests, goldens, smoke, short bench for both arms, ncu launch list.
cd "$GRAFT_REPO_ROOT"
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total --format=csv > gpurun_out/gpu.txt
nproc >> gpurun_out/gpu.txt
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -25 gpurun_out/pytest_gpu.log
timeout 300 python tests/golden/make_raster_golden.py gpurun_out/golden > gpurun_out/golden.log 2>&1; tail -3 gpurun_out/golden.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -3 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_native.json; tail -5 gpurun_out/bench_native.err
timeout 900 python bench.py --steps 10 --warmup 3 --mode fused --no-cpu-baseline > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err; echo "fused rc=$?"; tail -c 1500 gpurun_out/bench_fused.json; tail -5 gpurun_out/bench_fused.err
timeout 900 python bench.py --impl reference --steps 10 --warmup 3 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 1500 gpurun_out/bench_ref.json; tail -5 gpurun_out/bench_ref.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base mangled -k regex:r3dgw -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/ncu.log 2>&1
echo "ncu rc=$?"; tail -3 gpurun_out/ncu.log
