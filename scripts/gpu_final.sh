#!/bin/bash
# Round-end pass: parity tests, smoke, both bench arms, ncu launch list of the bench command, ncu --set full of the blend kernels.
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 20 --warmup 5 > gpurun_out/bench_native.json 2> gpurun_out/bench_native.err; echo "bench rc=$?"; tail -4 gpurun_out/bench_native.err
timeout 900 python bench.py --steps 20 --warmup 5 --mode fused --no-cpu-baseline > gpurun_out/bench_fused.json 2> gpurun_out/bench_fused.err; echo "fused rc=$?"
timeout 900 python bench.py --impl reference --steps 20 --warmup 5 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -2 gpurun_out/bench_ref.err
timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:r3dgw -c 600 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-profile > gpurun_out/ncu.log 2>&1
echo "ncu list rc=$?"
timeout 300 python scripts/prof_u2.py 3 > gpurun_out/plain_u2.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:blend_ -s 24 -c 12 -o gpurun_out/prof_u2_blend -f python scripts/prof_u2.py 3 > gpurun_out/ncu_u2.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_u2.log
