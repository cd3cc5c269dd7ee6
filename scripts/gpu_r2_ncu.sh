#!/bin/bash
# ncu evidence for round 2: launch list of the bench command (durations only) + `--set full` of the step's kernels.
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --quick > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:r3dgw -c 700 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --quick > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/launches.csv
timeout 300 python scripts/prof_u2.py 3 > gpurun_out/plain_u2.log 2>&1 && \
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"blend_|shade_|geometry_backward|gaussian_reg|normal_tail|normal_loss|ssim" -s 60 -c 40 -o gpurun_out/prof_u2_r02 -f python scripts/prof_u2.py 3 > gpurun_out/ncu_u2.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_u2.log
