#!/bin/bash
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
echo "cpu.max: $(cat /sys/fs/cgroup/cpu.max 2>/dev/null)"; echo "nproc $(nproc)"; cat /sys/fs/cgroup/cpu.stat 2>/dev/null | tr '\n' ' '; echo
cat /proc/self/status | grep -i "cpus_allowed_list"
python bench.py --steps 60 --warmup 5 --quick 2>&1 >/dev/null | grep "step ms (resident)" | cut -c1-600
cat /sys/fs/cgroup/cpu.stat 2>/dev/null | tr '\n' ' '; echo
python scripts/diag_timeline.py > gpurun_out/timeline.txt 2>&1; head -12 gpurun_out/timeline.txt; grep -n "idle/step" -A30 gpurun_out/timeline.txt | head -50
