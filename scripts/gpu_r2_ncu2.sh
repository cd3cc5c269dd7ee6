#!/bin/bash
# ncu evidence, final state of round 2: launch list of the bench command (durations only), `--set full` of one whole
# training step's kernels (summary + DRAM traffic + warp-stall breakdown), source-level dumps of two kernels.
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 600 python bench.py --steps 2 --warmup 3 --quick > gpurun_out/plain.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none --kernel-name-base demangled -k regex:r3dgw -c 700 --csv --log-file gpurun_out/launches.csv python bench.py --steps 2 --warmup 3 --quick > gpurun_out/ncu_list.log 2>&1
echo "ncu list rc=$?"; wc -l gpurun_out/launches.csv
timeout 300 python scripts/prof_u2.py 3 > gpurun_out/plain_u2.log 2>&1 && \
timeout 1800 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:r3dgw -s 98 -c 56 -o gpurun_out/prof_u2_r02c -f python scripts/prof_u2.py 3 > gpurun_out/ncu_u2.log 2>&1
echo "ncu full rc=$?"; tail -2 gpurun_out/ncu_u2.log
python scripts/ncu_summarize.py gpurun_out/prof_u2_r02c.ncu-rep gpurun_out/ncu_u2_summary.txt gpurun_out/ncu_u2_traffic.json > /dev/null
ncu -i gpurun_out/prof_u2_r02c.ncu-rep --page raw --csv > gpurun_out/ncu_u2_raw.csv
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/ncu_u2_raw.csv')))
hdr=next(i for i,r in enumerate(rows) if 'Kernel Name' in r)
names=rows[hdr]; col={n:i for i,n in enumerate(names)}
stall=[n for n in names if n.startswith('smsp__average_warps_issue_stalled') and n.endswith('_per_issue_active.ratio')]
seen=set(); out=[]
for r in rows[hdr+2:]:
    k=r[col['Kernel Name']].split('(')[0].replace('void ','').replace('r3dgw::','')
    if k in seen: continue
    seen.add(k)
    vals=sorted(((float(r[col[s]].replace(',','')) if r[col[s]] not in ('','n/a') else 0.0, s.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')) for s in stall), reverse=True)[:6]
    out.append(f"{k}: dur {r[col['gpu__time_duration.sum']]} " + ' '.join(f'{n}={v:.2f}' for v,n in vals))
open('gpurun_out/ncu_u2_stalls.txt','w').write('\n'.join(out)+'\n'); print('\n'.join(out))
PY
for k in blend_backward_geometry_cached_kernel shade_backward_kernel blend_backward_color_cached_kernel blend_forward_cached_kernel; do
  ncu -i gpurun_out/prof_u2_r02c.ncu-rep --page source --csv --kernel-name-base demangled -k regex:$k -c 1 > gpurun_out/src_$k.csv 2>/dev/null
  python scripts/ncu_src_summary.py gpurun_out/src_$k.csv 0.02 > gpurun_out/src_$k.txt 2>&1; head -30 gpurun_out/src_$k.txt
done
python scripts/diag_timeline.py > gpurun_out/timeline.txt 2>&1
# the report itself is > 64 MiB (gpurun_out/ would not travel back): keep the summaries, drop the report
rm -f gpurun_out/prof_u2_r02c.ncu-rep gpurun_out/ncu_u2_raw.csv gpurun_out/src_*.csv
