#!/bin/bash
# ncu --set full of the U1 forward chain (one launch per kernel), with the warp-stall breakdown
cd "$GRAFT_REPO_ROOT"; mkdir -p gpurun_out
timeout 300 python scripts/prof_u1.py B 3 > gpurun_out/plain_u1.log 2>&1 || { echo plain failed; tail -5 gpurun_out/plain_u1.log; exit 1; }
timeout 1500 ncu --set full --clock-control none --import-source on -k regex:"preprocess|duplicate|onesweep|radix_hist|ranges|blend_forward|blend_backward_kernel|geometry_backward" -s 30 -c 20 -o gpurun_out/prof_u1_r02b -f python scripts/prof_u1.py B 3 > gpurun_out/ncu_u1.log 2>&1
echo "ncu rc=$?"; tail -2 gpurun_out/ncu_u1.log
python scripts/ncu_summarize.py gpurun_out/prof_u1_r02b.ncu-rep gpurun_out/ncu_u1_summary.txt gpurun_out/ncu_u1_traffic.json > /dev/null
ncu -i gpurun_out/prof_u1_r02b.ncu-rep --page raw --csv > gpurun_out/ncu_u1_raw.csv
python - <<'PY'
import csv
rows=list(csv.reader(open('gpurun_out/ncu_u1_raw.csv')))
hdr=next(i for i,r in enumerate(rows) if 'Kernel Name' in r)
names=rows[hdr]; col={n:i for i,n in enumerate(names)}
stall=[n for n in names if n.startswith('smsp__average_warps_issue_stalled') and n.endswith('_per_issue_active.ratio')]
seen=set()
out=[]
for r in rows[hdr+2:]:
    k=r[col['Kernel Name']].split('(')[0].replace('void ','').replace('r3dgw::','')
    if k in seen: continue
    seen.add(k)
    vals=sorted(((float(r[col[s]].replace(',','')) if r[col[s]] not in ('','n/a') else 0.0, s.replace('smsp__average_warps_issue_stalled_','').replace('_per_issue_active.ratio','')) for s in stall), reverse=True)[:6]
    out.append(f"{k}: dur {r[col['gpu__time_duration.sum']]} " + ' '.join(f'{n}={v:.2f}' for v,n in vals))
open('gpurun_out/ncu_u1_stalls.txt','w').write('\n'.join(out)+'\n'); print('\n'.join(out))
PY
