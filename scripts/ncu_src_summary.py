"""Summarise an `ncu --page source --csv` dump: runs of SASS with equal execution counts (usage: file [min_share])."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
thr = float(sys.argv[2]) if len(sys.argv) > 2 else 0.01
hdr = rows[1]
ia, ie, ist = hdr.index('Source'), hdr.index('Instructions Executed'), hdr.index('# Samples')
body = []
for r in rows[2:]:
    if len(r) <= ie or r[0] == 'Kernel Name':
        break
    body.append(r)
tot = sum(int(r[ie]) for r in body); tots = sum(int(r[ist]) for r in body)
print('kernel:', rows[0][1][:80]); print('total warp-instr', tot, 'SASS lines', len(body), 'samples', tots)
prev = None; start = 0; acc = 0; sacc = 0; runs = []
for k, r in enumerate(body):
    n = int(r[ie]); key = round(n / 1e6, 1)
    if key != prev:
        if prev is not None: runs.append((start, k - 1, prev, acc, sacc))
        prev = key; start = k; acc = 0; sacc = 0
    acc += n; sacc += int(r[ist])
runs.append((start, len(body) - 1, prev, acc, sacc))
for s, e, key, acc, sacc in runs:
    if acc > thr * tot or sacc > thr * tots:
        print(f"lines {s:4d}-{e:4d} ({e-s+1:3d} instr) x {key:8.1f}M = {acc/1e6:8.1f}M ({acc/tot:6.1%}) samples {sacc/tots:6.1%}  first: {body[s][ia].strip()[:50]}")
