"""Aggregate an `ncu --page source --csv --print-source cuda,sass` export by source line: instruction and sample shares."""
import csv, sys
from collections import defaultdict
rows = list(csv.reader(open(sys.argv[1])))
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur = None; hdr = None; ci = {}
agg = defaultdict(lambda: [0, 0]); tot = [0, 0]; stall = defaultdict(int)
def num(x):
    try: return int(x)
    except Exception: return 0
for r in rows:
    if not r: continue
    if r[0] == 'File Path': cur = r[1].split('/')[-1]; continue
    if r[0] == 'Function Name': continue
    if r[0] == 'Line No':
        hdr = r; ci = {}
        for i, h in enumerate(r): ci.setdefault(h, i)
        continue
    if hdr is None or len(r) < len(hdr): continue
    if not r[0].strip().isdigit(): continue
    if r[ci['Address']].strip() not in ('', '-'): continue
    s, i = num(r[ci['# Samples']]), num(r[ci['Instructions Executed']])
    key = (cur, int(r[0]), r[1].strip()[:100])
    agg[key][0] += s; agg[key][1] += i; tot[0] += s; tot[1] += i
    for h in hdr:
        if h.startswith('stall_') and 'Not Issued' not in h: stall[h] += num(r[ci[h]])
print('total samples', tot[0], 'warp-inst', tot[1])
for k, v in sorted(stall.items(), key=lambda x: -x[1])[:8]: print(f'  {k:26s}{100*v/max(tot[0],1):5.1f}%')
for (f, ln, src), (s, i) in sorted(agg.items(), key=lambda x: -x[1][1])[:topn]:
    print(f'{100*i/max(tot[1],1):5.1f}% inst {100*s/max(tot[0],1):5.1f}% smp  {f}:{ln}  {src}')
