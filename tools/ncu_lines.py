#!/usr/bin/env python
"""Per-source-line instruction and stall shares from `ncu --page source --csv --print-source sass,cuda` output."""
import csv, sys
def num(x):
    try: return int(x)
    except ValueError: return 0
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 50
h = None; cur = None; agg = {}
for r in rows:
    if r and r[0] == 'Line No': h = r; continue
    if not h or len(r) < len(h): continue
    if r[0] != '':
        cur = (int(r[0]), r[1]); agg.setdefault(cur, [0, 0, 0]); 
        if r[2] in ('-', ''): continue
    ie = h.index('Instructions Executed'); ns = h.index('# Samples'); te = h.index('Thread Instructions Executed')
    if cur and r[2] not in ('-', ''):
        agg[cur][0] += num(r[ie]); agg[cur][1] += num(r[ns]); agg[cur][2] += num(r[te])
ti = sum(v[0] for v in agg.values()) or 1; ts = sum(v[1] for v in agg.values()) or 1
print("total warp instr", ti, "samples", ts)
for (ln, src), v in sorted(agg.items(), key=lambda x: -x[1][1])[:top]:
    print(f"{ln:4d} inst {100*v[0]/ti:5.1f}% samp {100*v[1]/ts:5.1f}% thr/inst {v[2]/max(v[0],1):4.1f} | {src.strip()[:100]}")
