#!/usr/bin/env python
"""Instruction / sample share per line range. usage: ncu_buckets.py src.csv [file=encode.cuh] name:a-b ..."""
import csv, sys
def num(x):
    try: return int(x)
    except ValueError: return 0
rows = list(csv.reader(open(sys.argv[1])))
h = None; cur = None; agg = {}; fn = None
for r in rows:
    if r and r[0] == 'File Path': fn = r[1]; continue
    if r and r[0] == 'Line No': h = r; continue
    if not h or len(r) < len(h): continue
    if r[0] != '':
        cur = (fn, int(r[0]))
        if r[2] in ('-', ''): continue
    if cur and r[2] not in ('-', ''):
        a = agg.setdefault(cur, [0, 0, 0]); a[0] += num(r[h.index('Instructions Executed')]); a[1] += num(r[h.index('# Samples')]); a[2] += num(r[h.index('Thread Instructions Executed')])
ti = sum(v[0] for v in agg.values()); ts = sum(v[1] for v in agg.values())
print("files:", sorted({k[0] for k in agg}))
specs = sys.argv[2:]
FILE = 'encode.cuh'
if specs and specs[0].startswith('file='): FILE = specs.pop(0)[5:]
for spec in specs:
    name, rng = spec.split(':'); a, b = map(int, rng.split('-'))
    sel = [v for k, v in agg.items() if k[0].endswith(FILE) and a <= k[1] <= b]
    i = sum(v[0] for v in sel); s = sum(v[1] for v in sel); t = sum(v[2] for v in sel)
    print(f"{name:18s} inst {100*i/ti:5.1f}% samples {100*s/ts:5.1f}% thr/inst {t/max(i,1):.1f}")
sel = [v for k, v in agg.items() if not k[0].endswith(FILE)]
i = sum(v[0] for v in sel); s = sum(v[1] for v in sel)
print(f"{'other files':18s} inst {100*i/ti:5.1f}% samples {100*s/ts:5.1f}%")
