#!/usr/bin/env python
"""SASS of a source line range from an `ncu --page source --csv --print-source cuda,sass` export, in address order.
usage: ncu_sass.py src.csv FIRST LAST [min_exec]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
a, b = int(sys.argv[2]), int(sys.argv[3])
mn = int(sys.argv[4]) if len(sys.argv) > 4 else 1000
h = None; fn = None; cur = None; out = []
for r in rows:
    if r and r[0] == 'File Path': fn = r[1]; continue
    if r and r[0] == 'Line No': h = r; continue
    if not h or len(r) < len(h): continue
    if r[0] != '': cur = (fn, int(r[0]))
    if cur and cur[0].endswith('encode.cuh') and a <= cur[1] <= b and r[2].startswith('0x'):
        try: ex = int(r[7])
        except ValueError: ex = 0
        if ex >= mn: out.append((int(r[2], 16), cur[1], r[3].strip(), ex, r[6]))
out.sort()
base = out[0][0] if out else 0
for ad, ln, sass, ex, smp in out: print(f"{ad-base:6x} L{ln:<4d} {ex/1e6:8.2f}M samp {smp:>6s}  {sass}")
print("total warp instr in range: %.1fM" % (sum(o[3] for o in out) / 1e6))
