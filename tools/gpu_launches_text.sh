#!/bin/bash
# Launch list (kernel name + duration) of the text path (file-to-file passes at C2). usage: gpurun -- 'bash tools/gpu_launches_text.sh tag'
tag=${1:-lt}; out=gpurun_out/$tag; mkdir -p $out
KREG='regex:^(encode_kernel|stitch_kernel|ingest_kernel|scan_onepass|scalars_init|words_out|op_text|line_begin|tx_|prefix_copy|ser_)'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" --launch-skip 160 -c 400 --csv --log-file $out/launches_text.csv \
    python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $out/ncu_launch.log 2>&1; echo "launch list exit $?"
python - $out/launches_text.csv <<'PY'
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
h = rows[0]; ni = h.index("Kernel Name"); vi = h.index("Metric Value"); agg = {}
for r in rows[1:]:
    k = r[ni].split("(")[0].replace("void ", "").replace("pafgpu::", ""); a = agg.setdefault(k, [0, 0.0, 0.0]); v = float(r[vi].replace(",", "")); a[0] += 1; a[1] += v; a[2] = max(a[2], v)
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]): print(f"{k:34s} n={a[0]:3d} mean {a[1]/a[0]/1e3:8.1f} us max {a[2]/1e3:8.1f} us share {100*a[1]/tot:5.1f}%")
PY
