#!/bin/bash
# Launch list (kernel name + duration) of the device-resident steps at C2. usage: gpurun -- 'bash tools/gpu_launches_resident.sh tag [ENV=..]'
tag=${1:-lr}; shift; out=gpurun_out/$tag; mkdir -p $out
KREG='regex:^(encode_kernel|stitch_kernel|ingest_|ig_|scan_onepass|scalars_init|words_out|tile_first_op)'
env "$@" timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" --launch-skip 100 -c 60 --csv --log-file $out/launches_resident_c2.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e --no-e2e-file > $out/ncu_launch.log 2>&1; echo "launch list exit $?"
python - $out/launches_resident_c2.csv <<'PY'
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 5]
h = rows[0]; ni = h.index("Kernel Name"); vi = h.index("Metric Value"); agg = {}
for r in rows[1:]:
    k = r[ni].split("(")[0].replace("void ", "").replace("pafgpu::", ""); a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += float(r[vi].replace(",", ""))
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]): print(f"{k:36s} n={a[0]:3d} mean {a[1]/a[0]/1e3:9.1f} us share {100*a[1]/tot:5.1f}%")
PY
