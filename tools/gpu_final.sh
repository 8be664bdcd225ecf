#!/bin/bash
# Final evidence of a round: GPU tests, the default bench line (with the CPU baseline), the reference arm, a launch list of
# the device-resident steps. usage: gpurun --timeout 2400 -- 'bash tools/gpu_final.sh tag'
tag=${1:-final}; out=gpurun_out/$tag; mkdir -p $out
timeout 600 python -m pytest tests -m gpu -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log; tail -3 $out/pytest_gpu.log
python -c "import __graft_entry__ as g; g.smoke()" > $out/smoke.log 2>&1; echo "smoke exit $?"; tail -1 $out/smoke.log
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > $out/bench_reference.json 2> $out/bench_reference.err; echo "reference arm exit $?"
timeout 900 python bench.py > $out/bench.json 2> $out/bench.err; echo "bench exit $?"; tail -3 $out/bench.err
KREG='regex:^(encode_kernel|stitch_kernel|ingest_kernel|scan_onepass|scalars_init|words_out|tile_first_op)'
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" --launch-skip 100 -c 49 --csv --log-file $out/launches_resident_c2.csv \
    python bench.py --steps 4 --warmup 3 --no-cpu-baseline --no-e2e --no-e2e-file > $out/ncu_launch.log 2>&1; echo "launch list exit $?"
python - $out <<'PY'
import json, sys, csv
o = sys.argv[1]
d = json.load(open(o + "/bench.json"))
print("value", d["value"], "ms/step", d["ms_per_step"], "frac", d["roofline"]["frac"], "traffic", d["roofline"].get("traffic"), "balance", d["roofline"].get("load_balance", {}).get("max_over_mean"))
print("e2e", d["e2e"]["value"], "e2e_file", d["e2e_file"].get("value"), d["e2e_file"].get("output_to_dev_null"))
print("cpu", {k: v for k, v in d["cpu_baseline"].items() if k != "reference_cli" and k != "sample"})
print("ref arm", open(o + "/bench_reference.json").read()[:400])
rows = [r for r in csv.reader(open(o + "/launches_resident_c2.csv")) if len(r) > 5]
h = rows[0]; ni = h.index("Kernel Name"); vi = h.index("Metric Value"); agg = {}
for r in rows[1:]:
    k = r[ni].split("(")[0]; a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += float(r[vi].replace(",", ""))
tot = sum(a[1] for a in agg.values())
for k, a in sorted(agg.items(), key=lambda x: -x[1][1]): print(f"{k:40s} n={a[0]:3d} mean {a[1]/a[0]/1e3:9.1f} us share {100*a[1]/tot:5.1f}%")
PY
