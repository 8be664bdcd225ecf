#!/bin/bash
# Sweep an env var over values: bash tools/gpu_sweep.sh tag VAR v1 v2 ... ; prints encode_kernel ms per value
tag=$1; var=$2; shift 2
out=gpurun_out/$tag; mkdir -p $out
for v in "$@"; do
  env $var=$v timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench_$v.json 2> $out/bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("$out/bench_$v.json")); print("$var=$v", d["device_ms"], "frac", round(d["roofline"]["frac"],4), d["parity"]["mismatch_counts_equal"], d["parity"]["output_op_counts_equal"])
except Exception as e:
    print("$var=$v failed", e); print(open("$out/bench_$v.err").read()[-800:])
PY
done
