#!/bin/bash
# bench under different environment settings. usage: bash tools/gpu_envsweep.sh tag "VAR=1" "VAR=2 OTHER=3" ...
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
i=0
for e in "$@"; do
  i=$((i+1))
  env $e timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench_$i.json 2> $out/bench_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$out/bench_$i.json")); print("[$e]", "encode_kernel ms", round(d["device_ms"]["encode_kernel"],3), "frac", round(d["roofline"]["frac"],4), d["parity"]["mismatch_counts_equal"], d["parity"]["output_op_counts_equal"])
except Exception as ex:
    print("[$e] failed", ex); print(open("$out/bench_$i.err").read()[-800:])
PY
done
