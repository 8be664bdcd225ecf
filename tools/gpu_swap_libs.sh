#!/bin/bash
# Time prebuilt variants of the library (tools/_var/*.so, built in the authoring container with different nvcc flags) on
# the GPU box: each is copied over paffy_b200/libpaffy_b200.so and benched at C2. usage: bash tools/gpu_swap_libs.sh tag
tag=${1:-swap}
out=gpurun_out/$tag; mkdir -p $out
for so in tools/_var/*.so; do
  v=$(basename $so .so)
  cp $so paffy_b200/libpaffy_b200.so
  timeout 200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench_$v.json 2> $out/bench_$v.err
  python - <<PY
import json
try:
    d=json.load(open("$out/bench_$v.json")); print("[$v]", "encode_kernel ms", round(d["device_ms"]["encode_kernel"],4), "ingest", round(d["device_ms"]["ingest"],4), "ms/step", round(d["ms_per_step"],4), "frac", round(d["roofline"]["frac"],4), d["parity"]["mismatch_counts_equal"], d["parity"]["output_op_counts_equal"])
except Exception as e:
    print("[$v] failed", e); print(open("$out/bench_$v.err").read()[-600:])
PY
done
