#!/bin/bash
# Rebuild the library on the GPU box with different -D sets and bench each. usage: bash tools/gpu_sweep_defs.sh tag "defs1" "defs2" ...
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
i=0
for d in "$@"; do
  i=$((i+1))
  PAFGPU_NVCC_DEFS="$d" python -m paffy_b200.build --force > $out/build_$i.log 2>&1 || { echo "build failed: $d"; tail -5 $out/build_$i.log; continue; }
  PAFGPU_DEBUG=1 timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench_$i.json 2> $out/bench_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$out/bench_$i.json")); print("[$d]", "encode_kernel ms", round(d["device_ms"]["encode_kernel"],3), "frac", round(d["roofline"]["frac"],4), d["parity"]["mismatch_counts_equal"], d["parity"]["output_op_counts_equal"])
except Exception as e:
    print("[$d] failed", e); print(open("$out/bench_$i.err").read()[-800:])
PY
done
