#!/bin/bash
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
i=0
for d in "$@"; do
  i=$((i+1))
  PAFGPU_NVCC_DEFS="$d" python -m paffy_b200.build --force > $out/build_$i.log 2>&1 || { echo "build failed: $d"; tail -5 $out/build_$i.log; continue; }
  timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench_$i.json 2> $out/bench_$i.err
  python - <<PY
import json
try:
    d=json.load(open("$out/bench_$i.json")); print("[$d]", d["device_ms"], round(d["value"]/1e12,3))
except Exception as e:
    print("[$d] failed", e); print(open("$out/bench_$i.err").read()[-800:])
PY
done
