#!/bin/bash
# e2e sweep. usage: bash tools/gpu_e2e.sh tag "subbatches:contexts" ...
out=gpurun_out/$1; mkdir -p $out; shift
for spec in "$@"; do
  n=${spec%%:*}; c=${spec##*:}
  timeout 400 python bench.py --steps 5 --warmup 3 --no-cpu-baseline --e2e-subbatches $n --e2e-contexts $c > $out/bench_${n}_$c.json 2> $out/bench_${n}_$c.err
  python - <<PY
import json
try:
    d=json.load(open("$out/bench_${n}_$c.json")); print("sub=$n ctx=$c", "e2e", round(d["e2e"]["value"]/1e9,1), "Gbp/s", round(d["e2e"]["ms_per_step"],2), "ms", d["e2e"]["host_wall_ms_per_step_summed_over_threads"])
except Exception as e:
    print("sub=$n failed", e); print(open("$out/bench_${n}_$c.err").read()[-1500:])
PY
done
