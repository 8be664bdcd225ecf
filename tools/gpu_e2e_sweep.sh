#!/bin/bash
# e2e (text interface) against the number of sibling contexts / chunks per step. usage: gpurun -- 'bash tools/gpu_e2e_sweep.sh tag "ctxs:chunks" ...'
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
for cs in "$@"; do
  c=${cs%%:*}; s=${cs##*:}
  timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e-file --e2e-contexts $c --e2e-subbatches $s > $out/bench_${c}_${s}.json 2> $out/bench_${c}_${s}.err
  python - $out/bench_${c}_${s}.json $c $s <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(f"ctxs {sys.argv[2]} chunks {sys.argv[3]}: e2e text {d['e2e']['ms_per_step']:.2f} ms/step ({d['e2e']['value']/1e9:.0f} Gbp/s), soa {d['e2e']['soa']['ms_per_step']:.2f} ms/step", {k: round(v, 2) for k, v in d['e2e']['wall_ms_per_step_summed_over_contexts'].items()})
except Exception as e:
    print("failed", sys.argv[2:], e)
PY
done
