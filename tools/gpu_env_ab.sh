#!/bin/bash
# Run-time switches A/B on the resident C2 step (no rebuild). usage: gpurun -- 'bash tools/gpu_env_ab.sh tag "VAR=1 VAR2=3" ...' ("-" = no switch)
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
i=0
for e in "$@"; do
  i=$((i+1)); [ "$e" == "-" ] && e=""
  env $e PAFGPU_DEBUG=1 timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-e2e --no-e2e-file ${BENCH_ARGS} > $out/bench_$i.json 2> $out/bench_$i.err
  python - $out/bench_$i.json "$e" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(f"[{sys.argv[2]}] ingest {d['device_ms']['ingest']:.3f} encode_kernel ms {d['device_ms']['encode_kernel']:.3f} frac {d['roofline']['frac']:.4f} step {d['ms_per_step']:.3f} parity {d['parity']['mismatch_counts_equal']} {d['parity']['output_op_counts_equal']}")
except Exception as e:
    print(f"[{sys.argv[2]}] failed: {e}")
PY
done
