#!/bin/bash
# Kernel iteration: GPU tests + C2 bench with the shipped build, then rebuild-and-bench sweeps of -D sets.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_r2_kernel.sh tag "defs1" "defs2" ...'
tag=$1; shift
out=gpurun_out/$tag; mkdir -p $out
timeout 600 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -4 $out/pytest_gpu.log
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
show() { python - "$1" "$2" <<'PY'
import json, sys
try:
    d = json.load(open(sys.argv[1]))
    print(f"[{sys.argv[2]}] encode_kernel ms {d['device_ms']['encode_kernel']:.3f} frac {d['roofline']['frac']:.4f} ingest {d['device_ms']['ingest']:.3f} "
          f"parity {d['parity']['mismatch_counts_equal']} {d['parity']['output_op_counts_equal']}")
except Exception as e:
    print(f"[{sys.argv[2]}] failed: {e}")
PY
}
PAFGPU_DEBUG=1 timeout 300 $B > $out/bench_0.json 2> $out/bench_0.err; show $out/bench_0.json shipped; tail -3 $out/bench_0.err
i=0
for d in "$@"; do
  i=$((i+1))
  if [[ "$d" == env:* ]]; then   # a run-time switch: no rebuild
    env ${d#env:} PAFGPU_DEBUG=1 timeout 300 $B > $out/bench_$i.json 2> $out/bench_$i.err; show $out/bench_$i.json "$d"
    continue
  fi
  PAFGPU_NVCC_DEFS="$d" timeout 600 python -m paffy_b200.build --force > $out/build_$i.log 2>&1 || { echo "build failed: $d"; tail -5 $out/build_$i.log; continue; }
  grep -A2 "encode_kernelILi8" $out/build_$i.log | grep -o "[0-9]* bytes spill stores\|Used [0-9]* registers" | tr '\n' ' '; echo
  PAFGPU_DEBUG=1 timeout 300 $B > $out/bench_$i.json 2> $out/bench_$i.err; show $out/bench_$i.json "$d"
done
