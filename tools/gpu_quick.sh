#!/bin/bash
# Quick GPU check: parity tests + one C2 bench line (no ncu). usage: gpurun --timeout 900 -- 'bash tools/gpu_quick.sh tag [bench args]'
tag=${1:-quick}; shift
out=gpurun_out/$tag; mkdir -p $out
timeout 300 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -30 $out/pytest_gpu.log
timeout 400 python bench.py --steps 5 --warmup 3 "$@" > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
tail -5 $out/bench.err; cat $out/bench.json
