#!/bin/bash
# Timeline of the file-to-file pipeline at C2 (PAFFY_TRACE=1). usage: gpurun --timeout 900 -- 'bash tools/gpu_file_trace.sh tag [extra env...]'
tag=${1:-trace}; shift
out=gpurun_out/$tag; mkdir -p $out
env "$@" PAFFY_TRACE=1 timeout 600 python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench.json 2> $out/trace.txt; echo "exit $?"
python - $out/bench.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1])); print("e2e_file", d.get("e2e_file"))
PY
grep -n "contexts ready" $out/trace.txt | tail -1 | cut -d: -f1 > $out/last.txt
tail -n +$(cat $out/last.txt) $out/trace.txt | head -60
