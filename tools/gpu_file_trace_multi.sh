#!/bin/bash
# Timeline of the file-to-file pipeline with rank 0 driving N GPUs (PAFFY_TRACE=1). usage: gpurun --gpus N -- 'bash tools/gpu_file_trace_multi.sh tag N'
tag=$1; n=$2; out=gpurun_out/$tag; mkdir -p $out
PAFFY_TRACE=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $n --steps 2 --warmup 3 --no-cpu-baseline --no-e2e > $out/bench.json 2> $out/trace.txt; echo "exit $?"
python - $out/bench.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1])); print("e2e_file", {k: v for k, v in d.get("e2e_file", {}).items() if k != "what"})
PY
grep -n "contexts ready" $out/trace.txt | tail -2 | head -1 | cut -d: -f1 > $out/last.txt
tail -n +$(cat $out/last.txt) $out/trace.txt | head -150 > $out/trace_tail.txt; rm -f $out/trace.txt
head -90 $out/trace_tail.txt
