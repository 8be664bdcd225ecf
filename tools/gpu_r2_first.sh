#!/bin/bash
# Round-2 baseline: GPU tests, a C2 bench line (no CPU baseline), one full ncu capture of encode_kernel with sources.
tag=${1:-r2a}; out=gpurun_out/$tag; mkdir -p $out
nvidia-smi -L > $out/gpus.txt 2>&1
timeout 600 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -5 $out/pytest_gpu.log
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 400 $B > $out/bench_c2.json 2> $out/bench_c2.err; echo "bench exit $?"
python -c "
import json; d=json.load(open('$out/bench_c2.json')); print('encode_kernel ms', d['device_ms']['encode_kernel'], 'frac', d['roofline']['frac'], 'ingest', d['device_ms']['ingest'], 'value', d['value'])"
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:^(encode_kernel)' -s 3 -c 1 -o $out/prof_c2 $B > $out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -2 $out/ncu_full.log; ls -la $out
