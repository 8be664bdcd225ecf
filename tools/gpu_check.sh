#!/bin/bash
# One gpurun call: GPU parity tests, the C2 bench line, the ncu launch list and one full capture of the two hot kernels.
# usage: gpurun --timeout 1500 -- 'bash tools/gpu_check.sh [tag]'
tag=${1:-run}
out=gpurun_out/$tag
mkdir -p $out
KREG='regex:^(encode_kernel|stitch_kernel|ingest_kernel|scan_|scalars_init|words_out|tile_first_op|op_text|prefix_copy|line_begin)'
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > $out/smi.csv
timeout 900 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -5 $out/pytest_gpu.log
timeout 600 python bench.py --steps 5 --warmup 3 > $out/bench_c2.json 2> $out/bench_c2.err; echo "bench exit $?"
tail -3 $out/bench_c2.err; cat $out/bench_c2.json
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $B > $out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" -c 400 --csv --log-file $out/launches_c2.csv $B > $out/ncu_launch.log 2>&1
echo "ncu launches exit $?"
timeout 600 $B > $out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:^(encode_kernel)' -s 3 -c 1 -o $out/prof_c2 $B > $out/ncu_full.log 2>&1
echo "ncu full exit $?"
ls -la $out
