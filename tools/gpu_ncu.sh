#!/bin/bash
# ncu launch list + one full capture of encode_kernel at C2. usage: gpurun --timeout 1500 -- 'bash tools/gpu_ncu.sh tag'
tag=${1:-ncu}; out=gpurun_out/$tag; mkdir -p $out
KREG='regex:^(encode_kernel|stitch_kernel|ingest_kernel|scan_|scalars_init|words_out|tile_first_op|op_text|prefix_copy|line_begin)'
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $B > $out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" -c 400 --csv --log-file $out/launches_c2.csv $B > $out/ncu_launch.log 2>&1
echo "ncu launches exit $?"
timeout 600 $B > $out/plain2.log 2>&1 &&
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:^(encode_kernel)' -s 3 -c 1 -o $out/prof_c2 $B > $out/ncu_full.log 2>&1
echo "ncu full exit $?"; tail -3 $out/ncu_full.log
