#!/bin/bash
# Round-2 profiles: (1) launch list (kernel name + duration) of one device-resident step and one file-to-file pass at C2,
# (2) one full capture of encode_kernel, (3) full captures of the other kernels of the path (ingest, text parse / render,
# serialiser). usage: gpurun --timeout 2400 -- 'bash tools/gpu_ncu_r2.sh tag'
tag=${1:-r2ncu}; out=gpurun_out/$tag; mkdir -p $out
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 600 $B > $out/plain.json 2> $out/plain.err || { echo "plain bench failed"; tail -5 $out/plain.err; exit 1; }
KREG='regex:^(encode_kernel|stitch_kernel|ingest_kernel|scan_onepass|scalars_init|words_out|tile_first_op|op_text|line_begin|tx_|prefix_copy)'
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -k "$KREG" --launch-skip 1400 -c 500 --csv --log-file $out/launches_c2.csv $B > $out/ncu_launch.log 2>&1
echo "launch list exit $?"
timeout 1200 ncu --set full --clock-control none --import-source on -k 'regex:^(encode_kernel)' -s 3 -c 1 -o $out/prof_encode $B > $out/ncu_full.log 2>&1
echo "encode capture exit $?"
timeout 1500 ncu --set full --clock-control none -k 'regex:^(ingest_kernel|tx_|op_text|line_begin|stitch_kernel)' -c 40 -o $out/prof_others $B > $out/ncu_others.log 2>&1
echo "others capture exit $?"
# the report of 40 full captures is tens of MB (gpurun brings back at most 64 MiB): keep its raw page as CSV only
ncu -i $out/prof_others.ncu-rep --page raw --csv > $out/others_raw.csv 2>/dev/null && rm -f $out/prof_others.ncu-rep
ls -la $out
