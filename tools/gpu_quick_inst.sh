#!/bin/bash
# Quick GPU check + instruction count of encode_kernel. usage: gpurun --timeout 900 -- 'bash tools/gpu_quick_inst.sh tag'
tag=${1:-quick}; shift
out=gpurun_out/$tag; mkdir -p $out
timeout 300 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -3 $out/pytest_gpu.log
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline --no-e2e"
timeout 400 $B "$@" > $out/bench.json 2> $out/bench.err; rc=$?; echo "bench exit $rc"
python - <<PY
import json
d=json.load(open("$out/bench.json")); print("device_ms", d["device_ms"], "frac", round(d["roofline"]["frac"],4), d["parity"]["mismatch_counts_equal"], d["parity"]["output_op_counts_equal"])
PY
[ $rc -eq 0 ] && timeout 600 ncu --metrics smsp__inst_executed.sum,smsp__thread_inst_executed.sum,smsp__issue_active.avg.pct_of_peak_sustained_active,gpu__time_duration.sum,smsp__warp_issue_stalled_barrier_per_warp_active.pct,smsp__warp_issue_stalled_long_scoreboard_per_warp_active.pct,smsp__warp_issue_stalled_no_instruction_per_warp_active.pct --clock-control none -k 'regex:^(encode_kernel)' -s 3 -c 1 --csv --log-file $out/inst.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e "$@" > $out/ncu.log 2>&1
grep -E "encode_kernel" $out/inst.csv | awk -F'","' '{print $(NF-2), $(NF)}' | tr -d '"'
