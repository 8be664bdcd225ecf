#!/bin/bash
# Full default bench line(s). usage: gpurun --timeout 1500 -- 'bash tools/gpu_bench.sh tag [bench args]'
tag=${1:-bench}; shift
out=gpurun_out/$tag; mkdir -p $out
timeout 1200 python bench.py "$@" > $out/bench.json 2> $out/bench.err; echo "bench exit $?"
tail -5 $out/bench.err
python - $out/bench.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print("value", d["value"], "ms/step", d["ms_per_step"], "kernel", d["device_ms"], "frac", d["roofline"]["frac"])
print("e2e", d["e2e"] and {k: d["e2e"][k] for k in ("value", "ms_per_step", "h2d_bytes_per_step", "d2h_bytes_per_step")})
print("e2e_file", d.get("e2e_file"))
print("cpu", d.get("cpu_baseline") and {k: v for k, v in d["cpu_baseline"].items() if k != "reference_cli"})
print("cli", d.get("cpu_baseline") and d["cpu_baseline"].get("reference_cli"))
print("parity", d["parity"])
PY
