#!/bin/bash
# N-GPU check: the GPU tests that use several devices, then the default bench under torchrun.
# usage: gpurun --gpus N --timeout 1800 -- 'bash tools/gpu_multi.sh tag N [bench args]'
tag=$1; n=$2; shift 2
out=gpurun_out/$tag; mkdir -p $out
nvidia-smi -L > $out/gpus.txt
timeout 600 python -m pytest tests -m gpu -q -k "sharded or clone or chunks" > $out/pytest_multi.log 2>&1; echo "pytest exit $?" >> $out/pytest_multi.log
tail -5 $out/pytest_multi.log
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n "$@" > $out/bench_n$n.json 2> $out/bench_n$n.err
echo "bench exit $?"; tail -5 $out/bench_n$n.err
python - $out/bench_n$n.json <<'PY'
import json, sys
d = json.load(open(sys.argv[1]))
print("value", d["value"], "ms/step", d["ms_per_step"], "scaling", d["scaling"], "n_gpus", d["n_gpus"])
print("workload", d["config"]["workload"])
print("shards", d["config"]["shards"])
print("e2e", d["e2e"] and {k: d["e2e"][k] for k in ("value", "ms_per_step")})
print("e2e_file", d.get("e2e_file"))
print("gen_s", d["gen_s"], "kernel", d["device_ms"], "frac", d["roofline"]["frac"])
PY
