#!/bin/bash
# GPU tests only (optionally a -k filter). usage: gpurun --timeout 900 -- 'bash tools/gpu_tests.sh tag [pytest args]'
tag=${1:-t}; shift
out=gpurun_out/$tag; mkdir -p $out
timeout 800 python -m pytest tests -m gpu -q "$@" > $out/pytest_gpu.log 2>&1; echo "pytest exit $?" >> $out/pytest_gpu.log
tail -60 $out/pytest_gpu.log
