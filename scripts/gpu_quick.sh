#!/bin/bash
# quick GPU pass: parity tests + bench (no ncu).  usage: bash scripts/gpu_quick.sh <tag> [extra bench args]
tag=${1:-quick}; shift
out=gpurun_out/$tag; mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 $out/pytest_gpu.log
timeout 900 python bench.py --no-cpu-baseline "$@" > $out/bench.log 2>&1; echo "bench rc=$?"
python scripts/bench_brief.py < $out/bench.log | tail -5
