#!/bin/bash
out=gpurun_out/r02d
mkdir -p $out
timeout 600 python scripts/gpu_configs.py C3 C5 > $out/configs.log 2>&1; echo "configs rc=$?" | tee -a $out/configs.log
cat $out/configs.log
timeout 1800 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_gpu.log
tail -5 $out/pytest_gpu.log
timeout 900 python bench.py > $out/bench.log 2>&1; echo "bench rc=$?"
python scripts/bench_brief.py < $out/bench.log | tail -5
tail -1 $out/bench.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('ms_per_step','value','e2e','result_check','exact_names_mode','stage_ms')})"
timeout 300 scripts/experiments/_build/dist_probes > $out/dist_probes.log 2>&1; echo "probes rc=$?"; cat $out/dist_probes.log
bash scripts/gpu_checked.sh r02d/checked
