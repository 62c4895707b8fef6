#!/bin/bash
out=gpurun_out/r02f
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_distance.py tests/test_gpu_linkage.py tests/test_gpu_cli.py tests/test_gpu_multidevice.py tests/test_gpu_cgmlst_dists.py -m gpu -x -q > $out/pytest_a.log 2>&1; echo "pytest a rc=$?"; tail -15 $out/pytest_a.log
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q > $out/pytest_full.log 2>&1; echo "pytest full rc=$?"; tail -5 $out/pytest_full.log
timeout 900 python bench.py --no-cpu-baseline > $out/bench.log 2>&1; echo "bench rc=$?"
python scripts/bench_brief.py < $out/bench.log | tail -3
tail -1 $out/bench.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print({k:d[k] for k in ('ms_per_step','value','result_check','stage_ms')}); print(d['e2e'])"
RTHC_LIB_PATH=$PWD/reportree_b200/build/lib_checked.so timeout 600 python -m pytest tests/test_gpu_distance.py -m gpu -x -q > $out/pytest_checked.log 2>&1; echo "pytest checked rc=$?"; tail -3 $out/pytest_checked.log
python - <<'PY'
import numpy as np, sys
sys.path.insert(0,'.')
from reportree_b200 import hc_backend as hc, synth
X = synth.workload("C4")
with hc.HCContext() as ctx:
    ctx.set_profiles(X); print("C4 info", ctx.info()); ctx.distance(); print("C4 timings", ctx.timings())
PY
