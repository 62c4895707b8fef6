#!/bin/bash
out=gpurun_out/r02k
mkdir -p $out
timeout 900 python -m pytest tests/test_gpu_linkage.py tests/test_gpu_exact_linkage.py tests/test_gpu_next.py tests/test_gpu_cli.py -m gpu -x -q > $out/pytest.log 2>&1; echo "pytest rc=$?"; tail -3 $out/pytest.log
RTHC_LIB_PATH=$PWD/reportree_b200/build/lib_checked.so timeout 600 python -m pytest tests/test_gpu_linkage.py -m gpu -x -q > $out/pytest_checked.log 2>&1; echo "pytest checked rc=$?"; tail -2 $out/pytest_checked.log
timeout 900 python bench.py --no-cpu-baseline > $out/bench.log 2>&1; echo "bench rc=$?"
python scripts/bench_brief.py < $out/bench.log | tail -3
head -1 $out/bench.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['host_wall_ms_rank0'], d['result_check']['oracle_match'], d['gpu_launches'])"
