#!/bin/bash
out=gpurun_out/r02g
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_exact_linkage.py tests/test_gpu_next.py -m gpu -x -q > $out/pytest_chain.log 2>&1; echo "pytest chain rc=$?"; tail -3 $out/pytest_chain.log
RTHC_LIB_PATH=$PWD/reportree_b200/build/lib_checked.so timeout 400 python -m pytest tests/test_gpu_exact_linkage.py -m gpu -x -q > $out/pytest_chain_checked.log 2>&1; echo "pytest chain checked rc=$?"; tail -3 $out/pytest_chain_checked.log
timeout 600 python scripts/gpu_configs.py C3 C5 --verify > $out/configs.log 2>&1; echo "configs rc=$?"; cat $out/configs.log
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "C3 or C5" > $out/pytest_full.log 2>&1; echo "pytest full rc=$?"; tail -3 $out/pytest_full.log
timeout 900 python bench.py --no-cpu-baseline > $out/bench.log 2>&1; echo "bench rc=$?"
python scripts/bench_brief.py < $out/bench.log | tail -3
tail -1 $out/bench.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['host_wall_ms_rank0'], d['result_check']['oracle_match'])"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:rthc_(rank32|pack_planes|escape_correct)" -c 3 -o $out/recode_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline > $out/ncu_recode.log 2>&1; echo "ncu recode rc=$?"
