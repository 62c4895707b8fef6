#!/bin/bash
out=gpurun_out/r02j
mkdir -p $out
timeout 400 python -m pytest tests/test_gpu_exact_linkage.py tests/test_gpu_next.py -m gpu -x -q > $out/pytest_chain.log 2>&1; echo "pytest chain rc=$?"; tail -3 $out/pytest_chain.log
timeout 600 python scripts/gpu_configs.py C3 C5 --verify > $out/configs.log 2>&1; echo "configs rc=$?"; cat $out/configs.log
timeout 900 python -m pytest tests/test_gpu_fullsize.py -m gpu -x -q -k "C3 or C5" > $out/pytest_full.log 2>&1; echo "pytest full rc=$?"; tail -3 $out/pytest_full.log
