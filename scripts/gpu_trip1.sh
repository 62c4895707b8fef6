#!/bin/bash
# round 2, trip 1: parity of the NN-cached merge loop, C3 golden, configs timing, sanitizers
out=gpurun_out/r02a
mkdir -p $out
timeout 1200 python -m pytest tests -m gpu -x -q --deselect "tests/test_gpu_fullsize.py::test_full_configs_equal_the_cpu_oracle[C4]" > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_gpu.log
tail -5 $out/pytest_gpu.log
timeout 600 python scripts/gpu_configs.py C3 C5 --verify > $out/configs.log 2>&1; echo "configs rc=$?" | tee -a $out/configs.log
cat $out/configs.log
bash scripts/gpu_sanitize.sh r02a_san
