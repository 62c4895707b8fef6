#!/bin/bash
out=gpurun_out/r02b
mkdir -p $out
timeout 1500 python -m pytest tests -m gpu -q --deselect "tests/test_gpu_fullsize.py::test_full_configs_equal_the_cpu_oracle[C4]" > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_gpu.log
tail -15 $out/pytest_gpu.log
