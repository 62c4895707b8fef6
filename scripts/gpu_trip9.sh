#!/bin/bash
out=gpurun_out/r02i
mkdir -p $out
RTHC_NNC_TIMING=1 timeout 600 python scripts/gpu_configs.py C3 C5 > $out/configs.log 2>&1; echo "configs rc=$?"; cat $out/configs.log
