#!/bin/bash
out=gpurun_out/r02h
mkdir -p $out
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:rthc_nn_chain -c 1 -o $out/nnchain_full -f \
    python scripts/gpu_configs.py C5 > $out/ncu_nnchain.log 2>&1; echo "ncu nnchain rc=$?"; tail -3 $out/ncu_nnchain.log
