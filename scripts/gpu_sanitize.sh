#!/bin/bash
# compute-sanitizer over every kernel of the library ("race detection / sanitizers").
# usage (on the GPU box, from the repo root): bash scripts/gpu_sanitize.sh <tag>
tag=${1:-san}
out=gpurun_out/$tag
mkdir -p $out
CS=/usr/local/cuda/bin/compute-sanitizer
timeout 300 python scripts/gpu_sanitize_driver.py > $out/driver_plain.log 2>&1; echo "plain rc=$?" | tee -a $out/driver_plain.log
for tool in memcheck racecheck synccheck initcheck; do
    arg=""
    [ $tool = racecheck ] && arg="quick"
    timeout 1500 $CS --tool $tool --error-exitcode 9 --print-limit 50 python scripts/gpu_sanitize_driver.py $arg \
        > $out/sanitizer_$tool.log 2>&1
    echo "$tool rc=$?" | tee -a $out/sanitizer_$tool.log
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|rc=" $out/sanitizer_$tool.log | tail -3
done
