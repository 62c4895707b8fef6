#!/bin/bash
# Memory / ordering evidence without compute-sanitizer (closed on this GPU pool: it answers rc 86, "runs under it
# have left GPUs needing a reset"): the driver that touches every kernel at small and ragged shapes is run
# (1) against the product library, results compared with scipy, and (2) against build/lib_checked.so
# (-DRTHC_CHECKED: device-side invariants that trap) -- see DESIGN.md section 9.
# usage (on the GPU box, from the repo root): bash scripts/gpu_checked.sh <tag>
tag=${1:-checked}
out=gpurun_out/$tag
mkdir -p $out
timeout 300 python scripts/gpu_sanitize_driver.py > $out/driver_plain.log 2>&1; echo "plain rc=$?" | tee -a $out/driver_plain.log
RTHC_LIB_PATH=$PWD/reportree_b200/build/lib_checked.so timeout 600 python scripts/gpu_sanitize_driver.py > $out/driver_checked.log 2>&1
echo "checked rc=$?" | tee -a $out/driver_checked.log
RTHC_LIB_PATH=$PWD/reportree_b200/build/lib_checked.so timeout 900 python -m pytest tests/test_gpu_distance.py tests/test_gpu_linkage.py tests/test_gpu_exact_linkage.py -m gpu -x -q > $out/pytest_checked.log 2>&1
echo "pytest checked rc=$?" | tee -a $out/pytest_checked.log
timeout 120 /usr/local/cuda/bin/compute-sanitizer --tool memcheck python -c "print(1)" > $out/compute_sanitizer_probe.log 2>&1; echo "compute-sanitizer rc=$?" | tee -a $out/compute_sanitizer_probe.log
for f in driver_plain driver_checked pytest_checked; do tail -n 2 $out/$f.log; done
