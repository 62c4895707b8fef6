#!/bin/bash
# One GPU-box pass: parity tests, smoke, bench (both arms), configs, checked build, launch list, ncu captures.
# usage (from the repo root on the box): bash scripts/gpu_round.sh <tag>
tag=${1:-run}
out=gpurun_out/$tag
mkdir -p $out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > $out/gpu.txt 2>&1
nproc > $out/nproc.txt; lscpu | grep -E "Model name|^CPU\(s\)" >> $out/nproc.txt; free -g | head -2 >> $out/nproc.txt
timeout 1800 python -m pytest tests -m gpu -x -q > $out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a $out/pytest_gpu.log
timeout 600 python __graft_entry__.py smoke > $out/smoke.log 2>&1; echo "smoke rc=$?" | tee -a $out/smoke.log
timeout 900 python bench.py > $out/bench.log 2>&1; echo "bench rc=$?" | tee -a $out/bench.log
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > $out/bench_ref.log 2>&1; echo "ref rc=$?" | tee -a $out/bench_ref.log
timeout 900 python scripts/gpu_configs.py C3 C5 C4prim --verify > $out/configs.log 2>&1; echo "configs rc=$?" | tee -a $out/configs.log
bash scripts/gpu_checked.sh $tag/checked > $out/checked.log 2>&1
timeout 300 python scripts/gpu_collide_probe.py C4 C5 C3 0 24576 > $out/collide_probe.log 2>&1; echo "collide probe rc=$?" | tee -a $out/collide_probe.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file $out/launches.csv \
    python bench.py --steps 1 --warmup 1 --no-cpu-baseline > $out/ncu_launch.log 2>&1; echo "ncu launches rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:rthc_distance -c 1 -o $out/dist_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline > $out/ncu_full.log 2>&1; echo "ncu full rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:rthc_(rank32_mark|rank32_code|collide|collide_fixup|pack_planes|boruvka_scan_tile|cut_leaders|cut_labels)" -c 10 -o $out/stages_full -f \
    python bench.py --steps 1 --warmup 0 --no-cpu-baseline > $out/ncu_stages.log 2>&1; echo "ncu stages rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:rthc_prim_cluster -c 1 -o $out/prim_full -f \
    python scripts/gpu_configs.py C4prim > $out/ncu_prim.log 2>&1; echo "ncu prim rc=$?"
timeout 1200 ncu --set full --clock-control none --import-source on -k regex:rthc_nn_chain -c 1 -o $out/nnchain_full -f \
    python scripts/gpu_configs.py C5 > $out/ncu_nnchain.log 2>&1; echo "ncu nnchain rc=$?"
tail -3 $out/pytest_gpu.log; tail -2 $out/bench.log | cut -c1-1500; cat $out/configs.log; cat $out/collide_probe.log; tail -4 $out/checked.log
