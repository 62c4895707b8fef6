#!/bin/bash
# scaling run on one 8-GPU box: N = 1, 2, 4, 8 back to back (same launch line the driver uses)
out=gpurun_out/${1:-scale}; mkdir -p $out
for N in ${NLIST:-1 2 4 8}; do
  if [ $N = 1 ]; then python bench.py --gpus 1 --steps 5 --warmup 3 --no-cpu-baseline > $out/bench_n$N.log 2>&1
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port $((29500+N)) bench.py --gpus $N --steps 5 --warmup 3 > $out/bench_n$N.log 2>&1; fi
  echo "N=$N rc=$?"; grep '^{' $out/bench_n$N.log | python scripts/bench_brief.py
done
