#!/bin/bash
# On-box: partitions of the C4 CLI run -> stability drop-in, wall time of the whole process (both modes)
set -e
python scripts/gpu_cli_e2e.py ${1:-C4} --skip-dist | grep -E "wall|partitions"
W=${E2E_DIR:-/tmp/rthc_e2e}; C=${1:-C4}
for extra in "--keep-redundants" ""; do
  s=$(date +%s.%N)
  python -m reportree_b200.comparing_partitions_v2 -i1 $W/${C}_partitions.tsv -o1 0 -a stability -t $W/${C} -n 5 -thr 0.99 -log $W/${C}_stab.log $extra > $W/stab_stdout.txt
  e=$(date +%s.%N)
  echo "stability CLI ($C, '$extra'): $(python -c "print('%.2f' % ($e - $s))") s wall; $(wc -l < $W/${C}_metrics.tsv) metric lines; regions:"; tail -n +3 $W/${C}_stableRegions.tsv | head -5
done
