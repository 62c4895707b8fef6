#!/usr/bin/env python3
"""Times the UNMODIFIED reference script (/root/reference/scripts/partitioning_HC.py: pandas + the external
cgmlst-dists through its 2,000-column chunk protocol + scipy linkage / fcluster + its naming loop) on a bounded
slice of the bench workload, in THIS container (the reference tree does not exist on the GPU box), with the CPU
oracle binary oracle/_build/cgmlst-dists standing in for the un-vendored program on PATH.

    python scripts/time_reference_script.py [n_samples] -> profiles/r02_reference_script.json
"""
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np                                       # noqa: E402
from oracle import oracle_lib                            # noqa: E402
from reportree_b200 import synth                         # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 3000
oracle_lib.build()
X = synth.allele_matrix(**synth.config("C4"))[:n]
L = X.shape[1]
tmp = tempfile.mkdtemp()
tsv = os.path.join(tmp, "c4_first_%d.tsv" % n)
with open(tsv, "w") as f:
    f.write("FILE\t" + "\t".join("L%05d" % (k + 1) for k in range(L)) + "\n")
    for i in range(n):
        f.write("S%07d\t" % (i + 1) + "\t".join(map(str, X[i].tolist())) + "\n")
env = dict(os.environ, PATH=os.path.dirname(oracle_lib.CGMLST_DISTS_BIN) + ":" + os.environ["PATH"])
cmd = [sys.executable, "/root/reference/scripts/partitioning_HC.py", "-a", tsv, "-o", os.path.join(tmp, "ref"),
       "--HC-threshold", "single-0-50"]
t0 = time.perf_counter()
out = subprocess.run(cmd, env=env, capture_output=True, text=True)
wall = time.perf_counter() - t0
if out.returncode != 0:
    raise SystemExit(out.stdout[-2000:] + out.stderr[-2000:])
cpu = [ln.split(":", 1)[1].strip() for ln in open("/proc/cpuinfo") if ln.startswith("model name")]
P = n * (n - 1) // 2
rec = {"what": "unmodified /root/reference/scripts/partitioning_HC.py -a <tsv> --HC-threshold single-0-50, "
               "oracle/_build/cgmlst-dists on PATH (1 thread, 2,000-column chunks, TSV round trips)",
       "sample": "first %d of 100000 samples x %d loci of the C4 workload" % (n, L),
       "wall_s": round(wall, 2), "comparisons": P * L, "value": P * L / wall, "unit": "comparisons/s",
       "cores": 1, "kind": "reference", "where": "build container (no GPU box access to /root/reference)",
       "host_cpu": cpu[0] if cpu else None, "nproc": os.cpu_count(),
       "python": sys.version.split()[0]}
dst = os.path.join(ROOT, "profiles", "r02_reference_script.json")
json.dump(rec, open(dst, "w"), indent=1)
print(json.dumps(rec))
