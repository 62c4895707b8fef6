"""On-box: where the _dist_hamming.tsv time goes at n samples (D2H of row blocks / formatting / copy into the file)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from reportree_b200 import fastio, hc_backend as hc, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
X = synth.allele_matrix(n, 600, 5)
os.environ["RTHIO_TIMING"] = "1"
with hc.HCContext() as ctx:
    ctx.set_profiles(X); ctx.distance()
    names = ["S%07d" % (k + 1) for k in range(n)]
    t = time.perf_counter()
    for r0 in range(0, n, 1024):
        ctx.square_rows(r0, min(1024, n - r0))
    print("D2H of all row blocks alone: %.2f s" % (time.perf_counter() - t), flush=True)
    for block in (1024, 4096):
        t = time.perf_counter()
        fastio.write_square_tsv("/tmp/sq_probe.tsv", names, ctx.square_rows, n, block=block)
        dt = time.perf_counter() - t
        print("write_square_tsv block=%d: %.2f s, %.2f GB/s" % (block, dt, os.path.getsize("/tmp/sq_probe.tsv") / dt / 1e9), flush=True)
        os.remove("/tmp/sq_probe.tsv")
