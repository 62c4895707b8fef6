"""On-box probe of the single-linkage stage alone (C4 matrix): linkage + cut timings."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from reportree_b200 import hc_backend as hc, synth

n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
cfg = synth.config("C4"); cfg["n"] = n
X = synth.allele_matrix(**cfg)
with hc.HCContext() as ctx:
    ctx.set_profiles(X); ctx.distance()
    for rep in range(3):
        t0 = time.perf_counter(); Z = ctx.linkage("single"); t1 = time.perf_counter()
        lab = ctx.fcluster(np.arange(51.0)); t2 = time.perf_counter()
        tm = ctx.timings()
        print("rep %d linkage %.1f ms (device %.1f ms, rounds %d) cut51 %.1f ms (device %.1f)" % (
            rep, 1e3 * (t1 - t0), tm["linkage_ms"], ctx.info()["boruvka_rounds"], 1e3 * (t2 - t1), tm["cut_ms"]))
    print("checksum", float(Z[:, 2].sum()), int(lab.sum()))
