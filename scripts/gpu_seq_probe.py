"""On-box: sequential linkage kernels (Prim-exact, NN-chain) vs RTHC_SEQ_BLOCKS."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from reportree_b200 import hc_backend as hc, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 20000
X = synth.allele_matrix(n, 3000, 5)
with hc.HCContext() as ctx:
    ctx.set_profiles(X); ctx.distance()
    ctx.set_single_algo(hc.SL_PRIM)
    ref = {}
    for nb in sys.argv[2:] or ["148"]:
        os.environ["RTHC_SEQ_BLOCKS"] = nb
        out = []
        for method in ("single", "complete", "average"):
            ctx.linkage(method)
            t0 = time.perf_counter(); Z = ctx.linkage(method); dt = time.perf_counter() - t0
            if method not in ref: ref[method] = Z
            out.append("%s %.3f s%s" % (method, dt, "" if np.array_equal(Z, ref[method]) else " MISMATCH"))
        print("blocks %4s: %s  (scans %d)" % (nb, "  ".join(out), ctx.info()["row_scans"]), flush=True)
