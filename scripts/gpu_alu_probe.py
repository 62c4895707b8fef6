"""On-box: ALU peak probes + distance kernel at both tile shapes (RTHC_DIST_NJ is read per set_profiles)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from reportree_b200 import hc_backend as hc, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 40000
X = synth.allele_matrix(n, 3000, 5)
with hc.HCContext() as ctx:
    pk = ctx.alu_peak()
    print("LOP3 stream peak %.3e lane-ops/s" % pk)
    for w in (8, 16, 24, 32):
        print("  block pattern @%d warps/SM: %.3e (%.1f%%)" % (w, ctx.alu_peak_block(w), 100 * ctx.alu_peak_block(w) / pk))
    for nj in (2, 1):
        os.environ["RTHC_DIST_NJ"] = str(nj)
        ctx.set_profiles(X)
        for rep in range(3):
            ctx.distance()
        ms = ctx.timings()["distance_ms"]; info = ctx.info()
        P = n * (n - 1) // 2
        print("NJ=%d: %.2f ms  %.2f Tcmp/s  alu_frac %.3f" % (nj, ms, P * 3000 / ms / 1e9, P * info["plane_rows"] / (ms * 1e-3) / pk))
