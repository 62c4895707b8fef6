"""On-box: distance kernel time only (n x 3000 synthetic), for A/B runs of library variants (RTHC_LIB_PATH)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from reportree_b200 import hc_backend as hc, synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 40000
X = synth.allele_matrix(n, 3000, 5)
with hc.HCContext() as ctx:
    ctx.set_profiles(X)
    for rep in range(3):
        ctx.distance()
    ms = ctx.timings()["distance_ms"]; info = ctx.info()
    P = n * (n - 1) // 2
    print("%-40s %.2f ms  %.2f Tcmp/s  lop3 lane-ops/s %.3e" % (os.environ.get("RTHC_LIB_PATH", "default")[-40:], ms, P * 3000 / ms / 1e9,
                                                  P * info["plane_rows"] / (ms * 1e-3)))
