"""Quick on-box probe: stage timings of the hot path at a few sizes (not the bench)."""
import sys, time, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from reportree_b200 import hc_backend as hc, synth

sizes = [(5000, 3002), (20000, 3002)] + ([(100000, 3000)] if "--big" in sys.argv else [])
for n, L in sizes:
    t0 = time.time(); X = synth.allele_matrix(n, L, 42); tg = time.time() - t0
    with hc.HCContext() as ctx:
        t0 = time.time(); ctx.set_profiles(X); t_set = time.time() - t0
        for rep in range(2):
            t0 = time.time(); ctx.distance(); t_d = time.time() - t0
        t0 = time.time(); Z = ctx.linkage("single"); t_l = time.time() - t0
        thr = np.arange(0, 51, dtype=np.float64)
        t0 = time.time(); lab = ctx.fcluster(thr); t_c = time.time() - t0
        info, tm = ctx.info(), ctx.timings()
    P = n * (n - 1) // 2
    print(f"n={n} L={L} gen={tg:.2f}s set={t_set:.3f}s dist={t_d:.3f}s link={t_l:.3f}s cut51={t_c:.3f}s")
    print("  info", info)
    print("  timings", tm)
    print(f"  distance kernel: {P*L/tm['distance_ms']*1e3/1e12:.2f} Tcmp/s ; lop3-equivalent planes/32loci={info['plane_rows']/max(info['groups'],1):.2f}")
    print("  clusters at t=5:", lab[5].max(), " maxdist:", Z[-1, 2])
