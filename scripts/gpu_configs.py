"""On-box timings of BASELINE.json configs C3 / C5 / C4-Prim through the backend (no TSV I/O),
with scipy verification where scipy fits (C3)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from reportree_b200 import hc_backend as hc, synth

which = sys.argv[1:] or ["C3", "C5", "C4prim"]

def t(f):
    t0 = time.perf_counter(); r = f(); return r, time.perf_counter() - t0

if "C3" in which:
    X = synth.allele_matrix(**synth.config("C3"))
    with hc.HCContext() as ctx:
        _, ts = t(lambda: ctx.set_profiles(X)); _, td = t(ctx.distance)
        Zs, t1 = t(lambda: ctx.linkage("single"))
        ctx.set_single_algo(hc.SL_PRIM); Zp, t2 = t(lambda: ctx.linkage("single"))
        Za, t3 = t(lambda: ctx.linkage("average")); scans_a = ctx.info()["row_scans"]
        print("C3 average: grid barriers %d, look-ups %d" % (ctx.info()["chain_barriers"], ctx.info()["chain_cache_hits"]))
        lab, t4 = t(lambda: ctx.fcluster(np.arange(0, 51.0)))
        Zc, t5 = t(lambda: ctx.linkage("complete")); scans_c = ctx.info()["row_scans"]
        Zw, t6 = t(lambda: ctx.linkage("ward")); scans_w = ctx.info()["row_scans"]
        y = ctx.condensed()
    print(f"C3 ward {t6:.3f}s ({scans_w} scans)")
    print(f"C3 20000x3002: set {ts:.3f}s dist {td:.3f}s boruvka {t1:.3f}s prim {t2:.3f}s average {t3:.3f}s ({scans_a} scans) "
          f"cut51(avg) {t4:.3f}s complete {t5:.3f}s ({scans_c} scans)")
    if "--verify" in sys.argv:
        from scipy.cluster.hierarchy import linkage, fcluster
        yf = y.astype(np.float64)
        Z, ta = t(lambda: linkage(yf, "average")); print("  scipy average", round(ta, 2), "s  Z equal:", np.array_equal(Z, Za))
        for k in (0, 5, 50):
            assert np.array_equal(lab[k], fcluster(Z, float(k), criterion="distance"))
        Z, ta = t(lambda: linkage(yf, "single")); print("  scipy single", round(ta, 2), "s  Z equal (prim):", np.array_equal(Z, Zp),
                                                        " heights equal (boruvka):", np.array_equal(Z[:, 2], Zs[:, 2]))
        Z, ta = t(lambda: linkage(yf, "complete")); print("  scipy complete", round(ta, 2), "s  Z equal:", np.array_equal(Z, Zc))
        Z, ta = t(lambda: linkage(yf, "ward")); print("  scipy ward", round(ta, 2), "s  Z equal:", np.array_equal(Z, Zw))

if "C5" in which:
    cfg = synth.config("C5")
    X = synth.allele_matrix(**cfg)
    rng = np.random.default_rng(1)
    cols = np.sort(rng.choice(cfg["L"], 3000, replace=False))        # --loci subset
    Xs = X[:, cols]
    keep = (Xs != 0).sum(1) / Xs.shape[1] >= 0.95                        # --loci-called 0.95
    Xs = np.ascontiguousarray(Xs[keep])
    with hc.HCContext() as ctx:
        _, ts = t(lambda: ctx.set_profiles(Xs)); _, td = t(ctx.distance)
        Zc, t5 = t(lambda: ctx.linkage("complete")); scans = ctx.info()["row_scans"]; inf = ctx.info()
        Zc2, t5b = t(lambda: ctx.linkage("complete"))
        print("C5 complete again (warm pool) %.3fs; grid barriers %d, look-ups %d" % (t5b, inf["chain_barriers"], inf["chain_cache_hits"]))
        lab, t4 = t(lambda: ctx.fcluster(np.arange(0, 51.0)))
    print(f"C5 {Xs.shape[0]}x{Xs.shape[1]} (of 50000x7600): set {ts:.3f}s dist {td:.3f}s complete {t5:.3f}s ({scans} scans) cut51 {t4:.3f}s"
          f"  clusters@t=20: {lab[20].max()}")

if "C4prim" in which:
    X = synth.allele_matrix(**synth.config("C4"))
    with hc.HCContext() as ctx:
        ctx.set_profiles(X); ctx.distance()
        Zb, t1 = t(lambda: ctx.linkage("single"))
        ctx.set_single_algo(hc.SL_PRIM); Zp, t2 = t(lambda: ctx.linkage("single"))
        thr = np.arange(0, Zp[-1, 2] + 1)
        lab, t3 = t(lambda: ctx.fcluster(thr[:200])); 
    print(f"C4 100000x3000: boruvka {t1:.3f}s prim-exact {t2:.3f}s  heights equal: {np.array_equal(Zb[:,2], Zp[:,2])}  cut200 {t3:.3f}s  maxdist {Zp[-1,2]}")
