"""Drives every kernel of libreportree_hc.so once at small sizes (C1 14 x 94, a C2-like 300 x 1,748 and the ragged
shape n = 129, L = 2,001) so that compute-sanitizer (memcheck / racecheck / synccheck / initcheck) sees all of them:
re-coder (shared-memory bitmap path and the wide-id path), v5 distance, direct-pairs kernel, row copies, Boruvka tile
scan + edge-list Boruvka (forest exchange), cluster Prim (mirrored and condensed), grid Prim (float64), NN-chain
(complete uint16, average, weighted, ward, float input), cut, zoom (closest + subset), stability tables.
Results are checked against scipy so that a sanitizer run is also a parity run.   usage: python scripts/gpu_sanitize_driver.py [quick]"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from scipy.cluster.hierarchy import fcluster, linkage

from reportree_b200 import hc_backend as hc

quick = "quick" in sys.argv
rng = np.random.default_rng(12)
shapes = [(14, 94, 4), (129, 2001, 6)] + ([] if quick else [(300, 1748, 60)])
kernels = []

for n, L, A in shapes:
    X = rng.integers(1, A + 1, size=(n, L)).astype(np.int32)
    X[rng.random((n, L)) < 0.03] = 0
    X[:, 3] = 7                                     # monomorphic locus
    X[1, :] = 0                                     # all-missing sample
    if n > 100:
        X[:, 5] = rng.integers(1 << 20, 1 << 30, size=n)        # wide-id path of the re-coder
    ref = None
    for prim_mode in ("1", "2", "0"):               # mirrored cluster Prim, condensed cluster Prim, grid Prim
        os.environ["RTHC_PRIM_CLUSTER"] = prim_mode
        with hc.HCContext() as ctx:
            ctx.set_profiles(X)
            ctx.distance()
            y = ctx.condensed()
            if ref is None:
                ref = y.copy()
                d = np.zeros((n, n), dtype=np.int64)
                for i in range(n):
                    d[i] = ((X[i] != X) & (X[i] != 0) & (X != 0)).sum(1)
                assert np.array_equal(y, d[np.triu_indices(n, 1)].astype(np.uint16)), "distance"
                assert np.array_equal(ctx.square_rows(0, n), d.astype(np.uint16))
                pi, pj = np.triu_indices(n, 1)
                assert np.array_equal(ctx.direct_pairs(pi.astype(np.int32), pj.astype(np.int32)), d[pi, pj])
            yf = ref.astype(np.float64)
            ctx.set_single_algo(hc.SL_PRIM)
            Zp = ctx.linkage("single")
            assert np.array_equal(Zp, linkage(yf, "single")), "prim " + prim_mode
            thr = np.arange(0, 12.0)
            lab = ctx.fcluster(thr)
            assert all(np.array_equal(lab[k], fcluster(Zp, t, criterion="distance")) for k, t in enumerate(thr))
            if prim_mode != "1":
                continue
            ctx.set_single_algo(hc.SL_BORUVKA)
            Zb = ctx.linkage("single")
            assert np.array_equal(Zb[:, 2], Zp[:, 2]), "boruvka heights"
            keys = ctx.local_forest()
            with hc.HCContext() as fin:
                Zf = fin.linkage_from_forest(np.concatenate([keys, keys[::-1]]), n)
            assert np.array_equal(Zf, Zb), "edge-list boruvka"
            for method in ("complete", "average", "weighted", "ward"):
                Z = ctx.linkage(method)
                assert np.array_equal(Z, linkage(yf, method)), method
                lab = ctx.fcluster(thr)
                assert all(np.array_equal(lab[k], fcluster(Z, t, criterion="distance")) for k, t in enumerate(thr))
            idx, dist, vmax = ctx.closest(2, min(5, n - 1))
            assert np.array_equal(np.sort(dist), dist)
            sub = np.sort(rng.choice(n, n // 2, replace=False))
            with ctx.subset(sub) as child:
                Zc = child.linkage("single")
                sq = np.zeros((n, n)); sq[np.triu_indices(n, 1)] = yf; sq += sq.T
                assert np.array_equal(Zc[:, 2], linkage(sq[np.ix_(sub, sub)][np.triu_indices(sub.size, 1)], "single")[:, 2])
            cells = ctx.contingency_adjacent(lab - 1)
            assert len(cells) == lab.shape[0] - 1 and all(c[2].sum() == n for c in cells)
    # external float64 matrix (-d_mx entry): grid Prim on doubles + NN-chain on doubles
    yext = np.round(rng.random(n * (n - 1) // 2) * 9, 1)
    with hc.HCContext() as ctx:
        ctx.set_single_algo(hc.SL_PRIM)
        ctx.set_condensed(yext)
        assert np.array_equal(ctx.linkage("single"), linkage(yext, "single"))
        assert np.array_equal(ctx.linkage("average"), linkage(yext, "average"))
    print("shape %d x %d ok" % (n, L), flush=True)
print("sanitize driver: all results equal scipy")
