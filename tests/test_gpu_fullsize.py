"""Parity at BASELINE.json's full sizes (-m gpu).  Where the CPU reference still fits (C3) the
result is compared with scipy / the oracle directly; at C4 / C5 sizes, through size-independent
properties and spot checks: an independent scalar kernel on the uint16 codes, the oracle on a
600-sample sub-matrix, symmetry, single linkage = connected components of the <= t graph,
complete linkage = every flat cluster has diameter <= t, Boruvka == Prim-exact heights."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _canon(oracle, x):
    return oracle.canonical(np.asarray(x))


def test_c3_full_parity_with_scipy(hc, oracle):
    """configs[2]: 20,000 x 3,002, single + average -- Z byte-identical to scipy, labels identical."""
    from scipy.cluster.hierarchy import fcluster, linkage
    from reportree_b200 import synth
    X = synth.allele_matrix(**synth.config("C3"))
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        y = ctx.condensed()
        # distance: oracle on a 500-sample sub-matrix
        rng = np.random.default_rng(1)
        sub = np.sort(rng.choice(X.shape[0], 500, replace=False))
        D = oracle.dist_square(X[sub])
        n = X.shape[0]
        iu = np.triu_indices(sub.size, 1)
        a, b = sub[iu[0]], sub[iu[1]]
        idx = n * a - a * (a + 1) // 2 + (b - a - 1)
        assert np.array_equal(y[idx].astype(np.int32), D[iu])
        yf = y.astype(np.float64)
        ctx.set_single_algo(hc.SL_PRIM)
        Zs = ctx.linkage("single")
        assert np.array_equal(Zs, linkage(yf, "single"))
        ctx.set_single_algo(hc.SL_BORUVKA)
        Zb = ctx.linkage("single")
        assert np.array_equal(Zb[:, 2], Zs[:, 2])
        lab_b = ctx.fcluster([0.0, 3.0, 10.0, 50.0])
        for k, t in enumerate([0.0, 3.0, 10.0, 50.0]):
            assert np.array_equal(_canon(oracle, lab_b[k]), _canon(oracle, fcluster(Zs, t, criterion="distance")))
        Za = ctx.linkage("average")
        Zref = linkage(yf, "average")
        assert np.array_equal(Za, Zref)
        lab = ctx.fcluster([0.0, 5.0, 17.5, 50.0])
        for k, t in enumerate([0.0, 5.0, 17.5, 50.0]):
            assert np.array_equal(lab[k], fcluster(Zref, t, criterion="distance"))


def test_c4_full_size_properties(hc, oracle):
    """configs[3] on one GPU: 100,000 x 3,000, single linkage."""
    from reportree_b200 import synth
    X = synth.allele_matrix(**synth.config("C4"))
    n, L = X.shape
    rng = np.random.default_rng(2)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        assert ctx.owned_pairs() == n * (n - 1) // 2
        rows = np.sort(rng.choice(n, 48, replace=False))
        sq = np.stack([ctx.square_rows(int(r), 1)[0] for r in rows])           # 48 x n
        # (1) independent scalar kernel on the uint16 codes, every pair of the 48 rows
        pi = np.repeat(rows, n).astype(np.int32)
        pj = np.tile(np.arange(n, dtype=np.int32), rows.size)
        direct = ctx.direct_pairs(pi, pj).reshape(rows.size, n)
        assert np.array_equal(direct, sq.astype(np.int32))
        # (2) symmetry + zero diagonal through the condensed indexing
        assert np.array_equal(sq[:, rows], sq[:, rows].T)
        assert (sq[np.arange(rows.size), rows] == 0).all()
        # (3) the CPU oracle on a 600-sample sub-matrix (first 300 + 300 random)
        sub = np.unique(np.concatenate([np.arange(300), rng.choice(n, 300, replace=False)]))
        D = oracle.dist_square(X[sub])
        got = np.stack([ctx.square_rows(int(r), 1)[0][sub] for r in sub[:64]])
        assert np.array_equal(got.astype(np.int32), D[:64])
        # (4) Boruvka and scipy-order Prim agree on every merge height; cuts have the same membership
        Zb = ctx.linkage("single")
        thr = [0.0, 1.0, 4.0, 15.0, 60.0, 400.0]
        lab_b = ctx.fcluster(thr)
        ctx.set_single_algo(hc.SL_PRIM)
        Zp = ctx.linkage("single")
        lab_p = ctx.fcluster(thr)
        assert np.array_equal(Zb[:, 2], Zp[:, 2])
        assert Zp[-1, 3] == n and (np.diff(Zp[:, 2]) >= 0).all()
        for k in range(len(thr)):
            assert np.array_equal(_canon(oracle, lab_b[k]), _canon(oracle, lab_p[k]))
        # (5) single linkage at t = components of the <= t graph: neighbours within t share the label,
        #     and the number of merges with height <= t fixes the number of clusters
        for k, t in enumerate(thr):
            for q, r in enumerate(rows):
                near = np.flatnonzero(sq[q] <= t)
                assert (lab_p[k][near] == lab_p[k][r]).all()
            assert lab_p[k].max() == n - int((Zp[:, 2] <= t).sum())


def test_c5_complete_linkage_properties(hc, oracle):
    """configs[4]: 50,000 x 7,600 wgMLST, --loci subset (3,000) + --loci-called 0.95, complete linkage."""
    from reportree_b200 import synth
    cfg = synth.config("C5")
    X = synth.allele_matrix(**cfg)
    rng = np.random.default_rng(1)
    cols = np.sort(rng.choice(cfg["L"], 3000, replace=False))
    Xs = X[:, cols]
    Xs = np.ascontiguousarray(Xs[(Xs != 0).sum(1) / Xs.shape[1] >= 0.95])
    n = Xs.shape[0]
    assert 40000 < n < 50000
    with hc.HCContext() as ctx:
        ctx.set_profiles(Xs)
        ctx.distance()
        Z = ctx.linkage("complete")
        assert Z.shape == (n - 1, 4) and Z[-1, 3] == n and (np.diff(Z[:, 2]) >= 0).all()
        thr = [0.0, 5.0, 20.0, 100.0]
        lab = ctx.fcluster(thr)
        for k, t in enumerate(thr):
            assert np.array_equal(lab[k], oracle.fcluster(Z, t))            # cut == scipy's rule on this Z
        # diameter property on sampled members: everything sharing a flat cluster is within t
        rows = np.sort(rng.choice(n, 40, replace=False))
        sq = np.stack([ctx.square_rows(int(r), 1)[0] for r in rows])
        for k, t in enumerate(thr):
            for q, r in enumerate(rows):
                same = lab[k] == lab[k][r]
                assert sq[q][same].max() <= t
        # and the oracle's distance rule on a sub-matrix
        sub = np.sort(rng.choice(n, 300, replace=False))
        D = oracle.dist_square(Xs[sub])
        got = np.stack([ctx.square_rows(int(r), 1)[0][sub] for r in sub[:32]])
        assert np.array_equal(got.astype(np.int32), D[:32])


# ------------------------------------------------------------------------------------------------
# BASELINE.json configs[2..4] against the CPU oracle's own answer (tests/golden/large/*.json, written
# by tests/golden/make_golden_large.py: oracle/cgmlst_dists.c + oracle/hclust.c on this container's
# host cores).  Distances: CRC32 of every 1,024-row block of the condensed matrix.  Z: byte-identical
# (CRC32 of the float64 array) for the scipy-order paths.  Labels: at every threshold 0..50 the
# fcluster ids (CRC32), the cluster count and the membership-only relabelling (CRC32).
# ------------------------------------------------------------------------------------------------
import json
import os

LARGE = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "large")


def _golden(name):
    with open(os.path.join(LARGE, name + ".json")) as f:
        return json.load(f)


def _check_distances(ctx, g):
    from reportree_b200 import checksums as C
    assert g["rows_per_block"] == C.ROWS_PER_BLOCK
    bounds = C.block_bounds(ctx.n)
    assert len(bounds) == len(g["condensed_block_crc32"])
    bad = [b for b, (r0, r1, p0, p1) in enumerate(bounds)
           if C.crc(ctx.condensed_range(p0, p1 - p0)) != g["condensed_block_crc32"][b]]
    assert not bad, "distance blocks (1,024 rows each) that differ from the oracle: %s" % bad[:10]


def _check_method(ctx, g, method, exact_z):
    from reportree_b200 import checksums as C
    gm = g["methods"][method]
    Z = ctx.linkage(method)
    zf = C.z_fingerprint(Z)
    assert zf["heights_crc32"] == gm["heights_crc32"], method
    if exact_z:
        assert zf["z_crc32"] == gm["z_crc32"], method
    lab = ctx.fcluster(np.asarray(g["thresholds"], dtype=np.float64))
    for k, t in enumerate(g["thresholds"]):
        f = C.labels_fingerprint(lab[k])
        assert f["clusters"] == gm["labels"][k]["clusters"], (method, t)
        assert f["membership_crc32"] == gm["labels"][k]["membership_crc32"], (method, t)
        if exact_z:
            assert f["ids_crc32"] == gm["labels"][k]["ids_crc32"], (method, t)


@pytest.mark.parametrize("name", ["C3", "C5", "C4"])
def test_full_configs_equal_the_cpu_oracle(hc, name):
    """configs[2] (single / average / complete), configs[4] (complete, + average), configs[3] (single):
    every distance block, Z and the partitions at all thresholds 0..50 equal the oracle's."""
    from reportree_b200 import checksums as C, synth
    g = _golden(name)
    X = synth.workload(name)
    assert X.shape == (g["n"], g["L"]) and C.crc(X) == g["input_crc32"]
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        _check_distances(ctx, g)
        for method in g["methods"]:
            if method == "single":
                ctx.set_single_algo(hc.SL_PRIM)                 # scipy order: Z and ids byte-identical
                _check_method(ctx, g, "single", exact_z=True)
                ctx.set_single_algo(hc.SL_BORUVKA)              # the benchmarked path: same heights + membership
                _check_method(ctx, g, "single", exact_z=False)
            else:
                _check_method(ctx, g, method, exact_z=True)
