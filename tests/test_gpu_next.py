"""GPU parity of the SURVEY 8(f) N3 / N4 components, all through the C ABI:
zoom-in on the resident matrix (rthc_closest, rthc_subset_from) and the adjacent-column contingency
tables behind the cluster-stability drop-in (rthc_contingency_adjacent)."""
import gzip
import json
import os
import shutil

import numpy as np
import pytest
from scipy.cluster.hierarchy import fcluster, linkage
from scipy.spatial.distance import squareform

from oracle import oracle_lib as O
from oracle import partition_metrics as pm
from oracle import zoom_oracle
from reportree_b200 import comparing_partitions_v2 as cp
from reportree_b200 import hc_backend, zoom

pytestmark = pytest.mark.gpu

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
NEXT = os.path.join(GOLD, "next")


def random_profiles(rng, n, L, alleles=6, missing=0.05):
    codes = rng.integers(1, alleles + 1, (n, L)).astype(np.int32)
    codes[rng.random((n, L)) < missing] = 0
    return codes


# ---- N3 ---------------------------------------------------------------------------------------------
def test_closest_equals_reference_function_on_c2():
    z = np.load(os.path.join(GOLD, "C2", "C2_dist_hamming.npz"))
    ids, dist = [str(x) for x in z["ids"]], z["dist"]
    with open(os.path.join(NEXT, "closest_C2.json")) as fh:
        gold = json.load(fh)
    with hc_backend.HCContext() as ctx:
        ctx.set_condensed(squareform(dist.astype(np.float64), checks=False))
        rm = zoom.ResidentMatrix(ctx, ids)
        for q in gold:
            if q["code"] == "IndexError":
                with pytest.raises(IndexError):
                    zoom.get_closest_samples(q["sample"], q["n"], rm, set(ids))
                continue
            names, code = zoom.get_closest_samples(q["sample"], q["n"], rm, set(ids))
            assert code == q["code"]
            assert sorted(names) == q["names"]
            if code == "run":
                pos = ids.index(q["sample"])
                assert sorted(int(dist[pos, ids.index(s)]) for s in names) == q["dists"]
        assert zoom.get_closest_samples(ids[0], 3, rm, set()) == ("", "not_in_hamming")


def test_closest_on_computed_distances_with_heavy_ties():
    rng = np.random.default_rng(11)
    codes = random_profiles(rng, 3000, 40, alleles=3)
    D = O.dist_square(codes)
    names = ["s%d" % k for k in range(codes.shape[0])]
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes)
        ctx.distance()
        for sample, k in [(0, 0), (1, 1), (77, 10), (2999, 500), (1500, 2999), (4, 2998)]:
            idx, d, vmax = ctx.closest(sample, k)
            want, wd, _ = zoom_oracle.closest_samples(D, names, names[sample], k)
            assert {names[i] for i in idx} == want
            assert d.tolist() == wd
            assert vmax == wd[-1]
            assert np.all(np.diff(d.astype(np.int64)) >= 0)
        with pytest.raises(hc_backend.HCBackendError):
            ctx.closest(0, 3000)
        with pytest.raises(hc_backend.HCBackendError):
            ctx.closest(3000, 1)


@pytest.mark.parametrize("method", ["single", "complete", "average", "weighted", "ward"])
def test_subset_clustering_equals_scipy_on_the_sub_matrix(method):
    rng = np.random.default_rng(5)
    codes = random_profiles(rng, 1500, 120)
    D = O.dist_square(codes)
    picks = rng.permutation(1500)[:400]
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes)
        ctx.distance()
        with ctx.subset(picks) as child:
            sub = zoom_oracle.sub_square(D, picks)
            y = squareform(sub.astype(np.float64), checks=False)
            assert np.array_equal(child.condensed().astype(np.float64), y)
            assert np.array_equal(child.square_rows(3, 5), sub[3:8].astype(np.uint16))
            child.set_single_algo(hc_backend.SL_PRIM)
            Z = child.linkage(method)
            Zs = linkage(y, method)
            assert Z.tobytes() == Zs.tobytes()
            ts = [0, 1, 5, 20, 60.5, 1e9]
            labels = child.fcluster(ts)
            for k, t in enumerate(ts):
                assert np.array_equal(labels[k], fcluster(Zs, t, criterion="distance"))
        # the parent is untouched and still serves queries
        assert np.array_equal(ctx.square_rows(10, 2), D[10:12].astype(np.uint16))


def test_subset_argument_errors_and_resident_wrapper():
    rng = np.random.default_rng(6)
    codes = random_profiles(rng, 300, 50)
    names = ["n%03d" % k for k in range(300)]
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes)
        ctx.distance()
        for bad in ([1, 1, 2], [0, 300], [-1, 2], [5]):
            with pytest.raises(hc_backend.HCBackendError):
                ctx.subset(bad)
        rm = zoom.ResidentMatrix(ctx, names)
        near, code = zoom.get_closest_samples("n007", 20, rm, set(names))
        assert code == "run" and "n007" in near and len(near) >= 21
        Z, labels = rm.cluster_subset(near, "single", [3, 10])
        D = O.dist_square(codes)
        sub = zoom_oracle.sub_square(D, [rm.position(s) for s in near])
        Zs = linkage(squareform(sub.astype(np.float64), checks=False), "single")
        assert np.array_equal(np.sort(Z[:, 2]), np.sort(Zs[:, 2]))
        for k, t in enumerate([3, 10]):
            assert same_partition(labels[k], fcluster(Zs, t, criterion="distance"))


def test_subset_refills_an_existing_child():
    rng = np.random.default_rng(8)
    codes = random_profiles(rng, 900, 80)
    D = O.dist_square(codes)
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes)
        ctx.distance()
        child = None
        for m in (50, 700, 2, 300):
            picks = rng.permutation(900)[:m]
            child = ctx.subset(picks, into=child)
            assert child.n == m
            y = squareform(zoom_oracle.sub_square(D, picks).astype(np.float64), checks=False)
            for method in ("complete", "single"):
                child.set_single_algo(hc_backend.SL_PRIM)
                assert child.linkage(method).tobytes() == linkage(y, method).tobytes()
            assert np.array_equal(child.fcluster([4.0])[0], fcluster(linkage(y, "single"), 4.0, criterion="distance"))
        child.close()


def same_partition(a, b):
    pairs = {}
    back = {}
    for x, y in zip(a.tolist(), b.tolist()):
        if pairs.setdefault(x, y) != y or back.setdefault(y, x) != x:
            return False
    return True


# ---- N4 ---------------------------------------------------------------------------------------------
def cells_by_numpy(codes, p):
    """Vectorised statement of the dict-of-dicts order: rows by first appearance, cells by first appearance."""
    r, c = codes[p + 1].astype(np.int64), codes[p].astype(np.int64)
    n = r.size
    _, cell_first, counts = np.unique(r * n + c, return_index=True, return_counts=True)
    row_first = np.full(n, n, dtype=np.int64)
    np.minimum.at(row_first, r, np.arange(n))
    order = np.lexsort((cell_first, row_first[r[cell_first]]))
    f = cell_first[order]
    return r[f].astype(np.int32), c[f].astype(np.int32), counts[order].astype(np.int32)


def dense(col):
    return np.unique(col, return_inverse=True)[1].astype(np.int32)


def label_matrix(rng, n, T, kind):
    if kind == "nested":                 # what a dendrogram cut gives: each column merges clusters of the previous
        cols = [rng.permutation(n).astype(np.int32)]
        for _ in range(T - 1):
            k = int(cols[-1].max()) + 1
            merge = rng.integers(0, max(1, (2 * k) // 3), k).astype(np.int32)
            cols.append(dense(merge[cols[-1]]))
        return np.stack(cols)
    if kind == "random":
        return np.stack([rng.integers(0, max(1, n // int(rng.integers(1, 50))), n).astype(np.int32) for _ in range(T)])
    if kind == "extremes":               # all singletons / one cluster / arbitrary ids below n
        cols = [np.arange(n, dtype=np.int32)[::-1].copy(), np.zeros(n, dtype=np.int32), np.full(n, n - 1, dtype=np.int32),
                rng.integers(0, 3, n).astype(np.int32), np.arange(n, dtype=np.int32)]
        return np.stack(cols[:max(2, min(T, 5))])
    raise ValueError(kind)


@pytest.mark.parametrize("n,T,kind", [(1, 2, "extremes"), (7, 5, "extremes"), (100, 6, "random"), (2048, 4, "nested"), (2049, 3, "random"),
                                      (5000, 12, "nested"), (70000, 5, "random"), (70000, 6, "nested"), (200, 400, "random")])
def test_contingency_cells_in_reference_order(n, T, kind):
    rng = np.random.default_rng(n * 31 + T)
    codes = label_matrix(rng, n, T, kind)
    with hc_backend.HCContext() as ctx:
        got = ctx.contingency_adjacent(codes)
        assert len(got) == codes.shape[0] - 1
        for p, (rows, cols, counts) in enumerate(got):
            wr, wc, wn = cells_by_numpy(codes, p)
            assert np.array_equal(rows, wr) and np.array_equal(cols, wc) and np.array_equal(counts, wn)
            assert int(counts.sum()) == n
        if n <= 5000:                    # the dict-of-dicts restatement itself (slow Python loops)
            for p in range(min(3, codes.shape[0] - 1)):
                table = pm.contingency(codes[p + 1].tolist(), codes[p].tolist())
                flat = [(r, c, v) for r, inner in table.items() for c, v in inner.items()]
                rows, cols, counts = got[p]
                assert flat == list(zip(rows.tolist(), cols.tolist(), counts.tolist()))
                assert ctx.adjusted_wallace_sums(rows, counts)[1] == sum(r * (r - 1.0) for r in (sum(i.values()) for i in table.values()))
        assert ctx.timings()["stability_ms"] > 0


def test_contingency_rejects_codes_out_of_range():
    with hc_backend.HCContext() as ctx:
        for bad in (np.array([[0, 1, 2], [0, 3, 1]]), np.array([[0, -1, 2], [0, 1, 1]])):
            with pytest.raises(hc_backend.HCBackendError):
                ctx.contingency_adjacent(bad.astype(np.int32))
        with pytest.raises(hc_backend.HCBackendError):
            ctx.contingency_adjacent(np.zeros((1, 4), dtype=np.int32))


@pytest.mark.parametrize("case,table,extra", [
    ("C2_keep", "C2", ["-o1", "0", "--keep-redundants"]),
    ("C1b_keep", "C1b", ["-o1", "0", "--keep-redundants"]),
    ("C2_keep_rev", "C2", ["--keep-redundants", "-o1", "1"]),
    ("C2_default", "C2", ["-o1", "0"]),
])
def test_stability_cli_files_equal_reference(tmp_path, case, table, extra):
    with gzip.open(os.path.join(GOLD, table, table + "_partitions.tsv.gz"), "rb") as fi, open(tmp_path / "partitions.tsv", "wb") as fo:
        shutil.copyfileobj(fi, fo)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        cp.main(["-i1", "partitions.tsv", "-t", "out", "-a", "stability", "-n", "5", "-thr", "0.99", "-log", "out.log"] + extra)
    finally:
        os.chdir(cwd)
    metrics = (tmp_path / "out_metrics.tsv").read_text()
    regions = (tmp_path / "out_stableRegions.tsv").read_text()
    with open(os.path.join(NEXT, "stability_%s_metrics.tsv" % case)) as fh:
        ref_metrics = fh.read()
    with open(os.path.join(NEXT, "stability_%s_stableRegions.tsv" % case)) as fh:
        ref_regions = fh.read()
    assert regions == ref_regions
    if case == "C2_default":             # the reference's representatives are random: same tables, sums in another order
        for a, b in zip(metrics.split("\n")[1:-1], ref_metrics.split("\n")[1:-1]):
            a, b = a.split("\t"), b.split("\t")
            assert a[:2] == b[:2] and a[9:] == b[9:]
            np.testing.assert_allclose([float(x) for x in a[2:9]], [float(x) for x in b[2:9]], rtol=1e-12, atol=1e-12)
    else:
        assert metrics == ref_metrics


def test_stability_on_a_device_cut_at_scale():
    """60,000 samples, 40 thresholds straight from rthc_cut: every table against the vectorised statement,
    and the metric rows against the oracle on a few pairs."""
    rng = np.random.default_rng(99)
    n = 60000
    founders = rng.integers(1, 30, (600, 64)).astype(np.int32)
    codes = founders[rng.integers(0, 600, n)]
    mut = rng.random(codes.shape) < 0.03
    codes = np.where(mut, rng.integers(1, 30, codes.shape), codes).astype(np.int32)
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes)
        ctx.distance()
        ctx.linkage("single")
        labels = ctx.fcluster(np.arange(40, dtype=np.float64)) - 1
        got = ctx.contingency_adjacent(labels)
        for p, (rows, cols, counts) in enumerate(got):
            wr, wc, wn = cells_by_numpy(labels, p)
            assert np.array_equal(rows, wr) and np.array_equal(cols, wc) and np.array_equal(counts, wn)
        stats = [cp.ColumnStats(row) for row in labels]
        for p in (0, 7, 38):
            rand, wallace, adj = cp.pair_metrics(ctx, got[p], stats[p], stats[p + 1])
            sub = slice(0, 3000)         # the oracle's loops are quadratic in the clusters: check a prefix separately
            assert 0.0 <= wallace <= 1.0 and -1.0 <= rand <= 1.0 and adj[1] <= adj[0] <= adj[2]
            icol, jcol = labels[p][sub].tolist(), labels[p + 1][sub].tolist()
            small = ctx.contingency_adjacent(np.stack([dense(labels[p][sub]), dense(labels[p + 1][sub])]))[0]
            st_i, st_j = cp.ColumnStats(dense(labels[p][sub])), cp.ColumnStats(dense(labels[p + 1][sub]))
            r2, w2, a2 = cp.pair_metrics(ctx, small, st_i, st_j)
            assert r2 == pm.adjusted_rand(icol, jcol)
            assert w2 == pm.wallace(pm.contingency(icol, jcol))
            assert a2 == pm.adjusted_wallace(pm.contingency(jcol, icol), icol)
            assert st_i.simpson == pm.simpson(icol)[0] and st_j.k == pm.simpson(jcol)[3]
