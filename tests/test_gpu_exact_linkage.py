"""GPU parity: scipy-identical Z (Prim-order single linkage; NN-chain complete/average/weighted)
and scipy-identical fcluster ids, on tie-heavy integer matrices and on float matrices."""
import numpy as np
import pytest
from scipy.cluster.hierarchy import fcluster, linkage

pytestmark = pytest.mark.gpu


def _rand_codes(rng, n, L, max_allele, missing):
    X = rng.integers(1, max_allele + 1, size=(n, L)).astype(np.int32)
    X[rng.random((n, L)) < missing] = 0
    return X


@pytest.mark.parametrize("n,L,max_allele", [(2, 3, 2), (3, 3, 2), (5, 4, 2), (33, 10, 2), (200, 16, 3), (1000, 30, 3), (2500, 12, 2)])
def test_prim_exact_z_equals_scipy(hc, n, L, max_allele):
    rng = np.random.default_rng(7 * n + L)
    X = _rand_codes(rng, n, L, max_allele, 0.05)
    with hc.HCContext() as ctx:
        ctx.set_single_algo(hc.SL_PRIM)
        ctx.set_profiles(X)
        ctx.distance()
        y = ctx.condensed().astype(np.float64)
        Z = ctx.linkage("single")
        Zs = linkage(y, "single")
        assert np.array_equal(Z, Zs)
        thr = np.arange(0, int(Zs[-1, 2]) + 2, dtype=np.float64)
        lab = ctx.fcluster(thr)
    for k, t in enumerate(thr):
        assert np.array_equal(lab[k], fcluster(Zs, t, criterion="distance"))


@pytest.mark.parametrize("method", ["complete", "average", "weighted", "ward"])
@pytest.mark.parametrize("n,L,max_allele", [(2, 3, 2), (3, 4, 2), (12, 6, 2), (150, 14, 3), (900, 40, 4)])
def test_nn_chain_z_equals_scipy(hc, method, n, L, max_allele):
    rng = np.random.default_rng(11 * n + L)
    X = _rand_codes(rng, n, L, max_allele, 0.05)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        y = ctx.condensed().astype(np.float64)
        Z = ctx.linkage(method)
        Zs = linkage(y, method)
        assert np.array_equal(Z, Zs), np.flatnonzero((Z != Zs).any(axis=1))[:5]
        thr = np.unique(np.concatenate([np.arange(0, 8.0), Zs[::max(1, n // 20), 2], [Zs[-1, 2] + 1]]))
        lab = ctx.fcluster(thr)
        info = ctx.info()
        assert info["row_scans"] + info["chain_cache_hits"] >= n - 1      # every chain step is a scan or a cache hit
    for k, t in enumerate(thr):
        assert np.array_equal(lab[k], fcluster(Zs, t, criterion="distance")), t


@pytest.mark.parametrize("method", ["single", "complete", "average", "weighted", "ward"])
def test_external_matrix_integer_and_float(hc, method):
    rng = np.random.default_rng(101)
    n = 300
    for y in (rng.integers(0, 9, size=n * (n - 1) // 2).astype(np.float64),
              np.round(rng.random(n * (n - 1) // 2) * 10, 1)):
        Zs = linkage(y, method)
        for algo in ((hc.SL_PRIM, hc.SL_BORUVKA) if method == "single" else (hc.SL_PRIM,)):
            with hc.HCContext() as ctx:
                ctx.set_single_algo(algo)
                ctx.set_condensed(y)
                Z = ctx.linkage(method)
                lab = ctx.fcluster([0.0, 0.5, 1.0, 2.0, 5.0])
            if method != "single" or algo == hc.SL_PRIM or y.max() > 9:
                assert np.array_equal(Z, Zs)
            else:
                assert np.array_equal(Z[:, 2], Zs[:, 2])
            for k, t in enumerate([0.0, 0.5, 1.0, 2.0, 5.0]):
                assert np.array_equal(lab[k], fcluster(Z, t, criterion="distance"))


def test_external_matrix_rejects_bad_values(hc):
    y = np.array([1.0, -1.0, 2.0])
    with hc.HCContext() as ctx:
        with pytest.raises(hc.HCBackendError):
            ctx.set_condensed(y)
        with pytest.raises(hc.HCBackendError):
            ctx.set_condensed(np.array([1.0, np.nan, 2.0]))
