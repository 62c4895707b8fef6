"""GPU parity: single linkage (Boruvka MST) + multi-threshold cut vs scipy / the oracle."""
import numpy as np
import pytest
from scipy.cluster.hierarchy import fcluster, linkage

pytestmark = pytest.mark.gpu


def _rand_codes(rng, n, L, max_allele, missing):
    X = rng.integers(1, max_allele + 1, size=(n, L)).astype(np.int32)
    X[rng.random((n, L)) < missing] = 0
    return X


def _same_membership(a, b, oracle):
    return np.array_equal(oracle.canonical(a), oracle.canonical(b))


@pytest.mark.parametrize("n,L,max_allele", [(2, 3, 2), (3, 4, 2), (17, 20, 3), (200, 30, 2), (777, 60, 4), (1500, 12, 2)])
def test_single_linkage_membership_every_threshold(hc, oracle, n, L, max_allele):
    rng = np.random.default_rng(n + L)
    X = _rand_codes(rng, n, L, max_allele, 0.05)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        y = ctx.condensed().astype(np.float64)
        Z = ctx.linkage("single")
        Zs = linkage(y, "single")
        # heights (sorted MST weights) are tie-independent
        assert np.array_equal(Z[:, 2], Zs[:, 2])
        assert np.array_equal(Z[:, 3].max(), n)
        thr = np.arange(0, int(Zs[-1, 2]) + 2, dtype=np.float64)
        labels = ctx.fcluster(thr)
    for k, t in enumerate(thr):
        ref = fcluster(Zs, t, criterion="distance")
        assert _same_membership(labels[k], ref, oracle), (t,)
        # and the ids are exactly scipy's numbering of OUR dendrogram
        assert np.array_equal(labels[k], fcluster(Z, t, criterion="distance")), (t,)
        assert np.array_equal(labels[k], oracle.fcluster(Z, t)), (t,)


def test_cut_threshold_order_and_float_thresholds(hc, oracle):
    rng = np.random.default_rng(99)
    X = _rand_codes(rng, 300, 40, 3, 0.02)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        Z = ctx.linkage("single")
        thr = np.array([7.0, 0.0, 3.5, 100.0, 2.0, 2.0, 1e-9])
        labels = ctx.fcluster(thr)
    for k, t in enumerate(thr):
        assert np.array_equal(labels[k], fcluster(Z, t, criterion="distance"))


def test_cut_with_more_thresholds_than_a_grid_dimension(hc, oracle):
    """70,000 thresholds on a small tree (`--HC-threshold ward` cuts at every integer up to the root
    height, which grows like sqrt(n/2)*d; -d_mx matrices can exceed 65,535 too): the batched cut must
    chunk below the gridDim.y limit instead of failing the launch."""
    rng = np.random.default_rng(4)
    X = _rand_codes(rng, 64, 50, 3, 0.02)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        Z = ctx.linkage("complete")
        thr = np.linspace(0.0, float(Z[-1, 2]) + 1.0, 70000)
        labels = ctx.fcluster(thr)
    for k in list(range(0, 70000, 997)) + [59999, 60000, 60001, 65535, 65536, 69999]:
        assert np.array_equal(labels[k], fcluster(Z, thr[k], criterion="distance")), k


def test_forest_union_equals_single_gpu(hc, oracle):
    # multi-GPU algebra on one GPU: per-shard forests, union, final MST == full MST
    rng = np.random.default_rng(5)
    n = 900
    X = _rand_codes(rng, n, 25, 3, 0.05)
    with hc.HCContext() as full:
        full.set_profiles(X)
        full.distance()
        Zfull = full.linkage("single")
    for world in (2, 4):
        keys = []
        for rank in range(world):
            with hc.HCContext(rank=rank, world=world) as ctx:
                ctx.set_profiles(X)
                ctx.distance()
                keys.append(ctx.local_forest())
        allk = np.concatenate(keys)
        with hc.HCContext() as ctx:
            Z = ctx.linkage_from_forest(allk, n)
            assert np.array_equal(Z, Zfull)
            lab = ctx.fcluster([0.0, 1.0, 2.0, 3.0])
        for k, t in enumerate([0.0, 1.0, 2.0, 3.0]):
            assert np.array_equal(lab[k], fcluster(Zfull, t, criterion="distance"))


def test_forest_exchange_on_device_equals_single_gpu(hc):
    # the NCCL path of bench.py without NCCL: forests written into torch buffers (+inf padded, unsorted),
    # concatenated on the device, final MST from the device pointer
    import torch
    rng = np.random.default_rng(6)
    n = 1100
    X = _rand_codes(rng, n, 30, 3, 0.05)
    with hc.HCContext() as full:
        full.set_profiles(X)
        full.distance()
        Zfull = full.linkage("single")
    for world in (2, 3):
        bufs, counts = [], []
        for rank in range(world):
            with hc.HCContext(rank=rank, world=world) as ctx:
                ctx.set_profiles(X)
                ctx.distance()
                t = torch.zeros(n - 1, dtype=torch.int64, device="cuda")
                counts.append(ctx.local_forest_device(t.data_ptr(), n - 1))
                host = ctx.local_forest()
                got = t.cpu().numpy().view(np.uint64)
                assert np.all(got[counts[-1]:] == np.uint64(0xFFFFFFFFFFFFFFFF))
                assert np.array_equal(np.sort(got[: counts[-1]]), host)          # same forest as the host entry point
                assert np.all(host[1:] > host[:-1])                              # which returns it ascending
                bufs.append(t)
        allk = torch.cat(bufs)
        torch.cuda.synchronize()
        assert sum(counts) >= n - 1
        with hc.HCContext() as ctx:
            Z = ctx.linkage_from_forest_device(allk.data_ptr(), allk.numel(), n)
            assert np.array_equal(Z, Zfull)
            lab = ctx.fcluster([0.0, 2.0])
        for k, t in enumerate([0.0, 2.0]):
            assert np.array_equal(lab[k], fcluster(Zfull, t, criterion="distance"))


def test_device_sort_of_forest_keys_all_sizes(hc):
    # the key sort (shared-memory bitonic chunks + global steps) through linkage_from_forest at sizes around
    # the chunk boundaries: a path graph with shuffled, duplicated and padded keys
    from reportree_b200 import sharded
    rng = np.random.default_rng(8)
    for n in (2, 3, 1025, 2049, 2050, 4097, 9000):
        i = np.arange(n - 1)
        w = rng.integers(0, 50, n - 1)
        keys = sharded.pack_keys(w, i, i + 1)
        extra = sharded.pack_keys(rng.integers(50, 60, n // 2), np.zeros(n // 2, dtype=np.int64), 1 + rng.integers(0, n - 1, n // 2))
        allk = np.concatenate([keys, extra, keys[: n // 3]])
        rng.shuffle(allk)
        with hc.HCContext() as ctx:
            Z = ctx.linkage_from_forest(allk, n)
        assert np.array_equal(Z[:, 2], np.sort(w).astype(np.float64))
        assert Z[-1, 3] == n


def test_event_marks_and_stage_timings(hc):
    """rthc_mark / rthc_mark_elapsed_ms (bench.py's device-side step clock) and rthc_timings."""
    from reportree_b200 import synth
    X = synth.allele_matrix(1500, 200, 3)
    with hc.HCContext() as ctx:
        ctx.mark(0)
        ctx.set_profiles(X)
        ctx.distance()
        ctx.linkage("single")
        ctx.fcluster([1.0, 2.0])
        ctx.mark(1)
        total = ctx.mark_elapsed_ms()
        tm = ctx.timings()
        assert total > 0
        assert 0 < tm["distance_ms"] < total and 0 < tm["linkage_ms"] < total and tm["launches"] > 5
        assert tm["recode_ms"] + tm["distance_ms"] + tm["linkage_ms"] + tm["cut_ms"] <= total * 1.05


@pytest.mark.parametrize("kind", ["ties", "clusters"])
def test_contraction_after_the_first_round_gives_the_same_forest(hc, oracle, monkeypatch, kind):
    """After Boruvka's first round the matrix is reduced to the lightest edge per pair of components and the remaining
    rounds scan that list (mst.cu, pair_min_contract; production: shards of 1e9 pairs and more, here forced for small n).  The strict
    (w, i, j) order makes the forest unique, so Z and every per-shard forest must be IDENTICAL with and without."""
    rng = np.random.default_rng(12)
    if kind == "ties":
        n = 1100
        X = _rand_codes(rng, n, 25, 3, 0.05)                       # few distinct distances: ties everywhere
    else:
        from reportree_b200 import synth
        n = 2600
        X = synth.allele_matrix(n, 120, seed=3, missing=0.02, p_mut=0.01, founders=40)

    def run():
        out = {}
        with hc.HCContext() as ctx:
            ctx.set_profiles(X)
            ctx.distance()
            out["Z"] = ctx.linkage("single")
            out["rounds"] = ctx.info()["boruvka_rounds"]
        for world in (2, 3):
            for rank in range(world):
                with hc.HCContext(rank=rank, world=world) as ctx:
                    ctx.set_profiles(X)
                    ctx.distance()
                    out[(world, rank)] = np.sort(ctx.local_forest())
        return out

    monkeypatch.setenv("RTHC_PAIRMIN_MIN_N", "1000000000")          # never
    plain = run()
    monkeypatch.setenv("RTHC_PAIRMIN_MIN_N", "0")                   # always (when at least 2 components are left)
    contracted = run()
    assert plain["rounds"] >= 2
    for k in plain:
        assert np.array_equal(plain[k], contracted[k]), k
    y = None
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        y = ctx.condensed().astype(np.float64)
    assert np.array_equal(contracted["Z"][:, 2], linkage(y, "single")[:, 2])
