"""GPU parity: bit-sliced distance kernel + device re-coder vs the CPU oracle (bit-exact)."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _rand_codes(rng, n, L, max_allele, missing):
    X = rng.integers(1, max_allele + 1, size=(n, L)).astype(np.int32)
    X[rng.random((n, L)) < missing] = 0
    return X


@pytest.mark.parametrize("n,L,max_allele,missing", [
    (2, 1, 2, 0.0),
    (3, 5, 3, 0.3),
    (14, 94, 4, 0.05),          # shape of tests/allele_matrix.tsv
    (127, 33, 9, 0.1),
    (128, 64, 300, 0.02),
    (129, 2001, 40, 0.02),      # > 2,000 loci: the reference would chunk and sum
    (300, 700, 70000, 0.02),    # ids above 2^16 but < 2^20 (few distinct per locus)
    (513, 257, 2, 0.5),
    (1000, 3002, 500, 0.02),
])
def test_distance_matches_oracle(hc, oracle, n, L, max_allele, missing):
    rng = np.random.default_rng(n * 1000 + L)
    X = _rand_codes(rng, n, L, max_allele, missing)
    ref = oracle.dist_condensed(X)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        got = ctx.condensed()
        info = ctx.info()
    assert got.dtype == np.uint16
    assert np.array_equal(got, ref), (info, np.flatnonzero(got != ref)[:10])


def test_chunk_sum_equivalence(hc, oracle):
    # the reference sums per-2,000-column chunks (scripts/partitioning_HC.py:489-516)
    rng = np.random.default_rng(7)
    X = _rand_codes(rng, 60, 4500, 30, 0.05)
    sq = oracle.dist_square_chunked(X, 2000)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        rows = ctx.square_rows(0, 60)
    assert np.array_equal(rows.astype(np.int32), sq)


def test_all_missing_rows_and_monomorphic_loci(hc, oracle):
    rng = np.random.default_rng(11)
    X = _rand_codes(rng, 200, 150, 6, 0.1)
    X[5] = 0                       # sample with every locus missing
    X[:, 10] = 0                   # locus missing everywhere
    X[:, 11] = 3                   # monomorphic locus
    X[:, 12] = np.where(rng.random(200) < 0.5, 0, 9)   # one allele + missing
    ref = oracle.dist_condensed(X)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        got = ctx.condensed()
        assert ctx.info()["loci_kept"] <= 147
    assert np.array_equal(got, ref)


def test_large_allele_ids_slow_path(hc, oracle):
    rng = np.random.default_rng(13)
    X = _rand_codes(rng, 150, 40, 50, 0.05)
    X[:, 3] = rng.integers(1, 2**31 - 1, size=150)          # hashed-profile style ids
    X[:, 7] = np.where(X[:, 7] > 25, X[:, 7] + (1 << 20), X[:, 7])
    ref = oracle.dist_condensed(X)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        got = ctx.condensed()
    assert np.array_equal(got, ref)


def test_sixteen_bit_loci_and_the_uint16_limit(hc, oracle):
    # a locus with > 32,767 distinct called alleles needs all 16 bit planes (the widest uint16 plane packer)
    rng = np.random.default_rng(19)
    n = 33500
    X = np.zeros((n, 3), dtype=np.int32)
    X[:, 0] = rng.permutation(n) + 1                       # every sample its own allele: 16 bits
    X[:, 1] = rng.integers(1, 40000, n)                    # ~ 22,000 distinct: 15 bits
    X[:, 2] = rng.integers(0, 4, n)                        # 3 alleles + missing
    X[rng.random(n) < 0.01, 0] = 0
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        info = ctx.info()
        assert info["plane_rows"] == 17 and info["max_alleles"] >= 32768      # 16 bit planes + the valid plane
        ctx.distance()
        rows = ctx.square_rows(n - 700, 700)
    ref = oracle.dist_square(np.ascontiguousarray(X[n - 2000:]))[-700:, :]
    assert np.array_equal(rows[:, n - 2000:].astype(np.int32), ref)
    head = (X[n - 700:, None, :] != X[None, :50, :]) & (X[n - 700:, None, :] != 0) & (X[None, :50, :] != 0)
    assert np.array_equal(rows[:, :50].astype(np.int32), head.sum(axis=2))


def test_more_than_65535_alleles_in_a_locus_take_the_uint32_codes(hc, oracle):
    """The reference has no limit on the alleles of a locus (scripts/partitioning_HC.py:293-298: any integer
    is an allele).  Past 65,535 distinct called alleles the re-coder switches the matrix to uint32 codes and
    the locus gets 17+ bit planes (generic group body of the distance kernel): n = 70,000 with one locus where
    every sample has its own allele, one with ~ 50,000 alleles incl. ids >= 2^20, and two ordinary loci."""
    rng = np.random.default_rng(23)
    n = 70000
    X = np.zeros((n, 4), dtype=np.int32)
    X[:, 0] = rng.permutation(n) + 1                                   # 70,000 alleles: 17 bits
    X[:, 1] = rng.integers(1, 90000, n) * 37 + (1 << 20)               # ~ 48,000 alleles, all ids >= 2^20 (wide-id path)
    X[:, 2] = rng.integers(0, 4, n)
    X[:, 3] = rng.integers(0, 300, n)
    X[rng.random(n) < 0.01, 0] = 0
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        info = ctx.info()
        assert info["max_alleles"] > 65535 and info["plane_rows"] == 18           # one group: 17 bit planes + valid
        ctx.distance()
        rows = ctx.square_rows(n - 600, 600)
        pi = np.repeat(np.arange(n - 40, n, dtype=np.int32), 500)
        pj = np.tile(np.arange(0, 500, dtype=np.int32), 40)
        direct = ctx.direct_pairs(pi, pj)                              # the scalar check kernel reads the uint32 codes too
        Z = ctx.linkage("single")
    ref = oracle.dist_square(np.ascontiguousarray(X[n - 1500:]))[-600:, :]
    assert np.array_equal(rows[:, n - 1500:].astype(np.int32), ref)
    head = (X[n - 600:, None, :] != X[None, :64, :]) & (X[n - 600:, None, :] != 0) & (X[None, :64, :] != 0)
    assert np.array_equal(rows[:, :64].astype(np.int32), head.sum(axis=2))
    assert np.array_equal(direct.reshape(40, 500), rows[-40:, :500].astype(np.int32))
    assert Z.shape == (n - 1, 4) and Z[-1, 3] == n


@pytest.mark.parametrize("n,L", [(5, 1), (130, 33), (700, 95), (1500, 260)])
def test_uint16_input_and_both_recoder_paths_agree(hc, oracle, n, L):
    """The re-coder has a fast path (ids < 2^14: 32 loci per CTA straight from the row-major input) and a general
    one (transposed copy, 2^20-bit bitmaps); the input may be int32 or uint16.  Dense ranking is equality-preserving,
    so the four combinations must give the oracle's distances on the same matrix (ids scaled by 40 leave the
    fast path but still fit uint16)."""
    rng = np.random.default_rng(n * 31 + L)
    X = rng.integers(0, 1500, size=(n, L)).astype(np.int32)
    X[rng.random((n, L)) < 0.05] = 0
    X[:, 0] = rng.integers(1, 3, n)
    ref = oracle.dist_condensed(X)
    for scale in (1, 40):
        for dtype in (np.int32, np.uint16):
            with hc.HCContext() as ctx:
                ctx.set_profiles((X * scale).astype(dtype))
                ctx.distance()
                assert np.array_equal(ctx.condensed(), ref), (scale, dtype)
    # negative int32 ids are missing on both paths
    Y = X.copy()
    Y[X == 0] = -7
    with hc.HCContext() as ctx:
        ctx.set_profiles(Y)
        ctx.distance()
        assert np.array_equal(ctx.condensed(), ref)


def _shared_code_matrix(kind):
    rng = np.random.default_rng(77)
    if kind == "singletons":
        # mixtures: no singletons, a few, hundreds, only singletons + one shared allele, singletons next to missing cells
        n, L = 3000, 40
        X = rng.integers(1, 6, size=(n, L)).astype(np.int32)
        for l in range(L):
            k = [0, 2, 3, 17, 60, 250, 700, 1500][l % 8]                    # singletons of this locus
            rows = rng.choice(n, k, replace=False)
            X[rows, l] = 1000 + np.arange(k)
            if l % 5 == 0:
                X[rng.choice(n, 40, replace=False), l] = 0
        X[:, 7] = 1
        X[rng.choice(n, 300, replace=False), 7] = 2000 + np.arange(300)      # one shared allele + 300 singletons
        return X
    if kind == "zipf":
        # founder alleles with a long tail of rare ones (the shape of the C3 / C4 / C5 workloads), 3 % missing
        from reportree_b200 import synth
        return synth.allele_matrix(2500, 96, seed=5, missing=0.03, p_mut=0.02, founders=60)
    # "flat": many alleles of similar, small counts -- every code of a locus receives (k = C), ties everywhere
    n, L = 2000, 64
    X = np.empty((n, L), dtype=np.int32)
    for l in range(L):
        A = [9, 20, 64, 130, 300, 700, 1500, 1999][l % 8]
        X[:, l] = 1 + rng.integers(0, A, n) * 3
    X[rng.random((n, L)) < 0.02] = 0
    return X


@pytest.mark.parametrize("kind,div,world", [("singletons", None, 1), ("singletons", None, 3), ("singletons", "1", 1),
                                            ("zipf", None, 1), ("zipf", "1", 3), ("zipf", "64", 1),
                                            ("flat", "1", 1), ("flat", "300", 1), ("flat", None, 2)])
def test_shared_codes_are_fixed_up_exactly(hc, oracle, monkeypatch, kind, div, world):
    """Rare alleles of a locus share codes (fewer bit planes, collide.cu) and the pairs of cells that hold two
    different alleles of one code get their mismatch back from rthc_collide_fixup.  RTHC_COLLIDE_DIV moves the
    device's cost rule: 1 = share wherever the per-locus cell cap allows (every branch of the plan: all codes
    receiving, level search, cap refusals), default = what production does at this size."""
    if div is None:
        monkeypatch.delenv("RTHC_COLLIDE_DIV", raising=False)
    else:
        monkeypatch.setenv("RTHC_COLLIDE_DIV", div)
    X = _shared_code_matrix(kind)
    n = X.shape[0]
    rng = np.random.default_rng(1)
    ref = oracle.dist_condensed(X)
    ctxs = [hc.HCContext(rank=r, world=world) for r in range(world)]
    try:
        got = np.zeros_like(ref)
        seen = np.zeros(ref.size, dtype=bool)
        from reportree_b200 import sharded
        pairs = 0
        for r, ctx in enumerate(ctxs):
            ctx.set_profiles(X)
            info = ctx.info()
            if div is not None or kind == "singletons":
                assert info["escape_loci"] >= 5 and info["escape_cells"] > 0, info
            assert info["max_alleles"] == max(np.unique(c[c > 0]).size for c in X.T)      # distinct alleles, not codes
            ctx.distance()
            pairs += ctx.info()["escape_pairs"]
            rows = sharded.owned_rows(n, r, world)
            own = ctx.owned_condensed()
            p = 0
            for i in rows:
                base = n * i - i * (i + 1) // 2
                cnt = n - 1 - i
                got[base:base + cnt] = own[p:p + cnt]
                seen[base:base + cnt] = True
                p += cnt
            if world == 1:                                                # the scalar check kernel reads the dense ranks
                pi = rng.integers(0, n, 4000).astype(np.int32)
                pj = rng.integers(0, n, 4000).astype(np.int32)
                sq = (X[pi] != X[pj]) & (X[pi] != 0) & (X[pj] != 0)
                assert np.array_equal(ctx.direct_pairs(pi, pj), sq.sum(1))
        assert seen.all() and np.array_equal(got, ref)
        if div is not None or kind == "singletons":
            assert pairs > 0
    finally:
        for c in ctxs:
            c.close()


def test_wide_snp_matrix_and_the_informative_loci_limit(hc, oracle):
    # SNP matrices (conv_nucl, scripts/partitioning_HC.py:28-47) can have far more than 65,535 columns; only
    # the columns with >= 2 called alleles can mismatch, and only their number is bounded by uint16 distances
    rng = np.random.default_rng(23)
    n, L = 150, 70001
    X = np.full((n, L), 2, dtype=np.int32)                 # monomorphic background
    info_cols = rng.choice(L, 4000, replace=False)
    X[:, info_cols] = rng.integers(1, 5, (n, info_cols.size))
    X[rng.random((n, L)) < 0.01] = 0
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        assert ctx.info()["loci_kept"] == info_cols.size
        ctx.distance()
        got = ctx.condensed()
    assert np.array_equal(got, oracle.dist_condensed(X))
    Y = rng.integers(1, 3, (40, 65600)).astype(np.int32)   # every column informative
    Y[0], Y[1] = 1, 2                                      # one pair differs everywhere: d = 65,535 at the limit
    with hc.HCContext() as ctx:
        with pytest.raises(hc.HCBackendError, match="65,535"):
            ctx.set_profiles(Y)
        ctx.set_profiles(np.ascontiguousarray(Y[:, :65535]))    # exactly at the limit: distances up to 65,535 fit
        ctx.distance()
        got = ctx.condensed()
    assert got[0] == 65535
    assert np.array_equal(got, oracle.dist_condensed(np.ascontiguousarray(Y[:, :65535])))


def test_negative_codes_are_missing(hc, oracle):
    rng = np.random.default_rng(17)
    X = _rand_codes(rng, 64, 64, 5, 0.0)
    Y = X.copy()
    Y[rng.random(X.shape) < 0.1] = -1
    X0 = np.where(Y < 0, 0, Y)
    ref = oracle.dist_condensed(X0)
    with hc.HCContext() as ctx:
        ctx.set_profiles(Y)
        ctx.distance()
        assert np.array_equal(ctx.condensed(), ref)


def test_direct_kernel_agrees_with_bitsliced_on_synthetic(hc):
    from reportree_b200 import synth
    X = synth.allele_matrix(3000, 3002, 5)
    rng = np.random.default_rng(3)
    pi = rng.integers(0, 3000, size=20000).astype(np.int32)
    pj = rng.integers(0, 3000, size=20000).astype(np.int32)
    keep = pi != pj
    pi, pj = np.minimum(pi, pj)[keep], np.maximum(pi, pj)[keep]
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        y = ctx.condensed()
        d = ctx.direct_pairs(pi, pj)
    n = 3000
    idx = n * pi.astype(np.int64) - pi.astype(np.int64) * (pi + 1) // 2 + (pj - pi - 1)
    assert np.array_equal(y[idx].astype(np.int32), d)


@pytest.mark.parametrize("world", [2, 3, 8])
def test_sharded_rows_cover_matrix(hc, oracle, world):
    rng = np.random.default_rng(23)
    n = 700
    X = _rand_codes(rng, n, 90, 12, 0.05)
    ref = oracle.dist_condensed(X)
    seen = np.zeros(ref.size, dtype=bool)
    B = (n + 127) // 128
    for rank in range(world):
        with hc.HCContext(rank=rank, world=world) as ctx:
            ctx.set_profiles(X)
            ctx.distance()
            part = ctx.owned_condensed()
        rows = [i for i in range(n) if min(i // 128, B - 1 - i // 128) % world == rank]
        pos = 0
        for i in rows:
            ln = n - 1 - i
            base = n * i - i * (i + 1) // 2
            assert np.array_equal(part[pos:pos + ln], ref[base:base + ln])
            seen[base:base + ln] = True
            pos += ln
        assert pos == part.size
    assert seen.all()


def _sharded_recode_distances(hc, X, world, dtype=np.int32):
    """rthc_recode_shard / rthc_export_shard / rthc_import_planes with `world` contexts on ONE device: the gathered
    buffers of sharded.PlaneExchange are then simply shared, no collective needed.  Returns the assembled matrix."""
    import torch
    from reportree_b200 import sharded
    n, L = X.shape
    chunk = sharded.shard_chunk(L, world)
    assert chunk == hc.load_library().rthc_shard_chunk(L, world)
    hdr_len = sharded.SHARD_HDR + chunk // 32
    dX = torch.from_numpy(np.ascontiguousarray(X.astype(dtype))).cuda() if dtype != np.uint16 else \
        torch.from_numpy(np.ascontiguousarray(X.astype(np.uint16)).view(np.int16)).cuda()
    ctxs = [hc.HCContext(rank=r, world=world) for r in range(world)]
    try:
        hdr_all = np.zeros((world, hdr_len), dtype=np.int32)
        for r, ctx in enumerate(ctxs):
            ctx.recode_shard(dX.data_ptr(), n, L, dtype, hdr_all[r])
        rows_max = max(1, int(hdr_all[:, 1].max()))
        n_blk = (n + 127) // 128 * 2
        planes = torch.zeros((world, n_blk, rows_max, 64), dtype=torch.int32, device="cuda")
        cnt = torch.zeros((world, chunk), dtype=torch.int32, device="cuda")
        cells = torch.zeros((world, chunk, sharded.LIST_CAP), dtype=torch.int64, device="cuda")
        for r, ctx in enumerate(ctxs):
            ctx.export_shard(planes[r].data_ptr(), rows_max, cnt[r].data_ptr(), cells[r].data_ptr())
        torch.cuda.synchronize()
        P = n * (n - 1) // 2
        got = np.zeros(P, dtype=np.uint16)
        seen = np.zeros(P, dtype=bool)
        infos = []
        for r, ctx in enumerate(ctxs):
            ctx.import_planes(planes.data_ptr(), rows_max, hdr_all, cnt.data_ptr(), cells.data_ptr())
            ctx.distance()
            infos.append(ctx.info())
            own = ctx.owned_condensed()
            p = 0
            for i in sharded.owned_rows(n, r, world):
                base = n * i - i * (i + 1) // 2
                c = n - 1 - i
                got[base:base + c] = own[p:p + c]
                seen[base:base + c] = True
                p += c
        assert seen.all()
        return got, infos, hdr_all
    finally:
        for c in ctxs:
            c.close()


@pytest.mark.parametrize("kind,world,div", [("zipf", 2, None), ("zipf", 3, "1"), ("singletons", 8, None),
                                            ("flat", 4, "300"), ("narrow", 3, None), ("wideids", 2, None)])
def test_recoder_sharded_by_loci_gives_the_same_distances(hc, oracle, monkeypatch, kind, world, div):
    """world > 1: every rank re-codes only its range of loci and the plane matrix is assembled from the ranks'
    parts (rthc_recode_shard / rthc_export_shard / rthc_import_planes).  Cases: rare alleles with shared codes and
    their fix-up lists crossing ranks, more ranks than 32-locus blocks (ranks without loci), uint16 input, a range
    that needs the general path (ids >= 2^14) next to ranges that do not."""
    if div is None:
        monkeypatch.delenv("RTHC_COLLIDE_DIV", raising=False)
    else:
        monkeypatch.setenv("RTHC_COLLIDE_DIV", div)
    dtype = np.int32
    if kind == "narrow":
        rng = np.random.default_rng(3)
        X = _rand_codes(rng, 300, 40, 9, 0.05)              # 40 loci over 3 ranks: chunk 32 -> 32 + 8 + 0
        dtype = np.uint16
    elif kind == "wideids":
        rng = np.random.default_rng(4)
        X = _rand_codes(rng, 500, 150, 30, 0.03)
        X[:, 100:] = np.where(X[:, 100:] > 0, X[:, 100:] * 40000, 0)     # second rank only: ids up to 1.2e6
    else:
        X = _shared_code_matrix(kind)
    ref = oracle.dist_condensed(X)
    got, infos, hdr_all = _sharded_recode_distances(hc, X, world, dtype)
    assert np.array_equal(got, ref)
    assert all(i["L"] == X.shape[1] for i in infos)
    assert infos[0]["plane_rows"] == int(hdr_all[:, 1].sum()) and infos[0]["groups"] == int(hdr_all[:, 0].sum())
    if kind == "narrow":
        assert hdr_all[2, 0] == 0 and hdr_all[2, 7] == 0            # the third rank had no loci
    if div is not None or kind == "singletons":
        assert infos[0]["escape_cells"] > 0


@pytest.mark.parametrize("div", [None, "1"])
def test_shared_codes_at_the_fast_path_limit_of_alleles(hc, oracle, monkeypatch, div):
    """A locus with 16,383 distinct allele ids (the most the fast re-coder path takes: ids < 2^14) next to ordinary
    loci: the shared-code plan then sorts 16,384 keys in shared memory and works with 14-bit dense codes."""
    if div is None:
        monkeypatch.delenv("RTHC_COLLIDE_DIV", raising=False)
    else:
        monkeypatch.setenv("RTHC_COLLIDE_DIV", div)
    rng = np.random.default_rng(99)
    n, L = 17000, 6
    X = rng.integers(1, 40, size=(n, L)).astype(np.int32)
    X[:, 2] = 1 + (np.arange(n) % 16383)                       # ids up to 16,383, 617 of them twice
    X[:, 4] = 1 + (rng.zipf(1.3, n) % 9000)                    # long tail
    X[rng.random((n, L)) < 0.01] = 0
    ref = oracle.dist_condensed(X)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        info = ctx.info()
        assert info["max_alleles"] == max(np.unique(c[c > 0]).size for c in X.T) > 16000     # (missing cells took a few)
        ctx.distance()
        assert np.array_equal(ctx.condensed(), ref)
        if div == "1":
            assert ctx.info()["escape_pairs"] > 0
