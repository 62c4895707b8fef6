"""CPU (gloo, world_size 2 and 3) test of the multi-GPU host logic in reportree_b200/sharded.py:
tile-row ownership, per-rank minimum spanning forests under the (w, i, j) order, the single
all_gather of padded edge lists, and the final merge -- with the per-rank forest computed by a
plain Kruskal on distances from the CPU oracle (the GPU kernels are covered by the -m gpu tests,
tests/test_gpu_linkage.py::test_forest_union_equals_single_gpu)."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _kruskal(keys, n):
    """Minimum spanning forest of the packed edges under the total key order."""
    parent = list(range(n))

    def find(v):
        while parent[v] != v:
            parent[v] = parent[parent[v]]
            v = parent[v]
        return v

    out = []
    for k in np.sort(keys):
        i = int((k >> np.uint64(24)) & np.uint64(0xFFFFFF))
        j = int(k & np.uint64(0xFFFFFF))
        a, b = find(i), find(j)
        if a != b:
            parent[a] = b
            out.append(k)
    return np.asarray(out, dtype=np.uint64)


def _worker(rank, world, port, n, L, seed, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from oracle import oracle_lib as O
    from reportree_b200 import sharded, synth
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    X = synth.allele_matrix(n, L, seed)
    D = O.dist_square(X)
    rows = sharded.owned_rows(n, rank, world)
    ii, jj = [], []
    for i in rows:
        j = np.arange(i + 1, n)
        ii.append(np.full(j.size, i)); jj.append(j)
    ii = np.concatenate(ii) if ii else np.zeros(0, np.int64)
    jj = np.concatenate(jj) if jj else np.zeros(0, np.int64)
    local = _kruskal(sharded.pack_keys(D[ii, jj], ii, jj), n)
    allk = sharded.all_gather_forests(local, n, dist)
    if rank == 0:
        q.put((int(ii.size), allk))
    else:
        q.put((int(ii.size), None))
    dist.barrier()
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("world", [2, 3])
def test_forest_union_over_gloo_equals_single_rank_mst(world, oracle):
    from scipy.cluster.hierarchy import fcluster, linkage
    from reportree_b200 import sharded, synth
    n, L, seed = 700, 60, 11
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, n, L, seed, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    # every pair is owned exactly once, and the ranks are balanced to within one tile-row pair
    owned = [g[0] for g in got]
    assert sum(owned) == n * (n - 1) // 2
    allk = next(g[1] for g in got if g[1] is not None)
    assert allk.size <= world * (n - 1)
    mst = _kruskal(allk, n)
    assert mst.size == n - 1
    # reference: MST of the complete graph on one rank, same total order
    X = synth.allele_matrix(n, L, seed)
    D = oracle.dist_square(X)
    iu = np.triu_indices(n, 1)
    full = _kruskal(sharded.pack_keys(D[iu], iu[0], iu[1]), n)
    assert np.array_equal(np.sort(mst), np.sort(full))
    # and it cuts like scipy's single linkage at every threshold
    w, i, j = sharded.unpack_keys(np.sort(mst))
    Z = linkage(D[iu].astype(np.float64), "single")
    assert np.array_equal(np.sort(w).astype(np.float64), Z[:, 2])
    for t in (0, 1, 3, 8, 20):
        parent = list(range(n))

        def find(v):
            while parent[v] != v:
                parent[v] = parent[parent[v]]
                v = parent[v]
            return v
        for ww, a, b in zip(w, i, j):
            if ww <= t:
                parent[find(int(a))] = find(int(b))
        mine = oracle.canonical(np.asarray([find(v) for v in range(n)]))
        assert np.array_equal(mine, oracle.canonical(fcluster(Z, t, criterion="distance")))


def test_ownership_is_a_balanced_partition():
    from reportree_b200 import sharded
    for n in (1, 127, 128, 129, 1000, 5000):
        for world in (1, 2, 4, 8):
            rows = [sharded.owned_rows(n, r, world) for r in range(world)]
            allr = np.concatenate(rows)
            assert np.array_equal(np.sort(allr), np.arange(n))
            if n >= 128 * 2 * world:
                pairs = [int(np.sum(n - 1 - r)) for r in rows]
                assert max(pairs) - min(pairs) <= 2 * 128 * n


class _FakeForestCtx:
    """Stands in for HCContext in the exchange helpers: 'device' memory is host memory here."""

    def __init__(self, keys):
        self.keys = np.ascontiguousarray(keys, dtype=np.uint64)
        self.seen = None

    def local_forest_device(self, ptr, capacity):
        import ctypes
        buf = np.full(capacity, np.uint64(0xFFFFFFFFFFFFFFFF), dtype=np.uint64)
        buf[: self.keys.size] = self.keys[::-1]                     # unsorted on purpose
        ctypes.memmove(ptr, buf.ctypes.data, buf.nbytes)
        return int(self.keys.size)

    def linkage_from_forest_device(self, ptr, m, n):
        import ctypes
        out = np.empty(m, dtype=np.uint64)
        ctypes.memmove(out.ctypes.data, ptr, out.nbytes)
        self.seen = out
        return out


def _exchange_worker(rank, world, port, n, L, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from reportree_b200 import sharded
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    # forests of different lengths per rank (rank r holds r + 2 edges)
    mine = sharded.pack_keys(np.arange(rank + 2) + 10 * rank, np.arange(rank + 2), np.arange(rank + 2) + 1)
    ctx = _FakeForestCtx(mine)
    ex = sharded.ForestExchange(n, dist, device="cpu")
    m = ex.local_forest(ctx)
    ex.gather()
    slots = ex.linkage(ctx)
    real = ex.gathered_keys()
    # host matrix -> every rank copies its own row block, one all_gather replicates it
    X = (np.arange(n * L, dtype=np.int32).reshape(n, L) * 7) % 1001
    ld = sharded.ShardedProfileLoader(n, L, dist, device="cpu")
    ptr = ld.load(X)                                   # ids < 65,536: narrowed to uint16 before the gather
    assert ld.dtype == np.uint16 and ptr == ld.full16.data_ptr()
    narrow = ld.replica().numpy()[:n].astype(np.int32)
    Y = X.copy()
    Y[0, 0] = 70000                                    # one wide id: every rank must fall back to int32
    Y[n - 1, 1] = -3                                   # negative = missing, kept as it is on the int32 path
    ptr = ld.load(Y)
    assert ld.dtype == np.int32 and ptr == ld.full.data_ptr()
    assert np.array_equal(ld.replica().numpy()[:n], Y)
    q.put((rank, m, slots.size, real, narrow, ld.own_bytes()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_device_resident_exchange_helpers_over_gloo(world):
    """ForestExchange / ShardedProfileLoader (the NCCL path of bench.py) with CPU tensors: padding, one
    all_gather_into_tensor each, ragged row blocks (n not divisible by world)."""
    from reportree_b200 import sharded
    n, L = 11, 5
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_exchange_worker, args=(r, world, port, n, L, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    X = (np.arange(n * L, dtype=np.int32).reshape(n, L) * 7) % 1001
    expect = np.sort(np.concatenate([sharded.pack_keys(np.arange(r + 2) + 10 * r, np.arange(r + 2), np.arange(r + 2) + 1)
                                     for r in range(world)]))
    for rank, m, slots, real, full, own in got:
        assert m == rank + 2
        assert slots == world * (n - 1)                      # fixed-size gather
        assert np.array_equal(np.sort(real), expect)         # every rank sees every forest, padding dropped
        assert np.array_equal(full, X)
    assert sum(g[5] for g in got) == n * L * 4               # the H2D bytes add up to one copy of the matrix


# ---- re-coder sharded by locus ranges: the exchange of sharded.PlaneExchange over gloo ------------------------------
class _StandInContext:
    """What PlaneExchange needs from hc_backend.HCContext, without a GPU: recode_shard reports a header for this
    rank's locus range, export_shard writes a recognisable pattern into the rank's slots, import_planes keeps what
    it was given.  (The kernels behind the real calls are covered by
    tests/test_gpu_distance.py::test_recoder_sharded_by_loci_gives_the_same_distances.)"""

    def __init__(self, rank, world):
        self.rank, self.world = rank, world
        self.imported = None

    def recode_shard(self, dev_ptr, n, L, dtype, hdr):
        from reportree_b200 import sharded
        chunk = sharded.shard_chunk(L, self.world)
        l0 = min(L, self.rank * chunk)
        loci = min(L, l0 + chunk) - l0
        groups = (loci + 31) // 32
        hdr[:] = 0
        hdr[0], hdr[7] = groups, loci
        for g in range(groups):
            hdr[sharded.SHARD_HDR + g] = 4 + (self.rank + g) % 5            # planes of group g
        hdr[1] = int(hdr[sharded.SHARD_HDR:sharded.SHARD_HDR + groups].sum())
        hdr[2] = int(hdr[sharded.SHARD_HDR:sharded.SHARD_HDR + groups].max()) if groups else 0
        hdr[3], hdr[4], hdr[5], hdr[6] = loci, 100 + self.rank, groups, 7 * self.rank
        self.rows = int(hdr[1])

    def export_shard(self, planes_ptr, rows_max, cnt_ptr, list_ptr):
        import ctypes
        assert rows_max >= self.rows
        self.rows_max = rows_max
        ctypes.memset(planes_ptr, 0x10 + self.rank, self.n_blk * rows_max * 64 * 4)
        ctypes.memset(cnt_ptr, 0x20 + self.rank, self.chunk * 4)
        ctypes.memset(list_ptr, 0x30 + self.rank, self.chunk * 4096 * 8)

    def import_planes(self, planes_ptr, rows_max, hdr_all, cnt_ptr, list_ptr):
        self.imported = (planes_ptr, rows_max, np.array(hdr_all), cnt_ptr, list_ptr)


def _plane_worker(rank, world, port, n, L, q):
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    from reportree_b200 import sharded
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    ex = sharded.PlaneExchange(n, L, dist, device="cpu")
    ctx = _StandInContext(rank, world)
    ctx.n_blk, ctx.chunk = ex.n_blk, ex.chunk
    ok = True
    for rep in range(2):                                     # the buffers are reused by a second matrix of the same shape
        hdr_all = ex.set_profiles(ctx, 0, np.int32)
        ok &= ctx.imported is not None and ctx.imported[0] == ex.planes_all.data_ptr() and ctx.imported[1] == ex.rows_max
        ok &= ex.rows_max == int(hdr_all[:, 1].max())
        for r in range(world):
            ok &= bool((ex.planes_all[r].view(-1).numpy().view(np.uint8) == 0x10 + r).all())
            ok &= bool((ex.cnt_all[r].numpy().view(np.uint8) == 0x20 + r).all())
            ok &= bool((ex.list_all[r].view(-1).numpy().view(np.uint8) == 0x30 + r).all())
            ok &= int(hdr_all[r, 4]) == 100 + r and int(hdr_all[r, 6]) == 7 * r
    q.put((rank, ok, hdr_all.tolist(), ex.chunk))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("world,L", [(2, 3000), (3, 40)])
def test_plane_exchange_over_gloo(world, L):
    from reportree_b200 import sharded
    n = 300
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_plane_worker, args=(r, world, port, n, L, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    chunk = sharded.shard_chunk(L, world)
    for rank, ok, hdr_all, ch in got:
        assert ok and ch == chunk
        assert hdr_all == got[0][2]                           # every rank holds the same gathered headers
        assert sum(h[7] for h in hdr_all) == L                # the ranks' locus ranges tile the matrix
    if world == 3:
        assert [h[7] for h in got[0][2]] == [32, 8, 0]        # more ranks than 32-locus blocks
