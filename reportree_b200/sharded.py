"""Host side of the multi-GPU single-linkage path.

The upper triangle is cut into 128-sample tile-rows; tile-row r of B belongs to rank
min(r, B-1-r) mod world (long and short rows pair up, so every rank owns the same number of
pairs to within one tile-row).  Each rank computes the minimum spanning forest of the edges it
owns under the strict total order (w, i, j); the union of the forests contains the global MST
(cycle property), so ONE fixed-size all_gather of <= n-1 packed 64-bit edges per rank is the
only exchange; rank 0 finishes Boruvka on the <= world*(n-1) gathered edges and cuts.

These helpers are backend-agnostic (`nccl` with CUDA tensors on the GPU box, `gloo` with CPU
tensors in the CPU tests); the distance and forest kernels themselves live in
libreportree_hc.so (rthc_set_shard / rthc_local_forest / rthc_linkage_from_forest).
"""
import numpy as np

TILE = 128
KEY_INF = np.uint64(0xFFFFFFFFFFFFFFFF)


def owner_of_tile_row(r, n_tile_rows, world):
    """Mirror of owner_of_tile_row() in csrc/api.cu."""
    return min(r, n_tile_rows - 1 - r) % world


def owned_rows(n, rank, world):
    """Sample rows whose condensed-matrix row this rank computes and scans."""
    B = (n + TILE - 1) // TILE
    rows = []
    for r in range(B):
        if owner_of_tile_row(r, B, world) == rank:
            rows.extend(range(r * TILE, min(n, (r + 1) * TILE)))
    return np.asarray(rows, dtype=np.int64)


def pack_keys(w, i, j):
    """(w, i, j) -> w << 48 | i << 24 | j, the total order every kernel uses (i < j < 2^24)."""
    return (np.asarray(w, dtype=np.uint64) << np.uint64(48)) | (np.asarray(i, dtype=np.uint64) << np.uint64(24)) \
        | np.asarray(j, dtype=np.uint64)


def unpack_keys(keys):
    keys = np.asarray(keys, dtype=np.uint64)
    return ((keys >> np.uint64(48)).astype(np.int64), ((keys >> np.uint64(24)) & np.uint64(0xFFFFFF)).astype(np.int64),
            (keys & np.uint64(0xFFFFFF)).astype(np.int64))


def all_gather_forests(local_keys, n, dist, device=None):
    """all_gather of every rank's forest, each padded to n-1 keys with +inf.

    Returns the concatenated real keys (uint64 numpy) on every rank.  `dist` is
    torch.distributed (initialised); `device` is the tensor device ('cuda' for nccl, None/cpu
    for gloo)."""
    import torch
    buf = np.full(max(n - 1, 1), KEY_INF, dtype=np.uint64)
    k = np.asarray(local_keys, dtype=np.uint64)
    if k.size > n - 1:
        raise ValueError("a forest over n vertices has at most n-1 edges")
    buf[: k.size] = k
    t = torch.from_numpy(buf.view(np.int64))
    if device is not None:
        t = t.to(device)
    out = [torch.empty_like(t) for _ in range(dist.get_world_size())]
    dist.all_gather(out, t)
    allk = torch.cat(out).cpu().numpy().view(np.uint64)
    return allk[allk != KEY_INF]


class ForestExchange:
    """The same exchange without leaving the device (NCCL): every rank's forest is written by the library
    straight into a torch buffer of n-1 slots (+inf padded, `rthc_local_forest_device`), ONE
    `all_gather_into_tensor` concatenates them, and rank 0 hands the gathered buffer to
    `rthc_linkage_from_forest_device`.  Buffers are allocated once and reused every step."""

    def __init__(self, n, dist, device="cuda"):
        import torch
        self.n, self.dist, self.torch = int(n), dist, torch
        self.cap = max(self.n - 1, 1)
        self.world = dist.get_world_size()
        self.local = torch.empty(self.cap, dtype=torch.int64, device=device)
        self.gathered = torch.empty(self.cap * self.world, dtype=torch.int64, device=device)

    def local_forest(self, ctx):
        """Boruvka over the owned tile-rows; returns this rank's edge count (synchronous)."""
        return ctx.local_forest_device(self.local.data_ptr(), self.cap)

    def gather(self):
        """The one collective of the path; returns when the gathered keys are visible to any stream."""
        self.dist.all_gather_into_tensor(self.gathered, self.local)
        if self.gathered.is_cuda:
            self.torch.cuda.current_stream().synchronize()
        return self.gathered

    def linkage(self, ctx):
        """rank 0: final MST over the <= world*(n-1) gathered edges -> Z."""
        return ctx.linkage_from_forest_device(self.gathered.data_ptr(), self.gathered.numel(), self.n)

    def gathered_keys(self):
        """Real (non-padding) keys of the last gather as uint64 numpy (tests / debugging)."""
        k = self.gathered.cpu().numpy().view(np.uint64)
        return k[k != KEY_INF]


class ShardedProfileLoader:
    """Host -> device load of the n x L int32 code matrix for world > 1: every rank copies only its own
    block of ceil(n / world) rows over PCIe and one all_gather over NVLink replicates the matrix, instead
    of `world` full copies competing for the host's memory bandwidth.  When every allele id of the matrix
    fits 16 bits (one tiny all_reduce(max) decides, so that all ranks take the same branch) the blocks are
    narrowed on the device first and the all_gather moves half the bytes; `load` then returns the uint16
    replica, which `rthc_set_profiles_u16_device` takes as it is."""

    def __init__(self, n, L, dist, device="cuda", narrow=True):
        import torch
        self.n, self.L, self.dist, self.torch = int(n), int(L), dist, torch
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.rows = (self.n + self.world - 1) // self.world
        self.block = torch.zeros((self.rows, self.L), dtype=torch.int32, device=device)
        self.full = torch.empty((self.rows * self.world, self.L), dtype=torch.int32, device=device)
        self.narrow = bool(narrow) and hasattr(torch, "uint16")
        if self.narrow:
            self.block16 = torch.zeros((self.rows, self.L), dtype=torch.uint16, device=device)
            self.full16 = torch.empty((self.rows * self.world, self.L), dtype=torch.uint16, device=device)
            self.vmax = torch.zeros(1, dtype=torch.int32, device=device)
        self.r0 = min(self.n, self.rank * self.rows)
        self.r1 = min(self.n, self.r0 + self.rows)
        self.dtype = np.int32                     # of the replica returned by the last load()

    def own_bytes(self):
        return (self.r1 - self.r0) * self.L * 4

    def load(self, host_codes):
        """host_codes: C-contiguous int32 [n, L] (page-locked for full PCIe speed).  Returns the device
        pointer of the replicated matrix (first n rows are the data); `self.dtype` says int32 or uint16."""
        t = self.torch
        if self.r1 > self.r0:
            self.block[: self.r1 - self.r0].copy_(t.from_numpy(host_codes[self.r0:self.r1]), non_blocking=True)
        if self.narrow:
            self.vmax[0] = self.block.max()
            self.dist.all_reduce(self.vmax, op=self.dist.ReduceOp.MAX)
            if int(self.vmax[0]) < 65536:                              # ids <= 0 are missing either way
                t.clamp(self.block, min=0, out=self.block)
                self.block16.copy_(self.block)
                self.dist.all_gather_into_tensor(self.full16.view(t.uint8), self.block16.view(t.uint8))
                if self.full16.is_cuda:
                    t.cuda.current_stream().synchronize()
                self.dtype = np.uint16
                return self.full16.data_ptr()
        self.dist.all_gather_into_tensor(self.full, self.block)
        if self.full.is_cuda:
            t.cuda.current_stream().synchronize()
        self.dtype = np.int32
        return self.full.data_ptr()

    def replica(self):
        """The replicated matrix of the last load() as a tensor (tests)."""
        return self.full16 if self.dtype == np.uint16 else self.full


SHARD_HDR = 8          # RTHC_SHARD_HDR
LIST_CAP = 4096        # RTHC_ESC_CAP: listed cells per locus


def shard_chunk(L, world):
    """Loci per rank of the sharded re-coder (rthc_shard_chunk): ceil(L / world) rounded up to 32."""
    return ((int(L) + world - 1) // world + 31) // 32 * 32


class PlaneExchange:
    """Re-coder sharded by locus ranges for world > 1 (include/rthc.h, "re-coder sharded by locus ranges"): every
    rank dense-ranks, re-maps and bit-slices only its chunk of the loci, then ONE small all_gather of the headers
    and three all_gather_into_tensor calls (planes, listed-cell counts, listed cells) give every rank the plane
    matrix of the whole matrix.  Replaces ctx.set_profiles_device(ptr, n, L, dtype) on every rank; the gathered
    buffers belong to this object and stay alive with it (the context borrows them)."""

    MAX_LIST_BYTES = 1 << 30

    def __init__(self, n, L, dist, device="cuda"):
        import torch
        self.n, self.L, self.dist, self.torch, self.device = int(n), int(L), dist, torch, device
        self.world, self.rank = dist.get_world_size(), dist.get_rank()
        self.chunk = shard_chunk(self.L, self.world)
        self.hdr_len = SHARD_HDR + self.chunk // 32
        self.n_blk = (self.n + 127) // 128 * 2
        self.hdr = np.zeros(self.hdr_len, dtype=np.int32)
        self.hdr_dev = torch.zeros(self.hdr_len, dtype=torch.int32, device=device)
        self.hdr_all_dev = torch.zeros((self.world, self.hdr_len), dtype=torch.int32, device=device)
        self.cnt_all = torch.zeros((self.world, self.chunk), dtype=torch.int32, device=device)
        self.list_all = torch.zeros((self.world, self.chunk, LIST_CAP), dtype=torch.int64, device=device)
        self.cnt_mine = torch.zeros(self.chunk, dtype=torch.int32, device=device)
        self.list_mine = torch.zeros((self.chunk, LIST_CAP), dtype=torch.int64, device=device)
        self.planes_all = self.planes_mine = None
        self.rows_max = 0

    @staticmethod
    def usable(L, world):
        """False for matrices so wide that the gathered cell lists would not be small (SNP matrices): the caller
        then re-codes the whole matrix on every rank as before."""
        return world > 1 and world * shard_chunk(L, world) * LIST_CAP * 8 <= PlaneExchange.MAX_LIST_BYTES

    def set_profiles(self, ctx, dev_ptr, dtype=np.int32):
        t = self.torch
        ctx.recode_shard(dev_ptr, self.n, self.L, dtype, self.hdr)
        self.hdr_dev.copy_(t.from_numpy(self.hdr))
        self.dist.all_gather_into_tensor(self.hdr_all_dev.view(-1), self.hdr_dev)
        hdr_all = self.hdr_all_dev.cpu().numpy()
        rows_max = max(1, int(hdr_all[:, 1].max()))
        if self.planes_all is None or rows_max > self.rows_max:
            self.rows_max = rows_max
            self.planes_all = t.empty((self.world, self.n_blk, rows_max, 64), dtype=t.int32, device=self.device)
            self.planes_mine = t.empty((self.n_blk, rows_max, 64), dtype=t.int32, device=self.device)
        ctx.export_shard(self.planes_mine.data_ptr(), self.rows_max, self.cnt_mine.data_ptr(), self.list_mine.data_ptr())
        self.dist.all_gather_into_tensor(self.planes_all.view(-1), self.planes_mine.view(-1))      # flat: gloo and NCCL agree on that
        self.dist.all_gather_into_tensor(self.cnt_all.view(-1), self.cnt_mine)
        self.dist.all_gather_into_tensor(self.list_all.view(-1), self.list_mine.view(-1))
        if self.planes_all.is_cuda:
            t.cuda.current_stream().synchronize()
        ctx.import_planes(self.planes_all.data_ptr(), self.rows_max, hdr_all, self.cnt_all.data_ptr(), self.list_all.data_ptr())
        return hdr_all
