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
