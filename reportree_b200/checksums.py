"""Result fingerprints (pure numpy + zlib) used to compare a run with the committed CPU-generated goldens
under tests/golden/large/ without shipping the 10 GB matrices: CRC32 of every 1,024-row block of the
condensed distance vector, CRC32 of Z, and per threshold the CRC32 of the fcluster ids, the cluster
count and the CRC32 of the membership-only (first-appearance) relabelling."""
import zlib

import numpy as np

ROWS_PER_BLOCK = 1024


def block_bounds(n, rows_per_block=ROWS_PER_BLOCK):
    """[(row0, row1, first condensed index, one-past-last condensed index)] of each block of rows."""
    out = []
    for r0 in range(0, n - 1, rows_per_block):
        r1 = min(r0 + rows_per_block, n - 1)
        out.append((r0, r1, n * r0 - r0 * (r0 + 1) // 2, n * r1 - r1 * (r1 + 1) // 2))
    return out


def crc(a):
    return zlib.crc32(np.ascontiguousarray(a).view(np.uint8).reshape(-1)) & 0xFFFFFFFF


def canonical_i32(labels):
    """Cluster ids replaced by the rank of the cluster's first member: equal iff the memberships are equal."""
    labels = np.asarray(labels)
    _, first, inv = np.unique(labels, return_index=True, return_inverse=True)
    rank = np.empty(first.size, dtype=np.int32)
    rank[np.argsort(first, kind="stable")] = np.arange(first.size, dtype=np.int32)
    return rank[inv.reshape(-1)]


def labels_fingerprint(row):
    row = np.ascontiguousarray(row, dtype=np.int32)
    return {"ids_crc32": crc(row), "clusters": int(row.max()), "membership_crc32": crc(canonical_i32(row))}


def z_fingerprint(Z):
    Z = np.ascontiguousarray(Z, dtype=np.float64)
    return {"z_crc32": crc(Z), "heights_crc32": crc(np.ascontiguousarray(Z[:, 2])), "sizes_crc32": crc(np.ascontiguousarray(Z[:, 3]))}
