"""Host-side text <-> array helpers of the drop-in CLI (vectorised replacements for the
reference's per-cell pandas / pure-Python loops; outputs are byte-identical).

    encode_cells     <- conv_nucl (scripts/partitioning_HC.py:28-47) + the cell parse of
                        `cgmlst-dists` (contract :293-298)
    write_square_tsv <- dist_df.to_csv(<out>_dist_hamming.tsv) (:517-518)
    name_partition   <- the naming loop (:651-670 and its three copies)
    newick           <- to_tree + recursive get_newick (:232-257, :629-633)
"""
import string

import numpy as np
from reportree_b200.lazy_pandas import pandas

_LETTERS = {c: i + 1 for i, c in enumerate(string.ascii_lowercase + string.ascii_uppercase)}


def _token_to_code(tok):
    """One cleaned cell -> allele code.  Single ASCII letter -> 1..52 (conv_nucl); clean
    non-negative decimal integer -> its value; anything else (NaN, 'LNF', '-', ...) -> 0."""
    if not isinstance(tok, str):
        return 0
    v = _LETTERS.get(tok)
    if v is not None:
        return v
    if tok.isascii() and tok.isdigit() and len(tok) <= 9:
        return int(tok)
    return 0


def encode_cells(frame):
    """String DataFrame of loci columns (no id column) -> n x L int32 codes, 0 = missing."""
    vals = frame.to_numpy(dtype=object)
    n, L = vals.shape
    codes, uniques = pandas.factorize(vals.ravel(), use_na_sentinel=True)
    lut = np.fromiter((_token_to_code(u) for u in uniques), dtype=np.int64, count=len(uniques))
    big = lut > np.iinfo(np.int32).max
    if big.any():
        lut[big] = 0
    lut = np.concatenate([lut, [0]]).astype(np.int32)       # slot -1 (NaN) -> 0
    return lut[codes].reshape(n, L)


def write_square_tsv(path, names, row_block_fn, n, block=512, corner="dists"):
    """Stream the square distance matrix as TSV: header = corner + names; one row per sample."""
    with open(path, "w") as fh:
        fh.write(corner + "\t" + "\t".join(names) + "\n")
        for r0 in range(0, n, block):
            nr = min(block, n - r0)
            rows = row_block_fn(r0, nr)
            txt = rows.astype("U5") if rows.dtype != object else rows
            lines = ["\t".join(r) for r in txt.tolist()]
            fh.write("".join(names[r0 + k] + "\t" + lines[k] + "\n" for k in range(nr)))


def name_partition(labels, sample_names):
    """fcluster ids -> ReporTree cluster names + composition in first-appearance order.

    Size-1 clusters become singleton_<running count in row order>, the others cluster_<id>.
    Returns (names list[str] per sample, composition list of (name, size, [samples]))."""
    labels = np.asarray(labels)
    n = labels.size
    counts = np.bincount(labels)
    size = counts[labels]
    single = size == 1
    names = np.empty(n, dtype=object)
    k = np.cumsum(single)
    names[single] = ["singleton_" + str(v) for v in k[single].tolist()]
    multi = ~single
    names[multi] = ["cluster_" + str(v) for v in labels[multi].tolist()]
    # composition: clusters in order of first appearance, members in row order
    order = np.argsort(labels, kind="stable")
    sorted_lab = labels[order]
    starts = np.flatnonzero(np.r_[True, sorted_lab[1:] != sorted_lab[:-1]])
    ends = np.r_[starts[1:], n]
    first_row = order[starts]                      # stable sort -> first member is the earliest row
    comp = []
    sample_arr = np.asarray(sample_names, dtype=object)
    for q in np.argsort(first_row, kind="stable").tolist():
        members = order[starts[q]:ends[q]]
        comp.append((names[members[0]], int(ends[q] - starts[q]), sample_arr[members].tolist()))
    return names.tolist(), comp


def newick(Z, leaf_names):
    """Newick string identical to the reference's get_newick(to_tree(Z), ...): children are
    emitted right-then-left, branch length = parent height - node height, '%.2f'."""
    n = Z.shape[0] + 1
    left = Z[:, 0].astype(np.int64)
    right = Z[:, 1].astype(np.int64)
    height = Z[:, 2]
    out = []
    root = 2 * n - 2
    # explicit stack of (node, parent_height, state); state 0 = open, 1 = between children, 2 = close
    stack = [(root, height[root - n], 0)]
    while stack:
        node, ph, state = stack.pop()
        if node < n:
            out.append("%s:%.2f" % (leaf_names[node], ph - 0.0))
            continue
        k = node - n
        if state == 0:
            out.append("(")
            stack.append((node, ph, 1))
            stack.append((int(right[k]), height[k], 0))
        elif state == 1:
            out.append(",")
            stack.append((node, ph, 2))
            stack.append((int(left[k]), height[k], 0))
        else:
            out.append(");" if node == root else "):%.2f" % (ph - height[k]))
    return "".join(out)


def write_partitions(path, columns):
    """<out>_partitions.tsv: same text as DataFrame(columns).to_csv(sep='\\t', index=None) (:779-780)."""
    pandas.DataFrame(data=columns).to_csv(path, sep="\t", index=None)
