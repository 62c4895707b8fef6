"""CPU restatement of the reference's zoom-in look-ups.

TEST INFRASTRUCTURE ONLY (imported by tests/): the product path is csrc/zoom.cu through
reportree_b200/zoom.py.

Follows, in the reference tree:
  * get_closest_samples, reportree.py:591-617 -- read the column of `sample` from
    <out>_dist_hamming.tsv, sort ascending, max_dist = distances[int(n)], keep d <= max_dist
    (codes "run" / "all_samples" / "not_in_hamming").
  * the zoom-in re-run, reportree.py:1965-2055 -- partitioning_HC.py on the subset with
    `--loci <out>_loci_used.txt`: with the loci fixed the subset's distance matrix is the
    sub-matrix of the full one.

Pinned: tests/golden/next/closest_C2.json holds the outputs of the reference function itself
(imported from /root/reference/reportree.py by tests/golden/make_golden_next.py) on the C2 golden
matrix; tests/test_oracle_next.py checks this restatement against them.
"""
import numpy as np


def closest_samples(square, names, sample, n, samples_in_partitions=None):
    """(set of names, ascending distances of that set, code) -- the order inside ties is
    unspecified in the reference (unstable sort), hence a set."""
    names = [str(x) for x in names]
    if samples_in_partitions is not None and sample not in samples_in_partitions:
        return set(), [], "not_in_hamming"
    if sample not in names:
        return set(), [], "not_in_hamming"
    col = np.asarray(square)[:, names.index(sample)].astype(np.int64)
    order = np.argsort(col, kind="stable")
    dist = col[order]
    if int(n) - 1 < len(dist):
        cut = dist[int(n)]                      # IndexError when int(n) == len(dist), as in the reference
        keep = order[dist <= cut]
        return {names[i] for i in keep}, sorted(int(col[i]) for i in keep), "run"
    return set(names), [int(x) for x in dist], "all_samples"


def sub_square(square, positions):
    p = np.asarray(positions, dtype=np.int64)
    return np.asarray(square)[np.ix_(p, p)]
