"""Zoom-in / subtree queries on a distance matrix that stays resident in HBM.

The reference serves both from text files: `get_closest_samples` (reportree.py:591-617) re-reads one
column of `<out>_dist_hamming.tsv` for every sample of interest, and the zoom-in loop
(reportree.py:1965-2055) re-invokes partitioning_HC.py -- `cgmlst-dists` included -- on the subset
with `--loci <out>_loci_used.txt`.  With the loci fixed, a pair's distance does not depend on which
other samples are present, so both are look-ups in the matrix the first run already computed:

    rm = ResidentMatrix(ctx, sample_names)            # ctx: HCContext after distance()
    names, code = get_closest_samples("S0000042", 50, rm, set(sample_names))
    Z, labels = rm.cluster_subset(names, "single", [4, 7, 14])

There is no CPU fallback: every number comes from `rthc_closest` / `rthc_subset_from`
(csrc/zoom.cu) through hc_backend.
"""
import numpy as np


class ResidentMatrix:
    """A context holding a full distance matrix + the sample names of its rows."""

    def __init__(self, ctx, names):
        names = [str(x) for x in names]
        if len(names) != ctx.n:
            raise ValueError("%d names for a %d-sample matrix" % (len(names), ctx.n))
        self.ctx = ctx
        self.names = names
        self._child = None                    # one child context, refilled by every cluster_subset()
        self._pos = {}
        for k, s in enumerate(names):
            self._pos.setdefault(s, k)        # duplicated ids: the first row wins, as a column look-up by name would

    def __contains__(self, sample):
        return str(sample) in self._pos

    def position(self, sample):
        return self._pos[str(sample)]

    def closest(self, sample, n):
        """(names, distances, max_dist) of every sample within the distance found at sorted position
        `n` of `sample`'s column (its own 0 included), ordered by (distance, row)."""
        idx, d, vmax = self.ctx.closest(self.position(sample), int(n))
        return [self.names[i] for i in idx], d, vmax

    def subset_context(self, samples):
        """Child context with the sub-matrix of `samples` (names), rows in the given order."""
        return self.ctx.subset(np.fromiter((self.position(s) for s in samples), dtype=np.int64))

    def cluster_subset(self, samples, method, thresholds):
        """linkage + fcluster of the subset without recomputing a distance: (Z, T x m labels)."""
        pos = np.fromiter((self.position(s) for s in samples), dtype=np.int64)
        self._child = self.ctx.subset(pos, into=self._child)
        Z = self._child.linkage(method)
        return Z, self._child.fcluster(thresholds)

    def close(self):
        if self._child is not None:
            self._child.close()
            self._child = None


def get_closest_samples(sample, n, resident, samples_in_partitions):
    """Mirror of get_closest_samples(sample, n, hamming, samples_in_partitions) (reportree.py:591-617)
    with the `_dist_hamming.tsv` path replaced by a ResidentMatrix; same (closest_samples, code):

        "run"            the samples with d <= distances[int(n)] (ascending column, self included)
        "all_samples"    n - 1 >= number of samples: every sample, ascending by distance
        "not_in_hamming" sample absent from the analysis: ""

    Ties: the reference sorts with pandas' default (unstable) quicksort, so the order inside a run of
    equal distances is unspecified there; here it is row order.  The SET is identical, and the caller
    only joins it into a `sample == a,b,c` filter (reportree.py:1959).  As in the reference,
    int(n) == number of samples indexes past the column and raises IndexError."""
    if sample in samples_in_partitions and sample in resident:
        total = resident.ctx.n
        if int(n) - 1 < total:
            if int(n) >= total:
                raise IndexError("list index out of range")
            names, _, _ = resident.closest(sample, int(n))
            return [str(s) for s in names], "run"
        names, _, _ = resident.closest(sample, total - 1)
        return names, "all_samples"
    return "", "not_in_hamming"
