"""Pins the N3 / N4 oracle restatements (oracle/zoom_oracle.py, oracle/partition_metrics.py) on the
outputs of the unmodified reference code stored under tests/golden/next/ (made by
tests/golden/make_golden_next.py), and -- where /root/reference exists -- on the reference's own
metrics.py imported directly.  CPU only."""
import gzip
import json
import os
import sys

import numpy as np
import pandas as pd
import pytest

from oracle import partition_metrics as pm
from oracle import zoom_oracle

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
NEXT = os.path.join(GOLD, "next")
REF_CP = "/root/reference/scripts/ComparingPartitions"


def read_partitions(case):
    with gzip.open(os.path.join(GOLD, case, case + "_partitions.tsv.gz"), "rt") as fh:
        return pd.read_table(fh)


def golden_lines(name):
    with open(os.path.join(NEXT, name)) as fh:
        return fh.read().split("\n")


def columns_of(df, reverse=False):
    names = list(df.columns[1:])
    if reverse:
        names = names[::-1]
    return names, [df[c].values.tolist() for c in names]


@pytest.mark.parametrize("case,table,reverse", [("C2_keep", "C2", False), ("C1b_keep", "C1b", False), ("C2_keep_rev", "C2", True)])
def test_stability_lines_equal_reference_files(case, table, reverse):
    names, cols = columns_of(read_partitions(table), reverse)
    lines = pm.stability_lines(names, cols)
    gold = golden_lines("stability_%s_metrics.tsv" % case)
    assert gold[-1] == ""
    assert lines == gold[:-1]
    regions = pm.stable_regions(lines, 0.99, 5)
    gold_regions = [l.split("\t") for l in golden_lines("stability_%s_stableRegions.tsv" % case)[3:] if l]
    assert [[b, f, l, str(k)] for b, f, l, k in regions] == gold_regions


def first_representatives(df):
    """rm_redundant (comparing_partitions_v2.py:33-62) with the first sample of every cluster of the
    most discriminatory column standing in for random.choice."""
    best, main = 0, None
    for c in df.columns[1:]:
        k = df[c].nunique()
        if k > best:
            best, main = k, c
    return df[~df[main].duplicated()]


def test_stability_default_mode_matches_up_to_the_random_representatives():
    # single-linkage columns are nested, so any representative of a cluster of the finest column gives
    # the same tables; only the order of the floating-point sums can differ
    names, cols = columns_of(first_representatives(read_partitions("C2")))
    lines = pm.stability_lines(names, cols)
    gold = golden_lines("stability_C2_default_metrics.tsv")[:-1]
    assert len(lines) == len(gold)
    for mine, ref in zip(lines[1:], gold[1:]):
        a, b = mine.split("\t"), ref.split("\t")
        assert a[:2] == b[:2] and a[9:] == b[9:]
        np.testing.assert_allclose([float(x) for x in a[2:9]], [float(x) for x in b[2:9]], rtol=1e-12, atol=1e-12)


def test_closest_samples_equal_reference_function():
    z = np.load(os.path.join(GOLD, "C2", "C2_dist_hamming.npz"))
    ids, dist = [str(x) for x in z["ids"]], z["dist"]
    with open(os.path.join(NEXT, "closest_C2.json")) as fh:
        gold = json.load(fh)
    assert len(gold) >= 10
    for q in gold:
        if q["code"] == "IndexError":
            with pytest.raises(IndexError):
                zoom_oracle.closest_samples(dist, ids, q["sample"], q["n"], set(ids))
            continue
        names, dists, code = zoom_oracle.closest_samples(dist, ids, q["sample"], q["n"], set(ids))
        assert code == q["code"]
        assert sorted(names) == q["names"]
        if code == "run":
            assert dists == q["dists"]


@pytest.mark.skipif(not os.path.isdir(REF_CP), reason="reference tree not present (GPU box)")
def test_functions_equal_reference_metrics_module_on_random_labels():
    sys.path.insert(0, REF_CP)
    try:
        import metrics as ref
    finally:
        sys.path.remove(REF_CP)
    rng = np.random.default_rng(7)
    for trial in range(6):
        n = int(rng.integers(5, 400))
        x = ["c%d" % v for v in rng.integers(0, max(2, n // 3), n)]
        y = ["k%d" % v for v in rng.integers(0, max(2, n // 5), n)]
        if trial == 0:
            y = list(x)
        if trial == 1:
            x = ["same"] * n
        cont = ref.getContTable(x, y)
        assert pm.contingency(x, y) == cont
        abcdn = ref.getMismatchMatrix(cont, x, y)
        a, a1 = pm.pair_counts(pm.contingency(x, y))
        assert a == abcdn[0] and a1 - a == abcdn[1]
        assert pm.wallace(pm.contingency(x, y)) == ref.getWallace(abcdn[0], abcdn[1], abcdn[2])[0]
        assert pm.simpson(x) == ref.getSimpsons(x)
        assert pm.adjusted_wallace(pm.contingency(y, x), x) == ref.getAdjustedWallace(ref.getContTable(y, x), y, x)
        assert pm.adjusted_rand(x, y) == ref.getAdjRand(x, y)
