"""CPU: the oracle is pinned on the reference's shipped golden pair and on outputs of the
unmodified reference script (tests/golden/make_golden.py)."""
import gzip
import io
import os

import numpy as np
import pandas as pd
import pytest
from scipy.cluster.hierarchy import fcluster, linkage, maxdists

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def test_distance_rule_on_shipped_listeria_pair(oracle):
    z = np.load(os.path.join(G, "Lm_shipped_pair.npz"))
    D = oracle.dist_square(z["codes"].astype(np.int32))
    assert D.shape == (861, 861)
    assert np.array_equal(D, z["dist"].astype(np.int32))
    # plain Hamming (missing counted as an allele) does NOT reproduce the golden
    c = z["codes"][:50].astype(np.int32)
    plain = (c[:, None, :] != c[None, :, :]).sum(-1)
    assert not np.array_equal(plain, z["dist"][:50, :50].astype(np.int32))


def test_chunked_sum_equals_single_pass(oracle):
    z = np.load(os.path.join(G, "Lm_shipped_pair.npz"))
    c = z["codes"][:120].astype(np.int32)
    assert np.array_equal(oracle.dist_square_chunked(c, 500), oracle.dist_square(c))


@pytest.mark.parametrize("case", ["C1", "C2"])
def test_reference_run_distances_match_oracle_via_hostio(oracle, case):
    from reportree_b200 import hostio
    inp = {"C1": "allele_matrix.tsv.gz", "C2": "Listeria_alleles.tsv.gz"}[case]
    mx = pd.read_table(os.path.join(G, "inputs", inp), dtype=str)
    codes = hostio.encode_cells(mx.iloc[:, 1:])
    z = np.load(os.path.join(G, case, case + "_dist_hamming.npz"))
    assert np.array_equal(oracle.dist_square(codes), z["dist"].astype(np.int32))
    assert list(z["ids"]) == mx.iloc[:, 0].astype(str).tolist()


def test_oracle_linkage_and_fcluster_equal_scipy_on_tie_heavy_matrices(oracle):
    rng = np.random.default_rng(2024)
    for _ in range(150):
        n = int(rng.integers(2, 45))
        y = rng.integers(0, int(rng.integers(2, 9)), size=n * (n - 1) // 2).astype(np.float64)
        for method in ("single", "complete", "average", "weighted", "ward"):
            Zs = linkage(y, method)
            Zo = oracle.linkage(y, method)
            assert np.array_equal(Zs, Zo), (method, n)
            assert np.array_equal(maxdists(Zs), oracle.maxdists(Zo))
            for t in (0, 1, 1.5, 2, 3, 7):
                assert np.array_equal(fcluster(Zs, t, criterion="distance"), oracle.fcluster(Zo, t))


def test_oracle_uint16_input_equals_scipy(oracle):
    """The uint16-condensed entry points (what lets the oracle run configs[3] / [4] in host memory)
    give the same Z as live scipy on the float64 copy of the same tie-heavy matrices."""
    rng = np.random.default_rng(77)
    for _ in range(120):
        n = int(rng.integers(2, 60))
        y16 = rng.integers(0, int(rng.integers(2, 12)), size=n * (n - 1) // 2).astype(np.uint16)
        if rng.random() < 0.2:
            y16[rng.integers(0, y16.size, size=max(1, y16.size // 7))] = 65535      # the storage maximum
        for method in ("single", "complete", "average", "weighted", "ward"):
            assert np.array_equal(linkage(y16.astype(np.float64), method), oracle.linkage(y16, method)), (method, n)


def test_oracle_reproduces_reference_partitions_listeria(oracle):
    """Z from the oracle on the golden distances -> the partitions.tsv the reference wrote."""
    from reportree_b200 import hostio
    z = np.load(os.path.join(G, "C2", "C2_dist_hamming.npz"))
    D = z["dist"].astype(np.float64)
    y = D[np.triu_indices(D.shape[0], 1)]
    Z = oracle.linkage(y, "single")
    ref = pd.read_table(gzip.open(os.path.join(G, "C2", "C2_partitions.tsv.gz")), dtype=str)
    names = list(z["ids"])
    assert ref["sequence"].tolist() == names
    for thr in (1, 7, 50):
        col, _ = hostio.name_partition(oracle.fcluster(Z, float(thr)), names)
        assert col == ref["single-%dx1.0" % thr].tolist()
