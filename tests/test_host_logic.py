"""CPU: host-side mirror of scripts/partitioning_HC.py (threshold grammar, naming, Newick,
cell encoding, filters) against scipy and the golden outputs of the reference script."""
import gzip
import os
import types

import numpy as np
import pandas as pd
import pytest
from scipy.cluster.hierarchy import fcluster, linkage, to_tree

from reportree_b200 import hostio, partitioning_HC as phc

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _ref_newick(node, parent_dist, leaf_names, newick=""):
    # behaviour of get_newick (scripts/partitioning_HC.py:232-257) restated for the test
    if node.is_leaf():
        return "%s:%.2f%s" % (leaf_names[node.id], parent_dist - node.dist, newick)
    newick = "):%.2f%s" % (parent_dist - node.dist, newick) if len(newick) > 0 else ");"
    newick = _ref_newick(node.get_left(), node.dist, leaf_names, newick)
    newick = _ref_newick(node.get_right(), node.dist, leaf_names, ",%s" % newick)
    return "(%s" % newick


def test_newick_matches_recursive_reference_and_golden():
    rng = np.random.default_rng(3)
    for n in (2, 3, 9, 40):
        y = rng.integers(0, 6, size=n * (n - 1) // 2).astype(float)
        for method in ("single", "average"):
            Z = linkage(y, method)
            names = ["s%d" % i for i in range(n)]
            tree = to_tree(Z, False)
            assert hostio.newick(Z, names) == _ref_newick(tree, tree.dist, names)
    z = np.load(os.path.join(G, "C1", "C1_dist_hamming.npz"))
    D = z["dist"].astype(float)
    Z = linkage(D[np.triu_indices(D.shape[0], 1)], "single")
    gold = gzip.open(os.path.join(G, "C1", "C1_single_HC.nwk.gz"), "rt").read()
    assert hostio.newick(Z, list(z["ids"])) + "\n" == gold


def test_deep_chain_newick_does_not_recurse():
    n = 30000
    Z = np.zeros((n - 1, 4))
    Z[0] = (0, 1, 1, 2)
    for k in range(1, n - 1):
        Z[k] = (k + 1, n + k - 1, k + 1, k + 2)
    s = hostio.newick(Z, [str(i) for i in range(n)])
    assert s.count("(") == n - 1 and s.endswith(");")


def _naming_loop(info_clusters, names):
    # behaviour of scripts/partitioning_HC.py:651-670 restated for the test
    counter = {c: info_clusters.count(c) for c in set(info_clusters)}
    out, details, k = [], {}, 0
    for i, c in enumerate(info_clusters):
        if counter[c] == 1:
            k += 1
            nm, size = "singleton_" + str(k), 1
        else:
            nm, size = "cluster_" + str(c), counter[c]
        details.setdefault(nm, {}).setdefault(size, []).append(names[i])
        out.append(nm)
    return out, details


def test_name_partition_matches_reference_loop():
    rng = np.random.default_rng(5)
    for _ in range(50):
        n = int(rng.integers(2, 60))
        labels = rng.integers(1, max(2, n // 2), size=n)
        _, labels = np.unique(labels, return_inverse=True)
        labels = labels + 1
        names = ["id%d" % i for i in range(n)]
        col, comp = hostio.name_partition(labels, names)
        rcol, rdet = _naming_loop(list(labels), names)
        assert col == rcol
        assert [(c, s, m) for c, s, m in comp] == [(c, s, m) for c, d in rdet.items() for s, m in d.items()]


def test_encode_cells_contract():
    df = pd.DataFrame({"a": ["1", "01", "A", "a", "LNF", None, "0", "-3", "12x", "T"],
                       "b": ["437", "z", "Z", "", "1.0", "7", "00", " 5", "99999999999", "c"]}, dtype=object)
    got = hostio.encode_cells(df)
    assert got[:, 0].tolist() == [1, 1, 27, 1, 0, 0, 0, 0, 0, 46]
    assert got[:, 1].tolist() == [437, 26, 52, 0, 0, 7, 0, 0, 0, 3]


def test_plan_and_expand_cuts_grammar():
    said = []
    say = lambda *p: said.append(" ".join(str(x) for x in p))
    runs, pct = phc.plan_runs("single-1-10,complete,single-7,average-3", "single-0.005,single-0.0051", 1000, say)
    assert runs == {"single": [("5", "pct"), ("1-10", "general"), ("7", "general")],
                    "complete": [("all", "general")], "average": [("3", "general")]}
    assert pct == {"5": ["0.005", "0.0051"]}
    cuts = phc.expand_cuts("single", runs["single"], pct, np.float64(8.0), 1.0, say)
    assert [c[0] for c in cuts] == ["single-0.005", "single-0.0051"] + ["single-%dx1.0" % t for t in range(1, 9)] + ["single-7x1.0"]
    assert cuts[0][1] == 5 and cuts[2][1] == 1.0
    cuts = phc.expand_cuts("complete", [("all", "general")], {}, np.float64(3.0), 2.0, say)
    assert cuts == [("complete-%dx2.0" % t, t * 2.0) for t in range(4)]
    assert phc.expand_cuts("single", [("20", "general"), ("30-40", "general")], {}, np.float64(8.0), 1.0, say) == []
    assert any("is higher than the maximum distance" in s for s in said)
    assert "\tCalculating clustering in range 0 3.0 with a distance of 2.0" in said


def test_column_and_row_filters_match_reference_outputs(tmp_path):
    say = lambda *p: None
    mx = pd.read_table(os.path.join(G, "inputs", "Listeria_alleles.tsv.gz"), dtype=str)
    args = types.SimpleNamespace(loci="none", samples_called="0.99", loci_called="0.95", out=str(tmp_path / "x"))
    mx2, nl = phc.filter_columns(mx, args, say)
    gold_loci = gzip.open(os.path.join(G, "C2b", "C2b_loci_used.txt.gz"), "rt").read().split("\n")[:-1]
    assert list(mx2.columns[1:]) == gold_loci and nl == len(gold_loci)
    mx3 = phc.filter_rows(mx2, args, say)
    gold_report = gzip.open(os.path.join(G, "C2b", "C2b_loci_report.tsv.gz"), "rt").read()
    assert open(str(tmp_path / "x_loci_report.tsv")).read() == gold_report
    z = np.load(os.path.join(G, "C2b", "C2b_dist_hamming.npz"))
    assert mx3.iloc[:, 0].tolist() == list(z["ids"])


def test_loci_list_errors(tmp_path):
    mx = pd.DataFrame({"id": ["a"], "l1": ["1"], "l2": ["2"]})
    p = tmp_path / "loci.txt"
    p.write_text("l1\nl1\n")
    with pytest.raises(SystemExit) as e:
        phc.read_loci_list(str(p), mx)
    assert "duplicated" in str(e.value)
    p.write_text("l9\n")
    with pytest.raises(SystemExit) as e:
        phc.read_loci_list(str(p), mx)
    assert "not present" in str(e.value)


def test_square_to_condensed_checks():
    a = np.array([[0, 1, 2], [1, 0, 3], [2, 3, 0.0]])
    assert phc.square_to_condensed(a).tolist() == [1, 2, 3]
    with pytest.raises(ValueError):
        phc.square_to_condensed(np.array([[0, 1], [2, 0.0]]))
    with pytest.raises(ValueError):
        phc.square_to_condensed(np.array([[1, 1], [1, 0.0]]))


def test_shard_chunk_matches_the_library():
    """Loci per rank of the sharded re-coder: the Python helper and the C entry point agree (no GPU call)."""
    from reportree_b200 import hc_backend, sharded
    L = hc_backend.load_library()
    for loci in (1, 31, 32, 33, 40, 3000, 3002, 7600, 65535, 70001):
        for world in (1, 2, 3, 4, 8):
            c = sharded.shard_chunk(loci, world)
            assert c == L.rthc_shard_chunk(loci, world)
            assert c % 32 == 0 and c * world >= loci and (c - 32) * world < loci + 32 * world
