"""GPU parity of the drop-in CLI against outputs of the UNMODIFIED reference script
(tests/golden/make_golden.py): byte-identical partitions, cluster composition, Newick,
loci report, loci list, distance matrix and log messages."""
import gzip
import io
import os
import shutil
import sys

import numpy as np
import pandas as pd
import pytest

pytestmark = pytest.mark.gpu

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = {
    "C1": ("allele_matrix.tsv.gz", ["--HC-threshold", "single-1-10"]),
    "C1b": ("allele_matrix.tsv.gz", ["--HC-threshold", "single,complete,average,weighted", "--loci-called", "0.9"]),
    "C2": ("Listeria_alleles.tsv.gz", ["--HC-threshold", "single-1-50"]),
    "C2b": ("Listeria_alleles.tsv.gz", ["--HC-threshold", "complete-1-30,average-5", "--pct-HC-threshold", "single-0.004",
                                         "--loci-called", "0.95", "--site-inclusion", "0.99"]),
    "C1c": ("allele_matrix.tsv.gz", ["--HC-threshold", "ward,complete-3"]),
    "C2c": ("Listeria_alleles.tsv.gz", ["--HC-threshold", "ward-1-60", "--pct-HC-threshold", "ward-0.01", "--loci-called", "0.95"]),
    "C1d": ("allele_matrix_dupids.tsv.gz", ["--HC-threshold", "single-1-10,complete-4"]),     # duplicated sample ids
}


def _gunzip(src, dst):
    with gzip.open(src, "rb") as fi, open(dst, "wb") as fo:
        shutil.copyfileobj(fi, fo)


@pytest.mark.parametrize("case", list(CASES))
def test_cli_outputs_byte_identical_to_reference(case, tmp_path, capsys, monkeypatch):
    from reportree_b200 import partitioning_HC as phc
    inp_gz, extra = CASES[case]
    inp = str(tmp_path / inp_gz[:-3])
    _gunzip(os.path.join(G, "inputs", inp_gz), inp)
    tag = str(tmp_path / case)
    argv = ["-a", inp, "-o", tag] + extra
    monkeypatch.setattr(sys, "argv", ["partitioning_HC.py"] + argv)
    monkeypatch.setenv("REPORTREE_HC_EXACT_NAMES", "1")
    phc.main(argv)
    out = capsys.readouterr().out
    gdir = os.path.join(G, case)
    for f in sorted(os.listdir(gdir)):
        if f.endswith(".gz"):
            produced = tag + f[len(case):-3] if f.startswith(case + "_") else None
            assert produced and os.path.exists(produced), f
            assert open(produced, "rb").read() == gzip.open(os.path.join(gdir, f), "rb").read(), f
    z = np.load(os.path.join(gdir, case + "_dist_hamming.npz"))
    d = pd.read_table(tag + "_dist_hamming.tsv")
    assert d.columns[0] == "dists"
    assert d.iloc[:, 0].astype(str).tolist() == list(z["ids"])
    assert np.array_equal(d.iloc[:, 1:].to_numpy(), z["dist"].astype(np.int64))
    # messages: same lines as the reference, apart from the argv echo (holds temp paths)
    gold = open(os.path.join(gdir, case + "_stdout.txt")).read().split("\n")
    mine = [l for l in out.split("\n") if str(tmp_path) not in l]
    assert mine == gold
    assert open(tag + ".log").read().count("partitioning_HC.py is done!") == 1


def test_cli_boruvka_mode_same_membership(tmp_path, monkeypatch, capsys):
    from reportree_b200 import partitioning_HC as phc
    inp = str(tmp_path / "Listeria_alleles.tsv")
    _gunzip(os.path.join(G, "inputs", "Listeria_alleles.tsv.gz"), inp)
    tag = str(tmp_path / "bor")
    argv = ["-a", inp, "-o", tag, "--HC-threshold", "single-1-50"]
    monkeypatch.setattr(sys, "argv", ["partitioning_HC.py"] + argv)
    monkeypatch.setenv("REPORTREE_HC_EXACT_NAMES", "0")
    monkeypatch.setenv("REPORTREE_HC_SKIP_DIST_TSV", "1")
    phc.main(argv)
    capsys.readouterr()
    assert not os.path.exists(tag + "_dist_hamming.tsv")
    mine = pd.read_table(tag + "_partitions.tsv", dtype=str)
    ref = pd.read_table(gzip.open(os.path.join(G, "C2", "C2_partitions.tsv.gz")), dtype=str)
    assert mine.columns.tolist() == ref.columns.tolist()
    for col in ref.columns[1:]:
        a = pd.factorize(mine[col])[0]
        b = pd.factorize(ref[col])[0]
        assert np.array_equal(a, b), col
        # singletons keep identical names; only cluster_<id> numbering may differ
        s = ref[col].str.startswith("singleton_")
        assert (mine[col][s] == ref[col][s]).all()


def test_cli_distance_matrix_entry(tmp_path, monkeypatch, capsys):
    from reportree_b200 import partitioning_HC as phc
    z = np.load(os.path.join(G, "C1", "C1_dist_hamming.npz"))
    ids = list(z["ids"])
    df = pd.DataFrame(z["dist"].astype(int), columns=ids)
    df.insert(0, "dists", ids)
    dm = str(tmp_path / "dm.tsv")
    df.to_csv(dm, sep="\t", index=False)
    tag = str(tmp_path / "dmx")
    argv = ["-d_mx", dm, "-o", tag, "--HC-threshold", "single-1-10"]
    monkeypatch.setattr(sys, "argv", ["partitioning_HC.py"] + argv)
    phc.main(argv)
    capsys.readouterr()
    assert open(tag + "_partitions.tsv", "rb").read() == gzip.open(os.path.join(G, "C1", "C1_partitions.tsv.gz"), "rb").read()
    assert open(tag + "_single_HC.nwk", "rb").read() == gzip.open(os.path.join(G, "C1", "C1_single_HC.nwk.gz"), "rb").read()
