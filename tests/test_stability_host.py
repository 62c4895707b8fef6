"""Host side of the stability drop-in (reportree_b200/comparing_partitions_v2.py) on CPU: the
contingency cells come from the fake backend, the floating-point finish from the library's host
function rthc_adjusted_wallace_sums, and the files must equal those written by the unmodified
reference script (tests/golden/next/, made by tests/golden/make_golden_next.py)."""
import gzip
import os
import shutil

import pytest

from reportree_b200 import comparing_partitions_v2 as cp
from tests.fake_backend import FakeContext

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
NEXT = os.path.join(GOLD, "next")


def run_case(tmp_path, table, extra, ctx):
    src = os.path.join(GOLD, table, table + "_partitions.tsv.gz")
    with gzip.open(src, "rb") as fi, open(tmp_path / "partitions.tsv", "wb") as fo:
        shutil.copyfileobj(fi, fo)
    cwd = os.getcwd()
    os.chdir(tmp_path)
    try:
        cp.main(["-i1", "partitions.tsv", "-t", "out", "-a", "stability", "-n", "5", "-thr", "0.99", "-log", "out.log"] + extra, ctx=ctx)
    finally:
        os.chdir(cwd)
    return (tmp_path / "out_metrics.tsv").read_text(), (tmp_path / "out_stableRegions.tsv").read_text()


def gold(name):
    with open(os.path.join(NEXT, name)) as fh:
        return fh.read()


@pytest.mark.parametrize("case,table,extra", [
    ("C2_keep", "C2", ["-o1", "0", "--keep-redundants"]),
    ("C1b_keep", "C1b", ["-o1", "0", "--keep-redundants"]),
    ("C2_keep_rev", "C2", ["--keep-redundants", "-o1", "1"]),
])
def test_files_equal_reference(tmp_path, capsys, case, table, extra):
    metrics, regions = run_case(tmp_path, table, extra, FakeContext())
    assert metrics == gold("stability_%s_metrics.tsv" % case)
    assert regions == gold("stability_%s_stableRegions.tsv" % case)
    out = capsys.readouterr().out.split("\n")
    ref = gold("stability_%s_stdout.txt" % case).split("\n")
    keep = lambda lines: [l for l in lines if "comparing_partitions_v2.py -i1" not in l and "partitions.tsv -t out" not in l]
    assert keep(out) == keep(ref)


def test_default_mode_keeps_first_sample_of_each_cluster(tmp_path):
    metrics, regions = run_case(tmp_path, "C2", ["-o1", "0"], FakeContext())
    ref = gold("stability_C2_default_metrics.tsv").split("\n")
    mine = metrics.split("\n")
    assert len(mine) == len(ref)
    for a, b in zip(mine[1:-1], ref[1:-1]):
        a, b = a.split("\t"), b.split("\t")
        assert a[:2] == b[:2] and a[9:] == b[9:]
        assert all(abs(float(x) - float(y)) <= 1e-12 for x, y in zip(a[2:9], b[2:9]))
    assert regions == gold("stability_C2_default_stableRegions.tsv")


def test_between_methods_is_refused(tmp_path):
    with pytest.raises(SystemExit) as e:
        cp.main(["-i1", "x", "-i2", "y", "-o1", "0", "-o2", "0", "-t", str(tmp_path / "o"), "-log", str(tmp_path / "l")], ctx=FakeContext())
    assert e.value.code == 2
