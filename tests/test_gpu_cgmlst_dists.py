"""Seam B2: reportree_b200/bin/cgmlst-dists (GPU) prints the same square TSV as the CPU oracle's
stand-in for the external program (oracle/_build/cgmlst-dists), byte for byte, incl. odd tokens,
blank lines and the 2,000-column chunk protocol of scripts/partitioning_HC.py:489-516."""
import os
import subprocess

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GPU_BIN = os.path.join(ROOT, "reportree_b200", "bin", "cgmlst-dists")


def _write(path, n, L, seed, odd=False):
    rng = np.random.default_rng(seed)
    X = rng.integers(0, 6, size=(n, L))
    toks = X.astype("U6")
    if odd:
        toks[rng.random((n, L)) < 0.05] = "LNF"
        toks[rng.random((n, L)) < 0.02] = ""
        toks[rng.random((n, L)) < 0.02] = "-3"
        toks[rng.random((n, L)) < 0.02] = "007"
    with open(path, "w") as fh:
        fh.write("FILE\t" + "\t".join("l%d" % k for k in range(L)) + "\n")
        for i in range(n):
            fh.write("id_%d\t" % i + "\t".join(toks[i].tolist()) + "\n")
            if odd and i == 3:
                fh.write("\n")
    return X


@pytest.mark.parametrize("n,L,odd", [(2, 3, False), (37, 50, True), (300, 2300, False), (1, 5, False)])
def test_binary_output_equals_oracle_binary(n, L, odd, tmp_path, oracle):
    assert os.path.exists(GPU_BIN), "run python -m reportree_b200.build"
    inp = str(tmp_path / "m.tsv")
    _write(inp, n, L, 4, odd)
    ours = subprocess.run([GPU_BIN, inp], capture_output=True)
    ref = subprocess.run([oracle.CGMLST_DISTS_BIN, inp], capture_output=True)
    assert ours.returncode == 0, ours.stderr.decode()
    assert ref.returncode == 0
    assert ours.stdout == ref.stdout


def test_chunked_calls_sum_to_single_call(tmp_path, oracle):
    """The reference calls the program per 2,000-column chunk and adds the matrices."""
    import pandas as pd
    n, L = 120, 4100
    inp = str(tmp_path / "m.tsv")
    X = _write(inp, n, L, 9)
    mx = pd.read_table(inp, dtype=str)
    total = None
    for c0 in range(1, L + 1, 2000):
        sub = pd.concat([mx.iloc[:, [0]], mx.iloc[:, c0:c0 + 2000]], axis=1)
        p = str(tmp_path / ("chunk%d.tsv" % c0))
        sub.to_csv(p, sep="\t", index=False)
        out = subprocess.run([GPU_BIN, p], capture_output=True, text=True, check=True).stdout
        rows = [l.split("\t") for l in out.strip().split("\n")]
        d = np.asarray([[int(v) for v in r[1:]] for r in rows[1:]])
        total = d if total is None else total + d
    assert np.array_equal(total, oracle.dist_square(X))
