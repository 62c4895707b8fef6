"""Single-process multi-device path (seam B0): MultiDeviceContext shards the distance tiles over
several contexts and assembles the full condensed matrix on the first one.  [0, 0] runs the
whole logic (sharded tiles, full-storage layout, peer copies, finish) on one GPU; real device
lists are used when the box has them."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _device_lists():
    import torch
    n = torch.cuda.device_count()
    lists = [[0, 0], [0, 0, 0]]
    if n >= 2:
        lists.append([0, 1])
    if n >= 4:
        lists.append([0, 1, 2, 3])
    if n >= 8:
        lists.append(list(range(8)))
    return lists


@pytest.mark.parametrize("n,L", [(300, 40), (1000, 333), (2177, 64)])
def test_multi_device_matrix_and_linkage_equal_single_context(hc, oracle, n, L):
    from scipy.cluster.hierarchy import linkage
    from reportree_b200 import synth
    X = synth.allele_matrix(n, L, 5, missing=0.05)
    ref = oracle.dist_condensed(X)
    for devices in _device_lists():
        with hc.MultiDeviceContext(devices) as ctx:
            ctx.set_single_algo(hc.SL_PRIM)
            ctx.set_profiles(X)
            ctx.distance()
            assert np.array_equal(ctx.condensed(), ref), devices
            assert np.array_equal(ctx.square_rows(5, 3)[:, :7].astype(np.int64),
                                  oracle.dist_square(X)[5:8, :7]), devices
            for method in ("single", "complete", "average"):
                assert np.array_equal(ctx.linkage(method), linkage(ref.astype(np.float64), method)), (devices, method)
            lab = ctx.fcluster([0.0, 2.0, 7.0])
            assert lab.shape == (3, n)


def test_cli_with_device_list_is_byte_identical(tmp_path, monkeypatch, capsys):
    import gzip
    import shutil
    from reportree_b200 import partitioning_HC as phc
    G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
    inp = str(tmp_path / "Listeria_alleles.tsv")
    with gzip.open(os.path.join(G, "inputs", "Listeria_alleles.tsv.gz"), "rb") as fi, open(inp, "wb") as fo:
        shutil.copyfileobj(fi, fo)
    tag = str(tmp_path / "C2")
    argv = ["-a", inp, "-o", tag, "--HC-threshold", "single-1-50"]
    monkeypatch.setattr(sys, "argv", ["partitioning_HC.py"] + argv)
    monkeypatch.setenv("REPORTREE_HC_DEVICES", ",".join(str(d) for d in _device_lists()[-1]))
    phc.main(argv)
    capsys.readouterr()
    gdir = os.path.join(G, "C2")
    for f in sorted(os.listdir(gdir)):
        if f.endswith(".gz"):
            assert open(tag + f[len("C2"):-3], "rb").read() == gzip.open(os.path.join(gdir, f), "rb").read(), f
