"""CPU tests of the drop-in CLI's host side through a fake numeric backend (oracle + scipy):
 (1) the C++ text I/O path (librthc_hostio.so) writes byte-identical files and messages to the
     pandas path on regular inputs, incl. INF-/'*' cleaning, NA tokens, letters, --missing-code,
     --loci / --site-inclusion / --loci-called filters, duplicated sample ids;
 (2) irregular inputs (quotes, CRLF, ragged rows) fall back to the pandas path;
 (3) both paths reproduce the golden outputs of the unmodified reference script."""
import gzip
import os
import shutil
import sys

import numpy as np
import pytest

G = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _run(monkeypatch, capsys, argv, fast):
    from reportree_b200 import hc_backend, partitioning_HC as phc
    from tests.fake_backend import FakeContext
    monkeypatch.setattr(hc_backend, "HCContext", FakeContext)
    monkeypatch.setenv("REPORTREE_HC_FASTIO", "1" if fast else "0")
    monkeypatch.setattr(sys, "argv", ["partitioning_HC.py"] + argv)
    try:
        phc.main(argv)
    except SystemExit:
        pass
    return capsys.readouterr().out


def _files(tag):
    d, base = os.path.dirname(tag), os.path.basename(tag)
    return {f[len(base):]: open(os.path.join(d, f), "rb").read() for f in sorted(os.listdir(d))
            if f.startswith(base + "_") or f == base + ".log"}


def _messy_matrix(path, n=60, L=40, seed=3, dup_ids=False):
    rng = np.random.default_rng(seed)
    toks = ["0", "1", "2", "3", "17", "INF-5", "*8", "INF-*12", "LNF", "NA", "", "nan", "a", "C", "T", "00", "007", "-", "N/A",
            "ININF-F-3", "4*", "PLOT5", "123456789", "1234567890", "NIPH"]
    p = np.array([8, 20, 20, 15, 10, 3, 3, 1, 2, 1, 1, 1, 1, 1, 1, 1, 1, 2, 1, 1, 1, 1, 1, 1, 1], dtype=float)
    cells = rng.choice(len(toks), size=(n, L), p=p / p.sum())
    with open(path, "w") as fh:
        fh.write("FILE\t" + "\t".join("loc%d" % k for k in range(L)) + "\n")
        for i in range(n):
            name = "s%03d" % (i if not (dup_ids and i % 17 == 5) else 3)
            if i == 7:
                name = "INF-x*7"
            if i == 11:
                name = "0"
            fh.write(name + "\t" + "\t".join(toks[c] for c in cells[i]) + "\n")
            if i == 20:
                fh.write("\n")                       # blank line: skipped by both parsers
    return path


CASES = [
    ["--HC-threshold", "single-1-6,complete-3"],
    ["--HC-threshold", "single-2", "--loci-called", "0.6", "--site-inclusion", "0.7"],
    ["--HC-threshold", "average-1-4", "--missing-code", "-", "--loci-called", "0.5"],
    ["--HC-threshold", "single-3", "--missing-code", "empty", "--site-inclusion", "0.8"],
    ["--HC-threshold", "single-1,single-1", "--pct-HC-threshold", "single-0.1"],
]


@pytest.mark.parametrize("case", range(len(CASES)))
def test_fast_text_path_equals_pandas_path(case, tmp_path, monkeypatch, capsys):
    inp = _messy_matrix(str(tmp_path / "in.tsv"), dup_ids=(case == 1))
    outs = {}
    for fast in (False, True):
        d = tmp_path / ("fast" if fast else "slow")
        d.mkdir()
        argv = ["-a", inp, "-o", str(d / "run")] + CASES[case]
        msgs = _run(monkeypatch, capsys, argv, fast)
        outs[fast] = (_files(str(d / "run")), [l for l in msgs.split("\n") if str(tmp_path) not in l])
    files_slow, msg_slow = outs[False]
    files_fast, msg_fast = outs[True]
    assert msg_fast == msg_slow
    assert sorted(files_fast) == sorted(files_slow)
    for k in files_slow:
        if k == ".log":
            continue
        assert files_fast[k] == files_slow[k], k
    assert "_partitions.tsv" in files_fast and "_dist_hamming.tsv" in files_fast


def test_loci_list_and_metadata_free_paths(tmp_path, monkeypatch, capsys):
    inp = _messy_matrix(str(tmp_path / "in.tsv"), seed=5)
    loci = tmp_path / "loci.txt"
    loci.write_text("".join("loc%d\n" % k for k in range(0, 40, 3)))
    outs = {}
    for fast in (False, True):
        d = tmp_path / ("f" if fast else "s")
        d.mkdir()
        argv = ["-a", inp, "-o", str(d / "run"), "--HC-threshold", "single-1-3", "-l", str(loci), "--site-inclusion", "0.9"]
        msgs = _run(monkeypatch, capsys, argv, fast)
        outs[fast] = (_files(str(d / "run")), [l for l in msgs.split("\n") if str(tmp_path) not in l])
    assert outs[True][1] == outs[False][1]
    for k in outs[False][0]:
        if k != ".log":
            assert outs[True][0][k] == outs[False][0][k], k


@pytest.mark.parametrize("kind", ["quote", "crlf", "ragged", "dup_header"])
def test_irregular_text_takes_the_pandas_path(kind, tmp_path):
    from reportree_b200 import fastio
    p = tmp_path / "x.tsv"
    rows = ["FILE\tA\tB", "s1\t1\t2", "s2\t1\t3"]
    if kind == "quote":
        rows[1] = 's1\t"1"\t2'
    if kind == "ragged":
        rows[2] = "s2\t1"
    if kind == "dup_header":
        rows[0] = "FILE\tA\tA"
    p.write_bytes((("\r\n" if kind == "crlf" else "\n").join(rows) + "\n").encode())
    assert fastio.open_profile(str(p)) is None


def test_hostio_library_exports_every_declared_symbol():
    import re
    from reportree_b200 import fastio
    lib = fastio.load_library()
    txt = open(os.path.join(os.path.dirname(G), "..", "include", "rthio.h")).read()
    declared = sorted(set(re.findall(r"\b(rthio_[a-z0-9_]+)\s*\(", txt)))
    assert declared == sorted(fastio.EXPORTED_SYMBOLS)
    for name in declared:
        getattr(lib, name)
    assert lib.rthio_version() >= 100


GOLD = {
    "C1": ("allele_matrix.tsv.gz", ["--HC-threshold", "single-1-10"]),
    "C1b": ("allele_matrix.tsv.gz", ["--HC-threshold", "single,complete,average,weighted", "--loci-called", "0.9"]),
    "C2b": ("Listeria_alleles.tsv.gz", ["--HC-threshold", "complete-1-30,average-5", "--pct-HC-threshold", "single-0.004",
                                         "--loci-called", "0.95", "--site-inclusion", "0.99"]),
    "C1c": ("allele_matrix.tsv.gz", ["--HC-threshold", "ward,complete-3"]),
    "C2c": ("Listeria_alleles.tsv.gz", ["--HC-threshold", "ward-1-60", "--pct-HC-threshold", "ward-0.01", "--loci-called", "0.95"]),
    "C1d": ("allele_matrix_dupids.tsv.gz", ["--HC-threshold", "single-1-10,complete-4"]),     # duplicated sample ids
}


@pytest.mark.parametrize("case", list(GOLD))
@pytest.mark.parametrize("fast", [False, True])
def test_cli_host_side_reproduces_reference_goldens(case, fast, tmp_path, monkeypatch, capsys):
    inp_gz, extra = GOLD[case]
    inp = str(tmp_path / inp_gz[:-3])
    with gzip.open(os.path.join(G, "inputs", inp_gz), "rb") as fi, open(inp, "wb") as fo:
        shutil.copyfileobj(fi, fo)
    tag = str(tmp_path / case)
    out = _run(monkeypatch, capsys, ["-a", inp, "-o", tag] + extra, fast)
    gdir = os.path.join(G, case)
    for f in sorted(os.listdir(gdir)):
        if f.endswith(".gz"):
            produced = tag + f[len(case):-3]
            assert os.path.exists(produced), f
            assert open(produced, "rb").read() == gzip.open(os.path.join(gdir, f), "rb").read(), f
    gold = open(os.path.join(gdir, case + "_stdout.txt")).read().split("\n")
    assert [l for l in out.split("\n") if str(tmp_path) not in l] == gold


@pytest.mark.parametrize("trailing_newline", [True, False])
def test_vector_cell_parser_equals_scalar_parser(tmp_path, monkeypatch, trailing_newline):
    """The AVX2 tab scan + 8-digit SWAR parse gives the codes of the byte-by-byte loop, token by token."""
    from reportree_b200 import fastio, hostio
    import pandas as pd
    rng = np.random.default_rng(12)
    toks = ["0", "1", "9", "10", "437", "65535", "12345678", "99999999", "00000000", "123456789", "00", "01", "1:", ":1", "3;",
            "<2", "/7", "1/", "12a", "a", "Z", "LNF", "", "NA", "*5", "INF-44", "4 ", " 4", "-1", "+1", "1.0", "1e3", "１"]
    n, L = 300, 97                                   # 97 cells: rows end at every offset of the 32-byte blocks
    cells = rng.choice(len(toks), size=(n, L), p=np.r_[[0.3, 0.2], np.full(len(toks) - 2, 0.5 / (len(toks) - 2))])
    path = str(tmp_path / "m.tsv")
    with open(path, "w", encoding="utf-8") as fh:
        fh.write("F\t" + "\t".join("l%d" % k for k in range(L)) + "\n")
        for i in range(n):
            fh.write("s%d\t" % i + "\t".join(toks[c] for c in cells[i]) + ("\n" if trailing_newline or i < n - 1 else ""))
    monkeypatch.delenv("RTHIO_NO_SIMD", raising=False)
    a = fastio.open_profile(path, "0")
    fast = np.array(a.codes, copy=True)
    a.close()
    monkeypatch.setenv("RTHIO_NO_SIMD", "1")
    b = fastio.open_profile(path, "0")
    slow = np.array(b.codes, copy=True)
    b.close()
    assert np.array_equal(fast, slow)
    # and both follow the specification: pandas strings -> hostio.encode_cells, with -1 folded to 0 (missing)
    mx = pd.read_table(path, dtype=str)
    mx = mx.replace({"INF-": "", r"\*": ""}, regex=True)
    spec = hostio.encode_cells(mx.iloc[:, 1:])
    assert np.array_equal(np.maximum(fast, 0), spec)


@pytest.mark.parametrize("use_map", ["1", "0"])
def test_square_writer_blocks_and_modes(tmp_path, monkeypatch, use_map):
    """Row blocks through the shared mapping (default) and through write(2) give the same bytes as a plain join."""
    from reportree_b200 import fastio
    monkeypatch.setenv("RTHIO_SQUARE_MMAP", use_map)
    rng = np.random.default_rng(4)
    n = 1231
    D = rng.integers(0, 65536, (n, n)).astype(np.uint16)
    D[rng.random((n, n)) < 0.5] = 0
    names = ["sample_%d" % (k * 7) for k in range(n)]
    calls = []

    def rows(r0, nr):
        calls.append((r0, nr))
        return D[r0:r0 + nr]
    path = tmp_path / "sq.tsv"
    fastio.write_square_tsv(str(path), names, rows, n, block=100)
    want = "dists\t" + "\t".join(names) + "\n" + "".join(
        names[i] + "\t" + "\t".join(str(int(v)) for v in D[i]) + "\n" for i in range(n))
    assert path.read_text() == want
    assert calls == [(r0, min(100, n - r0)) for r0 in range(0, n, 100)]
    # short rows of one-digit cells, many threads: the 8-byte cell stores must never reach into a neighbour's row
    for n2 in (2, 3, 5, 6, 40):
        Z = (rng.integers(0, 12, (n2, n2)) * (rng.random((n2, n2)) < 0.3)).astype(np.uint16)
        want2 = "dists\t" + "\t".join("x" for _ in range(n2)) + "\n" + "".join(
            "x\t" + "\t".join(str(int(v)) for v in Z[i]) + "\n" for i in range(n2))
        for _ in range(25):
            fastio.write_square_tsv(str(path), ["x"] * n2, lambda r0, nr: Z[r0:r0 + nr], n2, threads=16)
            assert path.read_text() == want2
    fastio.write_square_tsv(str(path), names[:1], lambda r0, nr: D[:1, :1], 1)
    assert path.read_text() == "dists\t" + names[0] + "\n" + names[0] + "\t" + str(int(D[0, 0])) + "\n"
