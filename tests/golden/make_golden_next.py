"""Regenerates tests/golden/next/*  from the UNMODIFIED reference code, run in
this container (pandas + sklearn are installed here; /root/reference does not travel to the GPU box):

    python tests/golden/make_golden_next.py

closest_C2.json      get_closest_samples (reportree.py:591-617), imported from /root/reference/reportree.py,
                     on the C2 golden distance matrix written back as <out>_dist_hamming.tsv
stability_<case>_*   files + stdout of scripts/ComparingPartitions/comparing_partitions_v2.py -a stability on
                     golden partition tables: C2 (single-1-50, 865 samples) with and without
                     --keep-redundants, C1b (four linkages, not nested across methods) with --keep-redundants
"""
import gzip
import importlib.util
import json
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference"
OUT = os.path.join(HERE, "next")

CLOSEST_QUERIES = [(0, 1), (0, 5), (3, 10), (17, 50), (100, 2), (400, 25), (864, 100), (5, 864), (7, 900), (12, 2000)]
STABILITY_CASES = {
    "C2_keep": ("C2/C2_partitions.tsv.gz", ["--keep-redundants"]),
    "C2_default": ("C2/C2_partitions.tsv.gz", []),
    "C1b_keep": ("C1b/C1b_partitions.tsv.gz", ["--keep-redundants"]),
    "C2_keep_rev": ("C2/C2_partitions.tsv.gz", ["--keep-redundants", "-o1", "1"]),
}


def load_reference_module():
    argv, sys.argv = sys.argv, ["reportree.py"]
    try:
        spec = importlib.util.spec_from_file_location("reference_reportree", REF + "/reportree.py")
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        sys.argv = argv
    return mod


def golden_closest():
    ref = load_reference_module()
    z = np.load(os.path.join(HERE, "C2", "C2_dist_hamming.npz"))
    ids, dist = [str(x) for x in z["ids"]], z["dist"]
    tmp = tempfile.mkdtemp()
    path = os.path.join(tmp, "C2_dist_hamming.tsv")
    with open(path, "w") as fh:
        fh.write("dists\t" + "\t".join(ids) + "\n")
        for k, name in enumerate(ids):
            fh.write(name + "\t" + "\t".join(str(int(v)) for v in dist[k]) + "\n")
    res = []
    for pos, n in CLOSEST_QUERIES:
        try:
            names, code = ref.get_closest_samples(ids[pos], n, path, set(ids))
            res.append({"sample": ids[pos], "n": n, "code": code, "names": sorted(str(s) for s in names),
                        "dists": sorted(int(dist[pos, ids.index(str(s))]) for s in names)})
        except IndexError:
            res.append({"sample": ids[pos], "n": n, "code": "IndexError", "names": [], "dists": []})
    names, code = ref.get_closest_samples("not_a_sample", 5, path, set(ids))
    res.append({"sample": "not_a_sample", "n": 5, "code": code, "names": list(names), "dists": []})
    shutil.rmtree(tmp)
    with open(os.path.join(OUT, "closest_C2.json"), "w") as fh:
        json.dump(res, fh, indent=0)
    print("golden closest_C2 ok (%d queries)" % len(res))


def golden_stability():
    script = REF + "/scripts/ComparingPartitions/comparing_partitions_v2.py"
    for case, (table, extra) in STABILITY_CASES.items():
        tmp = tempfile.mkdtemp()
        with gzip.open(os.path.join(HERE, table), "rb") as fi, open(os.path.join(tmp, "partitions.tsv"), "wb") as fo:
            shutil.copyfileobj(fi, fo)
        cmd = [sys.executable, script, "-i1", "partitions.tsv", "-t", "out", "-a", "stability", "-n", "5", "-thr", "0.99",
               "-log", "out.log"]
        if "-o1" not in extra:
            cmd += ["-o1", "0"]
        run = subprocess.run(cmd + extra, cwd=tmp, capture_output=True, text=True, env=dict(os.environ, PYTHONHASHSEED="0"))
        if run.returncode != 0:
            raise SystemExit(run.stdout + run.stderr)
        for suffix in ("_metrics.tsv", "_stableRegions.tsv"):
            shutil.copy(os.path.join(tmp, "out" + suffix), os.path.join(OUT, "stability_%s%s" % (case, suffix)))
        with open(os.path.join(OUT, "stability_%s_stdout.txt" % case), "w") as fh:
            fh.write(run.stdout)
        shutil.rmtree(tmp)
        print("golden stability", case, "ok")


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    golden_closest()
    golden_stability()
