"""Regenerates tests/golden/* by running the UNMODIFIED reference script
/root/reference/scripts/partitioning_HC.py (pandas + scipy, in this container) with the CPU oracle
binary oracle/_build/cgmlst-dists standing in for the un-vendored `cgmlst-dists` on PATH.

    python tests/golden/make_golden.py

C1: tests/allele_matrix.tsv, --HC-threshold single-1-10           (BASELINE.json configs[0])
C1b: same matrix, single,complete,average,weighted all thresholds + --loci-called 0.9
C2: examples/Listeria/input/Listeria_alleles.tsv, single-1-50     (BASELINE.json configs[1])
C2b: Listeria with --loci-called 0.95 --site-inclusion 0.99, complete-1-30,average-5
C1c: tests/allele_matrix.tsv, ward at all thresholds + complete-3    (ward = scipy nn_chain, float heights)
C2c: Listeria with --loci-called 0.95, ward-1-60 + --pct-HC-threshold ward-0.01
C1d: tests/allele_matrix.tsv with two sample ids duplicated (derived input, written by this script):
     the reference names samples from the re-read TSV header, so duplicates come back as "id.1"
Case names on the command line regenerate only those cases.
Also stores the Listeria golden distance pair shipped with the reference
(examples/Listeria/output/Lm_flt_samples_matrix.tsv -> Lm_dist_hamming.tsv) as compact arrays.
"""
import gzip
import os
import shutil
import subprocess
import sys
import tempfile

import numpy as np
import pandas as pd

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = "/root/reference"
sys.path.insert(0, ROOT)

CASES = {
    "C1": (REF + "/tests/allele_matrix.tsv", ["--HC-threshold", "single-1-10"]),
    "C1b": (REF + "/tests/allele_matrix.tsv", ["--HC-threshold", "single,complete,average,weighted", "--loci-called", "0.9"]),
    "C2": (REF + "/examples/Listeria/input/Listeria_alleles.tsv", ["--HC-threshold", "single-1-50"]),
    "C2b": (REF + "/examples/Listeria/input/Listeria_alleles.tsv",
            ["--HC-threshold", "complete-1-30,average-5", "--pct-HC-threshold", "single-0.004",
             "--loci-called", "0.95", "--site-inclusion", "0.99"]),
    "C1c": (REF + "/tests/allele_matrix.tsv", ["--HC-threshold", "ward,complete-3"]),
    "C2c": (REF + "/examples/Listeria/input/Listeria_alleles.tsv",
            ["--HC-threshold", "ward-1-60", "--pct-HC-threshold", "ward-0.01", "--loci-called", "0.95"]),
}
DERIVED = os.path.join(HERE, "inputs", "allele_matrix_dupids.tsv.gz")
CASES["C1d"] = (DERIVED, ["--HC-threshold", "single-1-10,complete-4"])
KEEP = ["_partitions.tsv", "_clusterComposition.tsv", "_loci_report.tsv", "_loci_used.txt",
        "_parcentage2threshold.tsv"]


def main():
    from oracle import oracle_lib
    oracle_lib.build()
    env = dict(os.environ, PATH=os.path.dirname(oracle_lib.CGMLST_DISTS_BIN) + ":" + os.environ["PATH"])
    only = sys.argv[1:]
    for name, (inp, extra) in CASES.items():
        if only and name not in only:
            continue
        tmp = tempfile.mkdtemp()
        tag = os.path.join(tmp, name)
        if inp == DERIVED:                      # duplicated sample ids: sample3_2 -> sample3, sample9_2 -> sample9
            text = open(REF + "/tests/allele_matrix.tsv").read().replace("sample3_2\t", "sample3\t").replace("sample9_2\t", "sample9\t")
            with gzip.GzipFile(DERIVED, "wb", mtime=0) as fo:
                fo.write(text.encode())
            inp = os.path.join(tmp, "allele_matrix_dupids.tsv")
            with open(inp, "w") as fo:
                fo.write(text)
        cmd = [sys.executable, REF + "/scripts/partitioning_HC.py", "-a", inp, "-o", tag] + extra
        out = subprocess.run(cmd, env=env, capture_output=True, text=True)
        if out.returncode != 0:
            raise SystemExit(out.stdout + out.stderr)
        dst = os.path.join(HERE, name)
        os.makedirs(dst, exist_ok=True)
        for suffix in KEEP:
            if os.path.exists(tag + suffix):
                with open(tag + suffix, "rb") as fi, gzip.GzipFile(os.path.join(dst, name + suffix + ".gz"), "wb", mtime=0) as fo:
                    shutil.copyfileobj(fi, fo)
        for f in os.listdir(tmp):
            if f.endswith("_HC.nwk"):
                with open(os.path.join(tmp, f), "rb") as fi, gzip.GzipFile(os.path.join(dst, f + ".gz"), "wb", mtime=0) as fo:
                    shutil.copyfileobj(fi, fo)
        d = pd.read_table(tag + "_dist_hamming.tsv")
        np.savez_compressed(os.path.join(dst, name + "_dist_hamming.npz"),
                            ids=np.array(d.iloc[:, 0].astype(str).tolist(), dtype="U"), dist=d.iloc[:, 1:].to_numpy().astype(np.uint16))
        # stdout with the argv line removed (it embeds temp paths)
        lines = [l for l in out.stdout.split("\n") if tmp not in l]
        with open(os.path.join(dst, name + "_stdout.txt"), "w") as fh:
            fh.write("\n".join(lines))
        shutil.rmtree(tmp)
        print("golden", name, "ok")
    if only:
        return
    # the reference's own shipped distance golden
    mx = pd.read_table(REF + "/examples/Listeria/output/Lm_flt_samples_matrix.tsv", dtype=str)
    g = pd.read_table(REF + "/examples/Listeria/output/Lm_dist_hamming.tsv")
    codes = mx.iloc[:, 1:].to_numpy()
    ci = np.array([[int(x) if isinstance(x, str) and x.isdigit() else 0 for x in r] for r in codes], dtype=np.int16)
    np.savez_compressed(os.path.join(HERE, "Lm_shipped_pair.npz"), codes=ci,
                        dist=g.iloc[:, 1:].to_numpy().astype(np.uint16), ids=np.array(g.iloc[:, 0].astype(str).tolist(), dtype="U"))
    print("golden Lm_shipped_pair ok")
    # inputs travel with the fixtures: /root/reference does not exist on the GPU box
    os.makedirs(os.path.join(HERE, "inputs"), exist_ok=True)
    for src in (REF + "/tests/allele_matrix.tsv", REF + "/examples/Listeria/input/Listeria_alleles.tsv"):
        with open(src, "rb") as fi, gzip.GzipFile(os.path.join(HERE, "inputs", os.path.basename(src) + ".gz"), "wb", mtime=0) as fo:
            shutil.copyfileobj(fi, fo)


if __name__ == "__main__":
    main()
