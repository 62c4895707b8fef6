#!/usr/bin/env python3
"""Drop-in for ReporTree's scripts/partitioning_HC.py on B200.

Same command line, same stdout/log messages, same exit behaviour and the same output files
(`_partitions.tsv`, `_clusterComposition.tsv`, `_<method>_HC.nwk`, `_dist_hamming.tsv`,
`_loci_report.tsv`, `_loci_used.txt`, `_clean_missing_matrix.tsv`, `_flt_samples_matrix.tsv`,
`_missing0.tsv`, `_subset_matrix.tsv`, `_flt_dist.tsv`, `_parcentage2threshold.tsv`) as the
reference script (argv at scripts/partitioning_HC.py:302-332, called by
reportree.py:180-225), but the numeric layer underneath -- the `cgmlst-dists` subprocess
loop (:489-516), squareform (:186), scipy linkage (:197) and fcluster (:649 ff.) -- runs on
the GPU through reportree_b200.hc_backend.  There is no CPU fallback for that layer.

Environment switches (argv stays frozen so reportree.py can call this unchanged):
  REPORTREE_HC_DEVICE=<int>          CUDA device (default 0)
  REPORTREE_HC_EXACT_NAMES=0|1       1 (default): scipy-order Prim, cluster_<id> names and Newick
                                     byte-identical to scipy; 0: Boruvka MST (same membership)
  REPORTREE_HC_SKIP_DIST_TSV=1       do not write <out>_dist_hamming.tsv (benchmark runs)
"""
import argparse
import os
import sys
import textwrap

import numpy as np
import pandas

from reportree_b200 import hc_backend
from reportree_b200 import hostio

version = "1.9.0"
last_updated = "2025-10-21"

GPU_METHODS = ("single", "complete", "average", "weighted")


class Tee:
    """Every message goes to stdout and to <out>.log (scripts/partitioning_HC.py:340-348)."""

    def __init__(self, path):
        self.fh = open(path, "a+")

    def __call__(self, *parts):
        print(*parts)
        print(*parts, file=self.fh)

    def close(self):
        self.fh.close()


def build_parser():
    parser = argparse.ArgumentParser(
        prog="partitioning_HC.py", formatter_class=argparse.RawDescriptionHelpFormatter,
        description=textwrap.dedent("""\
            partitioning_HC.py (B200 backend): genetic clusters at any distance threshold(s) of an
            allele/SNP profile or pairwise distance matrix using hierarchical clustering.
            Profile cells may be integers or IUPAC letters; alleles starting with "*" or "INF-"
            count as new alleles; every other value is missing data."""))
    g = parser.add_argument_group("Partitioning with Hierarchical Clustering",
                                  "Specifications to get partitions with HC methods")
    g.add_argument("-d_mx", "--distance_matrix", dest="distance_matrix", required=False, type=str,
                   help="[OPTIONAL] Input pairwise distance matrix")
    g.add_argument("-a", "--allele-profile", dest="allele_profile", required=False, type=str,
                   help="[OPTIONAL] Input allele/SNP profile matrix")
    g.add_argument("-l", "--loci", dest="loci", required=False, type=str, default="none",
                   help="[OPTIONAL] List of loci that must be used (complemented by --site-inclusion if > 0)")
    g.add_argument("-o", "--output", dest="out", required=True, type=str, help="[MANDATORY] Tag for output file name")
    g.add_argument("--HC-threshold", dest="method_threshold", required=False, default="single",
                   help="[OPTIONAL] Comma-separated HC methods and thresholds: method, method-T or method-A-B "
                        "(inclusive). Default: single")
    g.add_argument("--pct-HC-threshold", dest="pct_HCmethod_threshold", required=False, default="none",
                   help="[OPTIONAL] Like --HC-threshold with the threshold given as a proportion of the loci in use")
    g.add_argument("--site-inclusion", dest="samples_called", required=False, default=0.0,
                   help="[OPTIONAL] Minimum proportion of samples per locus without missing data [default: 0.0]")
    g.add_argument("--loci-called", dest="loci_called", required=False, default=0.0,
                   help="[OPTIONAL] Minimum proportion of loci called per sample [default: 0.0]")
    g.add_argument("--missing-code", dest="missing_code", required=False, type=str, default="0",
                   help="[OPTIONAL] Code representing missing data [default: 0]")
    g.add_argument("-m", "--metadata", dest="metadata", required=False, default="", type=str,
                   help="[OPTIONAL] Metadata .tsv used with --filter to select samples")
    g.add_argument("-f", "--filter", dest="filter_column", required=False, default="",
                   help="[OPTIONAL] Metadata filter, 'column< >operation< >condition' separated by ';'")
    g.add_argument("-d", "--dist", dest="dist", required=False, default=1.0, type=float,
                   help="Distance unit by which partition thresholds are multiplied")
    return parser


# --------------------------------------------------------------------------------------------
# profile cleaning and filters (host side; string semantics of scripts/partitioning_HC.py:362-477)
# --------------------------------------------------------------------------------------------

def load_profile(path, missing_code, out, say):
    """Read + clean codes (:362-373).  Returns the string DataFrame."""
    mx = pandas.read_table(path, dtype=str)
    mx = mx.replace("INF-", "", regex=True)
    mx = mx.replace(r"\*", "", regex=True)
    if missing_code != "0":
        ids = mx[mx.columns[0]]
        if missing_code == "empty":
            mx = mx.fillna("0")
        else:
            mx = mx.replace({missing_code: "0"})
        mx[mx.columns[0]] = ids
        mx.to_csv(out + "_missing0.tsv", index=False, header=True, sep="\t")
    return mx


def read_loci_list(path, mx):
    """--loci file (:259-276): names split on '\\n' only; unknown or duplicated names abort."""
    wanted = []
    seen = set()
    columns = set(mx.columns)
    with open(path) as fh:
        for line in fh.readlines():
            name = line.split("\n")[0]
            if name in seen:
                sys.exit("Locus " + str(name) + " is duplicated in the loci list. Please revise your list before we proceed!!")
            if name not in columns:
                sys.exit("Locus " + str(name) + " is not present in the provided allele matrix. Please revise your list before we proceed!!")
            seen.add(name)
            wanted.append(name)
    return wanted


def called_fraction_per_locus(frame):
    """(col != "0").sum() / n in float64, exactly as :414-415 / :428-429 (NaN counts as called)."""
    vals = frame.to_numpy(dtype=object)
    called = (vals != "0").sum(axis=0)
    return called / len(frame)


def filter_columns(mx, args, say):
    """--loci / --site-inclusion (:400-443)."""
    first = mx.columns[0]
    n_before = len(mx.columns) - 1
    site = float(args.samples_called)
    n_after = n_before
    if args.loci != "none":
        wanted = set(read_loci_list(args.loci, mx))
        if site == 0.0:
            say("Keeping the sites/loci present in the loci list.")
            keep = [first] + [c for c in mx.columns[1:] if c in wanted]
        else:
            say("Keeping the sites/loci present in the loci list and those with information in >= " + str(site * 100) + "%% of the samples...")
            others = [c for c in mx.columns[1:] if c not in wanted]
            frac = called_fraction_per_locus(mx[others]) if others else np.zeros(0)
            extra = {c for c, f in zip(others, frac) if f >= site}
            keep = [c for c in mx.columns if c == first or c in extra or c in wanted]
        mx = mx[keep]
        n_after = len(mx.columns) - 1
        say("\tFrom the " + str(n_before) + " loci/positions, " + str(n_after) + " were kept in the matrix.")
    elif site > 0.0:
        say("Keeping the sites/loci with information in >= " + str(site * 100) + "%% of the samples...")
        cols = list(mx.columns[1:])
        frac = called_fraction_per_locus(mx[cols])
        keep = [first] + [c for c, f in zip(cols, frac) if f >= site]
        mx = mx[keep]
        n_after = len(mx.columns) - 1
        say("\tFrom the " + str(n_before) + " loci/positions, " + str(n_after) + " were kept in the matrix.")
    return mx, n_after


def filter_rows(mx, args, say):
    """--loci-called (:447-477).  Returns the kept DataFrame or None when <= 1 sample passes."""
    thr = float(args.loci_called)
    if thr < 0.0:
        return mx
    say("Cleaning the profile matrix using a threshold of >" + str(args.loci_called) + " alleles/positions called per sample...")
    len_schema = len(mx.columns) - 1
    vals = mx.to_numpy(dtype=object)
    missing = (vals == "0").sum(axis=1)            # all columns incl. the id column, like isin(["0"])
    called = len_schema - missing
    report = pandas.DataFrame({"samples": mx[mx.columns[0]], "missing": missing, "called": called,
                               "pct_called": called / len_schema})
    passing = report[report["pct_called"] >= thr]["samples"].values.tolist()
    report.to_csv(args.out + "_loci_report.tsv", index=False, header=True, sep="\t")
    say("\tFrom the " + str(len(mx)) + " samples, " + str(len(passing)) + " were kept in the profile matrix.")
    if len(passing) <= 1:
        say("\nCannot proceed because " + str(len(passing)) + " samples were kept in the matrix!")
        return None
    mx = mx[mx[mx.columns[0]].isin(passing)]
    mx.to_csv(args.out + "_flt_samples_matrix.tsv", index=False, header=True, sep="\t")
    return mx
