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
  REPORTREE_HC_DEVICES=0,1,..        several GPUs from this ONE process: distance tiles sharded, shards
                                     copied device-to-device into the first GPU, which does the rest
  REPORTREE_HC_EXACT_NAMES=0|1       1 (default): scipy-order Prim, cluster_<id> names and Newick
                                     byte-identical to scipy; 0: Boruvka MST (same membership)
  REPORTREE_HC_SKIP_DIST_TSV=1       do not write <out>_dist_hamming.tsv (benchmark runs)
"""
import os
import sys

if __name__ == "__main__" and os.environ.get("REPORTREE_HC_PREWARM", "1") != "0":   # CUDA start-up overlaps everything below, imports included
    from reportree_b200 import early_cuda
    early_cuda.start(int(os.environ.get("REPORTREE_HC_DEVICES", os.environ.get("REPORTREE_HC_DEVICE", "0")).split(",")[0] or 0))

import argparse
import textwrap

import numpy as np
from reportree_b200.lazy_pandas import pandas

from reportree_b200 import hc_backend
from reportree_b200 import hostio

version = "1.9.0"
last_updated = "2025-10-21"

GPU_METHODS = ("single", "complete", "average", "weighted", "ward")


class Phases:
    """REPORTREE_HC_TIMING=1: wall time of every phase of the run, printed to stderr at the end
    (stdout and <out>.log stay byte-identical to the reference's messages)."""

    def __init__(self):
        import time
        self.clock = time.perf_counter
        self.on = os.environ.get("REPORTREE_HC_TIMING", "0") == "1"
        self.t0 = self.last = self.clock()
        self.rows = []

    def mark(self, name):
        now = self.clock()
        self.rows.append((name, now - self.last))
        self.last = now

    def report(self):
        if self.on:
            total = self.clock() - self.t0
            for name, dt in self.rows:
                sys.stderr.write("[hc-timing] %-28s %9.3f s\n" % (name, dt))
            sys.stderr.write("[hc-timing] %-28s %9.3f s\n" % ("TOTAL (main)", total))


PH = Phases()


class Background:
    """Run fn(*args) on a helper thread (the C++ writers and the CUDA context set-up release the GIL);
    join() re-raises whatever it raised."""

    def __init__(self, fn, *args):
        import threading
        self.exc = None
        self.result = None

        def run():
            try:
                self.result = fn(*args)
            except BaseException as e:          # noqa: BLE001 -- re-raised in join()
                self.exc = e
        self.thread = threading.Thread(target=run, daemon=True)
        self.thread.start()

    def join(self):
        self.thread.join()
        if self.exc is not None:
            raise self.exc
        return self.result


class LazyContext:
    """The GPU context is created on a helper thread while the profile is being read (CUDA
    start-up is ~1.3 s); the first use waits for it.  A failure surfaces at that first use."""

    def __init__(self, device, devices=None, single_algo=None):
        def create():
            ctx = hc_backend.MultiDeviceContext(devices) if devices and len(devices) > 1 else hc_backend.HCContext(device=device)
            if single_algo is not None:          # set here, not by the caller: touching the proxy would join the thread
                ctx.set_single_algo(single_algo)
            return ctx
        self._bg = Background(create)
        self._ctx = None

    def __getattr__(self, name):
        if self._ctx is None:
            self._ctx = self._bg.join()
        return getattr(self._ctx, name)


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


# --------------------------------------------------------------------------------------------
# threshold planning (scripts/partitioning_HC.py:575-615) -- pure host logic, unit-tested on CPU
# --------------------------------------------------------------------------------------------

def plan_runs(method_threshold, pct_method_threshold, total_size, say):
    """Returns (runs: {method: [(threshold_str, kind)]}, pct_correspondence: {thr: [pct_str]})."""
    runs, pct_corr = {}, {}
    if pct_method_threshold != "none":
        say("\n\tCorrespondence between percentage and number of differences:")
        say("\n\t#METHOD\tPERCENTAGE\tDIFFERENCES")
        for item in pct_method_threshold.split(","):
            method = item.split("-")[0]
            pct = item.split("-", 1)[1]
            thr = str(int(int(total_size) * float(pct)))
            runs.setdefault(method, [])
            if (thr, "pct") not in runs[method]:
                runs[method].append((thr, "pct"))
            pct_corr.setdefault(thr, [])
            if str(pct) not in pct_corr[thr]:
                pct_corr[thr].append(str(pct))
                say("\t" + str(float(pct) * 100) + "\t" + thr)
    for item in method_threshold.split(","):
        if "-" not in item:
            runs.setdefault(item, []).append(("all", "general"))
        else:
            method = item.split("-")[0]
            runs.setdefault(method, []).append((str(item.split("-", 1)[1]), "general"))
    return runs, pct_corr


def expand_cuts(method, requests, pct_corr, max_dist, dist_unit, say):
    """One linkage method -> ordered list of (partition column name, cutoff) following
    :641-772 literally (range clipping, skip messages, pct cutoffs are not scaled by -d)."""
    cuts = []
    for threshold, kind in requests:
        if threshold == "all":
            say("\tCalculating clustering in range", str(0), str(max_dist), "with a distance of", str(dist_unit))
            for thr in range(0, int(max_dist) + 1):
                cuts.append((method + "-" + str(thr) + "x" + str(dist_unit), int(thr) * dist_unit))
        elif "-" in threshold:
            lo = int(threshold.split("-")[0])
            hi = int(threshold.split("-")[1]) + 1
            if lo < max_dist and hi > max_dist:
                hi = int(max_dist) + 1
            elif lo > max_dist:
                say("\tThe requested partition, " + str(lo) + ", is higher than the maximum distance. Clustering will not be computed.")
                continue
            say("\tCalculating clustering in range", str(lo), str(hi), "with a distance of", str(dist_unit))
            for thr in range(lo, hi):
                cuts.append((method + "-" + str(thr) + "x" + str(dist_unit), thr * dist_unit))
        elif kind == "general":
            if int(threshold) > max_dist:
                say("\tThe requested partition, " + str(threshold) + ", is higher than the maximum distance. Clustering will not be computed.")
                continue
            say("\tCalculating clustering for threshold", str(threshold), "with a distance of", str(dist_unit))
            cuts.append((method + "-" + str(threshold) + "x" + str(dist_unit), int(threshold) * dist_unit))
        elif kind == "pct":
            for pct in pct_corr[threshold]:
                say("\tCalculating clustering for threshold " + method + "-" + str(threshold) + ", which corresponds to the pct threshold of: " + str(pct))
                cuts.append((method + "-" + str(pct), int(threshold)))
    return cuts


def read_distance_matrix(path):
    """-d_mx entry (:522-525): square TSV with ids in the first column."""
    return pandas.read_table(path)


def square_to_condensed(square):
    """scipy.spatial.distance.squareform(checks=True) semantics for an n x n array (:186)."""
    a = np.ascontiguousarray(square, dtype=np.float64)
    if a.ndim != 2 or a.shape[0] != a.shape[1]:
        raise ValueError("The matrix argument must be square.")
    if not (a == a.T).all():                     # is_valid_dm(tol=0): exact symmetry, as scipy checks it
        raise ValueError("Distance matrix 'X' must be symmetric.")
    if not (np.diag(a) == 0).all():
        raise ValueError("Distance matrix 'X' diagonal must be zero.")
    iu = np.triu_indices(a.shape[0], 1)
    return a[iu]


# --------------------------------------------------------------------------------------------
# the pipeline
# --------------------------------------------------------------------------------------------

def _exit(log, code=None):
    log.close()
    sys.exit(code)


def profile_to_distances(args, say, log, ctx):
    """`-a` branch: load, clean, filter, encode, distance on the GPU, artefact files.
    Returns (sample names, total_size)."""
    say("Profile matrix provided... pairwise distance will be calculated!")
    from reportree_b200 import fastio
    if fastio.enabled():
        prof = fastio.open_profile(args.allele_profile, str(args.missing_code))
        if prof is not None:                      # None: irregular text, pandas decides what it means
            return profile_to_distances_fast(args, say, log, ctx, prof)
    mx = load_profile(args.allele_profile, str(args.missing_code), args.out, say)
    PH.mark("load + clean profile")

    if args.metadata != "" and args.filter_column != "":
        from reportree_b200 import metadata_filter
        say("Filtering the allele matrix...")
        meta = pandas.read_table(args.metadata, dtype=str)
        meta.columns = [c.replace(" ", "_") for c in meta.columns]
        meta = meta.replace(" ", "_", regex=True)
        n0 = len(mx)
        say("\tFiltering metadata for the following parameters: " + " & ".join(args.filter_column.split(";")))
        keep_ids = metadata_filter.select_samples(meta, args.filter_column)
        mx = mx[mx[mx.columns[0]].isin(keep_ids)]
        say("\tFrom the " + str(n0) + " samples, " + str(len(mx)) + " were kept in the matrix...")
        if len(mx) <= 1:
            say("\nCannot proceed because " + str(len(mx)) + " samples were kept in the matrix!")
            _exit(log)
        mx.to_csv(args.out + "_subset_matrix.tsv", sep="\t", index=None)

    mx, n_loci = filter_columns(mx, args, say)
    if n_loci <= 1:
        say("\nCannot proceed because " + str(n_loci) + " sites/loci were kept in the matrix!")
        _exit(log)
    with open(args.out + "_loci_used.txt", "w+") as fh:
        for locus in mx.columns[1:]:
            print(locus, file=fh)
    PH.mark("column filter")
    mx.to_csv(args.out + "_clean_missing_matrix.tsv", index=False, header=True, sep="\t")
    PH.mark("write clean_missing_matrix")

    mx = filter_rows(mx, args, say)
    if mx is None:
        _exit(log)
    PH.mark("row filter + flt_samples")

    say("Getting the pairwise distance matrix with cgmlst-dists (if your profile matrix is too big, this will be done in chunks of 2000 alleles/positions)...")
    total_size = len(mx.columns) - 1
    # sample ids travel through TSV text in the reference; NaN ids would print as empty strings
    names = ["" if not isinstance(s, str) else s for s in mx[mx.columns[0]].tolist()]
    codes = hostio.encode_cells(mx.iloc[:, 1:])
    PH.mark("encode cells")
    ctx.set_profiles(codes)
    ctx.distance()
    PH.mark("GPU recode + distance")
    if os.environ.get("REPORTREE_HC_SKIP_DIST_TSV", "0") != "1":
        write_dist_tsv(args.out + "_dist_hamming.tsv", names, ctx)
        PH.mark("write dist_hamming.tsv")
    return names_as_read_back(names), total_size


def _loci_selection_fast(prof, rows, args, say):
    """filter_columns() on a fastio.Profile: returns the kept locus indices (ascending)."""
    loci = prof.header[1:]
    n_before = len(loci)
    site = float(args.samples_called)
    keep = np.arange(n_before)

    def called_fraction(cols):
        if len(cols) == 0:
            return np.zeros(0)
        sub = prof.codes[np.ix_(rows, cols)] if len(rows) != prof.n else prof.codes[:, cols]
        return (sub != 0).sum(axis=0) / len(rows)           # (col != "0").sum() / n, float64 (:414-415)

    if args.loci != "none":
        wanted = set(read_loci_list(args.loci, _Columns(prof.header)))
        in_list = np.fromiter((c in wanted for c in loci), dtype=bool, count=n_before)
        if site == 0.0:
            say("Keeping the sites/loci present in the loci list.")
            keep = np.flatnonzero(in_list)
        else:
            say("Keeping the sites/loci present in the loci list and those with information in >= " + str(site * 100) + "%% of the samples...")
            others = np.flatnonzero(~in_list)
            extra = np.zeros(n_before, dtype=bool)
            extra[others[called_fraction(others) >= site]] = True
            keep = np.flatnonzero(in_list | extra)
        say("\tFrom the " + str(n_before) + " loci/positions, " + str(len(keep)) + " were kept in the matrix.")
    elif site > 0.0:
        say("Keeping the sites/loci with information in >= " + str(site * 100) + "%% of the samples...")
        keep = np.flatnonzero(called_fraction(np.arange(n_before)) >= site)
        say("\tFrom the " + str(n_before) + " loci/positions, " + str(len(keep)) + " were kept in the matrix.")
    return keep


class _Columns:
    """Just enough of a DataFrame for read_loci_list(): `.columns`."""

    def __init__(self, header):
        self.columns = header


def _selected_codes(prof, rows, cols):
    """int32 codes of the surviving samples x loci; the parsed matrix itself when nothing was dropped."""
    if len(rows) == prof.n and len(cols) == prof.n_loci:       # both are ascending and distinct: the identity
        return prof.codes
    return prof.gather(rows, cols)


def profile_to_distances_fast(args, say, log, ctx, prof):
    """`-a` branch on a profile parsed by librthc_hostio.so: same filters, messages and artefact
    files as profile_to_distances(), without a string DataFrame."""
    out = args.out
    all_cols = np.arange(prof.n_loci)
    rows = np.arange(prof.n)
    if str(args.missing_code) != "0":
        prof.write(out + "_missing0.tsv", rows, all_cols)
    PH.mark("load + clean profile")

    if args.metadata != "" and args.filter_column != "":
        from reportree_b200 import metadata_filter
        say("Filtering the allele matrix...")
        meta = pandas.read_table(args.metadata, dtype=str)
        meta.columns = [c.replace(" ", "_") for c in meta.columns]
        meta = meta.replace(" ", "_", regex=True)
        say("\tFiltering metadata for the following parameters: " + " & ".join(args.filter_column.split(";")))
        keep_ids = set(metadata_filter.select_samples(meta, args.filter_column))
        rows = np.asarray([i for i, s in enumerate(prof.ids) if s is not None and s in keep_ids], dtype=np.int64)
        say("\tFrom the " + str(prof.n) + " samples, " + str(len(rows)) + " were kept in the matrix...")
        if len(rows) <= 1:
            say("\nCannot proceed because " + str(len(rows)) + " samples were kept in the matrix!")
            _exit(log)
        prof.write(out + "_subset_matrix.tsv", rows, all_cols)

    cols = _loci_selection_fast(prof, rows, args, say)
    if len(cols) <= 1:
        say("\nCannot proceed because " + str(len(cols)) + " sites/loci were kept in the matrix!")
        _exit(log)
    with open(out + "_loci_used.txt", "w+") as fh:
        fh.write("".join(prof.header[c + 1] + "\n" for c in cols.tolist()))
    PH.mark("column filter")
    writers = [Background(prof.write, out + "_clean_missing_matrix.tsv", rows, cols)]     # overlaps the row filter

    thr = float(args.loci_called)
    if thr >= 0.0:
        say("Cleaning the profile matrix using a threshold of >" + str(args.loci_called) + " alleles/positions called per sample...")
        len_schema = len(cols)
        id_is_zero = np.fromiter((prof.ids[i] == "0" for i in rows.tolist()), dtype=bool, count=len(rows))
        missing = prof.count_eq(rows, cols, 0) + id_is_zero   # isin(["0"]) over every column incl. the id (:451)
        called = len_schema - missing
        pct_called = called / len_schema                      # float64, as the reference's Series division (:453)
        passed = pct_called >= thr
        # DataFrame.to_csv without pandas (its import costs ~1 s, more than the whole parse): ints as str, float64 as
        # the shortest round-trip repr, NaN ids as the empty field -- byte-identical on the goldens
        with open(out + "_loci_report.tsv", "w") as fh:
            fh.write("samples\tmissing\tcalled\tpct_called\n")
            fh.write("".join("%s\t%d\t%d\t%r\n" % ("" if prof.ids[i] is None else prof.ids[i], m, c, p)
                             for i, m, c, p in zip(rows.tolist(), missing.tolist(), called.tolist(), pct_called.tolist())))
        say("\tFrom the " + str(len(rows)) + " samples, " + str(int(passed.sum())) + " were kept in the profile matrix.")
        if int(passed.sum()) <= 1:
            say("\nCannot proceed because " + str(int(passed.sum())) + " samples were kept in the matrix!")
            for w in writers:
                w.join()
            _exit(log)
        # isin(passing): every row whose id equals the id of a passing row stays (duplicated ids)
        passing_ids = {prof.ids[i] for i, ok in zip(rows.tolist(), passed.tolist()) if ok}
        sel = np.fromiter((prof.ids[i] in passing_ids for i in rows.tolist()), dtype=bool, count=len(rows))
        rows = rows[sel]
        codes = _selected_codes(prof, rows, cols)
        writers.append(Background(prof.write, out + "_flt_samples_matrix.tsv", rows, cols))   # overlaps the GPU stages
    else:
        codes = _selected_codes(prof, rows, cols)
    PH.mark("row filter")

    say("Getting the pairwise distance matrix with cgmlst-dists (if your profile matrix is too big, this will be done in chunks of 2000 alleles/positions)...")
    total_size = len(cols)
    names = ["" if prof.ids[i] is None else prof.ids[i] for i in rows.tolist()]
    PH.mark("encode cells")
    try:
        ctx.set_profiles(np.ascontiguousarray(codes))
        ctx.distance()
        PH.mark("GPU context + recode + distance")
    finally:
        for w in writers:
            w.join()
        prof.close()
    PH.mark("wait for matrix artefacts")
    if os.environ.get("REPORTREE_HC_SKIP_DIST_TSV", "0") != "1":
        write_dist_tsv(out + "_dist_hamming.tsv", names, ctx)
        PH.mark("write dist_hamming.tsv")
    return names_as_read_back(names), total_size


def names_as_read_back(names):
    """The reference names the samples from the HEADER of `<out>_dist_hamming.tsv` as pandas reads it back
    (`dist.columns` after read_table, :518-520, :184), not from the profile's id column: duplicated ids
    come back mangled ("A", "A.1", ...), an empty id as "Unnamed: k".  Plain distinct ids are unchanged, so
    only those cases take the round trip through pandas' own writer and header parser."""
    if len(set(names)) == len(names) and "" not in names and "dists" not in names:
        return names
    import io
    text = pandas.DataFrame(columns=["dists"] + list(names)).to_csv(sep="\t", index=False)
    return [str(c) for c in pandas.read_table(io.StringIO(text), nrows=0).columns[1:]]


def write_dist_tsv(path, names, ctx):
    from reportree_b200 import fastio
    if fastio.enabled() and fastio.plain_text(names):
        fastio.write_square_tsv(path, names, ctx.square_rows, len(names))
    else:
        hostio.write_square_tsv(path, names, ctx.square_rows, len(names))


def matrix_to_distances(args, say, log, ctx):
    """`-d_mx` branch (:522-567): distance matrix supplied."""
    say("Distance matrix provided... pairwise distance will not be calculated!")
    dist = read_distance_matrix(args.distance_matrix)
    if args.metadata != "" and args.filter_column != "":
        from reportree_b200 import metadata_filter
        say("Filtering the distance matrix...")
        meta = pandas.read_table(args.metadata, dtype=str)
        meta.columns = [c.replace(" ", "_") for c in meta.columns]
        n0 = len(dist)
        say("\tFiltering metadata for the following parameters: " + " & ".join(args.filter_column.split(";")))
        keep_ids = set(metadata_filter.select_samples(meta, args.filter_column))
        sq = dist.set_index(dist.columns[0], drop=True)
        cols = [c for c in sq.columns if c in keep_ids]
        dist = sq.loc[cols, cols].reset_index()
        say("\tFrom the " + str(n0) + " samples, " + str(len(dist)) + " were kept in the matrix...")
        if len(dist) <= 1:
            say("\nCannot proceed because " + str(len(dist)) + " samples were kept in the matrix!")
            _exit(log)
        dist.to_csv(args.out + "_flt_dist.tsv", sep="\t", index=None)
    elif args.metadata != "" and args.filter_column == "":
        say("Metadata file was provided but no filter was found... I am confused :-(")
        _exit(log)
    elif args.metadata == "" and args.filter_column != "":
        say("Metadata file was not provided but a filter was found... I am confused :-(")
        _exit(log)
    sq = dist.set_index(dist.columns[0], drop=True)
    names = [str(c) for c in sq.columns.tolist()]
    ctx.set_condensed(square_to_condensed(sq.to_numpy()))
    return names, None


def main(argv=None):
    args = build_parser().parse_args(argv)
    from reportree_b200 import fastio, lazy_pandas
    if not (fastio.enabled() and args.allele_profile and args.metadata == ""):
        lazy_pandas.preload()                     # ~1 s of import on a helper thread; the C++ text path never needs it
    log = Tee(args.out + ".log")
    say = log
    say("\n-------------------- partitioning_HC.py --------------------\n")
    say("version", version, "last updated on", last_updated, "\n")
    say(" ".join(sys.argv), "\n")

    if not args.allele_profile and not args.distance_matrix:
        say("Could not find a profile or a distance matrix... One of them needs to be specified!!")
        _exit(log)

    PH.__init__()
    devices = [int(d) for d in os.environ.get("REPORTREE_HC_DEVICES", "").split(",") if d.strip() != ""]
    exact = os.environ.get("REPORTREE_HC_EXACT_NAMES", "1") != "0"
    ctx = LazyContext(int(os.environ.get("REPORTREE_HC_DEVICE", "0")), devices,
                      single_algo=hc_backend.SL_PRIM if exact else hc_backend.SL_BORUVKA)

    if args.allele_profile:
        names, total_size = profile_to_distances(args, say, log, ctx)
    else:
        names, total_size = matrix_to_distances(args, say, log, ctx)

    say("\tGetting condensed distance matrix...")
    if args.pct_HCmethod_threshold != "none" and total_size is None:
        raise NameError("name 'total_size' is not defined")      # same failure as the reference (:586)
    runs, pct_corr = plan_runs(args.method_threshold, args.pct_HCmethod_threshold, total_size, say)

    from reportree_b200 import fastio
    cut_names, cut_labels = [], []              # every partition column in output order
    for method, requests in runs.items():
        say("Hierarchical clustering with method: " + method + "...")
        if method not in GPU_METHODS:
            raise hc_backend.HCBackendError(
                "HC method %r is not implemented on the CUDA backend (available: %s)" % (method, ", ".join(GPU_METHODS)))
        PH.mark("plan")
        Z = ctx.linkage(method)
        PH.mark("GPU linkage " + method)
        max_dist = Z[-1, 2]
        say("\tMaximum distance " + str(max_dist) + "...")
        say("\tGenerating newick file...")
        with open(args.out + "_" + method + "_HC.nwk", "w+") as fh:
            print(hostio.newick(Z, names), file=fh)
        PH.mark("newick " + method)
        say("\tDefining clusters...")
        cuts = expand_cuts(method, requests, pct_corr, max_dist, args.dist, say)
        if cuts:
            labels = ctx.fcluster([c[1] for c in cuts])
            PH.mark("GPU cut " + method)
            cut_names.extend(c[0] for c in cuts)
            cut_labels.append(labels)

    # The C++ writer needs distinct partition names (a repeated name makes the reference extend the
    # composition of the first one, :217-230) and text that needs no CSV quoting; else the Python path.
    fast = (fastio.enabled() and len(set(cut_names)) == len(cut_names) and fastio.plain_text(names)
            and fastio.plain_text(cut_names))
    if fast:
        say("Creating sample partitions file...")
        say("Creating cluster composition file...")
        all_labels = np.concatenate(cut_labels, axis=0) if cut_labels else np.zeros((0, len(names)), dtype=np.int32)
        fastio.write_partitions(args.out + "_partitions.tsv", args.out + "_clusterComposition.tsv", names, all_labels, cut_names)
        PH.mark("write partitions + composition")
    else:
        columns = {"sequence": names}
        composition = {}
        k = 0
        for labels in cut_labels:
            for row in labels:
                partition = cut_names[k]
                k += 1
                col, comp = hostio.name_partition(row, names)
                columns[partition] = col
                bucket = composition.setdefault(partition, {})
                for cname, csize, members in comp:
                    bucket.setdefault(cname, {}).setdefault(csize, []).extend(members)
        PH.mark("name partitions")
        say("Creating sample partitions file...")
        hostio.write_partitions(args.out + "_partitions.tsv", columns)
        PH.mark("write partitions.tsv")
        say("Creating cluster composition file...")
        with open(args.out + "_clusterComposition.tsv", "w+") as fh:
            print("#partition\tcluster\tcluster_length\tsamples", file=fh)
            for partition, clusters in composition.items():
                for cname, by_size in clusters.items():
                    for csize, members in by_size.items():
                        print(partition + "\t" + cname + "\t" + str(csize) + "\t" + ",".join(members), file=fh)
        PH.mark("write clusterComposition")
    if len(pct_corr) > 0:
        with open(args.out + "_parcentage2threshold.tsv", "w+") as fh:
            print("#percentage\tthreshold", file=fh)
            for thr, pcts in pct_corr.items():
                for pct in pcts:
                    print(str(pct) + "\t" + str(thr), file=fh)
    say("\npartitioning_HC.py is done!")
    log.close()
    ctx.close()               # explicit release: leaving tens of GB to the driver's process teardown measured slower
    PH.mark("release device")
    PH.report()


if __name__ == "__main__":
    main()
    # every artefact is closed and every helper thread joined at this point: skip the interpreter's module
    # teardown and the CUDA runtime's atexit handlers (0.1-0.2 s of a 2.5 s run at 100,000 samples)
    sys.stdout.flush()
    sys.stderr.flush()
    if os.environ.get("REPORTREE_HC_FAST_EXIT", "1") != "0":
        os._exit(0)
