#!/usr/bin/env python3
"""Drop-in for the `-a stability` mode of scripts/ComparingPartitions/comparing_partitions_v2.py
(SURVEY.md 8f N4): same argv, messages, log lines and output files (`<tag>_metrics.tsv`,
`<tag>_stableRegions.tsv`), as invoked by run_stability (reportree.py:31-45).

What runs where
  * device (csrc/stability.cu, `rthc_contingency_adjacent`): the contingency table of every pair of
    neighbouring partition columns -- the reference's getContTable / getContTableTotals /
    getMismatchMatrix Python loops (metrics.py:23-91), which are quadratic in the number of clusters --
    returned cell by cell in the reference's dict order;
  * host C (`rthc_adjusted_wallace_sums`): the order-sensitive floating-point row sums of
    getAdjustedWallace (metrics.py:147-165);
  * here: the closed-form scalars per column pair (Simpson, Wallace, Adjusted Wallace + CI, Adjusted
    Rand from sklearn's pair-confusion integers) and the text.

Differences, both deliberate:
  * without --keep-redundants the reference keeps one RANDOM sample per cluster of the most
    discriminatory column (random.choice, unseeded; comparing_partitions_v2.py:47-52); this keeps the
    FIRST sample of each cluster -- one of the reference's possible outcomes, and reproducible;
  * `-a between_methods` ("still under development" upstream, not used by reportree.py) is refused.

There is no CPU fallback: without the CUDA library / a GPU this exits with the backend's error.
"""
import argparse
import math
import sys

import numpy as np

from reportree_b200.lazy_pandas import pandas

METRIC_COLUMNS = ["col2", "col1", "AdjRand", "Simpson_col2", "Simpson_col1", "Wallace", "AdjWallace", "AdjWallaceIClow",
                  "AdjWallaceIChigh", "col2_cluster", "col1_cluster"]


def say(log, *parts):
    print(*parts)
    print(*parts, file=log)


def encode_columns(matrix):
    """(names, T x n int32 codes): every partition column as dense cluster ids (first appearance order)."""
    names = list(matrix.columns[1:])
    codes = np.empty((len(names), len(matrix)), dtype=np.int32)
    for k, col in enumerate(names):
        codes[k] = pandas.factorize(matrix[col].to_numpy(), use_na_sentinel=False)[0]
    return names, codes


def rm_redundant(names, codes, log):
    """One sample per cluster of the most discriminatory column (first maximum, columns in file order);
    the surviving rows keep their order.  Codes are made dense again (the device wants ids < n)."""
    best, main = 0, 0
    for k in range(len(names)):
        size = int(codes[k].max()) + 1 if codes.shape[1] else 0
        if size > best:
            best, main = size, k
    say(log, "\tThe most discriminatory column found was " + (names[main] if names else ""))
    keep = np.sort(np.unique(codes[main], return_index=True)[1])
    sub = codes[:, keep]
    for k in range(sub.shape[0]):
        sub[k] = np.unique(sub[k], return_inverse=True)[1]
    return sub


class ColumnStats:
    """Cluster sizes of one partition column, from its dense codes."""

    def __init__(self, codes):
        sizes = np.bincount(codes).astype(np.int64)
        self.n = int(codes.size)
        self.k = int(np.count_nonzero(sizes))
        self.sum_sq = int(np.dot(sizes, sizes))
        self.pairs2 = self.sum_sq - self.n                    # sum of size * (size - 1)
        self.simpson = 1.0
        if self.n * (self.n - 1) > 0:
            self.simpson = 1.0 - (self.pairs2 / float(self.n * (self.n - 1)))


def pair_metrics(ctx, cells, st_i, st_j):
    """The seven numbers of one <tag>_metrics.tsv row.  cells = (rows, cols, counts) of
    getContTable(jcol, icol); st_i / st_j = ColumnStats of column i (col1) and i+1 (col2)."""
    rows, _, counts = cells
    n = st_i.n
    c64 = counts.astype(np.int64)
    ssq = int(np.dot(c64, c64))
    a = (ssq - n) / 2                                         # pairs together in both columns
    a1_i, a1_j = st_i.pairs2 / 2, st_j.pairs2 / 2
    wallace = a / float(a1_i) if a1_i > 0 else float(0)       # col1 -> col2 (metrics.py:100-108)
    # Adjusted Wallace col2 -> col1 with its confidence interval (metrics.py:142-193)
    w1sum, w2sum = ctx.adjusted_wallace_sums(rows, counts)
    var_w = 0.0
    if w2sum > 0:
        var_w = float(w1sum / float(float(w2sum) ** 2.0))
    w = a / float(a1_j) if a1_j > 0 else float(0)
    wi = 1 - st_i.simpson
    if float(1 - wi) > 0:
        aw = float(w - wi) / float(1 - wi)
        ci = 2.0 * (1.0 / float(1.0 - wi)) * math.sqrt(var_w)
        adj = (aw, aw - ci, aw + ci)
    else:
        adj = (0, 0, 0)
    # sklearn.metrics.adjusted_rand_score from the pair-confusion integers
    tp, fp, fn = ssq - n, st_j.sum_sq - ssq, st_i.sum_sq - ssq
    tn = n * n - fp - fn - ssq
    if fn == 0 and fp == 0:
        rand = 1.0
    else:
        rand = 2.0 * (tp * tn - fn * fp) / ((tp + fn) * (fn + tn) + (tp + fp) * (fp + tn))
    return rand, wallace, adj


def metrics4stability(names, codes, tag, log, ctx):
    stats = [ColumnStats(row) for row in codes]
    tables = ctx.contingency_adjacent(codes) if len(names) > 1 else []
    with open(tag + "_metrics.tsv", "w+") as out:
        print("\t".join(METRIC_COLUMNS), file=out)
        for k in range(len(names) - 1):
            col1, col2 = names[k], names[k + 1]
            if col1 == col2:
                continue
            print("\t\t", col2, " => ", col1)
            rand, wallace, adj = pair_metrics(ctx, tables[k], stats[k], stats[k + 1])
            say(log, "\t\t\tAdjusted Wallace ", col2, " vs ", col1, "\t%f\t%f\t%f" % adj)
            print("\t".join([str(col2), str(col1), str(rand), str(stats[k + 1].simpson), str(stats[k].simpson), str(wallace),
                             str(adj[0]), str(adj[1]), str(adj[2]), str(stats[k + 1].k), str(stats[k].k)]), file=out)


def get_stability(infile, thr, n, tag, log):
    say(log, "\t\tDetermining stability regions with threshold: ", str(thr), " and minimum n of ", str(n))
    blocks, open_block = [], False
    with open(infile) as fh:
        for line in fh.readlines()[1:]:
            cells = line.split("\t")
            if float(cells[6]) >= float(thr):
                if not open_block:
                    blocks.append([])
                    open_block = True
                blocks[-1].append(cells[0] + "->" + cells[1])
            else:
                open_block = False
    with open(tag + "_stableRegions.tsv", "w+") as out:
        print("#Stability regions of the Adjusted Wallace coefficient for ", infile, "\n#using a threshold of ", str(thr),
              " and minimum number of required observations of ", str(n), file=out)
        print("#block_id\tfirst_partition\tlast_partition\tlen_block", file=out)
        for k, members in enumerate(blocks):
            if len(members) >= int(n):
                block = "block_" + str(k + 1)
                print("\t".join([block, members[0], members[-1], str(len(members))]), file=out)
                say(log, "\t\t", block, members[0], members[-1], str(len(members)))


def build_parser():
    p = argparse.ArgumentParser(prog="comparing_partitions_v2.py", description="cluster stability regions (Adjusted Wallace "
                                "over neighbouring partition thresholds), CUDA backend")
    p.add_argument("-i1", "--input1", dest="input1", action="store", required=True, help="[MANDATORY] Input matrix 1 (table with partitions)")
    p.add_argument("-i2", "--input2", dest="input2", action="store", help="Input matrix 2 (table with partitions)")
    p.add_argument("-o1", "--order1", dest="order1", action="store", default=0, required=True,
                   help="Partitions order in matrix 1 (0: min -> max; 1: max -> min) [0]")
    p.add_argument("-o2", "--order2", dest="order2", action="store", help="Partitions order in matrix 2 (0: min -> max; 1: max -> min) [0]")
    p.add_argument("-a", "--analysis", dest="analysis", action="store", default="between_methods",
                   help="Type of analysis (options: stability or between_methods) [between_methods]")
    p.add_argument("-t", "--tag", dest="tag", action="store", required=True, help="[MANDATORY] Tag for output file name")
    p.add_argument("-n", "--n_obs", dest="n_obs", action="store", default=5,
                   help="Minimum number of sequencial observations to consider an interval for method stability analysis [5]")
    p.add_argument("-thr", "--threshold", dest="threshold", action="store", default=0.99,
                   help="Threshold of Adjusted Wallace score to consider an observation for method stability analysis [0.99]")
    p.add_argument("--keep-redundants", dest="keep_redundants", action="store_true",
                   help="Set ONLY if you want to keep all samples of each cluster of the most discriminatory partition")
    p.add_argument("-log", "--log", dest="log", action="store", default="log", help="Log file")
    return p


def main(argv=None, ctx=None):
    args = build_parser().parse_args(argv)
    log = open(args.log, "a+")
    say(log, "\n-------------------- comparing_partitions_v2.py --------------------\n")
    say(log, " ".join(sys.argv if argv is None else ["comparing_partitions_v2.py"] + list(argv)))

    if args.analysis == "stability":
        if args.input2 or args.order2:
            say(log, "Provided options are incompatible! 'stability' analysis only accepts input1 and order1!")
            sys.exit()
        import threading
        box = {}
        if ctx is None:                       # CUDA context creation overlaps the pandas parse
            def create():
                try:
                    import os
                    from reportree_b200 import hc_backend
                    box["ctx"] = hc_backend.HCContext(device=int(os.environ.get("REPORTREE_HC_DEVICE", "0")))
                except BaseException as e:    # noqa: BLE001 -- re-raised on the main thread
                    box["err"] = e
            maker = threading.Thread(target=create)
            maker.start()
        matrix = pandas.read_table(args.input1)
        say(log, "Preparing for stability analysis...")
        names, codes = encode_columns(matrix)
        if not args.keep_redundants:
            say(log, "Removing redundant samples...")
            codes = rm_redundant(names, codes, log)
        say(log, "Ordering matrix...")
        if int(args.order1) != 0:
            if int(args.order1) == 1:
                names, codes = names[::-1], np.ascontiguousarray(codes[::-1])
            else:
                say(log, "\tInvalid order!")
                sys.exit()
        say(log, "Cleaning matrix...")
        say(log, "Calculating metrics...")
        own = ctx is None
        if own:
            maker.join()
            if "err" in box:
                raise box["err"]
            ctx = box["ctx"]
        try:
            metrics4stability(names, codes, args.tag, log, ctx)
        finally:
            if own:
                ctx.close()
        say(log, "Get stability...")
        get_stability(args.tag + "_metrics.tsv", args.threshold, args.n_obs, args.tag, log)
    elif args.analysis == "between_methods":
        say(log, "'between_methods' is not available on the CUDA backend (upstream marks it as under development and "
                 "reportree.py never calls it); use -a stability")
        sys.exit(2)
    else:
        say(log, "Invalid analysis type!")
        sys.exit()

    say(log, "\ncomparing_partitions_v2.py is done!")
    log.close()


if __name__ == "__main__":
    main()
