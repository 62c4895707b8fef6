"""CPU restatement of the reference's cluster-stability metrics.

TEST INFRASTRUCTURE ONLY (imported by tests/ and bench-like scripts as the checker / timed CPU
baseline): the product path is csrc/stability.cu through reportree_b200/comparing_partitions_v2.py.

Follows, in the reference tree (scripts/ComparingPartitions/):
  * metrics.py:23-35   getContTable       -> contingency()
  * metrics.py:38-91   getContTableTotals / getMismatchMatrix -> pair_counts() (only a and a+b are
                       ever used by the stability output; the reference's `c` takes just the LAST
                       row total because metrics.py:86 sits outside its loop -- it feeds w2 and d,
                       neither of which is printed)
  * metrics.py:100-108 getWallace (first component)
  * metrics.py:111-139 getSimpsons        -> simpson()
  * metrics.py:142-193 getAdjustedWallace -> adjusted_wallace()
  * metrics.py:95-97   getAdjRand = sklearn adjusted_rand_score -> adjusted_rand() (sklearn's
                       pair-confusion integers, sklearn/metrics/cluster/_supervised.py:255-270, :454-461)
  * comparing_partitions_v2.py:74-113  metrics4stability -> stability_lines()
  * comparing_partitions_v2.py:218-257 get_stability     -> stable_regions()

Dict iteration order (= first appearance while scanning the samples) fixes the order of every
floating-point sum, so plain dicts and Python floats are used on purpose: `**` must stay libm pow.

Pinned: tests/golden/next/stability_* hold the files written by the UNMODIFIED reference script
(tests/golden/make_golden_next.py); tests/test_oracle_next.py compares this restatement with them
byte for byte, and, where /root/reference exists, with metrics.py imported directly.
"""
import math


def contingency(rows, cols):
    table = {}
    for r, c in zip(rows, cols):
        inner = table.setdefault(r, {})
        inner[c] = inner.get(c, 0) + 1
    return table


def _sizes(col):
    d = {}
    for x in col:
        d[x] = d.get(x, 0) + 1
    return d


def simpson(col):
    """(sid, low, high, number of clusters)"""
    n = len(col)
    d = _sizes(col)
    pairs = 0
    f2 = 0
    f3 = 0
    for v in d.values():
        pairs += v * (v - 1)
        f2 += (v / float(n)) ** 2.0
        f3 += (v / float(n)) ** 3.0
    sid = 1.0
    if n * (n - 1) > 0:
        sid = 1.0 - (pairs / float(n * (n - 1)))
    s2 = (4.0 / float(n)) * float(f3 - f2 ** 2.0)
    if s2 <= 0:
        return sid, "-", "-", len(d)
    return sid, sid - 2 * math.sqrt(s2), sid + 2 * math.sqrt(s2), len(d)


def pair_counts(table):
    """(a, a1): pairs together in both partitions / pairs together in the ROW partition."""
    a = 0
    a1 = 0
    for inner in table.values():
        tot = 0
        for v in inner.values():
            a += (v * (v - 1)) / 2
            tot += v
        a1 += (tot * (tot - 1)) / 2
    return a, a1


def wallace(table):
    a, a1 = pair_counts(table)
    return a / float(a1) if a1 > 0 else float(0)


def adjusted_wallace(table, cols):
    """getAdjustedWallace(table, rows, cols): (AW rows->cols, low, high)"""
    w1sum = 0
    w2sum = 0
    for inner in table.values():
        r = sum(inner.values())
        f2 = 0
        f3 = 0
        for v in inner.values():
            f2 += (v / float(r)) ** 2.0
            f3 += (v / float(r)) ** 3.0
        sq = f2 ** 2.0
        var = 0.0
        if r > 1:
            var = float(4.0 * r * (r - 1.0) * (r - 2.0) * f3 + 2.0 * r * (r - 1.0) * f2
                        - 2.0 * r * (r - 1.0) * (2.0 * r - 3.0) * sq) / float((r * (r - 1.0)) ** 2.0)
        w1sum += float((r * (r - 1.0)) ** 2.0) * var
        w2sum += r * (r - 1.0)
    var_w = 0.0
    if w2sum > 0:
        var_w = float(w1sum / float(float(w2sum) ** 2.0))
    w = wallace(table)
    wi = 1 - simpson(cols)[0]
    if float(1 - wi) > 0:
        aw = float(w - wi) / float(1 - wi)
        ci = 2.0 * (1.0 / float(1.0 - wi)) * math.sqrt(var_w)
        return aw, aw - ci, aw + ci
    return 0, 0, 0


def adjusted_rand(x, y):
    n = len(x)
    table = contingency(x, y)
    ssq = sum(v * v for inner in table.values() for v in inner.values())
    sx = sum(v * v for v in _sizes(x).values())
    sy = sum(v * v for v in _sizes(y).values())
    tp = ssq - n
    fp = sy - ssq
    fn = sx - ssq
    tn = n * n - fp - fn - ssq
    if fn == 0 and fp == 0:
        return 1.0
    return 2.0 * (tp * tn - fn * fp) / ((tp + fn) * (fn + tn) + (tp + fp) * (fp + tn))


HEADER = ["col2", "col1", "AdjRand", "Simpson_col2", "Simpson_col1", "Wallace", "AdjWallace", "AdjWallaceIClow",
          "AdjWallaceIChigh", "col2_cluster", "col1_cluster"]


def stability_lines(names, columns):
    """Text lines of <tag>_metrics.tsv for the ordered partition columns (lists of labels)."""
    out = ["\t".join(HEADER)]
    for k in range(len(names) - 1):
        if names[k] == names[k + 1]:
            continue
        icol, jcol = columns[k], columns[k + 1]
        s1, s2 = simpson(icol), simpson(jcol)
        w = wallace(contingency(icol, jcol))
        aw = adjusted_wallace(contingency(jcol, icol), icol)
        out.append("\t".join([str(names[k + 1]), str(names[k]), str(adjusted_rand(icol, jcol)), str(s2[0]), str(s1[0]), str(w),
                              str(aw[0]), str(aw[1]), str(aw[2]), str(s2[3]), str(s1[3])]))
    return out


def stable_regions(metric_lines, thr, n_obs):
    """[(block id, first, last, length)] from the lines of <tag>_metrics.tsv (header included)."""
    blocks = []
    open_block = False
    for line in metric_lines[1:]:
        cells = line.split("\t")
        if float(cells[6]) >= float(thr):
            if not open_block:
                blocks.append([])
                open_block = True
            blocks[-1].append(cells[0] + "->" + cells[1])
        else:
            open_block = False
    return [("block_%d" % (k + 1), b[0], b[-1], len(b)) for k, b in enumerate(blocks) if len(b) >= int(n_obs)]
