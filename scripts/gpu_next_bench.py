"""Measurement of the SURVEY 8(f) N3 / N4 components on one B200 at C4 size (100,000 x 3,000):
zoom-in look-ups on the resident matrix and the cluster-stability tables, each next to a CPU
timing of the reference's route on a bounded sample (oracle ports; /root/reference is not on the box).

    python bench.py --next [--next-workload C4|C3] > gpurun_out/next_bench.json      # with the CPU baselines
    python scripts/gpu_next_bench.py [C4|C3]                                            # GPU numbers only

The oracle (test infrastructure) is never imported here: bench.py's cpu_baseline leg passes its ports in as `cpu`.
"""
import json
import os
import sys
import tempfile
import time

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from reportree_b200 import comparing_partitions_v2 as cp       # noqa: E402
from reportree_b200 import fastio, hc_backend, synth           # noqa: E402


def wall(fn, reps=1):
    t0 = time.perf_counter()
    for _ in range(reps):
        out = fn()
    return (time.perf_counter() - t0) / reps * 1e3, out


def run(name="C4", cpu=None):
    """cpu: None, or an object with stability_lines(names, columns) and dist_square(codes) (bench.py's oracle ports)."""
    cfg = synth.config(name)
    X = synth.allele_matrix(**cfg)
    n, L = X.shape
    res = {"workload": "%s synthetic %dx%d" % (name, n, L)}
    peaks = {}
    if os.path.exists("MEASURED_PEAKS.json"):
        peaks = json.load(open("MEASURED_PEAKS.json"))
    hbm = float(peaks.get("hbm_gbs", 6547.5))
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        ctx.linkage("single")
        thresholds = np.arange(51, dtype=np.float64)
        labels = ctx.fcluster(thresholds) - 1

        # ---- N4: stability tables -----------------------------------------------------------
        ctx.contingency_adjacent(labels)                                   # warm-up (memory pool, code load)
        l0 = ctx.timings()["launches"]
        ms_tables, tables = wall(lambda: ctx.contingency_adjacent(labels))
        dev_ms = ctx.timings()["stability_ms"]
        launches = ctx.timings()["launches"] - l0
        stats = [cp.ColumnStats(r) for r in labels]
        ms_finish, rows = wall(lambda: [cp.pair_metrics(ctx, tables[p], stats[p], stats[p + 1]) for p in range(len(tables))])
        cells = int(sum(t[0].size for t in tables))
        N = 2048
        while N < n:
            N <<= 1
        H = 2 * N
        # bytes the kernels move per pair: table + key initialisation (16 H + 8 N), insert (8 n codes + 3 atomics of 4-8 B per
        # sample ~ 20 n), compact (12 H read + 8 cells), sort (local + tail: 2 x 16 N per stage above 2048; global steps 16 N each)
        stages = max(0, int(np.log2(N)) - 11)
        sort_bytes = 16 * N * (1 + stages + stages * (stages + 1) // 2)
        per_pair = 16 * H + 8 * N + 28 * n + 12 * H + sort_bytes + 24 * n
        res["stability"] = {
            "pairs": len(tables), "samples": n, "cells": cells, "device_ms": dev_ms, "wall_ms_tables": ms_tables,
            "wall_ms_host_finish": ms_finish, "gpu_launches": int(launches),
            "roofline": {"bound": "hbm", "achieved": per_pair * len(tables) / (dev_ms * 1e-3) / 1e9, "peak": hbm, "unit": "GB/s",
                         "frac": per_pair * len(tables) / (dev_ms * 1e-3) / 1e9 / hbm,
                         "note": "bytes per pair = fills + insert + compact + bitonic passes (formula in scripts/gpu_next_bench.py); "
                                 "device_ms includes the H2D of the label matrix and the D2H of the cells"},
            "first_rows": [[float(x) if not isinstance(x, tuple) else [float(y) for y in x] for x in r] for r in rows[:3]],
        }
        # CPU baseline: the oracle port of metrics4stability on a bounded prefix of the same label matrix
        ns = 4000
        small = np.stack([np.unique(labels[t][:ns], return_inverse=True)[1].astype(np.int32) for t in range(labels.shape[0])])
        ms_small, tb = wall(lambda: ctx.contingency_adjacent(small))
        res["stability"]["same_sample_gpu_ms"] = ms_small
        if cpu is not None:
            cols = [labels[t][:ns].tolist() for t in range(labels.shape[0])]
            names = ["single-%dx1.0" % t for t in range(labels.shape[0])]
            t0 = time.perf_counter()
            lines = cpu.stability_lines(names, cols)
            cpu_s = time.perf_counter() - t0
            res["stability"]["cpu_baseline"] = {
                "kind": "port", "cores": 1, "seconds": cpu_s, "sample": "first %d of %d samples, %d column pairs" % (ns, n, len(lines) - 1),
                "note": "oracle/partition_metrics.py iterates table cells (linear); the reference's getContTableTotals / "
                        "getMismatchMatrix walk set(col1) x set(col2) (metrics.py:38-91), quadratic in the clusters"}
            st = [cp.ColumnStats(r) for r in small]
            mine = [cp.pair_metrics(ctx, tb[p], st[p], st[p + 1]) for p in range(len(tb))]
            same = all(str(m[0]) == l.split("\t")[2] and str(m[2][0]) == l.split("\t")[6] for m, l in zip(mine, lines[1:]))
            res["stability"]["same_sample_rows_equal_oracle"] = bool(same)

        # ---- N3: zoom-in --------------------------------------------------------------------
        rng = np.random.default_rng(1)
        picks = rng.integers(0, n, 20)
        ctx.closest(int(picks[0]), 50)
        ms_closest, _ = wall(lambda: [ctx.closest(int(s), 50) for s in picks])
        idx, d, vmax = ctx.closest(int(picks[0]), 2000)
        m = int(idx.size)

        def zoom_once():
            with ctx.subset(idx) as child:
                Z = child.linkage("single")
                return Z, child.fcluster(thresholds)
        zoom_once()
        ms_zoom, (Z, lab) = wall(zoom_once, reps=3)
        reuse = ctx.subset(idx)

        def zoom_reuse():
            ctx.subset(idx, into=reuse)
            return reuse.linkage("single"), reuse.fcluster(thresholds)
        zoom_reuse()
        ms_reuse, _ = wall(zoom_reuse, reps=5)
        tm = reuse.timings()
        reuse.close()
        res["zoom"] = {"closest_ms_per_query": ms_closest / len(picks), "k": 50, "subset_samples": m,
                       "subset_gather_linkage_cut_ms_new_context": ms_zoom, "subset_gather_linkage_cut_ms": ms_reuse,
                       "subset_device_ms": {"linkage": tm["linkage_ms"], "cut": tm["cut_ms"]}}
        # CPU baselines of the reference's route at a bounded size
        if cpu is not None:
            from scipy.cluster.hierarchy import fcluster, linkage
            from scipy.spatial.distance import squareform
            t0 = time.perf_counter()
            D = cpu.dist_square(np.ascontiguousarray(X[idx]))
            Zs = linkage(squareform(D.astype(np.float64), checks=False), "single")
            for t in thresholds:
                fcluster(Zs, t, criterion="distance")
            cpu_zoom = time.perf_counter() - t0
            ok = bool(np.array_equal(np.sort(Z[:, 2]), np.sort(Zs[:, 2])))
            res["zoom"]["cpu_baseline_rerun"] = {"kind": "port", "cores": 1, "seconds": cpu_zoom, "heights_equal": ok,
                                                 "sample": "distance (oracle C) + scipy single linkage + 51 cuts on the same %d samples x %d loci; "
                                                           "TSV round trips of the reference re-run not included" % (m, L)}
        # get_closest_samples reads one column of <out>_dist_hamming.tsv with pandas: time it on a 10,000-sample sub-matrix
        import pandas as pd
        ns = min(10000, n)
        with ctx.subset(np.arange(ns)) as child:
            tmp = tempfile.mkdtemp()
            path = os.path.join(tmp, "sub_dist_hamming.tsv")
            ids = ["S%07d" % (k + 1) for k in range(ns)]
            fastio.write_square_tsv(path, ids, child.square_rows, ns)
            size = os.path.getsize(path)
            t0 = time.perf_counter()
            col = pd.read_table(path, usecols=["dists", ids[17]])
            col.sort_values(by=[ids[17]], inplace=True)
            cpu_closest = time.perf_counter() - t0
            ms_sub, _ = wall(lambda: child.closest(17, 50), reps=5)
            os.remove(path)
        res["zoom"]["cpu_baseline_closest"] = {"kind": "port", "cores": 1, "seconds": cpu_closest, "gpu_ms_same_matrix": ms_sub,
                                               "sample": "pandas.read_table(usecols) + sort of one column of a %d x %d _dist_hamming.tsv "
                                                         "(%.2f GB); the file grows with n^2" % (ns, ns, size / 1e9)}
    return res


if __name__ == "__main__":
    print(json.dumps(run(sys.argv[1] if len(sys.argv) > 1 else "C4")))
