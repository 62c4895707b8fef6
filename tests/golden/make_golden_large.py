#!/usr/bin/env python3
"""Generates tests/golden/large/<config>.json: the CPU oracle's answer for BASELINE.json configs[2..4]
(C3 20,000 x 3,002; C4 100,000 x 3,000; C5 50,000 x 7,600 after --loci / --loci-called) as fingerprints.

CPU only (no GPU, no product code on the path): oracle/cgmlst_dists.c for the distances (row blocks on
all host threads), oracle/hclust.c (uint16-input entry points) for linkage + fcluster.  What is stored
per config: CRC32 of every 1,024-row block of the condensed uint16 distance vector, and per linkage
method the CRC32 of Z and, at every threshold 0..50, the CRC32 of the fcluster ids, the cluster count
and the CRC32 of the membership-only relabelling (reportree_b200/checksums.py defines them).

    python tests/golden/make_golden_large.py C3 C5 C4      # ~3 min, ~6 min, ~12 min on 8 cores
"""
import json
import os
import sys
import time
from concurrent.futures import ThreadPoolExecutor

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import oracle_lib as O                      # noqa: E402
from reportree_b200 import checksums as C, synth        # noqa: E402

METHODS = {"C3": ["single", "average", "complete"], "C4": ["single"], "C5": ["complete", "average"]}
THRESHOLDS = list(range(0, 51))
OUT = os.path.join(ROOT, "tests", "golden", "large")


def oracle_condensed(X, threads):
    n = X.shape[0]
    y = np.empty(n * (n - 1) // 2, dtype=np.uint16)
    bounds = C.block_bounds(n)

    def job(b):
        r0, r1, p0, p1 = b
        y[p0:p1] = O.dist_condensed_rows(X, r0, r1)
        return C.crc(y[p0:p1])
    with ThreadPoolExecutor(threads) as ex:
        crcs = list(ex.map(job, bounds))
    return y, crcs


def main():
    threads = os.cpu_count() or 1
    for name in sys.argv[1:]:
        t0 = time.time()
        X = synth.workload(name)
        n, L = X.shape
        t1 = time.time()
        y, crcs = oracle_condensed(X, threads)
        t2 = time.time()
        rec = {"config": name, "n": n, "L": L, "input_crc32": C.crc(X), "rows_per_block": C.ROWS_PER_BLOCK,
               "condensed_block_crc32": crcs, "condensed_crc32": C.crc(y), "max_distance": int(y.max()),
               "thresholds": THRESHOLDS, "methods": {},
               "generator": "tests/golden/make_golden_large.py (oracle/cgmlst_dists.c + oracle/hclust.c, %d threads)" % threads,
               "seconds": {"synthesise": round(t1 - t0, 1), "distance_all_threads": round(t2 - t1, 1)}}
        for method in METHODS[name]:
            t3 = time.time()
            Z = O.linkage(y, method)
            t4 = time.time()
            m = C.z_fingerprint(Z)
            m["labels"] = [C.labels_fingerprint(O.fcluster(Z, float(t))) for t in THRESHOLDS]
            rec["methods"][method] = m
            rec["seconds"]["linkage_" + method] = round(t4 - t3, 1)
            rec["seconds"]["cut_" + method] = round(time.time() - t4, 1)
            print(name, method, "linkage %.1f s" % (t4 - t3), flush=True)
        os.makedirs(OUT, exist_ok=True)
        with open(os.path.join(OUT, name + ".json"), "w") as f:
            json.dump(rec, f, indent=1)
        print(name, "done in %.1f s" % (time.time() - t0), rec["seconds"], flush=True)


if __name__ == "__main__":
    main()
