"""Largest single-GPU shape the layout is sized for: n x 3,000 with the uint16 condensed matrix in
HBM (n = 250,000 -> 62.5 GB of distances, 3.1e10 pairs, 9.4e13 comparisons).  Properties only --
the CPU oracle sees a few hundred samples of it:
  distances of sampled rows == the independent scalar kernel on the uint16 codes (every column),
  == the CPU oracle on a sub-matrix, symmetric, zero diagonal;
  Boruvka single linkage: n-1 merges, heights sorted, last size == n;
  cut: #clusters at t == n - #(heights <= t), neighbours within t share the label.
usage (GPU box): python scripts/gpu_maxsize.py [n] [L]"""
import sys
import time

import numpy as np

sys.path.insert(0, ".")
from oracle import oracle_lib as O                      # noqa: E402  (checker only)
from reportree_b200 import hc_backend as hc, synth      # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 250000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 3000


def t(f):
    t0 = time.perf_counter(); r = f(); return r, time.perf_counter() - t0


X, tg = t(lambda: synth.allele_matrix(n=n, L=L, seed=20260306, missing=0.02))
print(f"synthetic {n} x {L}: generated in {tg:.1f} s, max allele id {int(X.max())}", flush=True)
rng = np.random.default_rng(3)
with hc.HCContext() as ctx:
    _, ts = t(lambda: ctx.set_profiles(X))
    _, td = t(ctx.distance)
    P = n * (n - 1) // 2
    assert ctx.owned_pairs() == P
    tm = ctx.timings()
    print(f"set_profiles {ts:.3f} s  distance {td:.3f} s (kernel {tm['distance_ms']:.1f} ms = "
          f"{P * L / tm['distance_ms'] / 1e9:.2f} Tcmp/s)  condensed {2 * P / 1e9:.1f} GB", flush=True)
    rows = np.sort(rng.choice(n, 24, replace=False))
    rows[0], rows[-1] = 0, n - 1                                    # first / last row of the triangle
    sq = np.stack([ctx.square_rows(int(r), 1)[0] for r in rows])
    pi = np.repeat(rows, n).astype(np.int32)
    pj = np.tile(np.arange(n, dtype=np.int32), rows.size)
    direct = ctx.direct_pairs(pi, pj).reshape(rows.size, n)
    assert np.array_equal(direct, sq.astype(np.int32)), "bit-sliced kernel != scalar kernel"
    assert np.array_equal(sq[:, rows], sq[:, rows].T) and (sq[np.arange(rows.size), rows] == 0).all()
    sub = np.unique(np.concatenate([np.arange(200), rng.choice(n, 200, replace=False), [n - 1]]))
    D = O.dist_square(X[sub])
    got = np.stack([ctx.square_rows(int(r), 1)[0][sub] for r in sub[-48:]])
    assert np.array_equal(got.astype(np.int32), D[-48:]), "kernel != CPU oracle"
    print("distances: 24 full rows == scalar kernel, 48 x 401 sub-matrix == CPU oracle", flush=True)
    Z, tl = t(lambda: ctx.linkage("single"))
    info = ctx.info()
    print(f"single linkage (Boruvka, {info.get('boruvka_rounds', '?')} rounds): {tl:.3f} s (device {ctx.timings()['linkage_ms']:.1f} ms)",
          flush=True)
    assert Z.shape == (n - 1, 4) and Z[-1, 3] == n and (np.diff(Z[:, 2]) >= 0).all()
    thr = [0.0, 1.0, 4.0, 15.0, 60.0, 400.0]
    lab, tc = t(lambda: ctx.fcluster(thr))
    for k, th in enumerate(thr):
        assert lab[k].max() == n - int((Z[:, 2] <= th).sum())
        for q, r in enumerate(rows):
            near = np.flatnonzero(sq[q] <= th)
            assert (lab[k][near] == lab[k][r]).all()
    print(f"cut at {len(thr)} thresholds: {tc * 1e3:.1f} ms; clusters "
          + ", ".join(f"t={int(th)}: {int(lab[k].max())}" for k, th in enumerate(thr)))
print("maxsize ok")
