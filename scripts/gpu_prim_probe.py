"""Prim-exact (scipy-order single linkage) timing: cluster kernel (mirror copy / condensed gather) vs the
cooperative-grid kernel.  RTHC_PRIM_CLUSTER is read once per process, so every mode is its own process.
usage: python scripts/gpu_prim_probe.py [n] [L]     (parent)"""
import os
import subprocess
import sys
import time

import numpy as np

sys.path.insert(0, ".")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 100000
L = int(sys.argv[2]) if len(sys.argv) > 2 else 3000

if os.environ.get("PRIM_CHILD"):
    from reportree_b200 import hc_backend as hc, synth
    X = synth.allele_matrix(n=n, L=L, seed=20260304, missing=0.02)
    with hc.HCContext() as ctx:
        ctx.set_profiles(X)
        ctx.distance()
        ctx.set_single_algo(hc.SL_PRIM)
        ts = []
        for _ in range(3):
            t0 = time.perf_counter(); Z = ctx.linkage("single"); ts.append(time.perf_counter() - t0)
    np.save(os.environ["PRIM_CHILD"], Z)
    print("mode %s: n=%d linkage(single, Prim-exact) %s s" % (os.environ.get("RTHC_PRIM_CLUSTER", "1"), n,
                                                             " ".join("%.3f" % t for t in ts)), flush=True)
else:
    outs = []
    for mode in ("0", "1", "2"):
        out = "/tmp/prim_Z_%s.npy" % mode
        subprocess.check_call([sys.executable, __file__, str(n), str(L)], env=dict(os.environ, RTHC_PRIM_CLUSTER=mode, PRIM_CHILD=out))
        outs.append(np.load(out))
    print("Z identical across modes:", all(a.tobytes() == outs[0].tobytes() for a in outs[1:]))
