"""On-box: end-to-end CLI wall time on a synthetic TSV (BASELINE metric 2), phases on stderr.
usage: python scripts/gpu_cli_e2e.py C3|C4|<n>x<L> [--skip-dist] [extra CLI args...]"""
import os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
from reportree_b200 import synth

what = sys.argv[1]
extra = [a for a in sys.argv[2:] if a != "--skip-dist"]
if "x" in what:
    n, L = (int(v) for v in what.split("x")); X = synth.allele_matrix(n, L, 7)
else:
    X = synth.allele_matrix(**synth.config(what))
n, L = X.shape
work = os.environ.get("E2E_DIR", "/tmp/rthc_e2e"); os.makedirs(work, exist_ok=True)
tsv = os.path.join(work, "%s.tsv" % what)
t0 = time.perf_counter()
if not os.path.exists(tsv):
    from reportree_b200 import fastio
    fastio.write_int_matrix(tsv, X)
print("TSV %s: %d x %d, %.1f MB, written in %.1f s" % (tsv, n, L, os.path.getsize(tsv) / 1e6, time.perf_counter() - t0), flush=True)
env = dict(os.environ, REPORTREE_HC_TIMING="1", PYTHONPATH=ROOT)
if "--skip-dist" in sys.argv:
    env["REPORTREE_HC_SKIP_DIST_TSV"] = "1"
args = extra or ["--HC-threshold", "single-1-50"]
if what == "C3" and not extra:
    args = ["--HC-threshold", "single-1-50,average-1-50"]                  # configs[2]: single + average
if what == "C5" and not extra:                                             # configs[4]: --loci subset + --loci-called 0.95, complete
    loci = os.path.join(work, "C5_loci.txt")
    rng = np.random.default_rng(1)
    cols = rng.permutation(L)[:3000]
    with open(loci, "w") as fh:
        fh.write("".join("L%05d\n" % (c + 1) for c in cols))
    args = ["--HC-threshold", "complete-1-50", "-l", loci, "--loci-called", "0.95"]
t0 = time.perf_counter()
rc = subprocess.call([sys.executable, "-m", "reportree_b200.partitioning_HC", "-a", tsv, "-o", os.path.join(work, what)] + args,
                     env=env, stdout=subprocess.DEVNULL)
print("CLI rc=%d wall %.2f s (process start -> exit)  args=%s skip_dist=%s" % (rc, time.perf_counter() - t0, args, "--skip-dist" in sys.argv), flush=True)
for f in sorted(os.listdir(work)):
    if f.startswith(what + "_") or f == what + ".log":
        print("   %-40s %12d bytes" % (f, os.path.getsize(os.path.join(work, f))))
