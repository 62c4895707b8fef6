"""On-box: what the shared-code plan (collide.cu) buys per workload and per cost divisor (RTHC_COLLIDE_DIV):
plane rows, re-code time, plane-kernel time, fix-up time, listed cells, fixed-up pairs."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from reportree_b200 import hc_backend as hc, synth
names = [a for a in sys.argv[1:] if not a.isdigit()] or ["C4"]
divs = [a for a in sys.argv[1:] if a.isdigit()] or ["0", "6144", "12288", "24576", "49152", "98304"]
for name in names:
    X = synth.workload(name)
    for d in divs:
        os.environ.pop("RTHC_NO_ESCAPE", None); os.environ.pop("RTHC_COLLIDE_DIV", None)
        if d == "0": os.environ["RTHC_NO_ESCAPE"] = "1"
        else: os.environ["RTHC_COLLIDE_DIV"] = d
        with hc.HCContext() as ctx:
            for rep in range(3):
                ctx.set_profiles(X)
                ctx.distance()
            t, i = ctx.timings(), ctx.info()
            print("%s div %6s  plane_rows %4d  recode %.2f ms  planes %.2f ms  fixup %.2f ms  sum %.2f  loci %d cells %d pairs %d" % (
                name, d, i["plane_rows"], t["recode_ms"], t["distance_ms"], t["escape_fixup_ms"],
                t["recode_ms"] + t["distance_ms"] + t["escape_fixup_ms"], i["escape_loci"], i["escape_cells"], i["escape_pairs"]), flush=True)
