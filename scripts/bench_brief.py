import sys, json
for l in sys.stdin:
    if l.startswith("{"):
        d = json.loads(l)
        print("value %.2f Tcmp/s  step %.1f ms  stages %s  kernel %.2f Tcmp/s alu_frac %.3f  e2e %.1f ms" % (
            d["value"] / 1e12, d["ms_per_step"], {k: round(v, 2) for k, v in d["stage_ms"].items()},
            d["roofline"]["achieved"], d["roofline"]["alu_pipe_frac"], d["e2e"]["ms_per_step"]))
    else:
        print(l, end="")
