"""Summarise an .ncu-rep (read on the CPU box): key raw metrics per kernel + top stall sites.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep > profiles/rNN_name.txt"""
import csv, subprocess, sys, io

rep = sys.argv[1]
json_out = sys.argv[sys.argv.index("--json") + 1] if "--json" in sys.argv else None      # {kernel, ms, dram bytes} of the first kernel
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
KEYS = ["gpu__time_duration.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__bytes_read.sum.per_second", "dram__bytes_write.sum.per_second",
        "lts__t_bytes.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smsp__inst_executed.sum"]
for r in rows[2:]:
    d = dict(zip(hdr, r))
    print("== kernel:", d.get("Kernel Name", "?")[:100])
    if json_out:
        import json

        def num(key, scale):
            unit = units[hdr.index(key)]
            mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, scale)
            return float(d[key].replace(",", "")) * mult
        json.dump({"kernel": d.get("Kernel Name", "?")[:60], "capture": rep.split("/")[-1], "kernel_ms": num("gpu__time_duration.sum", 1e-6),
                   "dram_bytes_read": num("dram__bytes_read.sum", 1), "dram_bytes_write": num("dram__bytes_write.sum", 1),
                   "how": "ncu --set full --clock-control none, one launch; read with ncu -i --page raw --csv"}, open(json_out, "w"), indent=1)
        json_out = None
    for k in KEYS:
        if k in d:
            print("  %-72s %s %s" % (k, d[k], units[hdr.index(k)]))
    for k in hdr:
        if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("per_issue_active.ratio") and float(d[k] or 0) > 0.05:
            print("  %-72s %s" % (k.replace("smsp__average_warps_issue_stalled_", "stall/issue: ").replace("_per_issue_active.ratio", ""), d[k]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
# the source page repeats "Kernel Name" + header per kernel
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "Kernel Name":
        name = rows[i][1][:80]
        hdr = rows[i + 1]
        ix = {h: k for k, h in enumerate(hdr)}
        j = i + 2
        data = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "Kernel Name"):
            if len(rows[j]) == len(hdr):
                data.append(rows[j])
            j += 1
        tot = sum(int(r[ix["# Samples"]]) for r in data) or 1
        print("== top stall sites:", name, "samples", tot)
        skeys = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
        agg = {k: sum(int(r[ix[k]]) for r in data) / tot for k in skeys}
        print("  overall:", {k[6:]: round(v, 3) for k, v in agg.items() if v > 0.01})
        for r in sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[:12]:
            st = {k[6:]: int(r[ix[k]]) for k in skeys if int(r[ix[k]]) > 0.1 * int(r[ix["# Samples"]])}
            print("  %6.2f%%  %-58s x%s %s" % (100 * int(r[ix["# Samples"]]) / tot, r[ix["Source"]].strip()[:58],
                                               r[ix["Instructions Executed"]], st))
        i = j
    else:
        i += 1
