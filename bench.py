#!/usr/bin/env python3
"""bench.py -- allele comparisons/s of the partitioning hot path (BASELINE.json metric).

One "step" = one full pass of the hot path over the synthetic 100,000 x 3,000 cgMLST matrix
(BASELINE.json configs[3]; fits one GPU): device re-coder -> bit-sliced distance kernel ->
condensed uint16 matrix -> Boruvka MST (single linkage) -> cut at thresholds 0..50.

  value : P*L / step time with the int32 code matrix already resident in HBM
  e2e   : same, through the host C-ABI call with PINNED HOST buffers: H2D of the codes and
          D2H of Z + labels inside the timed region
  N > 1 : strong scaling -- 128-sample tile-rows of the upper triangle sharded over the ranks,
          per-rank Boruvka forest, one NCCL all_gather of <= n-1 packed edges per rank, final
          MST + cut on rank 0 (launched by torchrun, one process per GPU)

  --impl reference : the reference's CPU path (oracle/ C port of cgmlst-dists + scipy
          linkage/fcluster) on a bounded sample of the same workload, all host threads.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WORKLOAD = dict(name="C4 synthetic cgMLST 100000x3000, 2% missing, single linkage, cut 0..50",
                cfg="C4", thresholds=51)


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=0, help="override sample count (debug only; invalidates the metric)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--next", action="store_true",
                    help="measure the SURVEY 8(f) N3 / N4 components (zoom-in, cluster stability) instead of the hot path")
    ap.add_argument("--next-workload", default="C4", choices=["C3", "C4"])
    return ap.parse_args()


def make_workload(args):
    from reportree_b200 import synth
    cfg = synth.config(WORKLOAD["cfg"])
    if args.n:
        cfg["n"] = args.n
    X = synth.allele_matrix(**cfg)
    return X, cfg


class ClockSampler:
    """SM clocks + throttle reasons during the timed region (B200_PROFILING.md clocks line), read
    through NVML in-process (the same counters nvidia-smi prints; spawning nvidia-smi 5x per second
    was measured to stall the launch-heavy MST stage by up to 8x through driver lock contention)."""

    def __init__(self, index, period=0.5):
        self.index, self.period = index, period
        self.sm, self.mx, self.reasons = [], [], set()
        self._stop = threading.Event()
        self._thr = None

    def _run(self):
        try:
            import pynvml as N
            N.nvmlInit()
            h = N.nvmlDeviceGetHandleByIndex(self.index)
            names = {N.nvmlClocksEventReasonHwSlowdown: "hw_slowdown",
                     N.nvmlClocksEventReasonHwThermalSlowdown: "hw_thermal_slowdown",
                     N.nvmlClocksEventReasonSwThermalSlowdown: "sw_thermal_slowdown",
                     N.nvmlClocksEventReasonSwPowerCap: "sw_power_cap"}
            while not self._stop.is_set():
                self.sm.append(N.nvmlDeviceGetClockInfo(h, N.NVML_CLOCK_SM))
                self.mx.append(N.nvmlDeviceGetMaxClockInfo(h, N.NVML_CLOCK_SM))
                r = N.nvmlDeviceGetCurrentClocksEventReasons(h)
                self.reasons |= {v for k, v in names.items() if r & k}
                self._stop.wait(self.period)
        except Exception as e:                      # no NVML: one nvidia-smi sample per 2 s
            self.err = repr(e)
            q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
                 "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
            names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
            while not self._stop.is_set():
                try:
                    out = subprocess.run(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + q,
                                          "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                    parts = [p.strip() for p in out.strip().split(",")]
                    self.sm.append(float(parts[0])); self.mx.append(float(parts[1]))
                    self.reasons |= {names[k] for k in range(4) if parts[2 + k].lower().startswith("active")}
                except Exception:
                    pass
                self._stop.wait(2.0)

    def start(self):
        self._thr = threading.Thread(target=self._run, daemon=True)
        self._thr.start()

    def stop(self):
        self._stop.set()
        if self._thr:
            self._thr.join(timeout=6)
        return {"sm_mhz": float(np.median(self.sm)) if self.sm else None, "sm_max_mhz": max(self.mx) if self.mx else None,
                "reasons": sorted(self.reasons), "samples": len(self.sm), "source": "nvml" if not hasattr(self, "err") else "nvidia-smi"}


# ------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle port of the reference's CPU path
# ------------------------------------------------------------------------------------------

def cpu_reference_step(X, n_sample, threads, thresholds):
    """cgmlst-dists rule (C port) + scipy linkage/fcluster on the first n_sample rows."""
    from concurrent.futures import ThreadPoolExecutor
    from scipy.cluster.hierarchy import fcluster, linkage
    from oracle import oracle_lib as O
    Xs = np.ascontiguousarray(X[:n_sample])
    n, L = Xs.shape
    t0 = time.perf_counter()
    # balanced row ranges (row i costs n-1-i pairs)
    cost = np.cumsum((n - 1 - np.arange(n)).astype(np.float64))
    cuts = [0] + [int(np.searchsorted(cost, cost[-1] * k / threads)) for k in range(1, threads)] + [n]
    cuts = sorted(set(cuts))
    with ThreadPoolExecutor(max_workers=threads) as ex:
        parts = list(ex.map(lambda ab: O.dist_condensed_rows(Xs, ab[0], ab[1]), zip(cuts[:-1], cuts[1:])))
    y = np.concatenate(parts).astype(np.float64)
    t_dist = time.perf_counter() - t0
    Z = linkage(y, "single")
    for t in range(thresholds):
        fcluster(Z, t, criterion="distance")
    t_all = time.perf_counter() - t0
    P = n * (n - 1) // 2
    return P * L, t_all, t_dist


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    X, cfg = make_workload(args)
    threads = os.cpu_count() or 1
    n_sample = 4000
    vals = []
    for k in range(args.warmup + args.steps):
        work, t_all, t_dist = cpu_reference_step(X, n_sample, threads, WORKLOAD["thresholds"])
        if k >= args.warmup:
            vals.append((work, t_all))
    tot_work = sum(v[0] for v in vals)
    tot_t = sum(v[1] for v in vals)
    value = tot_work / tot_t
    sample = "first %d of %d samples x %d loci per step (%.2e comparisons), distance + scipy single linkage + %d cuts" % (
        n_sample, cfg["n"], cfg["L"], vals[0][0], WORKLOAD["thresholds"])
    line = {"impl": "reference", "metric": "allele comparisons/sec (pairs x loci / s), whole hot path",
            "value": value, "unit": "comparisons/s", "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * tot_t / len(vals), "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "int32", "data": "synthetic",
            "config": {"workload": WORKLOAD["name"], "n": cfg["n"], "L": cfg["L"], "sample": sample},
            "cpu_baseline": {"value": value, "unit": "comparisons/s", "cores": threads, "kind": "port", "sample": sample},
            "e2e": {"value": value, "unit": "comparisons/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    print(json.dumps(line))


# ------------------------------------------------------------------------------------------
# our arm
# ------------------------------------------------------------------------------------------

def run_ours(args):
    import torch
    import torch.distributed as dist
    from reportree_b200 import hc_backend as hc, sharded

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device; the product path has no CPU fallback")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    X, cfg = make_workload(args)
    n, L = X.shape
    P = n * (n - 1) // 2
    pinned = torch.empty((n, L), dtype=torch.int32, pin_memory=True)
    pinned.numpy()[:] = X
    host_codes = pinned.numpy()
    dX = pinned.cuda(non_blocking=False)
    thr = np.arange(WORKLOAD["thresholds"], dtype=np.float64)
    labels_host = torch.empty((thr.size, n), dtype=torch.int32, pin_memory=True).numpy()   # result buffer, reused every step
    ctx = hc.HCContext(device=local, rank=rank, world=world)
    state = {"dist_ms": [], "link_ms": [], "cut_ms": [], "recode_ms": [], "h2d_ms": [], "d2h_bytes": 0, "Z": None}
    exch = sharded.ForestExchange(n, dist) if world > 1 else None                 # device-resident forest gather
    loader = sharded.ShardedProfileLoader(n, L, dist) if world > 1 else None      # 1/world of the rows per rank + all_gather
    planes_x = sharded.PlaneExchange(n, L, dist) if sharded.PlaneExchange.usable(L, world) else None   # re-coder sharded by loci

    seg = {}

    def lap(name, t0):
        t1 = time.perf_counter()
        seg.setdefault(name, []).append(1e3 * (t1 - t0))
        return t1

    def step(resident):
        t = time.perf_counter()
        if resident:
            if planes_x is not None:
                planes_x.set_profiles(ctx, dX.data_ptr(), np.int32)
            else:
                ctx.set_profiles_device(dX.data_ptr(), n, L)
        elif world > 1:
            ptr = loader.load(host_codes)
            if planes_x is not None:
                planes_x.set_profiles(ctx, ptr, loader.dtype)
            else:
                ctx.set_profiles_device(ptr, n, L, loader.dtype)
        else:
            ctx.set_profiles(host_codes)
        t = lap("set_profiles", t)
        ctx.distance()
        t = lap("distance", t)
        d2h = 0
        if world == 1:
            Z = ctx.linkage("single")
            t = lap("linkage", t)
            labels = ctx.fcluster(thr, out=labels_host)
            t = lap("cut", t)
            state["Z"] = Z
            d2h = Z.nbytes + labels.nbytes + (n - 1) * 8
        else:
            exch.local_forest(ctx)
            t = lap("local_forest", t)
            exch.gather()                                                        # the one NCCL exchange
            t = lap("all_gather", t)
            labels = None
            if rank == 0:
                Z = exch.linkage(ctx)
                state["Z"] = Z
                t = lap("final_merge", t)
                labels = ctx.fcluster(thr, out=labels_host)
                t = lap("cut", t)
                d2h = (n - 1) * 8 + Z.nbytes + labels.nbytes
        tm = ctx.timings()
        state["dist_ms"].append(tm["distance_ms"]); state["link_ms"].append(tm["linkage_ms"])
        state["cut_ms"].append(tm["cut_ms"]); state["recode_ms"].append(tm["recode_ms"])
        state["h2d_ms"].append(tm["h2d_ms"]); state["d2h_bytes"] = d2h
        return labels

    def sync():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize()

    def timed(resident, steps, warmup):
        for _ in range(warmup):
            step(resident)
        for k in state:
            if isinstance(state[k], list):
                state[k].clear()
        seg.clear()
        l0 = ctx.timings()["launches"]
        sync()
        ctx.mark(0)                       # CUDA events on the library's own stream
        t0 = time.perf_counter()
        for _ in range(steps):
            labels = step(resident)
        ctx.mark(1)
        sync()
        wall = time.perf_counter() - t0
        dt = ctx.mark_elapsed_ms() * 1e-3
        state["wall_over_device"] = wall / dt if dt > 0 else None
        launches = ctx.timings()["launches"] - l0
        if world > 1:
            t = torch.tensor([dt, launches], dtype=torch.float64, device="cuda")
            tmax = t.clone()
            dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
            tsum = t.clone()
            dist.all_reduce(tsum, op=dist.ReduceOp.SUM)
            dt, launches = float(tmax[0]), float(tsum[1])
        return dt, launches, labels

    alu_peak = ctx.alu_peak()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    dt, launches, labels = timed(True, args.steps, args.warmup)
    check = {"resident": result_check(labels, state["Z"])} if rank == 0 else {}
    kern = {k: float(np.mean(v)) for k, v in state.items() if isinstance(v, list) and v}
    kern_steps = {k: [round(float(x), 3) for x in v] for k, v in state.items() if isinstance(v, list) and v}
    host_seg = {k: round(float(np.mean(v)), 3) for k, v in seg.items()}
    host_seg_all = None
    if world > 1:                                   # where each rank spends its step (a slow rank shows up as waiting time of the others)
        host_seg_all = [None] * world
        dist.all_gather_object(host_seg_all, dict(host_seg, dist_kernel=round(float(np.mean(state["dist_ms"])), 3),
                                                  recode=round(float(np.mean(state["recode_ms"])), 3)))
    info = ctx.info()
    owned_pairs = ctx.owned_pairs()
    clocks = sampler.stop() if rank == 0 else None
    dt_e2e, _, labels_e2e = timed(False, args.steps, 1)
    if rank == 0:
        check["e2e"] = result_check(labels_e2e, state["Z"])
        check["oracle_match"] = bool(check["resident"]["oracle_match"] and check["e2e"]["oracle_match"]) \
            if check["resident"]["oracle_match"] is not None else None
    d2h_bytes = state["d2h_bytes"]
    exact = None
    if world == 1:
        # the drop-in CLI's default (REPORTREE_HC_EXACT_NAMES=1) is the scipy-order Prim, whose Z and cluster ids are
        # byte-identical to scipy's; the headline number above is Boruvka mode (same partitions, ids may differ on ties)
        ctx.set_single_algo(hc.SL_PRIM)
        ks = min(args.steps, 3)
        dt_p, _, labels_p = timed(True, ks, 1)
        ctx.set_single_algo(hc.SL_BORUVKA)
        exact = {"ms_per_step": 1e3 * dt_p / ks, "value": float(P) * L * ks / dt_p, "unit": "comparisons/s", "steps": ks,
                 "linkage_ms": float(np.mean(state["link_ms"])), "what": "same step with the scipy-order Prim linkage (CLI default)"}
        gpath = os.path.join(ROOT, "tests", "golden", "large", WORKLOAD["cfg"] + ".json")
        if os.path.exists(gpath) and n == synth_n():
            from reportree_b200 import checksums as C
            g = json.load(open(gpath))["methods"]["single"]
            exact["oracle_ids_match"] = bool(C.crc(state["Z"]) == g["z_crc32"] and all(
                C.crc(labels_p[k]) == g["labels"][k]["ids_crc32"] for k in range(labels_p.shape[0])))

    if rank == 0:
        work = float(P) * L
        value = work * args.steps / dt
        e2e_value = work * args.steps / dt_e2e
        t_k = kern["dist_ms"] * 1e-3
        cmp_rate = owned_pairs * float(L) / t_k
        lop3_rate = owned_pairs * float(info["plane_rows"]) / t_k
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        hbm_peak = peaks.get("hbm_gbs", 6650.0)
        alg_bytes = 2.0 * n * L + 2.0 * owned_pairs
        roofline = {
            "kernel": "rthc_distance_ws",
            "bound": "int_alu",
            "achieved": cmp_rate / 1e12, "peak": alu_peak / 2 / 1e12, "unit": "Tcmp/s",
            "frac": cmp_rate / (alu_peak / 2),
            "note": "peak = measured LOP3 lane-ops/s / 2 (SURVEY 8d: 2 lane-ops per comparison for a direct "
                    "compare+accumulate); the bit-sliced kernel issues plane_rows/L_kept*... LOP3 per 32 loci, "
                    "so frac can exceed 1; alu_pipe_frac is the true pipe utilisation",
            "alu_pipe_frac": lop3_rate / alu_peak,
            "alu_peak_lane_ops_per_s": alu_peak, "alu_peak_source": "measured live (rthc_alu_peak LOP3 stream)",
            "lop3_per_32_loci": info["plane_rows"] / max(info["groups"], 1),
            "kernel_ms": kern["dist_ms"],
            # dram__bytes_read.sum + dram__bytes_write.sum of this kernel: not measurable from inside the run (needs
            # the profiler), so it is READ from the summary that scripts/gpu_round.sh extracts from this round's
            # `ncu --set full` capture of the same command (profiles/<round>_ncu_distance_C4.json) and labelled
            **ncu_traffic(n, world),
            "algorithmic_bytes": alg_bytes,
            "hbm": {"achieved": alg_bytes / t_k / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "frac": alg_bytes / t_k / 1e9 / hbm_peak,
                    "peak_source": "MEASURED_PEAKS.json" if peaks else "fallback"},
        }
        line = {
            "metric": "allele comparisons/sec (pairs x loci / s), whole hot path",
            "value": value, "unit": "comparisons/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "u16 codes / u32 bit planes (integer)", "data": "synthetic",
            "config": {"workload": WORKLOAD["name"], "n": n, "L": L, "pairs": P, "thresholds": WORKLOAD["thresholds"],
                       "l2": "inputs larger than L2 (bit planes %.0f MB + condensed output %.1f GB per step)" % (
                           info["plane_rows"] * ctx_npad(n) * 4 / 1e6, 2.0 * owned_pairs / 1e9),
                       "sharding": "tile-rows of the upper triangle over %d rank(s)" % world + (
                           "; re-coder sharded by locus ranges, planes gathered over NCCL" if planes_x is not None else ""),
                       "single_linkage": "Boruvka MST (north_star; same partitions as scipy, cluster ids may differ on ties); "
                                         "exact_names_mode times the CLI-default scipy-order Prim"},
            "stage_ms": kern,
            "stage_ms_per_step": kern_steps,
            "host_wall_ms_rank0": host_seg, "host_wall_ms_all_ranks": host_seg_all,
            "e2e_stage_ms_per_step": {k: [round(float(x), 3) for x in v] for k, v in state.items() if isinstance(v, list) and v},
            "roofline": roofline,
            "e2e": {"value": e2e_value, "unit": "comparisons/s", "ms_per_step": 1e3 * dt_e2e / args.steps,
                    "h2d_bytes_per_step": int(n) * int(L) * 4, "d2h_bytes_per_step": int(d2h_bytes),
                    "h2d": "rank 0 copies the whole int32 matrix" if world == 1 else
                           "summed over ranks: each copies n/%d rows (int32) from page-locked memory, narrows them to uint16 on the device when "
                           "every id fits, one NCCL all_gather replicates" % world},
            "result_check": check,
            "gpu_launches": int(launches),
            "timing": "CUDA events on the library stream around the K steps (max over ranks); wall/device = %.3f" % (
                state.get("wall_over_device") or 0.0),
            "clocks": clocks,
        }
        if world == 1 and not args.no_cpu_baseline:
            ns = 7000
            work_s, t_all, t_dist = cpu_reference_step(X, ns, 1, WORKLOAD["thresholds"])
            line["cpu_baseline"] = {"value": work_s / t_all, "unit": "comparisons/s", "cores": 1, "kind": "port",
                                    "sample": "first %d samples x %d loci, C port of the cgmlst-dists rule + scipy "
                                              "single linkage + %d cuts (%.1f s)" % (ns, L, WORKLOAD["thresholds"], t_all)}
        if exact:
            line["exact_names_mode"] = exact
        ref_script = reference_script_timing()
        if ref_script:
            line["cpu_baseline_script"] = ref_script
        print(json.dumps(line))
    ctx.close()
    if world > 1:
        dist.destroy_process_group()


def result_check(labels, Z=None):
    """Fingerprint of the T x n label matrix (taken outside the timed region) and its comparison with the CPU
    oracle's answer for this workload, committed as tests/golden/large/C4.json (tests/golden/make_golden_large.py:
    oracle distances + oracle single linkage + oracle fcluster at thresholds 0..50).  The benchmarked path is
    Boruvka mode, whose cluster numbering may differ from scipy's on ties, so what is compared per threshold is
    the cluster count and the CRC32 of the membership-only relabelling, plus the CRC32 of the merge heights."""
    import zlib
    from reportree_b200 import checksums as C
    out = {"labels_crc32": zlib.crc32(np.ascontiguousarray(labels).tobytes()),
           "clusters_at_first_mid_last_threshold": [int(labels[k].max()) for k in (0, labels.shape[0] // 2, labels.shape[0] - 1)],
           "oracle_match": None}
    gpath = os.path.join(ROOT, "tests", "golden", "large", WORKLOAD["cfg"] + ".json")
    if os.path.exists(gpath) and labels.shape[1] == synth_n():
        g = json.load(open(gpath))
        gm = g["methods"]["single"]
        ok = len(g["thresholds"]) == labels.shape[0]
        bad = []
        for k in range(labels.shape[0] if ok else 0):
            f = C.labels_fingerprint(labels[k])
            if f["clusters"] != gm["labels"][k]["clusters"] or f["membership_crc32"] != gm["labels"][k]["membership_crc32"]:
                bad.append(k)
        heights_ok = None
        if Z is not None:
            heights_ok = C.crc(np.ascontiguousarray(Z[:, 2])) == gm["heights_crc32"]
        out["oracle_match"] = bool(ok and not bad and heights_ok is not False)
        out["oracle"] = {"golden": "tests/golden/large/%s.json" % WORKLOAD["cfg"], "thresholds_compared": labels.shape[0],
                         "thresholds_differing": bad[:5], "merge_heights_equal": heights_ok}
    return out


def synth_n():
    from reportree_b200 import synth
    return synth.config(WORKLOAD["cfg"])["n"]


def ncu_traffic(n, world):
    """{"traffic": bytes, "traffic_source": file} from the newest profiles/*_ncu_distance_C4.json, else nulls."""
    import glob
    if n != synth_n() or world != 1:
        return {"traffic": None, "traffic_source": None}
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_ncu_distance_C4.json")))
    if not files:
        return {"traffic": None, "traffic_source": None}
    rec = json.load(open(files[-1]))
    return {"traffic": rec["dram_bytes_read"] + rec["dram_bytes_write"],
            "traffic_source": "ncu --set full capture, " + os.path.relpath(files[-1], ROOT)}


def reference_script_timing():
    """Wall time of the UNMODIFIED reference script on a bounded TSV (scripts/time_reference_script.py, run in the
    build container because /root/reference does not travel to the GPU box), newest profiles/*_reference_script.json."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "r*_reference_script.json")))
    return json.load(open(files[-1])) if files else None


def ctx_npad(n):
    return (n + 127) // 128 * 128


def run_next(args):
    """zoom-in + stability timings (scripts/gpu_next_bench.py) with this file's cpu_baseline ports handed in."""
    import importlib.util
    spec = importlib.util.spec_from_file_location("gpu_next_bench", os.path.join(os.path.dirname(os.path.abspath(__file__)),
                                                                                 "scripts", "gpu_next_bench.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    cpu = None
    if not args.no_cpu_baseline:
        from oracle import oracle_lib, partition_metrics

        class cpu:                                # noqa: N801 -- a namespace of the oracle ports
            stability_lines = staticmethod(partition_metrics.stability_lines)
            dist_square = staticmethod(oracle_lib.dist_square)
    print(json.dumps(mod.run(args.next_workload, cpu)))


if __name__ == "__main__":
    a = parse_args()
    if a.next:
        run_next(a)
    elif a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
