"""ctypes binding of libreportree_hc.so (include/rthc.h) with the call shapes that
scripts/partitioning_HC.py uses today:

    cgmlst-dists subprocess + chunk loop  (scripts/partitioning_HC.py:489-516)  -> HCContext.set_profiles / distance
    squareform(dist)                      (:186)                                 -> HCContext.condensed
    linkage(y, method)                    (:197)                                 -> HCContext.linkage(method)
    maxdists(Z)[-1]                       (:198)                                 -> Z[-1, 2] (monotone methods)
    fcluster(Z, t, criterion="distance")  (:649, :689, :722, :751)               -> HCContext.fcluster(thresholds)

There is no CPU fallback: if the CUDA library is missing or no B200 is visible, every
entry point raises HCBackendError.
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RTHC_LIB_PATH") or os.path.join(_HERE, "libreportree_hc.so")   # override: kernel experiments only

METHODS = {"single": 0, "complete": 1, "average": 2, "weighted": 3, "ward": 4}
SL_BORUVKA, SL_PRIM = 0, 1


class HCBackendError(RuntimeError):
    pass


_lib = None


def load_library():
    """dlopen the in-tree CUDA library and declare every symbol of include/rthc.h."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise HCBackendError(
            "CUDA library %s not built (run `python -m reportree_b200.build`); there is no CPU fallback" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, i64, i32, ci = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int32, ctypes.c_int
    sig = {
        "rthc_create": (ci, [ctypes.POINTER(vp), ci]),
        "rthc_destroy": (None, [vp]),
        "rthc_last_error": (ctypes.c_char_p, [vp]),
        "rthc_version": (ci, []),
        "rthc_set_shard": (ci, [vp, ci, ci]),
        "rthc_set_shard_full_storage": (ci, [vp, ci, ci]),
        "rthc_import_shard": (ci, [vp, vp]),
        "rthc_finish_import": (ci, [vp]),
        "rthc_set_profiles_i32": (ci, [vp, vp, i64, i64]),
        "rthc_set_profiles_i32_device": (ci, [vp, vp, i64, i64]),
        "rthc_set_profiles_u16": (ci, [vp, vp, i64, i64]),
        "rthc_set_profiles_u16_device": (ci, [vp, vp, i64, i64]),
        "rthc_shard_chunk": (i64, [i64, ci]),
        "rthc_recode_shard": (ci, [vp, vp, ci, i64, i64, vp, i64]),
        "rthc_export_shard": (ci, [vp, vp, i64, vp, vp]),
        "rthc_import_planes": (ci, [vp, vp, i64, vp, i64, vp, vp]),
        "rthc_distance": (ci, [vp]),
        "rthc_owned_pairs": (i64, [vp]),
        "rthc_get_condensed_u16": (ci, [vp, vp]),
        "rthc_get_owned_condensed_u16": (ci, [vp, vp]),
        "rthc_get_condensed_range_u16": (ci, [vp, i64, i64, vp]),
        "rthc_copy_rows_u16": (ci, [vp, i64, i64, vp]),
        "rthc_direct_pairs": (ci, [vp, vp, vp, i64, vp]),
        "rthc_set_condensed_f64": (ci, [vp, vp, i64]),
        "rthc_set_single_algo": (ci, [vp, ci]),
        "rthc_linkage": (ci, [vp, ci, vp]),
        "rthc_local_forest": (ci, [vp, vp, ctypes.POINTER(i64)]),
        "rthc_linkage_from_forest": (ci, [vp, vp, i64, i64, vp]),
        "rthc_local_forest_device": (ci, [vp, vp, i64, ctypes.POINTER(i64)]),
        "rthc_linkage_from_forest_device": (ci, [vp, vp, i64, i64, vp]),
        "rthc_cut": (ci, [vp, vp, ci, vp]),
        "rthc_info": (ci, [vp, vp, ci]),
        "rthc_timings": (ci, [vp, vp, ci]),
        "rthc_mark": (ci, [vp, ci]),
        "rthc_mark_elapsed_ms": (ci, [vp, ctypes.POINTER(ctypes.c_double)]),
        "rthc_alu_peak": (ci, [vp, ctypes.POINTER(ctypes.c_double)]),
        "rthc_alu_peak_block": (ci, [vp, ci, ctypes.POINTER(ctypes.c_double)]),
        "rthc_closest": (ci, [vp, i64, i64, vp, vp, ctypes.POINTER(i64), ctypes.POINTER(ci)]),
        "rthc_closest_noalloc": (ci, [vp, i64, i64, vp, vp, ctypes.POINTER(i64), ctypes.POINTER(ci)]),
        "rthc_subset_from": (ci, [vp, vp, vp, i64]),
        "rthc_contingency_adjacent": (ci, [vp, vp, ci, i64, vp, vp, vp, vp]),
        "rthc_adjusted_wallace_sums": (ci, [vp, vp, i64, ctypes.POINTER(ctypes.c_double), ctypes.POINTER(ctypes.c_double)]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


EXPORTED_SYMBOLS = [
    "rthc_create", "rthc_prewarm", "rthc_destroy", "rthc_last_error", "rthc_version", "rthc_set_shard",
    "rthc_set_shard_full_storage", "rthc_import_shard", "rthc_finish_import",
    "rthc_set_profiles_i32", "rthc_set_profiles_i32_device", "rthc_set_profiles_u16", "rthc_set_profiles_u16_device",
    "rthc_shard_chunk", "rthc_recode_shard", "rthc_export_shard", "rthc_import_planes",
    "rthc_distance", "rthc_owned_pairs",
    "rthc_get_condensed_u16", "rthc_get_owned_condensed_u16", "rthc_get_condensed_range_u16", "rthc_copy_rows_u16", "rthc_direct_pairs",
    "rthc_set_condensed_f64", "rthc_set_single_algo", "rthc_linkage", "rthc_local_forest",
    "rthc_linkage_from_forest", "rthc_local_forest_device", "rthc_linkage_from_forest_device", "rthc_cut", "rthc_info", "rthc_timings", "rthc_mark", "rthc_mark_elapsed_ms", "rthc_alu_peak",
    "rthc_alu_peak_block", "rthc_closest", "rthc_closest_noalloc", "rthc_subset_from",
    "rthc_contingency_adjacent", "rthc_adjusted_wallace_sums",
]


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


class HCContext:
    """One device context (one GPU).  Not thread-safe; calls are synchronous."""

    def __init__(self, device=0, rank=0, world=1, full_storage=False):
        self._L = load_library()
        h = ctypes.c_void_p()
        rc = self._L.rthc_create(ctypes.byref(h), int(device))
        if rc != 0:
            raise HCBackendError("rthc_create failed (%d): %s" % (rc, self._L.rthc_last_error(None).decode()))
        self._h = h
        self.device = int(device)
        self.n = 0
        self.L = 0
        if full_storage:
            self._check(self._L.rthc_set_shard_full_storage(self._h, int(rank), int(world)))
        elif world != 1:
            self._check(self._L.rthc_set_shard(self._h, int(rank), int(world)))
        self.rank, self.world = rank, world

    def _check(self, rc):
        if rc != 0:
            raise HCBackendError("rthc error %d: %s" % (rc, self._L.rthc_last_error(self._h).decode()))

    def close(self):
        if getattr(self, "_h", None):
            self._L.rthc_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- distance half ---------------------------------------------------------------
    def set_profiles(self, codes):
        """codes: n x L integer allele codes (0 = missing), host numpy array.  A uint16 array is sent as it
        is (half the host-to-device bytes); every other dtype goes through int32."""
        codes = np.asarray(codes)
        if codes.ndim != 2:
            raise HCBackendError("allele code matrix must be 2-D")
        self.n, self.L = codes.shape
        if codes.dtype == np.uint16:
            codes = np.ascontiguousarray(codes)
            self._check(self._L.rthc_set_profiles_u16(self._h, _ptr(codes), self.n, self.L))
        else:
            codes = np.ascontiguousarray(codes, dtype=np.int32)
            self._check(self._L.rthc_set_profiles_i32(self._h, _ptr(codes), self.n, self.L))

    def set_profiles_device(self, dev_ptr, n, L, dtype=np.int32):
        """The matrix is already in device memory (row-major int32, or uint16 with dtype=np.uint16)."""
        self.n, self.L = int(n), int(L)
        fn = self._L.rthc_set_profiles_u16_device if np.dtype(dtype) == np.uint16 else self._L.rthc_set_profiles_i32_device
        self._check(fn(self._h, ctypes.c_void_p(int(dev_ptr)), self.n, self.L))

    # ---- re-coder sharded by locus ranges (world > 1; driven by sharded.PlaneExchange) ----
    def recode_shard(self, dev_ptr, n, L, dtype, hdr):
        """This rank's loci of the device-resident n x L matrix -> planes; fills hdr (int32 numpy, see rthc.h)."""
        self.n, self.L = int(n), int(L)
        self._check(self._L.rthc_recode_shard(self._h, ctypes.c_void_p(int(dev_ptr)), int(np.dtype(dtype) == np.uint16),
                                              self.n, self.L, _ptr(hdr), hdr.size))

    def export_shard(self, planes_slot_ptr, rows_max, cnt_slot_ptr, list_slot_ptr):
        self._check(self._L.rthc_export_shard(self._h, ctypes.c_void_p(int(planes_slot_ptr)), int(rows_max),
                                              ctypes.c_void_p(int(cnt_slot_ptr)), ctypes.c_void_p(int(list_slot_ptr))))

    def import_planes(self, planes_ptr, rows_max, hdr_all, cnt_ptr, list_ptr):
        hdr_all = np.ascontiguousarray(hdr_all, dtype=np.int32)
        self._check(self._L.rthc_import_planes(self._h, ctypes.c_void_p(int(planes_ptr)), int(rows_max), _ptr(hdr_all),
                                               hdr_all.shape[1], ctypes.c_void_p(int(cnt_ptr)), ctypes.c_void_p(int(list_ptr))))

    def distance(self):
        self._check(self._L.rthc_distance(self._h))

    def import_shard(self, other):
        """Copy another context's owned rows (same matrix, any device) into this full-storage context."""
        self._check(self._L.rthc_import_shard(self._h, other._h))

    def finish_import(self):
        self._check(self._L.rthc_finish_import(self._h))
        self.rank, self.world = 0, 1

    def owned_pairs(self):
        return int(self._L.rthc_owned_pairs(self._h))

    def condensed(self):
        """Full condensed uint16 distance vector in scipy pdist order (squareform-compatible)."""
        out = np.empty(self.n * (self.n - 1) // 2, dtype=np.uint16)
        self._check(self._L.rthc_get_condensed_u16(self._h, _ptr(out)))
        return out

    def condensed_range(self, first, count):
        """Entries [first, first + count) of the (owned) condensed vector."""
        out = np.empty(int(count), dtype=np.uint16)
        self._check(self._L.rthc_get_condensed_range_u16(self._h, int(first), int(count), _ptr(out)))
        return out

    def owned_condensed(self):
        out = np.empty(self.owned_pairs(), dtype=np.uint16)
        self._check(self._L.rthc_get_owned_condensed_u16(self._h, _ptr(out)))
        return out

    def square_rows(self, row0, nrows):
        out = np.empty((nrows, self.n), dtype=np.uint16)
        self._check(self._L.rthc_copy_rows_u16(self._h, int(row0), int(nrows), _ptr(out)))
        return out

    def direct_pairs(self, pi, pj):
        pi = np.ascontiguousarray(pi, dtype=np.int32)
        pj = np.ascontiguousarray(pj, dtype=np.int32)
        out = np.empty(pi.size, dtype=np.int32)
        self._check(self._L.rthc_direct_pairs(self._h, _ptr(pi), _ptr(pj), pi.size, _ptr(out)))
        return out

    def set_condensed(self, y):
        y = np.ascontiguousarray(y, dtype=np.float64)
        n = int(round((1 + (1 + 8 * y.size) ** 0.5) / 2))
        if n * (n - 1) // 2 != y.size:
            raise HCBackendError("condensed vector has an invalid length")
        self.n = n
        self._check(self._L.rthc_set_condensed_f64(self._h, _ptr(y), n))

    # ---- clustering half -------------------------------------------------------------
    def set_single_algo(self, algo):
        self._check(self._L.rthc_set_single_algo(self._h, int(algo)))

    def linkage(self, method="single"):
        if method not in METHODS:
            raise HCBackendError("method %r is not available on the CUDA backend" % (method,))
        Z = np.empty((self.n - 1, 4), dtype=np.float64)
        self._check(self._L.rthc_linkage(self._h, METHODS[method], _ptr(Z)))
        return Z

    def local_forest(self):
        keys = np.empty(max(self.n - 1, 1), dtype=np.uint64)
        m = ctypes.c_int64(0)
        self._check(self._L.rthc_local_forest(self._h, _ptr(keys), ctypes.byref(m)))
        return keys[: m.value].copy()

    def linkage_from_forest(self, keys, n):
        keys = np.ascontiguousarray(keys, dtype=np.uint64)
        self.n = int(n)
        Z = np.empty((self.n - 1, 4), dtype=np.float64)
        self._check(self._L.rthc_linkage_from_forest(self._h, _ptr(keys), keys.size, self.n, _ptr(Z)))
        return Z

    def local_forest_device(self, device_ptr, capacity):
        """Forest of the owned edges written (unsorted, +inf padded to `capacity` slots) into caller-owned
        device memory, e.g. `tensor.data_ptr()` of an int64 torch tensor; returns the edge count."""
        m = ctypes.c_int64(0)
        self._check(self._L.rthc_local_forest_device(self._h, ctypes.c_void_p(int(device_ptr)), int(capacity), ctypes.byref(m)))
        return m.value

    def linkage_from_forest_device(self, device_ptr, m, n):
        """Z from `m` gathered key slots in device memory (+inf slots are ignored)."""
        self.n = int(n)
        Z = np.empty((self.n - 1, 4), dtype=np.float64)
        self._check(self._L.rthc_linkage_from_forest_device(self._h, ctypes.c_void_p(int(device_ptr)), int(m), self.n, _ptr(Z)))
        return Z

    def fcluster(self, thresholds, out=None):
        """labels[k] == scipy fcluster(Z, thresholds[k], criterion='distance') of the last linkage.
        out: optional preallocated C-contiguous int32 [T, n] array (e.g. a view of page-locked memory)."""
        t = np.ascontiguousarray(np.atleast_1d(thresholds), dtype=np.float64)
        if out is None:
            out = np.empty((t.size, self.n), dtype=np.int32)
        elif out.dtype != np.int32 or out.shape != (t.size, self.n) or not out.flags.c_contiguous:
            raise HCBackendError("fcluster: `out` must be a C-contiguous int32 array of shape (%d, %d)" % (t.size, self.n))
        self._check(self._L.rthc_cut(self._h, _ptr(t), t.size, _ptr(out)))
        return out

    # ---- zoom-in on the resident matrix ----------------------------------
    def closest(self, sample, k, noalloc=False):
        """(indices, distances, max_dist): every sample within the distance found at sorted position
        k of `sample`'s row (own 0 included), ordered by (distance, index).  noalloc=True: the variant
        that works in the library's static device scratch (rthc_closest_noalloc; n <= 262,144, no device
        allocation, calls on one device must not overlap)."""
        idx = np.empty(self.n, dtype=np.int64)
        d = np.empty(self.n, dtype=np.uint16)
        cnt, vmax = ctypes.c_int64(0), ctypes.c_int(0)
        fn = self._L.rthc_closest_noalloc if noalloc else self._L.rthc_closest
        self._check(fn(self._h, int(sample), int(k), _ptr(idx), _ptr(d), ctypes.byref(cnt), ctypes.byref(vmax)))
        idx, d = idx[: cnt.value], d[: cnt.value]
        order = np.lexsort((idx, d))
        return idx[order].copy(), d[order].copy(), int(vmax.value)

    def subset(self, indices, into=None):
        """Context on the same device holding the sub-matrix of `indices` (distinct sample positions),
        ready for linkage() / fcluster() / square_rows().  `into`: an earlier child to refill (its
        memory pool is warm, which is most of the cost of a small zoom-in); default a new context."""
        indices = np.ascontiguousarray(indices, dtype=np.int64)
        child = into if into is not None else HCContext(device=self.device)
        try:
            child._check(self._L.rthc_subset_from(child._h, self._h, _ptr(indices), indices.size))
        except Exception:
            if into is None:
                child.close()
            raise
        child.n, child.L = int(indices.size), self.L
        return child

    # ---- cluster stability -------------------------------------------------
    def contingency_adjacent(self, codes):
        """codes: T x n int32 cluster ids (0 <= id < n).  Returns one (rows, cols, counts) triple per
        adjacent column pair (p, p+1): the cells of getContTable(col[p+1], col[p]) in the reference's
        dict iteration order."""
        codes = np.ascontiguousarray(codes, dtype=np.int32)
        T, n = codes.shape
        off = np.zeros(T, dtype=np.int64)
        cap = max((T - 1) * n, 1)
        rows, cols, cnts = (np.empty(cap, dtype=np.int32) for _ in range(3))
        self._check(self._L.rthc_contingency_adjacent(self._h, _ptr(codes), T, n, _ptr(off), _ptr(rows), _ptr(cols), _ptr(cnts)))
        return [(rows[off[p]:off[p + 1]], cols[off[p]:off[p + 1]], cnts[off[p]:off[p + 1]]) for p in range(T - 1)]

    def adjusted_wallace_sums(self, rows, counts):
        """(rSumW1, rSumW2) of getAdjustedWallace for one table (host arithmetic, libm pow)."""
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        a, b = ctypes.c_double(0), ctypes.c_double(0)
        rc = self._L.rthc_adjusted_wallace_sums(_ptr(rows), _ptr(counts), rows.size, ctypes.byref(a), ctypes.byref(b))
        if rc != 0:
            raise HCBackendError("rthc_adjusted_wallace_sums failed (%d)" % rc)
        return a.value, b.value

    # ---- introspection -----------------------------------------------------------------
    def info(self):
        a = np.zeros(14, dtype=np.int64)
        self._check(self._L.rthc_info(self._h, _ptr(a), 14))
        keys = ["n", "L", "loci_kept", "groups", "plane_rows", "tiles", "boruvka_rounds", "max_alleles", "row_scans",
                "chain_barriers", "chain_cache_hits", "escape_loci", "escape_cells", "escape_pairs"]
        return dict(zip(keys, (int(x) for x in a)))

    def timings(self):
        a = np.zeros(8, dtype=np.float64)
        self._check(self._L.rthc_timings(self._h, _ptr(a), 8))
        keys = ["h2d_ms", "recode_ms", "distance_ms", "linkage_ms", "cut_ms", "launches", "stability_ms", "escape_fixup_ms"]
        return dict(zip(keys, (float(x) for x in a)))

    def mark(self, slot):
        """Record CUDA event `slot` (0 or 1) on the library stream."""
        self._check(self._L.rthc_mark(self._h, int(slot)))

    def mark_elapsed_ms(self):
        v = ctypes.c_double(0)
        self._check(self._L.rthc_mark_elapsed_ms(self._h, ctypes.byref(v)))
        return v.value

    def alu_peak(self):
        """Measured int32 ALU lane-ops/s of this GPU (LOP3 stream microbenchmark)."""
        v = ctypes.c_double(0)
        self._check(self._L.rthc_alu_peak(self._h, ctypes.byref(v)))
        return v.value

    def alu_peak_block(self, warps_per_sm=8):
        """LOP3 lane-ops/s with the distance kernel's 8x8 register-block operand pattern."""
        v = ctypes.c_double(0)
        self._check(self._L.rthc_alu_peak_block(self._h, int(warps_per_sm), ctypes.byref(v)))
        return v.value


class MultiDeviceContext:
    """One process, several GPUs (SURVEY.md 8b seam B0: reportree.py spawns exactly one
    partitioning_HC.py): the distance tiles are sharded over `devices` (one context and one host
    thread per device, tile-rows paired long + short), the shards are copied device-to-device into
    the full condensed matrix of devices[0], and everything after the distance stage (any linkage,
    cut, row blocks for the TSV writer) runs there exactly as on one GPU."""

    def __init__(self, devices):
        if len(devices) < 2:
            raise HCBackendError("MultiDeviceContext needs at least two devices")
        w = len(devices)
        self._ctx = [HCContext(device=devices[0], rank=0, world=w, full_storage=True)]
        self._ctx += [HCContext(device=d, rank=k, world=w) for k, d in enumerate(devices) if k > 0]
        self._algo = None

    def _each(self, fn):
        import threading
        errs = []

        def run(c):
            try:
                fn(c)
            except BaseException as e:      # noqa: BLE001 -- re-raised below
                errs.append(e)
        th = [threading.Thread(target=run, args=(c,)) for c in self._ctx]
        for t in th:
            t.start()
        for t in th:
            t.join()
        if errs:
            raise errs[0]

    def set_profiles(self, codes):
        codes = np.ascontiguousarray(codes, dtype=np.int32)
        self._each(lambda c: c.set_profiles(codes))

    def distance(self):
        self._each(lambda c: c.distance())
        head = self._ctx[0]
        for c in self._ctx[1:]:
            head.import_shard(c)
            c.close()
        head.finish_import()
        self._ctx = [head]

    def set_condensed(self, y):
        for c in self._ctx[1:]:
            c.close()
        self._ctx = self._ctx[:1]
        self._ctx[0].set_condensed(y)

    def close(self):
        for c in self._ctx:
            c.close()

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __getattr__(self, name):            # linkage, fcluster, square_rows, condensed, info, ... of the full context
        return getattr(self._ctx[0], name)
