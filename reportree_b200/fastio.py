"""ctypes binding of librthc_hostio.so (include/rthio.h): multithreaded C++ text ingest and
artefact writers with the string semantics of the reference's pandas code
(scripts/partitioning_HC.py:362-372, :443, :477, :517-518, :651-670, :779-787).

`open_profile` returns None for files whose pandas parse is not the plain rectangular case
(quotes, CR, ragged rows, duplicated header cells ...): the caller then takes the pandas path in
partitioning_HC.py, which remains the specification.  Set REPORTREE_HC_FASTIO=0 to force the
pandas path everywhere (used by the equivalence tests).
"""
import ctypes
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "librthc_hostio.so")
EXPORTED_SYMBOLS = [
    "rthio_last_error", "rthio_version", "rthio_profile_open", "rthio_profile_close", "rthio_profile_rows",
    "rthio_profile_cols", "rthio_profile_header", "rthio_profile_sample", "rthio_profile_codes",
    "rthio_profile_gather", "rthio_profile_write", "rthio_square_open", "rthio_square_append_u16",
    "rthio_square_close", "rthio_write_partitions", "rthio_write_int_matrix", "rthio_profile_names",
    "rthio_profile_count_eq",
]
_lib = None


class FastIOError(RuntimeError):
    pass


def enabled():
    return os.environ.get("REPORTREE_HC_FASTIO", "1") != "0"


def load_library():
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise FastIOError("host I/O library %s not built (run `python -m reportree_b200.build`)" % LIB_PATH)
    L = ctypes.CDLL(LIB_PATH)
    vp, i64, ci, cp = ctypes.c_void_p, ctypes.c_int64, ctypes.c_int, ctypes.c_char_p
    pp = ctypes.POINTER
    sig = {
        "rthio_last_error": (cp, []),
        "rthio_version": (ci, []),
        "rthio_profile_open": (ci, [cp, cp, ci, pp(vp)]),
        "rthio_profile_close": (None, [vp]),
        "rthio_profile_rows": (i64, [vp]),
        "rthio_profile_cols": (i64, [vp]),
        "rthio_profile_header": (ci, [vp, i64, pp(vp), pp(i64)]),
        "rthio_profile_sample": (ci, [vp, i64, pp(vp), pp(i64), pp(ci)]),
        "rthio_profile_codes": (vp, [vp]),
        "rthio_profile_names": (ci, [vp, pp(vp), pp(i64), pp(vp), pp(i64), pp(vp)]),
        "rthio_profile_count_eq": (ci, [vp, vp, i64, vp, i64, ctypes.c_int32, vp]),
        "rthio_profile_gather": (ci, [vp, vp, i64, vp, i64, vp]),
        "rthio_profile_write": (ci, [vp, cp, vp, i64, vp, i64]),
        "rthio_square_open": (ci, [cp, cp, vp, vp, i64, ci, pp(vp)]),
        "rthio_square_append_u16": (ci, [vp, i64, i64, vp]),
        "rthio_square_close": (ci, [vp]),
        "rthio_write_partitions": (ci, [cp, cp, vp, vp, i64, vp, i64, vp, vp, ci]),
        "rthio_write_int_matrix": (ci, [cp, cp, cp, cp, vp, i64, i64, ci]),
    }
    for name, (res, args) in sig.items():
        fn = getattr(L, name)
        fn.restype = res
        fn.argtypes = args
    _lib = L
    return L


def _err(L):
    return (L.rthio_last_error() or b"").decode("utf-8", "replace")


def _ptr(a):
    return ctypes.c_void_p(a.ctypes.data)


def _blob(strings):
    """list[str] -> (bytes blob, int64 offsets[n+1]); None when a string cannot be written verbatim."""
    enc = [s.encode("utf-8") for s in strings]
    offs = np.zeros(len(enc) + 1, dtype=np.int64)
    if enc:
        np.cumsum([len(b) for b in enc], out=offs[1:])
    return b"".join(enc), offs


def plain_text(strings):
    """True when no string needs CSV quoting (tab, quote, CR, LF) -- the fast writers copy verbatim."""
    return not any(("\t" in s) or ('"' in s) or ("\n" in s) or ("\r" in s) for s in strings)


class Profile:
    """An allele/SNP profile parsed by rthio_profile_open."""

    def __init__(self, L, handle):
        self._L, self._h = L, handle
        self.n = int(L.rthio_profile_rows(handle))
        self.n_loci = int(L.rthio_profile_cols(handle))
        hb, hl, ib, il, nf = ctypes.c_void_p(), ctypes.c_int64(), ctypes.c_void_p(), ctypes.c_int64(), ctypes.c_void_p()
        if L.rthio_profile_names(handle, ctypes.byref(hb), ctypes.byref(hl), ctypes.byref(ib), ctypes.byref(il), ctypes.byref(nf)) != 0:
            raise FastIOError(_err(L))
        self.header = ctypes.string_at(hb.value, hl.value).decode("utf-8").split("\t")
        self.ids = ctypes.string_at(ib.value, il.value).decode("utf-8").split("\n") if self.n else []   # str, or None where pandas has NaN
        if len(self.header) != self.n_loci + 1 or len(self.ids) != self.n:
            raise FastIOError("name tables do not match the profile shape")
        if self.n:
            flags = np.frombuffer(ctypes.string_at(nf.value, self.n), dtype=np.int8)
            for r in np.flatnonzero(flags).tolist():
                self.ids[r] = None
        addr = L.rthio_profile_codes(handle)
        if self.n and self.n_loci:
            buf = (ctypes.c_int32 * (self.n * self.n_loci)).from_address(addr)
            self.codes = np.frombuffer(buf, dtype=np.int32).reshape(self.n, self.n_loci)
        else:
            self.codes = np.zeros((self.n, self.n_loci), dtype=np.int32)

    def gather(self, rows, cols):
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        cols = np.ascontiguousarray(cols, dtype=np.int64)
        out = np.empty((rows.size, cols.size), dtype=np.int32)
        if self._L.rthio_profile_gather(self._h, _ptr(rows), rows.size, _ptr(cols), cols.size, _ptr(out)) != 0:
            raise FastIOError(_err(self._L))
        return out

    def count_eq(self, rows, cols, value=0):
        """out[r] = number of the selected loci whose code equals `value` in row rows[r]."""
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        cols = np.ascontiguousarray(cols, dtype=np.int64)
        out = np.empty(rows.size, dtype=np.int64)
        if self._L.rthio_profile_count_eq(self._h, _ptr(rows), rows.size, _ptr(cols), cols.size, int(value), _ptr(out)) != 0:
            raise FastIOError(_err(self._L))
        return out

    def write(self, path, rows, cols):
        rows = np.ascontiguousarray(rows, dtype=np.int64)
        cols = np.ascontiguousarray(cols, dtype=np.int64)
        if self._L.rthio_profile_write(self._h, os.fsencode(path), _ptr(rows), rows.size, _ptr(cols), cols.size) != 0:
            raise FastIOError(_err(self._L))

    def close(self):
        if self._h:
            self.codes = None
            self._L.rthio_profile_close(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def open_profile(path, missing_code="0", threads=0):
    """Profile, or None when the file needs the pandas path."""
    L = load_library()
    h = ctypes.c_void_p()
    rc = L.rthio_profile_open(os.fsencode(path), str(missing_code).encode("utf-8"), int(threads), ctypes.byref(h))
    if rc == 1:
        return None
    if rc != 0:
        raise FastIOError(_err(L))
    try:
        return Profile(L, h)
    except UnicodeDecodeError:
        L.rthio_profile_close(h)
        return None


def write_square_tsv(path, names, row_block_fn, n, block=1024, corner="dists", threads=0):
    """<out>_dist_hamming.tsv streamed from uint16 row blocks (row_block_fn(row0, nrows) -> nrows x n)."""
    L = load_library()
    blob, offs = _blob(names)
    h = ctypes.c_void_p()
    if L.rthio_square_open(os.fsencode(path), corner.encode("utf-8"), blob, _ptr(offs), n, int(threads), ctypes.byref(h)) != 0:
        raise FastIOError(_err(L))
    # the next row block is fetched (device -> host) on a helper thread while this one is formatted and written
    from concurrent.futures import ThreadPoolExecutor

    def fetch(r0):
        return np.ascontiguousarray(row_block_fn(r0, min(block, n - r0)), dtype=np.uint16)
    try:
        with ThreadPoolExecutor(max_workers=1) as pool:
            nxt = pool.submit(fetch, 0) if n > 0 else None
            for r0 in range(0, n, block):
                rows = nxt.result()
                nxt = pool.submit(fetch, r0 + block) if r0 + block < n else None
                if L.rthio_square_append_u16(h, r0, rows.shape[0], _ptr(rows)) != 0:
                    raise FastIOError(_err(L))
    finally:
        if L.rthio_square_close(h) != 0:
            raise FastIOError(_err(L))


def write_partitions(path_partitions, path_composition, names, labels, partition_names, threads=0):
    """partitions.tsv + clusterComposition.tsv from the T x n fcluster id matrix."""
    L = load_library()
    labels = np.ascontiguousarray(labels, dtype=np.int32)
    T, n = labels.shape
    nb, no = _blob(names)
    cb, co = _blob(partition_names)
    if L.rthio_write_partitions(os.fsencode(path_partitions), os.fsencode(path_composition), nb, _ptr(no), n,
                                _ptr(labels), T, cb, _ptr(co), int(threads)) != 0:
        raise FastIOError(_err(L))


def write_int_matrix(path, codes, corner="FILE", row_prefix="S", col_prefix="L", threads=0):
    """Synthetic profile TSV from an int32 matrix (ids S0000001.., loci L00001..)."""
    L = load_library()
    codes = np.ascontiguousarray(codes, dtype=np.int32)
    n, nl = codes.shape
    if L.rthio_write_int_matrix(os.fsencode(path), corner.encode(), row_prefix.encode(), col_prefix.encode(), _ptr(codes), n, nl,
                                int(threads)) != 0:
        raise FastIOError(_err(L))
