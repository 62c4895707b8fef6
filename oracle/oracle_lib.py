"""ctypes loader for the CPU oracle (oracle/_build/liboracle.so).

TEST INFRASTRUCTURE ONLY: imported by tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs.  The product package
(reportree_b200/) never imports this module.
"""
import ctypes
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "liboracle.so")
CGMLST_DISTS_BIN = os.path.join(_HERE, "_build", "cgmlst-dists")

_lib = None


def build():
    subprocess.check_call(["make", "-s", "-C", _HERE])


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(_SO):
            build()
        L = ctypes.CDLL(_SO)
        i64, p = ctypes.c_int64, ctypes.c_void_p
        L.oracle_dist_square_i32.argtypes = [p, i64, i64, p]
        L.oracle_dist_condensed_rows_u16.argtypes = [p, i64, i64, i64, i64, p]
        L.oracle_dist_square_chunked_i32.argtypes = [p, i64, i64, i64, p]
        L.oracle_linkage_single.argtypes = [p, i64, p]
        L.oracle_prim_edges.argtypes = [p, i64, p]
        L.oracle_linkage_nn_chain.argtypes = [p, i64, ctypes.c_int, p]
        L.oracle_linkage_single_u16.argtypes = [p, i64, p]
        L.oracle_linkage_nn_chain_u16.argtypes = [p, i64, ctypes.c_int, p]
        L.oracle_maxdists.argtypes = [p, i64, p]
        L.oracle_fcluster_distance.argtypes = [p, i64, ctypes.c_double, p]
        for f in ("oracle_dist_square_i32", "oracle_dist_condensed_rows_u16",
                  "oracle_dist_square_chunked_i32", "oracle_linkage_single", "oracle_prim_edges",
                  "oracle_linkage_nn_chain", "oracle_linkage_single_u16", "oracle_linkage_nn_chain_u16",
                  "oracle_maxdists", "oracle_fcluster_distance"):
            getattr(L, f).restype = None
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(ctypes.c_void_p)


def dist_square(codes):
    codes = np.ascontiguousarray(codes, dtype=np.int32)
    n, L = codes.shape
    out = np.empty((n, n), dtype=np.int32)
    lib().oracle_dist_square_i32(_ptr(codes), n, L, _ptr(out))
    return out


def dist_square_chunked(codes, chunk=2000):
    codes = np.ascontiguousarray(codes, dtype=np.int32)
    n, L = codes.shape
    out = np.empty((n, n), dtype=np.int32)
    lib().oracle_dist_square_chunked_i32(_ptr(codes), n, L, chunk, _ptr(out))
    return out


def dist_condensed_rows(codes, row0, row1):
    codes = np.ascontiguousarray(codes, dtype=np.int32)
    n, L = codes.shape
    cnt = sum(n - 1 - i for i in range(row0, row1))
    out = np.empty(cnt, dtype=np.uint16)
    lib().oracle_dist_condensed_rows_u16(_ptr(codes), n, L, row0, row1, _ptr(out))
    return out


def dist_condensed(codes):
    return dist_condensed_rows(codes, 0, codes.shape[0])


METHODS = {"single": 0, "complete": 1, "average": 2, "weighted": 3, "ward": 4}


def linkage(y, method):
    """scipy linkage(y, method).  A uint16 condensed vector (the distance kernel's storage type) is
    consumed as it is -- no float64 copy for single / complete -- so configs[3] / [4] fit the host."""
    y = np.asarray(y)
    if y.dtype == np.uint16:
        y = np.ascontiguousarray(y)
        n = int(round((1 + (1 + 8 * y.size) ** 0.5) / 2))
        assert n * (n - 1) // 2 == y.size
        Z = np.empty((n - 1, 4), dtype=np.float64)
        if method == "single":
            lib().oracle_linkage_single_u16(_ptr(y), n, _ptr(Z))
        else:
            lib().oracle_linkage_nn_chain_u16(_ptr(y), n, METHODS[method], _ptr(Z))
        return Z
    y = np.ascontiguousarray(y, dtype=np.float64)
    n = int(round((1 + (1 + 8 * y.size) ** 0.5) / 2))
    assert n * (n - 1) // 2 == y.size
    Z = np.empty((n - 1, 4), dtype=np.float64)
    if method == "single":
        lib().oracle_linkage_single(_ptr(y), n, _ptr(Z))
    else:
        lib().oracle_linkage_nn_chain(_ptr(y), n, METHODS[method], _ptr(Z))
    return Z


def prim_edges(y):
    y = np.ascontiguousarray(y, dtype=np.float64)
    n = int(round((1 + (1 + 8 * y.size) ** 0.5) / 2))
    E = np.empty((n - 1, 3), dtype=np.float64)
    lib().oracle_prim_edges(_ptr(y), n, _ptr(E))
    return E


def maxdists(Z):
    Z = np.ascontiguousarray(Z, dtype=np.float64)
    n = Z.shape[0] + 1
    MD = np.empty(n - 1, dtype=np.float64)
    lib().oracle_maxdists(_ptr(Z), n, _ptr(MD))
    return MD


def fcluster(Z, t):
    Z = np.ascontiguousarray(Z, dtype=np.float64)
    n = Z.shape[0] + 1
    T = np.empty(n, dtype=np.int32)
    lib().oracle_fcluster_distance(_ptr(Z), n, float(t), _ptr(T))
    return T


def canonical(labels):
    """Relabel clusters by order of first appearance (membership-only comparison)."""
    labels = np.asarray(labels)
    _, first, inv = np.unique(labels, return_index=True, return_inverse=True)
    rank = np.empty(first.size, dtype=np.int64)
    rank[np.argsort(first, kind="stable")] = np.arange(first.size)
    return rank[inv]
