"""TEST INFRASTRUCTURE: a stand-in for hc_backend.HCContext built on the CPU oracle + scipy, so
that the host side of the CLI (filters, messages, artefact writers, both text I/O paths) can be
exercised without a GPU.  The product never imports this."""
import numpy as np
from scipy.cluster.hierarchy import fcluster, linkage
from scipy.spatial.distance import squareform

from oracle import oracle_lib as O


class FakeContext:
    def __init__(self, device=0, rank=0, world=1):
        self.n = 0

    def set_single_algo(self, algo):
        pass

    def set_profiles(self, codes):
        self.codes = np.maximum(np.ascontiguousarray(codes, dtype=np.int32), 0)
        self.n = self.codes.shape[0]

    def distance(self):
        self.D = O.dist_square(self.codes)
        self.y = squareform(self.D.astype(np.float64), checks=False)

    def set_condensed(self, y):
        self.y = np.asarray(y, dtype=np.float64)
        self.D = squareform(self.y)
        self.n = self.D.shape[0]

    def square_rows(self, row0, nrows):
        return self.D[row0:row0 + nrows].astype(np.uint16)

    def linkage(self, method="single"):
        self.Z = linkage(self.y, method)
        return self.Z

    def fcluster(self, thresholds):
        return np.stack([fcluster(self.Z, float(t), criterion="distance") for t in np.atleast_1d(thresholds)]).astype(np.int32)

    def close(self):
        pass

    # ---- N3 / N4 stand-ins ------------------------------------------------------------------
    def closest(self, sample, k):
        col = self.D[:, sample].astype(np.int64)
        cut = np.sort(col, kind="stable")[k]
        idx = np.nonzero(col <= cut)[0]
        order = np.lexsort((idx, col[idx]))
        return idx[order], col[idx][order].astype(np.uint16), int(cut)

    def contingency_adjacent(self, codes):
        out = []
        for p in range(codes.shape[0] - 1):
            table = {}
            for r, c in zip(codes[p + 1].tolist(), codes[p].tolist()):
                inner = table.setdefault(r, {})
                inner[c] = inner.get(c, 0) + 1
            cells = [(r, c, v) for r, inner in table.items() for c, v in inner.items()]
            out.append(tuple(np.array([x[k] for x in cells], dtype=np.int32) for k in range(3)))
        return out

    def adjusted_wallace_sums(self, rows, counts):
        # the real host-side C function: it needs the library, not a GPU
        import ctypes
        from reportree_b200 import hc_backend
        L = hc_backend.load_library()
        rows = np.ascontiguousarray(rows, dtype=np.int32)
        counts = np.ascontiguousarray(counts, dtype=np.int32)
        a, b = ctypes.c_double(0), ctypes.c_double(0)
        assert L.rthc_adjusted_wallace_sums(rows.ctypes.data, counts.ctypes.data, rows.size, ctypes.byref(a), ctypes.byref(b)) == 0
        return a.value, b.value
