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
