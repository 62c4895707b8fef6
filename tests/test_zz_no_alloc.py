"""rthc_closest_noalloc: the allocation-free variant of rthc_closest (scratch in a file-scope static device array
of the library, include/rthc.h).  No counterpart in the reference -- added
for callers that cannot call cudaMalloc.  Checked through the C ABI against the existing entry point (exact:
integer distances and indices), on two calls in a row and on a second matrix, alone or after the other tests.
Collected last on purpose."""
import numpy as np
import pytest

from oracle import oracle_lib as O
from oracle import zoom_oracle
from reportree_b200 import hc_backend

pytestmark = pytest.mark.gpu


def profiles(rng, n, L, alleles=3, missing=0.05):
    codes = rng.integers(1, alleles + 1, (n, L)).astype(np.int32)
    codes[rng.random((n, L)) < missing] = 0
    return codes


def same(a, b):
    return np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]) and a[2] == b[2]


def test_noalloc_closest_equals_the_allocating_entry_point():
    import torch
    rng = np.random.default_rng(5)
    codes = profiles(rng, 3000, 40)
    D = O.dist_square(codes)
    names = ["s%d" % k for k in range(codes.shape[0])]
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes)
        ctx.distance()
        queries = [(0, 0), (0, 5), (17, 1), (2999, 2999), (1500, 300), (42, 2998)]
        want = [ctx.closest(s, k) for s, k in queries]
        free0 = torch.cuda.mem_get_info()[0]
        for rep in range(2):                       # two calls in a row per query: the scratch keeps the last call's contents
            for (s, k), w in zip(queries, want):
                got = ctx.closest(s, k, noalloc=True)
                assert same(got, w), (rep, s, k)
        assert torch.cuda.mem_get_info()[0] >= free0          # nothing was taken from the device
        # and against the CPU oracle, not only against the sibling entry point
        for s, k in queries[:4]:
            idx, d, vmax = ctx.closest(s, k, noalloc=True)
            wn, wd, _ = zoom_oracle.closest_samples(D, names, names[s], k)
            assert {names[i] for i in idx} == wn and d.tolist() == wd and vmax == wd[-1]
        # bad arguments: same codes as rthc_closest
        with pytest.raises(hc_backend.HCBackendError):
            ctx.closest(3000, 1, noalloc=True)
        with pytest.raises(hc_backend.HCBackendError):
            ctx.closest(0, 3000, noalloc=True)
    # a second, smaller matrix in a new context: stale rows / histogram of the first must not leak in
    codes2 = profiles(rng, 257, 33, alleles=2, missing=0.2)
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes2)
        ctx.distance()
        for s, k in [(0, 256), (256, 0), (100, 50), (100, 50)]:
            assert same(ctx.closest(s, k, noalloc=True), ctx.closest(s, k)), (s, k)


def test_noalloc_closest_needs_distances_and_rejects_more_samples_than_its_scratch_holds():
    with hc_backend.HCContext() as ctx:
        with pytest.raises(hc_backend.HCBackendError):
            ctx.closest(0, 0, noalloc=True)                   # no resident matrix
    n = 262144 + 128                                           # RTHC_NOALLOC_MAX_N + 128: 68.8 GB of uint16 distances
    rng = np.random.default_rng(6)
    codes = profiles(rng, n, 4, alleles=4, missing=0.0)
    with hc_backend.HCContext() as ctx:
        ctx.set_profiles(codes)
        ctx.distance()
        with pytest.raises(hc_backend.HCBackendError, match="exceed the static scratch"):
            ctx.closest(0, 1, noalloc=True)
        idx, d, vmax = ctx.closest(0, 1)                      # the allocating entry point still answers
        assert idx.size >= 2 and int(d[0]) == 0
