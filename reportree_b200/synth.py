"""Deterministic synthetic cg/wgMLST allele matrices.

Population model so that clusters exist at low thresholds: K = n/50 founder profiles, allele
of founder f at locus l drawn Zipf-like from 1..A_l with A_l ~ U{5..400}; every sample copies
a sub-lineage of a founder (geometric founder weights) and both levels mutate each locus
independently with probability p (a mutation creates a brand-new allele id for that locus);
cells are then set to 0 (missing) iid.  Returns int32 codes, 0 = missing.
"""
import numpy as np

SEEDS = {"C3": 20260303, "C4": 20260304, "C5": 20260305}


def allele_matrix(n, L, seed, missing=0.02, p_mut=0.002, founders=None, sublineages_per_founder=5,
                  bad_sample_frac=0.0, bad_missing=0.10):
    rng = np.random.default_rng(seed)
    K = founders or max(2, n // 50)
    A = rng.integers(5, 401, size=L)                       # alleles per locus among founders
    # Zipf-like founder alleles: floor(A * u^3) + 1 concentrates mass on low ids
    u = rng.random((K, L), dtype=np.float32)
    founder = (np.floor(A[None, :] * u ** 3).astype(np.int32) + 1)
    next_id = A.astype(np.int64) + 1                       # next unused allele id per locus

    def mutate(parent_rows, parents):
        nonlocal next_id
        m = parents.shape[0]
        out = parent_rows[parents]                         # m x L copy
        nmut = rng.binomial(m * L, p_mut)
        if nmut:
            flat = rng.integers(0, m * L, size=nmut)
            flat = np.unique(flat)
            cols = flat % L
            order = np.argsort(cols, kind="stable")
            flat, cols = flat[order], cols[order]
            # running index of each mutation within its locus
            first = np.searchsorted(cols, cols, side="left")
            within = np.arange(cols.size) - first
            out.reshape(-1)[flat] = (next_id[cols] + within).astype(np.int32)
            next_id = next_id + np.bincount(cols, minlength=L)
        return out

    S = K * sublineages_per_founder
    sub_parent = rng.integers(0, K, size=S)
    sub = mutate(founder, sub_parent)
    # geometric-ish weights over sub-lineages
    w = 0.97 ** np.arange(S) + 1.0 / S
    w /= w.sum()
    sample_parent = rng.choice(S, size=n, p=w)
    X = mutate(sub, sample_parent)
    # missing data
    if missing > 0:
        nmiss = rng.binomial(n * L, missing)
        X.reshape(-1)[rng.integers(0, n * L, size=nmiss)] = 0
    if bad_sample_frac > 0:
        bad = np.flatnonzero(rng.random(n) < bad_sample_frac)
        for r in bad:
            X[r, rng.random(L) < bad_missing] = 0
    return X


def config(name):
    """Shapes of BASELINE.json's synthetic configs."""
    if name == "C3":
        return dict(n=20000, L=3002, seed=SEEDS["C3"], missing=0.02)
    if name == "C4":
        return dict(n=100000, L=3000, seed=SEEDS["C4"], missing=0.02)
    if name == "C5":
        return dict(n=50000, L=7600, seed=SEEDS["C5"], missing=0.03, bad_sample_frac=0.05, bad_missing=0.10)
    raise KeyError(name)


def workload(name):
    """The matrix that reaches the distance kernel for a BASELINE.json config: C3 / C4 as generated;
    C5 = configs[4] after the reference's filters (scripts/partitioning_HC.py:400-477): a --loci subset
    of 3,000 of the 7,600 columns, then --loci-called 0.95 on the rows (pct_called >= 0.95)."""
    X = allele_matrix(**config(name))
    if name != "C5":
        return X
    rng = np.random.default_rng(1)
    cols = np.sort(rng.choice(X.shape[1], 3000, replace=False))
    Xs = X[:, cols]
    return np.ascontiguousarray(Xs[(Xs != 0).sum(1) / Xs.shape[1] >= 0.95])
