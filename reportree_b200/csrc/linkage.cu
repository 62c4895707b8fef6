// linkage.cu -- rthc_linkage / rthc_cut entry points and the O(n) dendrogram bookkeeping.
//
// The device produces the n-1 merges (Boruvka MST edges, Prim-order edges, or the NN-chain
// merge list); this file turns them into scipy's Z exactly as scipy's `label` does
// (stable order by height, union-find ids n+k, sizes) -- the same O(n alpha(n)) pass scipy
// runs after mst_single_linkage / nn_chain (scipy _hierarchy_impl.py:960-963) -- and
// derives the traversal positions the device cut needs (see cut.cu).
#include <algorithm>
#include <cmath>
#include <numeric>

#include "rthc_internal.cuh"

// merges (x_k, y_k, w_k) in final (sorted) order -> Z, representatives, traversal events.
// Union-find over the n leaves (union by size, path halving) with the current cluster id of every root kept
// beside it -- scipy's LinkageUnionFind; all scratch arrays live in the context and the two tables the cut
// needs are built in page-locked memory, so that a repeated call allocates nothing.
static int finish_linkage(rthc_ctx *ctx, int64_t n, const int32_t *ex, const int32_t *ey, const double *ew, size_t count,
                          double *Z)
{
    const int64_t m = n - 1;
    if ((int64_t)count != m)
        return rthc_fail(ctx, RTHC_E_STATE, "linkage produced %zu merges, expected %lld", count, (long long)m);
    const size_t nodes = (size_t)(2 * n - 1);
    int32_t *tpar = (int32_t *)rthc_pinned(ctx, 1, sizeof(int32_t) * nodes);
    int32_t *event = (int32_t *)rthc_pinned(ctx, 2, sizeof(int32_t) * nodes);
    if (!tpar || !event) return RTHC_E_NOMEM;
    ctx->h_uf.resize((size_t)n); ctx->h_cid.resize((size_t)n); ctx->h_nsize.resize(nodes);
    int32_t *uf = ctx->h_uf.data(), *cid = ctx->h_cid.data(), *nsize = ctx->h_nsize.data();
    for (int64_t v = 0; v < n; ++v) { uf[v] = (int32_t)v; cid[v] = (int32_t)v; nsize[v] = 1; }
    auto find = [&](int32_t v) {
        while (uf[v] != v) { uf[v] = uf[uf[v]]; v = uf[v]; }
        return v;
    };
    for (int64_t k = 0; k < m; ++k) {
        int32_t a = find(ex[k]), b = find(ey[k]);
        if (a == b) return rthc_fail(ctx, RTHC_E_STATE, "merge %lld joins a cluster with itself", (long long)k);
        int32_t rx = cid[a], ry = cid[b];
        if (nsize[rx] < nsize[ry]) std::swap(a, b);          // b hangs under a
        if (rx > ry) std::swap(rx, ry);
        const int32_t id = (int32_t)(n + k), sz = nsize[rx] + nsize[ry];
        Z[4 * k + 0] = rx; Z[4 * k + 1] = ry; Z[4 * k + 2] = ew[k]; Z[4 * k + 3] = sz;
        nsize[id] = sz;
        tpar[rx] = id; tpar[ry] = id;                        // the dendrogram itself
        uf[b] = a; cid[a] = id;
    }
    tpar[nodes - 1] = 0x7FFFFFFF;
    // Position of every node in scipy's numbering walk (cluster_dist / cluster_monocrit): a node is
    // numbered when first visited, then its left and right internal subtrees, then its leaf
    // children (left before right).  A subtree of s leaves holds 2 s - 1 events, so the positions
    // follow top-down (parents have larger ids than children): no stack, no recursion.
    event[nodes - 1] = 0;
    for (int64_t k = m - 1; k >= 0; --k) {
        const int32_t lc = (int32_t)Z[4 * k], rc = (int32_t)Z[4 * k + 1];
        int32_t cursor = event[n + k] + 1;
        if (lc >= n) { event[lc] = cursor; cursor += 2 * nsize[lc] - 1; }
        if (rc >= n) { event[rc] = cursor; cursor += 2 * nsize[rc] - 1; }
        if (lc < n) event[lc] = cursor++;
        if (rc < n) event[rc] = cursor++;
    }
    ctx->h_heights.assign(ew, ew + m);
    for (int64_t q = 1; q < m; ++q)
        if (ew[q] < ew[q - 1])
            return rthc_fail(ctx, RTHC_E_STATE, "merge heights are not monotone at row %lld", (long long)q);
    int rcode;
    if ((rcode = rthc_dev_alloc(ctx, &ctx->d_tpar, nodes))) return rcode;
    if ((rcode = rthc_dev_alloc(ctx, &ctx->d_event, nodes))) return rcode;
    rthc_dev_free(ctx, &ctx->d_up);
    ctx->up_levels = 0;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_tpar, tpar, sizeof(int32_t) * nodes, cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_event, event, sizeof(int32_t) * nodes, cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));     // the staging slots are reused by the next call
    ctx->m_edges = m;
    ctx->n = n;
    ctx->have_linkage = true;
    return 0;
}

static int finish_linkage(rthc_ctx *ctx, int64_t n, const std::vector<int32_t> &ex, const std::vector<int32_t> &ey,
                          const std::vector<double> &ew, double *Z)
{
    return finish_linkage(ctx, n, ex.data(), ey.data(), ew.data(), ex.size(), Z);
}

// keys: ascending packed (w, i, j) edges in host memory
static int finish_from_keys(rthc_ctx *ctx, int64_t n, const uint64_t *keys, size_t count, double *Z)
{
    ctx->h_ex.resize(count); ctx->h_ey.resize(count); ctx->h_ew.resize(count);
    for (size_t k = 0; k < count; ++k) {
        ctx->h_ex[k] = (int32_t)rthc_key_i(keys[k]);
        ctx->h_ey[k] = (int32_t)rthc_key_j(keys[k]);
        ctx->h_ew[k] = (double)rthc_key_w(keys[k]);
    }
    return finish_linkage(ctx, n, ctx->h_ex.data(), ctx->h_ey.data(), ctx->h_ew.data(), count, Z);
}

// stable order by height (scipy: np.argsort(Z[:, 2], kind="mergesort"))
static void stable_sort_merges(std::vector<int32_t> &ex, std::vector<int32_t> &ey, std::vector<double> &ew)
{
    std::vector<int64_t> order(ew.size());
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int64_t a, int64_t b) { return ew[(size_t)a] < ew[(size_t)b]; });
    std::vector<int32_t> x2(ex.size()), y2(ey.size());
    std::vector<double> w2(ew.size());
    for (size_t k = 0; k < order.size(); ++k) {
        x2[k] = ex[(size_t)order[k]]; y2[k] = ey[(size_t)order[k]]; w2[k] = ew[(size_t)order[k]];
    }
    ex.swap(x2); ey.swap(y2); ew.swap(w2);
}

extern "C" int rthc_linkage(rthc_ctx *ctx, int method, double *Z_out)
{
    if (!ctx || !Z_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_dist) return rthc_fail(ctx, RTHC_E_STATE, "rthc_linkage before a distance matrix exists");
    if (ctx->world != 1)
        return rthc_fail(ctx, RTHC_E_STATE, "rthc_linkage needs a full shard; use rthc_local_forest + "
                                            "rthc_linkage_from_forest for sharded single linkage");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    rthc_dev_free(ctx, &ctx->d_tpar); rthc_dev_free(ctx, &ctx->d_event); rthc_dev_free(ctx, &ctx->d_up);
    ctx->have_linkage = false;
    int rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    std::vector<int32_t> ex, ey;
    std::vector<double> ew;
    const bool boruvka = method == RTHC_SINGLE && ctx->single_algo == RTHC_SL_BORUVKA && ctx->d_cond;
    const uint64_t *h_keys = nullptr;
    int64_t m = 0;
    if (boruvka) {
        uint64_t *d_keys = nullptr;
        if ((rc = rthc_boruvka_owned(ctx, true, &d_keys, &m))) return rc;
        RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
        rc = rthc_keys_to_host(ctx, d_keys, m, &h_keys);
        rthc_dev_free(ctx, &d_keys);
        if (rc) return rc;
    } else if (method == RTHC_SINGLE) {
        if ((rc = rthc_prim_exact(ctx, ex, ey, ew))) return rc;
        stable_sort_merges(ex, ey, ew);
    } else if (method == RTHC_COMPLETE || method == RTHC_AVERAGE || method == RTHC_WEIGHTED || method == RTHC_WARD) {
        if ((rc = rthc_nn_chain(ctx, method, ex, ey, ew))) return rc;
        stable_sort_merges(ex, ey, ew);
    } else {
        return rthc_fail(ctx, RTHC_E_ARG, "unsupported linkage method %d", method);
    }
    if (!boruvka) RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[3] = ms;
    if (boruvka) return finish_from_keys(ctx, ctx->n, h_keys, (size_t)m, Z_out);
    return finish_linkage(ctx, ctx->n, ex, ey, ew, Z_out);
}

// forest of the owned edges, left on the device: timing + the state checks shared by both entry points
static int local_forest_device(rthc_ctx *ctx, bool sorted, uint64_t **d_keys, int64_t *m)
{
    if (!ctx->have_dist || !ctx->d_cond) return rthc_fail(ctx, RTHC_E_STATE, "no uint16 distance matrix yet");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    int rc = rthc_boruvka_owned(ctx, sorted, d_keys, m);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[3] = ms;
    return 0;
}

extern "C" int rthc_local_forest(rthc_ctx *ctx, uint64_t *keys_out, int64_t *m_out)
{
    if (!ctx || !keys_out || !m_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    uint64_t *d_keys = nullptr;
    int64_t m = 0;
    int rc = local_forest_device(ctx, true, &d_keys, &m);
    if (rc) return rc;
    if (m > 0)
        RTHC_CUDA(ctx, cudaMemcpyAsync(keys_out, d_keys, sizeof(uint64_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_keys);
    *m_out = m;
    return 0;
}

extern "C" int rthc_local_forest_device(rthc_ctx *ctx, uint64_t *d_keys_out, int64_t capacity, int64_t *m_out)
{
    if (!ctx || !d_keys_out || !m_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    uint64_t *d_keys = nullptr;
    int64_t m = 0;
    int rc = local_forest_device(ctx, false, &d_keys, &m);      // the gather's consumer does not need an order
    if (rc) return rc;
    if (m > capacity) {
        rthc_dev_free(ctx, &d_keys);
        return rthc_fail(ctx, RTHC_E_ARG, "forest of %lld edges does not fit %lld slots", (long long)m, (long long)capacity);
    }
    if (m > 0)
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_keys_out, d_keys, sizeof(uint64_t) * (size_t)m, cudaMemcpyDeviceToDevice, ctx->stream));
    if (capacity > m)                                            // 0xFF bytes = RTHC_KEY_INF
        RTHC_CUDA(ctx, cudaMemsetAsync(d_keys_out + m, 0xFF, sizeof(uint64_t) * (size_t)(capacity - m), ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_keys);
    *m_out = m;
    return 0;
}

static int linkage_from_forest_device(rthc_ctx *ctx, const uint64_t *d_in, int64_t m_in, int64_t n, double *Z_out)
{
    rthc_dev_free(ctx, &ctx->d_tpar); rthc_dev_free(ctx, &ctx->d_event); rthc_dev_free(ctx, &ctx->d_up);
    ctx->have_linkage = false;
    uint64_t *d_mst = nullptr;
    int64_t m = 0;
    int rc = rthc_boruvka_edge_list(ctx, d_in, m_in, n, &d_mst, &m);
    if (rc) return rc;
    const uint64_t *h_keys = nullptr;
    rc = rthc_keys_to_host(ctx, d_mst, m, &h_keys);
    rthc_dev_free(ctx, &d_mst);
    if (rc) return rc;
    return finish_from_keys(ctx, n, h_keys, (size_t)m, Z_out);
}

extern "C" int rthc_linkage_from_forest(rthc_ctx *ctx, const uint64_t *keys, int64_t m, int64_t n, double *Z_out)
{
    if (!ctx || !keys || !Z_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (n < 2 || m < n - 1) return rthc_fail(ctx, RTHC_E_ARG, "forest union has %lld edges for n=%lld", (long long)m, (long long)n);
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    uint64_t *d_in = nullptr;
    int rc = rthc_dev_alloc(ctx, &d_in, (size_t)m);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_in, keys, sizeof(uint64_t) * (size_t)m, cudaMemcpyHostToDevice, ctx->stream));
    rc = linkage_from_forest_device(ctx, d_in, m, n, Z_out);
    rthc_dev_free(ctx, &d_in);
    return rc;
}

extern "C" int rthc_linkage_from_forest_device(rthc_ctx *ctx, const uint64_t *d_keys, int64_t m, int64_t n, double *Z_out)
{
    if (!ctx || !d_keys || !Z_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (n < 2 || m < n - 1) return rthc_fail(ctx, RTHC_E_ARG, "forest union has %lld slots for n=%lld", (long long)m, (long long)n);
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    return linkage_from_forest_device(ctx, d_keys, m, n, Z_out);
}

extern "C" int rthc_cut(rthc_ctx *ctx, const double *t, int T, int32_t *labels_out)
{
    if (!ctx || !t || !labels_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_linkage) return rthc_fail(ctx, RTHC_E_STATE, "rthc_cut before rthc_linkage");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    return rthc_run_cut(ctx, t, T, labels_out);
}
