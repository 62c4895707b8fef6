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

// merges (x_k, y_k, w_k) in final (sorted) order -> Z, representatives, traversal events
static int finish_linkage(rthc_ctx *ctx, int64_t n, const std::vector<int32_t> &ex,
                          const std::vector<int32_t> &ey, const std::vector<double> &ew, double *Z)
{
    const int64_t m = n - 1;
    if ((int64_t)ex.size() != m)
        return rthc_fail(ctx, RTHC_E_STATE, "linkage produced %zu merges, expected %lld", ex.size(), (long long)m);
    std::vector<int32_t> parent((size_t)(2 * n), -1), size((size_t)(2 * n), 1), tpar((size_t)(2 * n), 0x7FFFFFFF);
    auto find = [&](int32_t v) {
        int32_t r = v;
        while (parent[(size_t)r] != -1) r = parent[(size_t)r];
        while (parent[(size_t)v] != -1) { int32_t t = parent[(size_t)v]; parent[(size_t)v] = r; v = t; }
        return r;
    };
    for (int64_t k = 0; k < m; ++k) {
        int32_t rx = find(ex[(size_t)k]), ry = find(ey[(size_t)k]);
        if (rx == ry) return rthc_fail(ctx, RTHC_E_STATE, "merge %lld joins a cluster with itself", (long long)k);
        if (rx > ry) std::swap(rx, ry);
        const int32_t id = (int32_t)(n + k);
        Z[4 * k + 0] = rx; Z[4 * k + 1] = ry; Z[4 * k + 2] = ew[(size_t)k];
        size[(size_t)id] = size[(size_t)rx] + size[(size_t)ry];
        Z[4 * k + 3] = size[(size_t)id];
        parent[(size_t)rx] = id; parent[(size_t)ry] = id;      // union-find (gets path-compressed)
        tpar[(size_t)rx] = id; tpar[(size_t)ry] = id;          // the dendrogram itself
    }
    // Position of every node in scipy's numbering walk (cluster_dist / cluster_monocrit): a node is
    // numbered when first visited, then its left and right internal subtrees, then its leaf
    // children (left before right).  A subtree of s leaves holds 2 s - 1 events, so the positions
    // follow top-down (parents have larger ids than children): no stack, no recursion.
    std::vector<int32_t> event((size_t)(2 * n - 1), 0);
    event[(size_t)(2 * n - 2)] = 0;
    for (int64_t k = m - 1; k >= 0; --k) {
        const int32_t lc = (int32_t)Z[4 * k], rc = (int32_t)Z[4 * k + 1];
        int32_t cursor = event[(size_t)(n + k)] + 1;
        if (lc >= n) { event[(size_t)lc] = cursor; cursor += 2 * size[(size_t)lc] - 1; }
        if (rc >= n) { event[(size_t)rc] = cursor; cursor += 2 * size[(size_t)rc] - 1; }
        if (lc < n) event[(size_t)lc] = cursor++;
        if (rc < n) event[(size_t)rc] = cursor++;
    }
    ctx->h_heights = ew;
    for (int64_t q = 1; q < m; ++q)
        if (ctx->h_heights[(size_t)q] < ctx->h_heights[(size_t)q - 1])
            return rthc_fail(ctx, RTHC_E_STATE, "merge heights are not monotone at row %lld", (long long)q);
    int rcode;
    if ((rcode = rthc_dev_alloc(ctx, &ctx->d_tpar, (size_t)(2 * n - 1)))) return rcode;
    if ((rcode = rthc_dev_alloc(ctx, &ctx->d_event, (size_t)(2 * n - 1)))) return rcode;
    rthc_dev_free(ctx, &ctx->d_up);
    ctx->up_levels = 0;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_tpar, tpar.data(), sizeof(int32_t) * (size_t)(2 * n - 1), cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_event, event.data(), sizeof(int32_t) * (size_t)(2 * n - 1), cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->m_edges = m;
    ctx->n = n;
    ctx->have_linkage = true;
    return 0;
}

static int finish_from_keys(rthc_ctx *ctx, int64_t n, const std::vector<uint64_t> &keys, double *Z)
{
    std::vector<int32_t> ex(keys.size()), ey(keys.size());
    std::vector<double> ew(keys.size());
    for (size_t k = 0; k < keys.size(); ++k) {
        ex[k] = (int32_t)rthc_key_i(keys[k]);
        ey[k] = (int32_t)rthc_key_j(keys[k]);
        ew[k] = (double)rthc_key_w(keys[k]);
    }
    return finish_linkage(ctx, n, ex, ey, ew, Z);
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
    if (method == RTHC_SINGLE && ctx->single_algo == RTHC_SL_BORUVKA && ctx->d_cond) {
        std::vector<uint64_t> keys;
        if ((rc = rthc_boruvka_owned(ctx, keys))) return rc;
        ex.resize(keys.size()); ey.resize(keys.size()); ew.resize(keys.size());
        for (size_t k = 0; k < keys.size(); ++k) {
            ex[k] = (int32_t)rthc_key_i(keys[k]); ey[k] = (int32_t)rthc_key_j(keys[k]);
            ew[k] = (double)rthc_key_w(keys[k]);
        }
    } else if (method == RTHC_SINGLE) {
        if ((rc = rthc_prim_exact(ctx, ex, ey, ew))) return rc;
        stable_sort_merges(ex, ey, ew);
    } else if (method == RTHC_COMPLETE || method == RTHC_AVERAGE || method == RTHC_WEIGHTED) {
        if ((rc = rthc_nn_chain(ctx, method, ex, ey, ew))) return rc;
        stable_sort_merges(ex, ey, ew);
    } else {
        return rthc_fail(ctx, RTHC_E_ARG, "unsupported linkage method %d", method);
    }
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[3] = ms;
    return finish_linkage(ctx, ctx->n, ex, ey, ew, Z_out);
}

extern "C" int rthc_local_forest(rthc_ctx *ctx, uint64_t *keys_out, int64_t *m_out)
{
    if (!ctx || !keys_out || !m_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_dist || !ctx->d_cond) return rthc_fail(ctx, RTHC_E_STATE, "no uint16 distance matrix yet");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    std::vector<uint64_t> keys;
    int rc = rthc_boruvka_owned(ctx, keys);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[3] = ms;
    std::copy(keys.begin(), keys.end(), keys_out);
    *m_out = (int64_t)keys.size();
    return 0;
}

extern "C" int rthc_linkage_from_forest(rthc_ctx *ctx, const uint64_t *keys, int64_t m, int64_t n, double *Z_out)
{
    if (!ctx || !keys || !Z_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (n < 2 || m < n - 1) return rthc_fail(ctx, RTHC_E_ARG, "forest union has %lld edges for n=%lld", (long long)m, (long long)n);
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    rthc_dev_free(ctx, &ctx->d_tpar); rthc_dev_free(ctx, &ctx->d_event); rthc_dev_free(ctx, &ctx->d_up);
    ctx->have_linkage = false;
    std::vector<uint64_t> mst;
    int rc = rthc_boruvka_edge_list(ctx, keys, m, n, mst);
    if (rc) return rc;
    return finish_from_keys(ctx, n, mst, Z_out);
}

extern "C" int rthc_cut(rthc_ctx *ctx, const double *t, int T, int32_t *labels_out)
{
    if (!ctx || !t || !labels_out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_linkage) return rthc_fail(ctx, RTHC_E_STATE, "rthc_cut before rthc_linkage");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    return rthc_run_cut(ctx, t, T, labels_out);
}
