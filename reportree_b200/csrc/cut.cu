// cut.cu -- multi-threshold cut with a lock-free union-find on the device.
//
// Replaces scipy's `_hierarchy.cluster_dist` behind fcluster(Z, t, criterion="distance")
// (scripts/partitioning_HC.py:649, :689, :722, :751; scipy _hierarchy_impl.py:2589-2590).
//
// Input (uploaded by rthc_linkage): the n-1 merges of the dendrogram in Z order (sorted by
// height) as (leaf representative a_k, leaf representative b_k, height h_k), plus the
// position of every dendrogram node in scipy's numbering traversal:
//     epre[k]  : when internal node k is first visited (where a leader takes its id)
//     eleaf[v] : when leaf v is numbered (at its parent's final visit, left before right)
// For a monotone dendrogram the flat clusters at cutoff t are the connected components of
// merges {k : h_k <= t}; the leader of a component is its last merge.  Thresholds are
// swept in ascending order, so every merge is united exactly once (atomicCAS hooking,
// path halving); a cluster's id is the rank of its leader/leaf event among the events of
// that threshold, which reproduces scipy's fcluster numbering exactly.
#include <algorithm>
#include <numeric>

#include "rthc_internal.cuh"

__device__ __forceinline__ int32_t uf_find(int32_t *parent, int32_t v)
{
    int32_t p = parent[v];
    while (p != v) {
        int32_t gp = parent[p];
        if (gp != p) parent[v] = gp;   // path halving (benign race)
        v = p;
        p = gp;
        if (p == v) break;
        p = parent[v];
    }
    return v;
}

__global__ void rthc_cut_init(int32_t *parent, int32_t *last, int64_t n)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) { parent[v] = (int32_t)v; last[v] = -1; }
}

__global__ void rthc_cut_union(int32_t *parent, const int32_t *__restrict__ ea,
                               const int32_t *__restrict__ eb, int64_t k0, int64_t k1)
{
    int64_t k = k0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= k1) return;
    int32_t a = ea[k], b = eb[k];
    for (;;) {
        a = uf_find(parent, a);
        b = uf_find(parent, b);
        if (a == b) return;
        if (a < b) { int32_t t = a; a = b; b = t; }          // hook larger root under smaller
        if (atomicCAS(&parent[a], a, b) == a) return;
    }
}

__global__ void rthc_cut_last(int32_t *parent, int32_t *last, const int32_t *__restrict__ ea,
                              int64_t k0, int64_t k1)
{
    int64_t k = k0 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= k1) return;
    atomicMax(&last[uf_find(parent, ea[k])], (int32_t)k);
}

__global__ void rthc_cut_keys(int32_t *parent, const int32_t *__restrict__ last,
                              const int32_t *__restrict__ epre, const int32_t *__restrict__ eleaf,
                              int32_t *__restrict__ key, int32_t *__restrict__ flags, int64_t n)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    const int32_t r = uf_find(parent, (int32_t)v);
    const int32_t lk = last[r];
    const int32_t e = lk < 0 ? eleaf[v] : epre[lk];
    key[v] = e;
    flags[e] = 1;
}

// inclusive scan of flags[0..m) (m = 2n - 1): every CTA scans 4,096 entries in place and
// publishes its total; one CTA then turns the totals into exclusive offsets, which the label
// kernel adds back.  (A single-CTA scan of the whole array was 85 % of the cut stage.)
#define CUT_SCAN_ELEMS 4096
__global__ void __launch_bounds__(1024) rthc_cut_scan_blocks(int32_t *flags, int64_t m, int32_t *aux)
{
    __shared__ int32_t wsum[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t idx = (int64_t)blockIdx.x * CUT_SCAN_ELEMS + 4 * tid;
    int4 v = make_int4(0, 0, 0, 0);
    if (idx + 3 < m) v = *reinterpret_cast<const int4 *>(flags + idx);
    else {
        if (idx < m) v.x = flags[idx];
        if (idx + 1 < m) v.y = flags[idx + 1];
        if (idx + 2 < m) v.z = flags[idx + 2];
    }
    v.y += v.x; v.z += v.y; v.w += v.z;
    int32_t inc = v.w;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int32_t t = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += t;
    }
    if (lane == 31) wsum[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        int32_t w = wsum[lane], winc = w;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int32_t t = __shfl_up_sync(0xffffffffu, winc, d);
            if (lane >= d) winc += t;
        }
        wsum[lane] = winc - w;
        if (lane == 31) aux[blockIdx.x] = winc;
    }
    __syncthreads();
    const int32_t before = inc - v.w + wsum[warp];
    v.x += before; v.y += before; v.z += before; v.w += before;
    if (idx + 3 < m) *reinterpret_cast<int4 *>(flags + idx) = v;
    else {
        if (idx < m) flags[idx] = v.x;
        if (idx + 1 < m) flags[idx + 1] = v.y;
        if (idx + 2 < m) flags[idx + 2] = v.z;
    }
}

// exclusive scan of the per-CTA totals in place (one CTA; nb is a few hundred)
__global__ void __launch_bounds__(1024) rthc_cut_scan_aux(int32_t *aux, int nb)
{
    __shared__ int32_t wsum[32];
    __shared__ int32_t carry;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) carry = 0;
    __syncthreads();
    for (int base = 0; base < nb; base += 1024) {
        const int idx = base + tid;
        const int32_t v = idx < nb ? aux[idx] : 0;
        int32_t inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            int32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        if (lane == 31) wsum[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            int32_t w = wsum[lane], winc = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                int32_t t = __shfl_up_sync(0xffffffffu, winc, d);
                if (lane >= d) winc += t;
            }
            wsum[lane] = winc - w;
        }
        __syncthreads();
        const int32_t c = carry;
        inc += wsum[warp] + c;
        if (idx < nb) aux[idx] = inc - v;
        __syncthreads();
        if (tid == 1023) carry = inc;
        __syncthreads();
    }
}

__global__ void rthc_cut_labels(const int32_t *__restrict__ key, const int32_t *__restrict__ scanned,
                                const int32_t *__restrict__ aux, int32_t *__restrict__ labels, int64_t n)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) {
        const int32_t e = key[v];
        labels[v] = scanned[e] + aux[e / CUT_SCAN_ELEMS];
    }
}

int rthc_run_cut(rthc_ctx *ctx, const double *t, int T, int32_t *labels_out)
{
    const int64_t n = ctx->n, m = ctx->m_edges;
    if (T <= 0) return 0;
    std::vector<int> order((size_t)T);
    std::iota(order.begin(), order.end(), 0);
    std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return t[a] < t[b]; });

    int32_t *d_parent = nullptr, *d_last = nullptr, *d_key = nullptr, *d_flags = nullptr, *d_labels = nullptr, *d_aux = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_parent, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_last, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_key, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_flags, (size_t)(2 * n + 4)))) return rc;
    const int scan_blocks = (int)((2 * n - 1 + CUT_SCAN_ELEMS - 1) / CUT_SCAN_ELEMS);
    if ((rc = rthc_dev_alloc(ctx, &d_aux, (size_t)scan_blocks))) return rc;
    // labels are produced in chunks of thresholds to bound device memory
    const int64_t chunkT = std::max<int64_t>(1, std::min<int64_t>(T, (int64_t)(1ll << 28) / std::max<int64_t>(n, 1)));
    if ((rc = rthc_dev_alloc(ctx, &d_labels, (size_t)(chunkT * n)))) return rc;

    const unsigned nb = (unsigned)((n + 255) / 256);
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    rthc_cut_init<<<nb, 256, 0, ctx->stream>>>(d_parent, d_last, n);
    RTHC_LAUNCH_CHECK(ctx);
    int64_t kprev = 0;
    int64_t in_chunk = 0, chunk_start = 0;
    auto flush = [&]() -> int {
        // thresholds order[chunk_start .. chunk_start + in_chunk) are in d_labels rows 0..in_chunk
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (int64_t q = 0; q < in_chunk; ++q) {
            int dst = order[(size_t)(chunk_start + q)];
            RTHC_CUDA(ctx, cudaMemcpyAsync(labels_out + (int64_t)dst * n, d_labels + q * n,
                                           sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost, ctx->stream));
        }
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        chunk_start += in_chunk;
        in_chunk = 0;
        return 0;
    };
    for (int q = 0; q < T; ++q) {
        const double cutoff = t[order[(size_t)q]];
        // merges with height <= cutoff (same float64 comparison as scipy's MD <= cutoff)
        int64_t knew = (int64_t)(std::upper_bound(ctx->h_heights.begin(), ctx->h_heights.end(), cutoff) -
                                 ctx->h_heights.begin());
        if (knew > m) knew = m;
        if (knew > kprev) {
            const unsigned eb = (unsigned)((knew - kprev + 255) / 256);
            rthc_cut_union<<<eb, 256, 0, ctx->stream>>>(d_parent, ctx->d_ea, ctx->d_eb, kprev, knew);
            RTHC_LAUNCH_CHECK(ctx);
            rthc_cut_last<<<eb, 256, 0, ctx->stream>>>(d_parent, d_last, ctx->d_ea, kprev, knew);
            RTHC_LAUNCH_CHECK(ctx);
            kprev = knew;
        }
        RTHC_CUDA(ctx, cudaMemsetAsync(d_flags, 0, sizeof(int32_t) * (size_t)(2 * n), ctx->stream));
        rthc_cut_keys<<<nb, 256, 0, ctx->stream>>>(d_parent, d_last, ctx->d_epre, ctx->d_eleaf, d_key, d_flags, n);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_cut_scan_blocks<<<scan_blocks, 1024, 0, ctx->stream>>>(d_flags, 2 * n - 1, d_aux);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_cut_scan_aux<<<1, 1024, 0, ctx->stream>>>(d_aux, scan_blocks);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_cut_labels<<<nb, 256, 0, ctx->stream>>>(d_key, d_flags, d_aux, d_labels + in_chunk * n, n);
        RTHC_LAUNCH_CHECK(ctx);
        if (++in_chunk == chunkT)
            if ((rc = flush())) return rc;
    }
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    if ((rc = flush())) return rc;
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[4] = ms;
    rthc_dev_free(ctx, &d_parent); rthc_dev_free(ctx, &d_last); rthc_dev_free(ctx, &d_key);
    rthc_dev_free(ctx, &d_flags); rthc_dev_free(ctx, &d_labels); rthc_dev_free(ctx, &d_aux);
    return 0;
}
