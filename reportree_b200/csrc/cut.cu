// cut.cu -- multi-threshold cut of the dendrogram, all thresholds in one batch.
//
// Replaces scipy's `_hierarchy.cluster_dist` behind fcluster(Z, t, criterion="distance")
// (scripts/partitioning_HC.py:649, :689, :722, :751; scipy _hierarchy_impl.py:2589-2590).
//
// Input (uploaded by rthc_linkage): the dendrogram as a parent array over its 2n-1 nodes
// (leaves 0..n-1, merge k = node n+k, merges sorted by height) and the position of every node in
// scipy's numbering walk (`event`).  For a monotone dendrogram the flat clusters at cutoff t are
// the subtrees of the merges {k : h_k <= t} = {k < K(t)}; ancestors have increasing ids, so the
// leader of leaf v is its highest ancestor with id < n + K(t), found by binary lifting in
// O(log n) independent of every other (threshold, leaf) pair.  A leaf whose parent is already
// beyond the cutoff is its own cluster.  scipy numbers clusters in the order its walk meets the
// leader (or the lone leaf), so the id of a cluster is the rank of its leader's event among the
// events present at that threshold: a flag array per threshold + an inclusive scan.
// Everything is batched over thresholds (grid.y): ~25 launches per call instead of 7 per threshold.
#include <algorithm>
#include <numeric>

#include "rthc_internal.cuh"

#define CUT_SCAN_ELEMS 4096
#define CUT_ROOT_PARENT 0x7FFFFFFF

__global__ void rthc_cut_lift(const int32_t *__restrict__ prev, int32_t *__restrict__ next, int64_t nodes)
{
    const int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x >= nodes) return;
    const int32_t p = prev[x];
    next[x] = p == CUT_ROOT_PARENT ? CUT_ROOT_PARENT : prev[p];
}

// key[t][v] = event of the leader of leaf v at threshold t; flags[t][event] = 1
__global__ void rthc_cut_leaders(const int32_t *__restrict__ up, int levels, int64_t nodes, int64_t n,
                                 const int32_t *__restrict__ limit, const int32_t *__restrict__ event,
                                 int32_t *__restrict__ key, int32_t *__restrict__ flags, int64_t flag_stride)
{
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int t = blockIdx.y;
    if (v >= n) return;
    const int32_t lim = limit[t];                  // n + K(t): first node id that is NOT merged at this cutoff
    int32_t x = (int32_t)v;
    for (int j = levels - 1; j >= 0; --j) {
        const int32_t a = up[(int64_t)j * nodes + x];
        if (a < lim) x = a;
    }
    const int32_t e = event[x];
    RTHC_DEV_ASSERT(x >= 0 && x < nodes && e >= 0 && e < nodes);
    key[(int64_t)t * n + v] = e;
    flags[(int64_t)t * flag_stride + e] = 1;
}

// inclusive scan of flags[t][0..m): every CTA scans 4,096 entries in place and publishes its total
__global__ void __launch_bounds__(1024) rthc_cut_scan_blocks(int32_t *flags_all, int64_t m, int64_t flag_stride,
                                                             int32_t *aux_all, int nb)
{
    __shared__ int32_t wsum[32];
    int32_t *flags = flags_all + (int64_t)blockIdx.y * flag_stride;
    int32_t *aux = aux_all + (int64_t)blockIdx.y * nb;
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

// exclusive scan of the per-CTA totals of threshold blockIdx.x, in place
__global__ void __launch_bounds__(1024) rthc_cut_scan_aux(int32_t *aux_all, int nb)
{
    __shared__ int32_t wsum[32];
    __shared__ int32_t carry;
    int32_t *aux = aux_all + (int64_t)blockIdx.x * nb;
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

__global__ void rthc_cut_labels(const int32_t *__restrict__ key, const int32_t *__restrict__ scanned, int64_t flag_stride,
                                const int32_t *__restrict__ aux, int nb, int32_t *__restrict__ labels, int64_t n)
{
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int t = blockIdx.y;
    if (v >= n) return;
    const int32_t e = key[(int64_t)t * n + v];
    RTHC_DEV_ASSERT(e >= 0 && e < 2 * n - 1);
    labels[(int64_t)t * n + v] = scanned[(int64_t)t * flag_stride + e] + aux[(int64_t)t * nb + e / CUT_SCAN_ELEMS];
}

static int build_lifting_table(rthc_ctx *ctx)
{
    const int64_t nodes = 2 * ctx->n - 1;
    int levels = 1;
    while (((int64_t)1 << levels) < nodes) ++levels;          // 2^levels >= any root path length
    int rc = rthc_dev_alloc(ctx, &ctx->d_up, (size_t)levels * (size_t)nodes);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_up, ctx->d_tpar, sizeof(int32_t) * (size_t)nodes, cudaMemcpyDeviceToDevice, ctx->stream));
    const unsigned nb = (unsigned)((nodes + 255) / 256);
    for (int j = 1; j < levels; ++j) {
        rthc_cut_lift<<<nb, 256, 0, ctx->stream>>>(ctx->d_up + (size_t)(j - 1) * nodes, ctx->d_up + (size_t)j * nodes, nodes);
        RTHC_LAUNCH_CHECK(ctx);
    }
    ctx->up_levels = levels;
    return 0;
}

int rthc_run_cut(rthc_ctx *ctx, const double *t, int T, int32_t *labels_out)
{
    const int64_t n = ctx->n, m = ctx->m_edges, nodes = 2 * n - 1;
    if (T <= 0) return 0;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    int rc;
    if (!ctx->d_up && (rc = build_lifting_table(ctx))) return rc;
    // merges with height <= cutoff (same float64 comparison as scipy's MD <= cutoff)
    std::vector<int32_t> limit((size_t)T);
    for (int q = 0; q < T; ++q) {
        int64_t K = (int64_t)(std::upper_bound(ctx->h_heights.begin(), ctx->h_heights.end(), t[q]) - ctx->h_heights.begin());
        if (K > m) K = m;
        limit[(size_t)q] = (int32_t)(n + K);
    }
    // thresholds are processed in chunks that bound device memory (labels + keys + flags)
    // ... and the threshold index rides on gridDim.y (and gridDim.x of the aux scan), limit 65,535
    const int64_t chunkT = std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(T, 60000),
                                                                  (int64_t)(1ll << 27) / std::max<int64_t>(n, 1)));
    const int64_t flag_stride = (nodes + 3) / 4 * 4;
    const int nbs = (int)((nodes + CUT_SCAN_ELEMS - 1) / CUT_SCAN_ELEMS);
    int32_t *d_limit = nullptr, *d_key = nullptr, *d_flags = nullptr, *d_labels = nullptr, *d_aux = nullptr;
    if ((rc = rthc_dev_alloc(ctx, &d_limit, (size_t)T))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_key, (size_t)(chunkT * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_labels, (size_t)(chunkT * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_flags, (size_t)(chunkT * flag_stride)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_aux, (size_t)(chunkT * nbs)))) return rc;
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_limit, limit.data(), sizeof(int32_t) * (size_t)T, cudaMemcpyHostToDevice, ctx->stream));
    const unsigned nb = (unsigned)((n + 255) / 256);
    if (!ctx->copy_stream) {
        RTHC_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking));
        for (int k = 0; k < 4; ++k) RTHC_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_piece[k], cudaEventDisableTiming));
    }
    for (int64_t q0 = 0; q0 < T; q0 += chunkT) {
        const int64_t tc = std::min<int64_t>(chunkT, T - q0);
        // a chunk is cut in up to 4 pieces of thresholds: the labels of piece p travel to the host (copy stream) while
        // the kernels of piece p + 1 run; small results stay in one piece (a launch set costs more than it hides)
        const int np = (tc * n * 4 >= (4ll << 20)) ? (int)std::min<int64_t>(4, tc) : 1;
        RTHC_CUDA(ctx, cudaMemsetAsync(d_flags, 0, sizeof(int32_t) * (size_t)(tc * flag_stride), ctx->stream));
        for (int p = 0; p < np; ++p) {
            const int64_t p0 = tc * p / np, p1 = tc * (p + 1) / np, pc = p1 - p0;
            if (pc == 0) continue;
            int32_t *key = d_key + p0 * n, *flags = d_flags + p0 * flag_stride, *aux = d_aux + p0 * nbs, *lab = d_labels + p0 * n;
            rthc_cut_leaders<<<dim3(nb, (unsigned)pc), 256, 0, ctx->stream>>>(ctx->d_up, ctx->up_levels, nodes, n, d_limit + q0 + p0,
                                                                              ctx->d_event, key, flags, flag_stride);
            RTHC_LAUNCH_CHECK(ctx);
            rthc_cut_scan_blocks<<<dim3((unsigned)nbs, (unsigned)pc), 1024, 0, ctx->stream>>>(flags, nodes, flag_stride, aux, nbs);
            RTHC_LAUNCH_CHECK(ctx);
            rthc_cut_scan_aux<<<(unsigned)pc, 1024, 0, ctx->stream>>>(aux, nbs);
            RTHC_LAUNCH_CHECK(ctx);
            rthc_cut_labels<<<dim3(nb, (unsigned)pc), 256, 0, ctx->stream>>>(key, flags, flag_stride, aux, nbs, lab, n);
            RTHC_LAUNCH_CHECK(ctx);
            if (q0 + tc >= T && p == np - 1) RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
            RTHC_CUDA(ctx, cudaEventRecord(ctx->ev_piece[p], ctx->stream));
            RTHC_CUDA(ctx, cudaStreamWaitEvent(ctx->copy_stream, ctx->ev_piece[p], 0));
            RTHC_CUDA(ctx, cudaMemcpyAsync(labels_out + (q0 + p0) * n, lab, sizeof(int32_t) * (size_t)(pc * n), cudaMemcpyDeviceToHost,
                                           ctx->copy_stream));
        }
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->copy_stream));      // the device buffers are reused by the next chunk
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[4] = ms;
    rthc_dev_free(ctx, &d_limit); rthc_dev_free(ctx, &d_key); rthc_dev_free(ctx, &d_flags);
    rthc_dev_free(ctx, &d_labels); rthc_dev_free(ctx, &d_aux);
    return 0;
}
