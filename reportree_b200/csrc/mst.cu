// mst.cu -- single linkage on the device-resident condensed matrix.
//
// Replaces scipy's `_hierarchy.mst_single_linkage` behind
// linkage(y, "single") (scripts/partitioning_HC.py:197; scipy _hierarchy_impl.py:960-961).
//
// (1) Boruvka minimum spanning forest over the condensed uint16 rows a context owns,
//     under the strict total order key = (w, i, j) packed in 64 bits, so the forest is
//     unique and identical for 1/2/4/8-way shards.  Each round is one streaming pass
//     over the owned part of the matrix (HBM-bound: 2 bytes per pair per round).
// (2) A bitonic sort of the <= 8 (n-1) packed edges.
// (3) Device Prim in scipy's visiting order (first strictly-smaller key wins) for the
//     mode that reproduces scipy's Z byte for byte.
#include <algorithm>

#include "rthc_internal.cuh"

#define SCAN_THREADS 256

__global__ void rthc_iota_i32(int32_t *a, int64_t n)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) a[v] = (int32_t)v;
}
__global__ void rthc_fill_u64(uint64_t *a, int64_t n, uint64_t val)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) a[v] = val;
}

__device__ __forceinline__ void consider(uint32_t w, uint32_t i, uint32_t j, int32_t ci,
                                         const int32_t *__restrict__ comp, uint64_t &mine)
{
    if (comp[j] != ci) {
        const uint64_t key = rthc_pack_key(w, i, j);
        if (key < mine) mine = key;
    }
}

// ROW PASS.  One CTA per owned row i: stream D[i, i+1..n-1] (16-byte loads), lightest edge
// leaving comp[i] through a column j > i; block reduction, one atomic per row.
__global__ void __launch_bounds__(SCAN_THREADS)
rthc_boruvka_scan(const uint16_t *__restrict__ cond, const int64_t *__restrict__ row_base,
                  const int32_t *__restrict__ owned_rows, int64_t n, const int32_t *__restrict__ comp,
                  uint64_t *best)
{
    const uint32_t i = (uint32_t)owned_rows[blockIdx.x];
    const int64_t len = n - 1 - (int64_t)i;
    if (len <= 0) return;
    const int64_t base = row_base[i];
    const int32_t ci = comp[i];
    const int tid = threadIdx.x;
    uint64_t mine = RTHC_KEY_INF;

    int64_t head = (8 - (base & 7)) & 7;
    if (head > len) head = len;
    if (tid < head) consider(cond[base + tid], i, i + 1 + (uint32_t)tid, ci, comp, mine);
    const int64_t nvec = (len - head) >> 3;
    const uint4 *vec = reinterpret_cast<const uint4 *>(cond + base + head);
    for (int64_t v = tid; v < nvec; v += SCAN_THREADS) {
        const uint4 q = vec[v];
        const uint32_t j = i + 1 + (uint32_t)head + (uint32_t)(v << 3);
        consider(q.x & 0xFFFFu, i, j + 0, ci, comp, mine);
        consider(q.x >> 16, i, j + 1, ci, comp, mine);
        consider(q.y & 0xFFFFu, i, j + 2, ci, comp, mine);
        consider(q.y >> 16, i, j + 3, ci, comp, mine);
        consider(q.z & 0xFFFFu, i, j + 4, ci, comp, mine);
        consider(q.z >> 16, i, j + 5, ci, comp, mine);
        consider(q.w & 0xFFFFu, i, j + 6, ci, comp, mine);
        consider(q.w >> 16, i, j + 7, ci, comp, mine);
    }
    const int64_t done = head + (nvec << 3);
    if (done + tid < len)
        consider(cond[base + done + tid], i, i + 1 + (uint32_t)(done + tid), ci, comp, mine);

    // block min
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        uint64_t o = __shfl_xor_sync(0xffffffffu, mine, s);
        if (o < mine) mine = o;
    }
    __shared__ uint64_t wmin[SCAN_THREADS / 32];
    if ((tid & 31) == 0) wmin[tid >> 5] = mine;
    __syncthreads();
    if (tid == 0) {
        for (int w = 1; w < SCAN_THREADS / 32; ++w)
            if (wmin[w] < mine) mine = wmin[w];
        if (mine != RTHC_KEY_INF && mine < best[ci])
            atomicMin((unsigned long long *)&best[ci], (unsigned long long)mine);
    }
}

// COLUMN PASS.  Edge (i, j), i < j, is also a candidate for comp[j].  A CTA owns 256
// consecutive columns and a chunk of COL_ROWS rows; thread t walks down column j: for a fixed
// row the 256 threads read 512 contiguous bytes of the condensed row, comp[i] and row_base[i]
// are warp-uniform.  One filtered atomic per thread per chunk.
#define COL_ROWS 512
__global__ void __launch_bounds__(256)
rthc_boruvka_scan_cols(const uint16_t *__restrict__ cond, const int64_t *__restrict__ row_base, int64_t n,
                       const int32_t *__restrict__ comp, uint64_t *best)
{
    const int64_t j = (int64_t)blockIdx.x * 256 + threadIdx.x;
    const int64_t r0 = (int64_t)blockIdx.y * COL_ROWS;
    const int64_t jmax = (int64_t)blockIdx.x * 256 + 255;
    if (r0 >= jmax || r0 >= n) return;
    int64_t r1 = r0 + COL_ROWS;
    if (r1 > jmax) r1 = jmax;            // rows i < j only
    if (r1 > n) r1 = n;
    const bool live = j < n;
    const int32_t cj = live ? comp[j] : -1;
    uint64_t mine = RTHC_KEY_INF;
#pragma unroll 4
    for (int64_t i = r0; i < r1; ++i) {
        const int64_t base = row_base[i];
        if (base < 0) continue;           // row not owned by this shard (uniform)
        if (live && i < j) {
            const uint32_t w = cond[base + (j - i - 1)];
            if (comp[i] != cj) {
                const uint64_t key = rthc_pack_key(w, (uint32_t)i, (uint32_t)j);
                if (key < mine) mine = key;
            }
        }
    }
    if (mine != RTHC_KEY_INF && mine < best[cj])
        atomicMin((unsigned long long *)&best[cj], (unsigned long long)mine);
}

// same round step over an explicit edge list (merge of the ranks' forests)
__global__ void rthc_boruvka_scan_edges(const uint64_t *__restrict__ keys, int64_t m,
                                        const int32_t *__restrict__ comp, uint64_t *best)
{
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= m) return;
    const uint64_t key = keys[e];
    const int32_t ci = comp[rthc_key_i(key)], cj = comp[rthc_key_j(key)];
    if (ci == cj) return;
    if (key < best[ci]) atomicMin((unsigned long long *)&best[ci], (unsigned long long)key);
    if (key < best[cj]) atomicMin((unsigned long long *)&best[cj], (unsigned long long)key);
}

// every component root hooks along its lightest edge; a mutual pair keeps the smaller id as root
__global__ void rthc_boruvka_hook(const uint64_t *__restrict__ best, const int32_t *__restrict__ comp,
                                  int32_t *__restrict__ parent, int64_t n, uint64_t *__restrict__ out_keys,
                                  unsigned int *out_count)
{
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    if (comp[c] != (int32_t)c) { parent[c] = comp[c]; return; }
    const uint64_t key = best[c];
    if (key == RTHC_KEY_INF) { parent[c] = (int32_t)c; return; }
    const int32_t ci = comp[rthc_key_i(key)], cj = comp[rthc_key_j(key)];
    const int32_t other = (ci == (int32_t)c) ? cj : ci;
    const bool mutual = best[other] == key;
    if (mutual && (int32_t)c < other) { parent[c] = (int32_t)c; return; }
    parent[c] = other;
    out_keys[atomicAdd(out_count, 1u)] = key;
}

__global__ void rthc_pointer_jump(int32_t *parent, int64_t n)
{
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n) return;
    const int32_t p = parent[c];
    const int32_t pp = parent[p];
    if (pp != p) parent[c] = pp;
}
// comp[v] = root of v in the hooked forest.  Every thread chases its own chain (read-only, so no
// ordering issues); chains are short because each round at least halves the components and a
// hooked root points straight at a previous root.
__global__ void rthc_relabel(int32_t *comp, const int32_t *__restrict__ parent, int64_t n)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n) return;
    int32_t r = parent[v];
    for (;;) {
        const int32_t p = parent[r];
        if (p == r) break;
        r = p;
    }
    comp[v] = r;
}

__global__ void rthc_bitonic_step(uint64_t *a, uint32_t j, uint32_t k, uint32_t N)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    uint32_t ixj = i ^ j;
    if (ixj > i) {
        uint64_t x = a[i], y = a[ixj];
        bool asc = (i & k) == 0;
        if ((x > y) == asc) { a[i] = y; a[ixj] = x; }
    }
}

static int sort_keys_device(rthc_ctx *ctx, uint64_t *d_keys, int64_t m, uint64_t **d_sorted_out)
{
    // pad to a power of two with +inf keys
    uint32_t N = 1;
    while ((int64_t)N < m) N <<= 1;
    uint64_t *d_pad = nullptr;
    int rc = rthc_dev_alloc(ctx, &d_pad, (size_t)N);
    if (rc) return rc;
    rthc_fill_u64<<<(N + 255) / 256, 256, 0, ctx->stream>>>(d_pad, N, RTHC_KEY_INF);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_pad, d_keys, sizeof(uint64_t) * (size_t)m, cudaMemcpyDeviceToDevice,
                                   ctx->stream));
    for (uint32_t k = 2; k <= N; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            rthc_bitonic_step<<<(N + 255) / 256, 256, 0, ctx->stream>>>(d_pad, j, k, N);
            RTHC_LAUNCH_CHECK(ctx);
        }
    *d_sorted_out = d_pad;
    return 0;
}

// shared round loop: `scan` launches the lightest-edge pass for the current components
template <typename ScanFn>
static int boruvka_rounds(rthc_ctx *ctx, int64_t n, ScanFn scan, std::vector<uint64_t> &keys_out)
{
    int32_t *d_comp = nullptr, *d_parent = nullptr;
    uint64_t *d_best = nullptr, *d_keys = nullptr, *d_sorted = nullptr;
    unsigned int *d_count = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_comp, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_parent, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_best, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_keys, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_count, 1))) return rc;
    const unsigned nb = (unsigned)((n + 255) / 256);
    rthc_iota_i32<<<nb, 256, 0, ctx->stream>>>(d_comp, n);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemsetAsync(d_count, 0, sizeof(unsigned int), ctx->stream));
    unsigned int total = 0;
    int rounds = 0;
    while ((int64_t)total < n - 1) {
        rthc_fill_u64<<<nb, 256, 0, ctx->stream>>>(d_best, n, RTHC_KEY_INF);
        RTHC_LAUNCH_CHECK(ctx);
        if ((rc = scan(d_comp, d_best))) return rc;
        rthc_boruvka_hook<<<nb, 256, 0, ctx->stream>>>(d_best, d_comp, d_parent, n, d_keys, d_count);
        RTHC_LAUNCH_CHECK(ctx);
        unsigned int now = 0;
        RTHC_CUDA(ctx, cudaMemcpyAsync(&now, d_count, sizeof(unsigned int), cudaMemcpyDeviceToHost, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        ++rounds;
        if (now == total) break;   // forest of a disconnected (sharded) edge set is complete
        total = now;
        for (int it = 0; it < 4; ++it) {      // 16x path compression bounds the chase below on adversarial chains
            rthc_pointer_jump<<<nb, 256, 0, ctx->stream>>>(d_parent, n);
            RTHC_LAUNCH_CHECK(ctx);
        }
        rthc_relabel<<<nb, 256, 0, ctx->stream>>>(d_comp, d_parent, n);
        RTHC_LAUNCH_CHECK(ctx);
    }
    ctx->boruvka_rounds = rounds;
    keys_out.resize(total);
    if (total > 0) {
        if ((rc = sort_keys_device(ctx, d_keys, (int64_t)total, &d_sorted))) return rc;
        RTHC_CUDA(ctx, cudaMemcpyAsync(keys_out.data(), d_sorted, sizeof(uint64_t) * total,
                                       cudaMemcpyDeviceToHost, ctx->stream));
    }
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(&d_sorted);
    rthc_dev_free(&d_comp); rthc_dev_free(&d_parent); rthc_dev_free(&d_best);
    rthc_dev_free(&d_keys); rthc_dev_free(&d_count);
    return 0;
}

int rthc_boruvka_owned(rthc_ctx *ctx, std::vector<uint64_t> &keys)
{
    const int64_t n = ctx->n;
    const int64_t n_owned = (int64_t)ctx->h_owned_rows.size();
    auto scan = [&](int32_t *d_comp, uint64_t *d_best) -> int {
        if (n_owned > 0) {
            rthc_boruvka_scan<<<(unsigned)n_owned, SCAN_THREADS, 0, ctx->stream>>>(
                ctx->d_cond, ctx->d_row_base, ctx->d_owned_rows, n, d_comp, d_best);
            RTHC_LAUNCH_CHECK(ctx);
            dim3 grd((unsigned)((n + 255) / 256), (unsigned)((n + COL_ROWS - 1) / COL_ROWS));
            rthc_boruvka_scan_cols<<<grd, 256, 0, ctx->stream>>>(ctx->d_cond, ctx->d_row_base, n, d_comp, d_best);
            RTHC_LAUNCH_CHECK(ctx);
        }
        return 0;
    };
    return boruvka_rounds(ctx, n, scan, keys);
}

int rthc_boruvka_edge_list(rthc_ctx *ctx, const uint64_t *h_keys, int64_t m, int64_t n,
                           std::vector<uint64_t> &keys)
{
    uint64_t *d_in = nullptr;
    int rc = rthc_dev_alloc(ctx, &d_in, (size_t)m);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_in, h_keys, sizeof(uint64_t) * (size_t)m, cudaMemcpyHostToDevice,
                                   ctx->stream));
    auto scan = [&](int32_t *d_comp, uint64_t *d_best) -> int {
        if (m > 0) {
            rthc_boruvka_scan_edges<<<(unsigned)((m + 255) / 256), 256, 0, ctx->stream>>>(d_in, m, d_comp, d_best);
            RTHC_LAUNCH_CHECK(ctx);
        }
        return 0;
    };
    rc = boruvka_rounds(ctx, n, scan, keys);
    rthc_dev_free(&d_in);
    return rc;
}
