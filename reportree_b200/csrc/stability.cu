// stability.cu -- SURVEY 8(f) N4: contingency tables of ADJACENT partition columns (the cluster
// stability analysis of scripts/ComparingPartitions/comparing_partitions_v2.py:74-113).
//
// The reference builds, for every pair of neighbouring threshold columns, a dict-of-dicts
// contingency table with a Python loop over the samples (metrics.py:23-35) and then walks
// set(col1) x set(col2) (metrics.py:38-58, :61-91): quadratic in the number of clusters.  All of its
// outputs are functions of the table's cells, and the floating-point sums of getAdjustedWallace
// (metrics.py:142-170) are taken in dict order = order of first appearance while scanning the
// samples.  The device produces exactly that: per pair, the cells (row cluster, column cluster,
// count) sorted by (first sample of the row cluster, first sample of the cell).
//
//   rthc_stab_first    first sample index of every cluster id of every column      (atomicMin)
//   rthc_stab_insert   open-addressing hash table per pair: key (row id, col id) -> count, first
//   rthc_stab_compact  occupied slots -> 64-bit sort keys (first of row << 32 | first of cell)
//   rthc_stab_sort_*   batched bitonic sort (2,048-key chunks in shared memory, global steps above)
//   rthc_stab_emit     sorted keys -> (row id, col id, count)
//
// rthc_adjusted_wallace_sums is the host-side finish of one table: the reference's sums in the
// reference's order, with libm pow() exactly where Python evaluates `x ** 2.0` / `x ** 3.0`.
#include <algorithm>
#include <cmath>

#include "rthc_internal.cuh"

#define STAB_EMPTY 0xFFFFFFFFFFFFFFFFull
#define STAB_SB 2048             // keys per CTA of the shared-memory sort phases

__device__ __forceinline__ uint32_t stab_hash(uint64_t k)
{
    k ^= k >> 33; k *= 0xff51afd7ed558ccdull;
    k ^= k >> 33; k *= 0xc4ceb9fe1a85ec53ull;
    k ^= k >> 33;
    return (uint32_t)k;
}

__global__ void rthc_stab_fill_u32(uint32_t *a, int64_t n, uint32_t v)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = v;
}
__global__ void rthc_stab_fill_u64(uint64_t *a, int64_t n, uint64_t v)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) a[i] = v;
}

// codes: [cols][n]; first: [cols][n]
__global__ void rthc_stab_first(const int32_t *__restrict__ codes, int64_t n, uint32_t *__restrict__ first, int *bad)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int64_t col = blockIdx.y;
    const int32_t c = codes[col * n + s];
    if (c < 0 || c >= n) { *bad = 1; return; }
    atomicMin(&first[col * n + c], (uint32_t)s);
}

// pair p = (column p+1 as ROW cluster, column p as COLUMN cluster), the orientation of
// getContTable(jcol, icol) in metrics4stability (comparing_partitions_v2.py:100)
__global__ void rthc_stab_insert(const int32_t *__restrict__ codes, int64_t n, uint32_t H, uint64_t *__restrict__ hkeys,
                                 uint32_t *__restrict__ hcnt, uint32_t *__restrict__ hfirst)
{
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n) return;
    const int64_t p = blockIdx.y;
    const uint32_t r = (uint32_t)codes[(p + 1) * n + s], c = (uint32_t)codes[p * n + s];
    const uint64_t key = ((uint64_t)r << 32) | c;
    uint64_t *keys = hkeys + p * (int64_t)H;
    uint32_t h = stab_hash(key) & (H - 1);
    for (;;) {
        const uint64_t old = atomicCAS((unsigned long long *)&keys[h], (unsigned long long)STAB_EMPTY, (unsigned long long)key);
        if (old == STAB_EMPTY || old == key) break;
        h = (h + 1) & (H - 1);
    }
    atomicAdd(&hcnt[p * (int64_t)H + h], 1u);
    atomicMin(&hfirst[p * (int64_t)H + h], (uint32_t)s);
}

__global__ void rthc_stab_compact(int64_t n, uint32_t H, const uint64_t *__restrict__ hkeys, const uint32_t *__restrict__ hfirst,
                                  const uint32_t *__restrict__ first, uint32_t N, uint64_t *__restrict__ skeys,
                                  unsigned int *__restrict__ m)
{
    const int64_t h = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (h >= H) return;
    const int64_t p = blockIdx.y;
    const uint64_t key = hkeys[p * (int64_t)H + h];
    if (key == STAB_EMPTY) return;
    const uint32_t r = (uint32_t)(key >> 32);
    const unsigned int pos = atomicAdd(&m[p], 1u);
    skeys[p * (int64_t)N + pos] = ((uint64_t)first[(p + 1) * n + r] << 32) | hfirst[p * (int64_t)H + h];
}

__device__ __forceinline__ void stab_cswap(uint64_t *s, uint32_t i, uint32_t l, bool asc)
{
    const uint64_t x = s[i], y = s[l];
    if ((x > y) == asc) { s[i] = y; s[l] = x; }
}

// every stage with k <= STAB_SB: each CTA sorts its 2,048-key chunk (direction from the global index)
__global__ void __launch_bounds__(STAB_SB / 2) rthc_stab_sort_local(uint64_t *a, uint32_t N)
{
    __shared__ uint64_t s[STAB_SB];
    uint64_t *base = a + (int64_t)blockIdx.y * N + (int64_t)blockIdx.x * STAB_SB;
    const uint32_t t = threadIdx.x, g0 = blockIdx.x * STAB_SB;
    s[t] = base[t]; s[t + STAB_SB / 2] = base[t + STAB_SB / 2];
    __syncthreads();
    for (uint32_t k = 2; k <= STAB_SB; k <<= 1)
        for (uint32_t j = k >> 1; j > 0; j >>= 1) {
            const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
            stab_cswap(s, i, i | j, ((g0 + i) & k) == 0);
            __syncthreads();
        }
    base[t] = s[t]; base[t + STAB_SB / 2] = s[t + STAB_SB / 2];
}
// one step (j >= STAB_SB) of stage k
__global__ void rthc_stab_sort_global(uint64_t *a, uint32_t N, uint32_t j, uint32_t k)
{
    const uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= N / 2) return;
    uint64_t *base = a + (int64_t)blockIdx.y * N;
    const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
    const uint64_t x = base[i], y = base[i | j];
    if ((x > y) == ((i & k) == 0)) { base[i] = y; base[i | j] = x; }
}
// the steps j = STAB_SB/2 .. 1 of stage k > STAB_SB
__global__ void __launch_bounds__(STAB_SB / 2) rthc_stab_sort_tail(uint64_t *a, uint32_t N, uint32_t k)
{
    __shared__ uint64_t s[STAB_SB];
    uint64_t *base = a + (int64_t)blockIdx.y * N + (int64_t)blockIdx.x * STAB_SB;
    const uint32_t t = threadIdx.x, g0 = blockIdx.x * STAB_SB;
    s[t] = base[t]; s[t + STAB_SB / 2] = base[t + STAB_SB / 2];
    __syncthreads();
    const bool asc = (g0 & k) == 0;                   // k > STAB_SB: one direction per chunk
    for (uint32_t j = STAB_SB >> 1; j > 0; j >>= 1) {
        const uint32_t i = ((t & ~(j - 1)) << 1) | (t & (j - 1));
        stab_cswap(s, i, i | j, asc);
        __syncthreads();
    }
    base[t] = s[t]; base[t + STAB_SB / 2] = s[t + STAB_SB / 2];
}

__global__ void rthc_stab_emit(const int32_t *__restrict__ codes, int64_t n, uint32_t H, const uint64_t *__restrict__ hkeys,
                               const uint32_t *__restrict__ hcnt, uint32_t N, const uint64_t *__restrict__ skeys,
                               const unsigned int *__restrict__ m, int32_t *__restrict__ out_row, int32_t *__restrict__ out_col,
                               int32_t *__restrict__ out_cnt)
{
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t p = blockIdx.y;
    if (q >= (int64_t)m[p]) return;
    const uint32_t s = (uint32_t)(skeys[p * (int64_t)N + q] & 0xFFFFFFFFu);       // first sample of the cell
    const uint32_t r = (uint32_t)codes[(p + 1) * n + s], c = (uint32_t)codes[p * n + s];
    const uint64_t key = ((uint64_t)r << 32) | c;
    uint32_t h = stab_hash(key) & (H - 1);
    while (hkeys[p * (int64_t)H + h] != key) h = (h + 1) & (H - 1);
    out_row[p * n + q] = (int32_t)r;
    out_col[p * n + q] = (int32_t)c;
    out_cnt[p * n + q] = (int32_t)hcnt[p * (int64_t)H + h];
}

extern "C" int rthc_contingency_adjacent(rthc_ctx *ctx, const int32_t *codes, int T, int64_t n, int64_t *cell_off,
                                         int32_t *cell_row, int32_t *cell_col, int32_t *cell_count)
{
    if (!ctx || !codes || !cell_off || !cell_row || !cell_col || !cell_count)
        return rthc_fail(ctx, RTHC_E_ARG, "rthc_contingency_adjacent: NULL argument");
    if (T < 2 || n < 1) return rthc_fail(ctx, RTHC_E_ARG, "need at least 2 partition columns and 1 sample");
    if (n >= (1ll << 30)) return rthc_fail(ctx, RTHC_E_RANGE, "n=%lld exceeds 2^30-1 samples", (long long)n);
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t N = STAB_SB;                    // sort length: power of two >= max(n, 2048)
    while ((int64_t)N < n) N <<= 1;
    const uint32_t H = 2 * N;                // hash slots per pair (load factor <= 0.5)
    // pairs per pass: (16 H + 8 N + 12 n + 4 n) bytes each, capped at ~1.5 GB
    const int64_t per_pair = 16ll * H + 8ll * N + 16ll * n;
    const int chunk = (int)std::max<int64_t>(1, std::min<int64_t>(std::min<int64_t>(T - 1, 60000), (1536ll << 20) / per_pair));
    int32_t *d_codes = nullptr, *d_or = nullptr, *d_oc = nullptr, *d_on = nullptr;
    uint32_t *d_first = nullptr, *d_hcnt = nullptr, *d_hfirst = nullptr;
    uint64_t *d_hkeys = nullptr, *d_skeys = nullptr;
    unsigned int *d_m = nullptr;
    int *d_bad = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_codes, (size_t)((chunk + 1) * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_first, (size_t)((chunk + 1) * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_hkeys, (size_t)chunk * H))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_hcnt, (size_t)chunk * H))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_hfirst, (size_t)chunk * H))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_skeys, (size_t)chunk * N))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_or, (size_t)(chunk * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_oc, (size_t)(chunk * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_on, (size_t)(chunk * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_m, (size_t)chunk))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_bad, 1))) return rc;
    RTHC_CUDA(ctx, cudaMemsetAsync(d_bad, 0, sizeof(int), ctx->stream));
    std::vector<unsigned int> h_m((size_t)chunk);
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    int64_t total = 0;
    cell_off[0] = 0;
    for (int p0 = 0; p0 < T - 1; p0 += chunk) {
        const int np = std::min(chunk, T - 1 - p0), ncol = np + 1;
        const unsigned nbn = (unsigned)((n + 255) / 256);
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_codes, codes + (int64_t)p0 * n, sizeof(int32_t) * (size_t)(ncol * n),
                                       cudaMemcpyHostToDevice, ctx->stream));
        rthc_stab_fill_u32<<<(unsigned)((ncol * n + 255) / 256), 256, 0, ctx->stream>>>(d_first, ncol * n, 0xFFFFFFFFu);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_stab_fill_u64<<<(unsigned)(((int64_t)np * H + 255) / 256), 256, 0, ctx->stream>>>(d_hkeys, (int64_t)np * H, STAB_EMPTY);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_stab_fill_u32<<<(unsigned)(((int64_t)np * H + 255) / 256), 256, 0, ctx->stream>>>(d_hfirst, (int64_t)np * H, 0xFFFFFFFFu);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_stab_fill_u64<<<(unsigned)(((int64_t)np * N + 255) / 256), 256, 0, ctx->stream>>>(d_skeys, (int64_t)np * N, STAB_EMPTY);
        RTHC_LAUNCH_CHECK(ctx);
        RTHC_CUDA(ctx, cudaMemsetAsync(d_hcnt, 0, sizeof(uint32_t) * (size_t)np * H, ctx->stream));
        RTHC_CUDA(ctx, cudaMemsetAsync(d_m, 0, sizeof(unsigned int) * (size_t)np, ctx->stream));
        rthc_stab_first<<<dim3(nbn, (unsigned)ncol), 256, 0, ctx->stream>>>(d_codes, n, d_first, d_bad);
        RTHC_LAUNCH_CHECK(ctx);
        int bad = 0;
        RTHC_CUDA(ctx, cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        if (bad) { rc = rthc_fail(ctx, RTHC_E_ARG, "partition codes must lie in [0, n)"); break; }
        rthc_stab_insert<<<dim3(nbn, (unsigned)np), 256, 0, ctx->stream>>>(d_codes, n, H, d_hkeys, d_hcnt, d_hfirst);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_stab_compact<<<dim3((H + 255) / 256, (unsigned)np), 256, 0, ctx->stream>>>(n, H, d_hkeys, d_hfirst, d_first, N, d_skeys, d_m);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_stab_sort_local<<<dim3(N / STAB_SB, (unsigned)np), STAB_SB / 2, 0, ctx->stream>>>(d_skeys, N);
        RTHC_LAUNCH_CHECK(ctx);
        for (uint32_t k = 2 * STAB_SB; k <= N && k != 0; k <<= 1) {
            for (uint32_t j = k >> 1; j >= STAB_SB; j >>= 1) {
                rthc_stab_sort_global<<<dim3((N / 2 + 255) / 256, (unsigned)np), 256, 0, ctx->stream>>>(d_skeys, N, j, k);
                RTHC_LAUNCH_CHECK(ctx);
            }
            rthc_stab_sort_tail<<<dim3(N / STAB_SB, (unsigned)np), STAB_SB / 2, 0, ctx->stream>>>(d_skeys, N, k);
            RTHC_LAUNCH_CHECK(ctx);
        }
        rthc_stab_emit<<<dim3(nbn, (unsigned)np), 256, 0, ctx->stream>>>(d_codes, n, H, d_hkeys, d_hcnt, N, d_skeys, d_m, d_or, d_oc, d_on);
        RTHC_LAUNCH_CHECK(ctx);
        RTHC_CUDA(ctx, cudaMemcpyAsync(h_m.data(), d_m, sizeof(unsigned int) * (size_t)np, cudaMemcpyDeviceToHost, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (int q = 0; q < np; ++q) {
            const size_t mq = h_m[(size_t)q];
            RTHC_CUDA(ctx, cudaMemcpyAsync(cell_row + total, d_or + (int64_t)q * n, sizeof(int32_t) * mq, cudaMemcpyDeviceToHost, ctx->stream));
            RTHC_CUDA(ctx, cudaMemcpyAsync(cell_col + total, d_oc + (int64_t)q * n, sizeof(int32_t) * mq, cudaMemcpyDeviceToHost, ctx->stream));
            RTHC_CUDA(ctx, cudaMemcpyAsync(cell_count + total, d_on + (int64_t)q * n, sizeof(int32_t) * mq, cudaMemcpyDeviceToHost, ctx->stream));
            total += (int64_t)mq;
            cell_off[p0 + q + 1] = total;
        }
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[6] = ms;
    rthc_dev_free(ctx, &d_codes); rthc_dev_free(ctx, &d_first); rthc_dev_free(ctx, &d_hkeys);
    rthc_dev_free(ctx, &d_hcnt); rthc_dev_free(ctx, &d_hfirst); rthc_dev_free(ctx, &d_skeys);
    rthc_dev_free(ctx, &d_or); rthc_dev_free(ctx, &d_oc); rthc_dev_free(ctx, &d_on);
    rthc_dev_free(ctx, &d_m); rthc_dev_free(ctx, &d_bad);
    return rc;
}

// ---- host finish of one table --------------------------------------------------------------------
// The per-row sums of getAdjustedWallace (metrics.py:147-165) in the reference's evaluation order.
// Python's `x ** 2.0` is libm pow(x, 2.0): the exponents live in volatiles so that the compiler
// cannot turn the calls into multiplications (pow is not guaranteed to round like x*x*x).
extern "C" int rthc_adjusted_wallace_sums(const int32_t *cell_row, const int32_t *cell_count, int64_t m, double *sum_w1,
                                          double *sum_w2)
{
    if ((!cell_row || !cell_count) && m > 0) return RTHC_E_ARG;
    if (!sum_w1 || !sum_w2) return RTHC_E_ARG;
    volatile double two = 2.0, three = 3.0;
    double w1 = 0.0, w2 = 0.0;
    int64_t q = 0;
    while (q < m) {
        int64_t e = q, total = 0;
        while (e < m && cell_row[e] == cell_row[q]) total += cell_count[e++];
        const double r = (double)total;
        double fc2 = 0.0, fc3 = 0.0;
        for (int64_t c = q; c < e; ++c) {
            const double f = (double)cell_count[c] / r;
            fc2 += pow(f, two);
            fc3 += pow(f, three);
        }
        const double sq = pow(fc2, two);
        double var = 0.0;
        const double rr = pow(r * (r - 1.0), two);
        if (total > 1)
            var = (4.0 * r * (r - 1.0) * (r - 2.0) * fc3 + 2.0 * r * (r - 1.0) * fc2 - 2.0 * r * (r - 1.0) * (2.0 * r - 3.0) * sq) / rr;
        w1 += rr * var;
        w2 += r * (r - 1.0);
        q = e;
    }
    *sum_w1 = w1;
    *sum_w2 = w2;
    return 0;
}
