// zoom.cu -- SURVEY 8(f) N3: zoom-in / subtree queries on the RESIDENT distance matrix.
//
// The reference answers both questions from text: get_closest_samples (reportree.py:591-617)
// re-reads one column of <out>_dist_hamming.tsv per sample of interest, and the zoom-in loop
// (reportree.py:1965-2055) re-runs the whole HC script (cgmlst-dists included) on the subset with
// `--loci <out>_loci_used.txt`.  With the same loci the distance of a pair does not depend on the
// other samples, so the subset's matrix is a sub-matrix of the one already in HBM:
//
//   rthc_closest      one row of the condensed matrix -> value at sorted position k (65,536-bin
//                     histogram) -> every sample with d <= that value
//   rthc_subset_from  gather of the m(m-1)/2 sub-matrix into a child context that then runs
//                     rthc_linkage / rthc_cut as usual
#include <algorithm>
#include <mutex>

#include "rthc_internal.cuh"

__device__ __forceinline__ int64_t cond_index(int64_t n, int64_t i, int64_t j)
{
    const int64_t a = i < j ? i : j, b = i < j ? j : i;
    return n * a - (a * (a + 1)) / 2 + (b - a - 1);
}

// ---- closest samples ---------------------------------------------------------------------------
__global__ void rthc_zoom_row_hist(const uint16_t *__restrict__ cond, int64_t n, int64_t s, uint16_t *__restrict__ row,
                                   unsigned int *__restrict__ hist)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint16_t v = j == s ? (uint16_t)0 : cond[cond_index(n, s, j)];
    row[j] = v;
    atomicAdd(&hist[v], 1u);
}

// one CTA of 1024 threads: smallest v with #(d <= v) > k  (= the value at 0-based sorted position k).
// Segment s = bins [64 s, 64 s + 64); warp w sums segments 32 w .. 32 w + 31 with coalesced loads.
__global__ void __launch_bounds__(1024) rthc_zoom_kth(const unsigned int *__restrict__ hist, int64_t k, int *__restrict__ out_v,
                                                      unsigned int *__restrict__ out_count)
{
    __shared__ unsigned int part[1024];
    const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
    for (int q = 0; q < 32; ++q) {
        const int seg = warp * 32 + q;
        unsigned int s = hist[seg * 64 + lane] + hist[seg * 64 + 32 + lane];
        for (int off = 16; off > 0; off >>= 1) s += __shfl_xor_sync(0xFFFFFFFFu, s, off);
        if (lane == 0) part[seg] = s;
    }
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {                // inclusive scan of the 1024 segment sums
        unsigned int v = t >= off ? part[t - off] : 0u;
        __syncthreads();
        part[t] += v;
        __syncthreads();
    }
    // the warp that owns the segment holding position k resolves it: lane l covers bins 2l, 2l+1 of the segment
    for (int q = 0; q < 32; ++q) {
        const int seg = warp * 32 + q;
        const unsigned int before = seg ? part[seg - 1] : 0u;
        if (!((int64_t)before <= k && (int64_t)part[seg] > k)) continue;      // warp-uniform
        const unsigned int h0 = hist[seg * 64 + 2 * lane], h1 = hist[seg * 64 + 2 * lane + 1];
        unsigned int inc = h0 + h1;
        for (int off = 1; off < 32; off <<= 1) {
            const unsigned int v = __shfl_up_sync(0xFFFFFFFFu, inc, off);
            if (lane >= off) inc += v;
        }
        const unsigned int c1 = before + inc, c0 = c1 - h1;                    // counts up to and including bin 2l / 2l+1
        const unsigned int hit = __ballot_sync(0xFFFFFFFFu, (int64_t)c1 > k);
        if (lane == __ffs(hit) - 1) {
            if ((int64_t)c0 > k) { *out_v = seg * 64 + 2 * lane; *out_count = c0; }
            else { *out_v = seg * 64 + 2 * lane + 1; *out_count = c1; }
        }
    }
}

__global__ void rthc_zoom_select(const uint16_t *__restrict__ row, int64_t n, const int *__restrict__ vmax,
                                 int64_t *__restrict__ out_idx, uint16_t *__restrict__ out_d, unsigned int *cursor)
{
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    const uint16_t v = row[j];
    if ((int)v <= *vmax) {
        const unsigned int p = atomicAdd(cursor, 1u);
        out_idx[p] = j;
        out_d[p] = v;
    }
}

extern "C" int rthc_closest(rthc_ctx *ctx, int64_t sample, int64_t k, int64_t *out_idx, uint16_t *out_dist,
                            int64_t *out_count, int *out_max_dist)
{
    if (!ctx || !out_idx || !out_dist || !out_count) return rthc_fail(ctx, RTHC_E_ARG, "rthc_closest: NULL argument");
    if (!ctx->have_dist || !ctx->d_cond) return rthc_fail(ctx, RTHC_E_STATE, "rthc_closest needs a resident uint16 distance matrix");
    if (ctx->world != 1) return rthc_fail(ctx, RTHC_E_STATE, "rthc_closest needs a full shard");
    const int64_t n = ctx->n;
    if (sample < 0 || sample >= n || k < 0 || k >= n) return rthc_fail(ctx, RTHC_E_ARG, "rthc_closest: sample / k out of range");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    uint16_t *d_row = nullptr, *d_od = nullptr;
    unsigned int *d_hist = nullptr;      // 65536 bins + [65536] count<=v + [65537] cursor
    int *d_v = nullptr;
    int64_t *d_oi = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_row, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_od, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_oi, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_hist, (size_t)65536 + 2))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_v, 1))) return rc;
    RTHC_CUDA(ctx, cudaMemsetAsync(d_hist, 0, sizeof(unsigned int) * (65536 + 2), ctx->stream));
    const unsigned nb = (unsigned)((n + 255) / 256);
    rthc_zoom_row_hist<<<nb, 256, 0, ctx->stream>>>(ctx->d_cond, n, sample, d_row, d_hist);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_zoom_kth<<<1, 1024, 0, ctx->stream>>>(d_hist, k, d_v, d_hist + 65536);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_zoom_select<<<nb, 256, 0, ctx->stream>>>(d_row, n, d_v, d_oi, d_od, d_hist + 65537);
    RTHC_LAUNCH_CHECK(ctx);
    unsigned int cnt = 0;
    int vmax = 0;
    RTHC_CUDA(ctx, cudaMemcpyAsync(&cnt, d_hist + 65536, sizeof(cnt), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(&vmax, d_v, sizeof(vmax), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(out_idx, d_oi, sizeof(int64_t) * cnt, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(out_dist, d_od, sizeof(uint16_t) * cnt, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_row); rthc_dev_free(ctx, &d_od); rthc_dev_free(ctx, &d_oi);
    rthc_dev_free(ctx, &d_hist); rthc_dev_free(ctx, &d_v);
    *out_count = cnt;
    if (out_max_dist) *out_max_dist = vmax;
    return 0;
}

// ---- closest samples without any allocation ------------------------------------------------------
// Same three kernels as rthc_closest, with the scratch (row copy, selected indices / distances, histogram, k-th
// value) in ONE file-scope device array instead of the context's pool, for callers that may not allocate (CUDA
// graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// Every GPU that loads the library holds the array (about 3.3 MiB) until exit.
// All calls share it and it keeps the last call's contents: the histogram / counter / cursor words are cleared at
// the start of every call, everything else is written before it is read.  Calls that overlap on the GPU (two
// contexts or two host threads on the same device) race on it -- serialise them.
#define RTHC_NOALLOC_HIST_WORDS (65536 + 2)                     // bins + [65536] count <= v + [65537] cursor
#define RTHC_NOALLOC_OFF_V      ((size_t)263168)                // hist: 262,152 B, rounded up to 256
#define RTHC_NOALLOC_OFF_IDX    (RTHC_NOALLOC_OFF_V + 256)
#define RTHC_NOALLOC_OFF_ROW    (RTHC_NOALLOC_OFF_IDX + 8 * (size_t)RTHC_NOALLOC_MAX_N)
#define RTHC_NOALLOC_OFF_DIST   (RTHC_NOALLOC_OFF_ROW + 2 * (size_t)RTHC_NOALLOC_MAX_N)
#define RTHC_NOALLOC_BYTES      (RTHC_NOALLOC_OFF_DIST + 2 * (size_t)RTHC_NOALLOC_MAX_N)
static_assert(RTHC_NOALLOC_HIST_WORDS * 4 <= RTHC_NOALLOC_OFF_V, "histogram does not fit its slot");
static __device__ __align__(256) unsigned char rthc_noalloc_scratch[RTHC_NOALLOC_BYTES];

// In host code `rthc_noalloc_scratch` is only a stand-in for the symbol, not a device address: the address is
// asked for once per device, with that device current, and cached.
#define RTHC_NOALLOC_MAX_DEVICES 64
static unsigned char *g_noalloc_ptr[RTHC_NOALLOC_MAX_DEVICES];
static std::mutex g_noalloc_mutex;

static int noalloc_scratch_for(rthc_ctx *ctx, unsigned char **out)
{
    if (ctx->device < 0 || ctx->device >= RTHC_NOALLOC_MAX_DEVICES)
        return rthc_fail(ctx, RTHC_E_RANGE, "rthc_closest_noalloc: device %d beyond the %d devices it keeps scratch for",
                         ctx->device, RTHC_NOALLOC_MAX_DEVICES);
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[ctx->device]) {
        RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
        void *p = nullptr;
        RTHC_CUDA(ctx, cudaGetSymbolAddress(&p, rthc_noalloc_scratch));
        if (!p) return rthc_fail(ctx, RTHC_E_CUDA, "rthc_closest_noalloc: cudaGetSymbolAddress returned NULL");
        g_noalloc_ptr[ctx->device] = static_cast<unsigned char *>(p);
    }
    *out = g_noalloc_ptr[ctx->device];
    return 0;
}

extern "C" int rthc_closest_noalloc(rthc_ctx *ctx, int64_t sample, int64_t k, int64_t *out_idx, uint16_t *out_dist,
                                    int64_t *out_count, int *out_max_dist)
{
    if (!ctx || !out_idx || !out_dist || !out_count) return rthc_fail(ctx, RTHC_E_ARG, "rthc_closest_noalloc: NULL argument");
    if (!ctx->have_dist || !ctx->d_cond)
        return rthc_fail(ctx, RTHC_E_STATE, "rthc_closest_noalloc needs a resident uint16 distance matrix");
    if (ctx->world != 1) return rthc_fail(ctx, RTHC_E_STATE, "rthc_closest_noalloc needs a full shard");
    const int64_t n = ctx->n;
    if (n > RTHC_NOALLOC_MAX_N)
        return rthc_fail(ctx, RTHC_E_RANGE, "rthc_closest_noalloc: n = %lld samples exceed the static scratch (at most %d); "
                         "use rthc_closest", (long long)n, RTHC_NOALLOC_MAX_N);
    if (sample < 0 || sample >= n || k < 0 || k >= n)
        return rthc_fail(ctx, RTHC_E_ARG, "rthc_closest_noalloc: sample / k out of range");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    unsigned char *base = nullptr;
    int rc;
    if ((rc = noalloc_scratch_for(ctx, &base))) return rc;
    unsigned int *d_hist = reinterpret_cast<unsigned int *>(base);
    int *d_v = reinterpret_cast<int *>(base + RTHC_NOALLOC_OFF_V);
    int64_t *d_oi = reinterpret_cast<int64_t *>(base + RTHC_NOALLOC_OFF_IDX);
    uint16_t *d_row = reinterpret_cast<uint16_t *>(base + RTHC_NOALLOC_OFF_ROW);
    uint16_t *d_od = reinterpret_cast<uint16_t *>(base + RTHC_NOALLOC_OFF_DIST);
    // the array keeps the previous call's contents: histogram, count and cursor start at zero, d_v is always written
    RTHC_CUDA(ctx, cudaMemsetAsync(d_hist, 0, sizeof(unsigned int) * RTHC_NOALLOC_HIST_WORDS, ctx->stream));
    const unsigned nb = (unsigned)((n + 255) / 256);
    rthc_zoom_row_hist<<<nb, 256, 0, ctx->stream>>>(ctx->d_cond, n, sample, d_row, d_hist);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_zoom_kth<<<1, 1024, 0, ctx->stream>>>(d_hist, k, d_v, d_hist + 65536);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_zoom_select<<<nb, 256, 0, ctx->stream>>>(d_row, n, d_v, d_oi, d_od, d_hist + 65537);
    RTHC_LAUNCH_CHECK(ctx);
    unsigned int cnt = 0;
    int vmax = 0;
    RTHC_CUDA(ctx, cudaMemcpyAsync(&cnt, d_hist + 65536, sizeof(cnt), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(&vmax, d_v, sizeof(vmax), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if ((int64_t)cnt > n) return rthc_fail(ctx, RTHC_E_STATE, "rthc_closest_noalloc: scratch corrupted (overlapping calls?)");
    RTHC_CUDA(ctx, cudaMemcpyAsync(out_idx, d_oi, sizeof(int64_t) * cnt, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(out_dist, d_od, sizeof(uint16_t) * cnt, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *out_count = cnt;
    if (out_max_dist) *out_max_dist = vmax;
    return 0;
}

// ---- sub-matrix ---------------------------------------------------------------------------------
__global__ void rthc_zoom_gather(const uint16_t *__restrict__ cond, int64_t n, const int64_t *__restrict__ idx, int64_t m,
                                 uint16_t *__restrict__ out, int *__restrict__ flag)
{
    // one CTA row per child row a; threads stride over b > a
    const int64_t a = blockIdx.x;
    const int64_t ia = idx[a];
    const int64_t base = m * a - (a * (a + 1)) / 2 - a - 1;
    for (int64_t b = a + 1 + threadIdx.x; b < m; b += blockDim.x) {
        const int64_t ib = idx[b];
        uint16_t v = 0;
        if (ia == ib) *flag = 1;
        else v = cond[cond_index(n, ia, ib)];
        out[base + b] = v;
    }
}

extern "C" int rthc_subset_from(rthc_ctx *child, rthc_ctx *parent, const int64_t *idx, int64_t m)
{
    if (!child || !parent || !idx) return rthc_fail(child, RTHC_E_ARG, "rthc_subset_from: NULL argument");
    if (child == parent) return rthc_fail(child, RTHC_E_ARG, "rthc_subset_from: child and parent must differ");
    if (child->device != parent->device) return rthc_fail(child, RTHC_E_ARG, "rthc_subset_from: contexts on different devices");
    if (!parent->have_dist || !parent->d_cond || parent->world != 1)
        return rthc_fail(child, RTHC_E_STATE, "parent has no resident full uint16 distance matrix");
    if (m < 2) return rthc_fail(child, RTHC_E_ARG, "need at least 2 samples (m=%lld)", (long long)m);
    for (int64_t a = 0; a < m; ++a)
        if (idx[a] < 0 || idx[a] >= parent->n) return rthc_fail(child, RTHC_E_ARG, "subset index %lld out of range", (long long)idx[a]);
    RTHC_CUDA(child, cudaSetDevice(child->device));
    RTHC_CUDA(child, cudaStreamSynchronize(parent->stream));
    rthc_free_distance_state(child);
    rthc_free_linkage_state(child);
    child->rank = 0; child->world = 1; child->full_storage = false;
    int rc = rthc_setup_geometry(child, m);
    if (rc) return rc;
    child->L = parent->L; child->L_kept = parent->L_kept;
    child->G = 0; child->plane_rows = 0;
    const int64_t P = m * (m - 1) / 2;
    if (!child->d_cond || child->cond_cap < P) {
        child->cond_cap = 0;
        if ((rc = rthc_dev_alloc(child, &child->d_cond, (size_t)P + RTHC_COND_PAD))) return rc;
        child->cond_cap = P;
    }
    int64_t *d_idx = nullptr;
    int *d_flag = nullptr;
    if ((rc = rthc_dev_alloc(child, &d_idx, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(child, &d_flag, 1))) return rc;
    RTHC_CUDA(child, cudaMemsetAsync(d_flag, 0, sizeof(int), child->stream));
    RTHC_CUDA(child, cudaMemcpyAsync(d_idx, idx, sizeof(int64_t) * (size_t)m, cudaMemcpyHostToDevice, child->stream));
    rthc_zoom_gather<<<(unsigned)(m - 1), 256, 0, child->stream>>>(parent->d_cond, parent->n, d_idx, m, child->d_cond, d_flag);
    RTHC_LAUNCH_CHECK(child);
    int flag = 0;
    RTHC_CUDA(child, cudaMemcpyAsync(&flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, child->stream));
    RTHC_CUDA(child, cudaStreamSynchronize(child->stream));
    rthc_dev_free(child, &d_idx); rthc_dev_free(child, &d_flag);
    if (flag) return rthc_fail(child, RTHC_E_ARG, "rthc_subset_from: a sample index is listed twice");
    child->ext_integer = true;
    child->have_dist = true;
    child->have_profiles = false;
    return 0;
}
