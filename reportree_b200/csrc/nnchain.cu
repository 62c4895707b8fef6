// nnchain.cu -- the sequential-order linkage loops, as persistent cooperative kernels.
//
//  * rthc_prim_exact : scipy `_hierarchy.mst_single_linkage` (scipy _hierarchy_impl.py:960-961,
//    called from scripts/partitioning_HC.py:197).  Prim from vertex 0; every step relaxes the
//    keys with the row of the vertex added last and picks the LOWEST index among the minimal
//    keys (scipy's "first strictly smaller wins").  The edge list therefore equals scipy's,
//    and Z is byte-identical after the stable sort + label pass (linkage.cu).
//  * rthc_nn_chain   : scipy `_hierarchy.nn_chain` (scipy _hierarchy_impl.py:962-963) for
//    complete / average / weighted / ward: nearest-neighbour chain with scipy's tie rule (the previous
//    chain element wins ties, otherwise the lowest index of the minimum), Lance-Williams row
//    update in HBM.  Works on a square working copy of the matrix so that every row scan and
//    row update is contiguous; uint16 for complete linkage on integer distances, float64
//    (no FMA contraction: the library is built with -fmad=false) for average / weighted.
//
// Both are inherently sequential (n-1 dependent steps).  One CTA per SM stays resident; a step
// is a slice-parallel scan + a grid-wide barrier (cooperative groups), no host round trips.
// Mutable data is read with ld.global.cg (L1 is not coherent across SMs).
#include <cooperative_groups.h>

#include <algorithm>
#include <cmath>

#include "rthc_internal.cuh"

namespace cg = cooperative_groups;

#define SEQ_THREADS 512

struct MinSlot {
    unsigned long long w;   // order-preserving bits of the distance
    unsigned int i;
    unsigned int pad;
};

__device__ __forceinline__ unsigned long long order_bits(uint16_t v) { return (unsigned long long)v; }
__device__ __forceinline__ unsigned long long order_bits(double v)
{
    return (unsigned long long)__double_as_longlong(v);   // distances are >= 0: bit order == value order
}
__device__ __forceinline__ double from_bits(unsigned long long b, uint16_t) { return (double)b; }
__device__ __forceinline__ double from_bits(unsigned long long b, double) { return __longlong_as_double((long long)b); }

__device__ __forceinline__ bool slot_less(unsigned long long w1, unsigned int i1, unsigned long long w2, unsigned int i2)
{
    return w1 < w2 || (w1 == w2 && i1 < i2);
}

// block-wide lexicographic min of (w, i); result valid in thread 0
__device__ __forceinline__ void block_min(unsigned long long &w, unsigned int &i, MinSlot *sh)
{
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        unsigned long long ow = __shfl_xor_sync(0xffffffffu, w, s);
        unsigned int oi = __shfl_xor_sync(0xffffffffu, i, s);
        if (slot_less(ow, oi, w, i)) { w = ow; i = oi; }
    }
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) { sh[warp].w = w; sh[warp].i = i; }
    __syncthreads();
    if (warp == 0) {
        unsigned long long bw = lane < SEQ_THREADS / 32 ? sh[lane].w : ~0ull;
        unsigned int bi = lane < SEQ_THREADS / 32 ? sh[lane].i : ~0u;
#pragma unroll
        for (int s = 16; s > 0; s >>= 1) {
            unsigned long long ow = __shfl_xor_sync(0xffffffffu, bw, s);
            unsigned int oi = __shfl_xor_sync(0xffffffffu, bi, s);
            if (slot_less(ow, oi, bw, bi)) { bw = ow; bi = oi; }
        }
        w = bw; i = bi;
    }
    __syncthreads();
}

// every block reduces the per-block slots of this step (redundantly): result in all threads
__device__ __forceinline__ void grid_min(const MinSlot *slots, int nblocks, unsigned long long &w, unsigned int &i,
                                         MinSlot *sh)
{
    w = ~0ull; i = ~0u;
    for (int b = threadIdx.x; b < nblocks; b += SEQ_THREADS) {
        const unsigned long long bw = __ldcg(&slots[b].w);
        const unsigned int bi = __ldcg(&slots[b].i);
        if (slot_less(bw, bi, w, i)) { w = bw; i = bi; }
    }
    block_min(w, i, sh);
    __shared__ MinSlot bc;
    if (threadIdx.x == 0) { bc.w = w; bc.i = i; }
    __syncthreads();
    w = bc.w; i = bc.i;
    __syncthreads();
}

// ---------------------------------------------------------------------------------------------
// Prim in scipy order over the condensed matrix (full shard)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(SEQ_THREADS, 1)
rthc_prim_kernel(const T *__restrict__ cond, int64_t n, int per_thread, MinSlot *slots, int32_t *ex, int32_t *ey,
                 double *ew)
{
    cg::grid_group grid = cg::this_grid();
    __shared__ MinSlot sh[SEQ_THREADS / 32];
    const int nblocks = gridDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * SEQ_THREADS + threadIdx.x;
    const int64_t gthreads = (int64_t)nblocks * SEQ_THREADS;
    // vertex v = gtid + q * gthreads is owned by this thread; its key lives in a register
    constexpr int MAXQ = 8;
    unsigned long long key[MAXQ];
    bool merged[MAXQ];
#pragma unroll
    for (int q = 0; q < MAXQ; ++q) { key[q] = ~0ull; merged[q] = false; }

    int64_t x = 0;
    for (int64_t k = 0; k < n - 1; ++k) {
        unsigned long long bw = ~0ull;
        unsigned int bi = ~0u;
        const int64_t xbase = n * x - (x * (x + 1)) / 2;
#pragma unroll
        for (int q = 0; q < MAXQ; ++q) {
            if (q >= per_thread) break;
            const int64_t v = gtid + (int64_t)q * gthreads;
            if (v >= n) break;
            if (v == x) merged[q] = true;
            if (merged[q]) continue;
            const int64_t idx = v > x ? xbase + (v - x - 1) : n * v - (v * (v + 1)) / 2 + (x - v - 1);
            const unsigned long long d = order_bits(cond[idx]);
            if (d < key[q]) key[q] = d;
            if (slot_less(key[q], (unsigned int)v, bw, bi)) { bw = key[q]; bi = (unsigned int)v; }
        }
        block_min(bw, bi, sh);
        MinSlot *cur = slots + (k & 1) * nblocks;
        if (threadIdx.x == 0) { cur[blockIdx.x].w = bw; cur[blockIdx.x].i = bi; }
        grid.sync();
        grid_min(cur, nblocks, bw, bi, sh);
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            ex[k] = (int32_t)x; ey[k] = (int32_t)bi; ew[k] = from_bits(bw, T());
        }
        x = bi;
    }
}

template <typename T>
static int run_prim(rthc_ctx *ctx, const T *d_cond, std::vector<int32_t> &ex, std::vector<int32_t> &ey,
                    std::vector<double> &ew)
{
    const int64_t n = ctx->n, m = n - 1;
    int nblocks = ctx->sm_count;
    int per_thread = (int)((n + (int64_t)nblocks * SEQ_THREADS - 1) / ((int64_t)nblocks * SEQ_THREADS));
    if (per_thread > 8)
        return rthc_fail(ctx, RTHC_E_RANGE, "Prim-exact supports up to %lld samples", (long long)nblocks * SEQ_THREADS * 8);
    MinSlot *d_slots = nullptr;
    int32_t *d_ex = nullptr, *d_ey = nullptr;
    double *d_ew = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_slots, (size_t)2 * nblocks))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ex, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ey, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ew, (size_t)m))) return rc;
    int64_t nn = n;
    void *args[] = {(void *)&d_cond, (void *)&nn, (void *)&per_thread, (void *)&d_slots, (void *)&d_ex, (void *)&d_ey,
                    (void *)&d_ew};
    RTHC_CUDA(ctx, cudaLaunchCooperativeKernel((void *)rthc_prim_kernel<T>, dim3(nblocks), dim3(SEQ_THREADS), args, 0,
                                               ctx->stream));
    ctx->launches++;
    ex.resize((size_t)m); ey.resize((size_t)m); ew.resize((size_t)m);
    RTHC_CUDA(ctx, cudaMemcpyAsync(ex.data(), d_ex, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ey.data(), d_ey, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ew.data(), d_ew, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_slots); rthc_dev_free(ctx, &d_ex); rthc_dev_free(ctx, &d_ey); rthc_dev_free(ctx, &d_ew);
    return 0;
}

int rthc_prim_exact(rthc_ctx *ctx, std::vector<int32_t> &ex, std::vector<int32_t> &ey, std::vector<double> &ew)
{
    if (ctx->d_cond) return run_prim<uint16_t>(ctx, ctx->d_cond, ex, ey, ew);
    if (ctx->d_cond_f64) return run_prim<double>(ctx, ctx->d_cond_f64, ex, ey, ew);
    return rthc_fail(ctx, RTHC_E_STATE, "no distance matrix on the device");
}

// ---------------------------------------------------------------------------------------------
// NN-chain merge loop on a square working matrix
// ---------------------------------------------------------------------------------------------
template <typename T>
__device__ __forceinline__ T lance_williams(int method, T dx, T dy, double dxy, int nx, int ny, int ni);
template <>
__device__ __forceinline__ uint16_t lance_williams<uint16_t>(int, uint16_t dx, uint16_t dy, double, int, int, int)
{
    return dx > dy ? dx : dy;                                   // complete linkage on integer distances
}
template <>
__device__ __forceinline__ double lance_williams<double>(int method, double dx, double dy, double dxy, int nx, int ny,
                                                         int ni)
{
    if (method == RTHC_COMPLETE) return dx > dy ? dx : dy;
    if (method == RTHC_AVERAGE)
        return __ddiv_rn(__dadd_rn(__dmul_rn((double)nx, dx), __dmul_rn((double)ny, dy)), __dadd_rn((double)nx, (double)ny));
    if (method == RTHC_WARD) {
        // scipy's association: t = 1/(nx+ny+ni); ((ni+nx)*t*dx)*dx + ((ni+ny)*t*dy)*dy - ((ni*t)*dxy)*dxy
        const double t = __ddiv_rn(1.0, (double)(nx + ny + ni));
        const double a = __dmul_rn(__dmul_rn(__dmul_rn((double)(ni + nx), t), dx), dx);
        const double b = __dmul_rn(__dmul_rn(__dmul_rn((double)(ni + ny), t), dy), dy);
        const double c = __dmul_rn(__dmul_rn(__dmul_rn((double)ni, t), dxy), dxy);
        return __dsqrt_rn(__dsub_rn(__dadd_rn(a, b), c));
    }
    return __dmul_rn(0.5, __dadd_rn(dx, dy));                   // weighted
}

template <typename T, typename S>
__global__ void rthc_expand_square(const S *__restrict__ cond, int64_t n, T *__restrict__ W)
{
    const int64_t i = blockIdx.y;
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= n) return;
    T v = 0;
    if (i != j) {
        const int64_t a = i < j ? i : j, b = i < j ? j : i;
        v = (T)cond[n * a - (a * (a + 1)) / 2 + (b - a - 1)];
    }
    W[i * n + j] = v;
}

// control block in global memory (written by block 0, thread 0 only)
struct ChainCtl {
    int64_t scans;
};

template <typename T>
__global__ void __launch_bounds__(SEQ_THREADS, 1)
rthc_nn_chain_kernel(T *W, int64_t n, int method, int32_t *size, int32_t *chain, MinSlot *slots, int32_t *ex,
                     int32_t *ey, double *ew, ChainCtl *ctl)
{
    cg::grid_group grid = cg::this_grid();
    __shared__ MinSlot sh[SEQ_THREADS / 32];
    const int nblocks = gridDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * SEQ_THREADS + threadIdx.x;
    const int64_t gthreads = (int64_t)nblocks * SEQ_THREADS;
    const bool leader = blockIdx.x == 0 && threadIdx.x == 0;

    // replicated scalar state (identical in every thread of the grid)
    int64_t len = 0, first_live = 0, step = 0, scans = 0;
    int32_t x = -1, prev = -1;
    int32_t dead = -1;                 // cluster retired by the previous merge: its size[] update may
    int32_t grown = -1, grown_size = 0; // not be visible yet, so it is excluded / patched explicitly

    for (int64_t k = 0; k < n - 1;) {
        if (len == 0) {
            while (first_live == dead || __ldcg(&size[first_live]) == 0) ++first_live;
            x = (int32_t)first_live;
            prev = -1;
            len = 1;
            if (leader) chain[0] = x;
        }
        // ---- scan row x: lowest index of the minimum over live i != x -------------------------
        unsigned long long bw = ~0ull;
        unsigned int bi = ~0u;
        const T *row = W + (int64_t)x * n;
        // D[x, prev] is needed right after the reduction: request it with the row, not after it
        const unsigned long long dprev = len > 1 ? order_bits(__ldcg(&row[prev])) : ~0ull;
        for (int64_t i = gtid; i < n; i += gthreads) {
            if (i == x || i == dead || __ldcg(&size[i]) == 0) continue;
            const unsigned long long d = order_bits(__ldcg(&row[i]));
            if (slot_less(d, (unsigned int)i, bw, bi)) { bw = d; bi = (unsigned int)i; }
        }
        block_min(bw, bi, sh);
        MinSlot *cur = slots + (step & 1) * nblocks;
        if (threadIdx.x == 0) { cur[blockIdx.x].w = bw; cur[blockIdx.x].i = bi; }
        if (leader && dead >= 0) { size[dead] = 0; size[grown] = grown_size; }   // publish last merge's sizes
        grid.sync();
        ++step; ++scans;
        dead = -1;
        grid_min(cur, nblocks, bw, bi, sh);
        int32_t y = (int32_t)bi;
        unsigned long long dmin = bw;
        if (len > 1 && dprev <= dmin) { y = prev; dmin = dprev; }  // previous chain element wins ties
        if (!(len > 1 && y == prev)) {
            // ---- extend the chain ---------------------------------------------------------------
            if (leader) chain[len] = y;
            prev = x; x = y; ++len;
            continue;
        }
        // ---- merge x and y (reciprocal nearest neighbours) ---------------------------------------
        int32_t a = x < y ? x : y, b = x < y ? y : x;             // a dies, b survives (scipy: x < y swap)
        const int32_t na = __ldcg(&size[a]), nb = __ldcg(&size[b]);
        if (leader) { ex[k] = a; ey[k] = b; ew[k] = from_bits(dmin, T()); }
        const T *ra = W + (int64_t)a * n;
        T *rb = W + (int64_t)b * n;
        const double dab = from_bits(dmin, T());
        for (int64_t i = gtid; i < n; i += gthreads) {
            int32_t ni;
            if (i == a || i == b || (ni = __ldcg(&size[i])) == 0) continue;
            const T v = lance_williams<T>(method, __ldcg(&ra[i]), __ldcg(&rb[i]), dab, na, nb, ni);
            rb[i] = v;                       // row b (contiguous)
            W[i * n + b] = v;                // column b (symmetric copy)
        }
        dead = a; grown = b; grown_size = na + nb;
        ++k;
        len -= 2;
        grid.sync();                          // row/column b visible before the next scan
        ++step;
        if (len > 0) {
            x = __ldcg(&chain[len - 1]);
            prev = len > 1 ? __ldcg(&chain[len - 2]) : -1;
        }
    }
    if (leader) ctl->scans = scans;
}

template <typename T, typename S>
static int run_nn_chain(rthc_ctx *ctx, const S *d_cond, int method, std::vector<int32_t> &ex,
                        std::vector<int32_t> &ey, std::vector<double> &ew)
{
    const int64_t n = ctx->n, m = n - 1;
    T *d_W = nullptr;
    int32_t *d_size = nullptr, *d_chain = nullptr, *d_ex = nullptr, *d_ey = nullptr;
    double *d_ew = nullptr;
    MinSlot *d_slots = nullptr;
    ChainCtl *d_ctl = nullptr;
    int nblocks = ctx->sm_count;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_W, (size_t)(n * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_size, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_chain, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_slots, (size_t)2 * nblocks))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ex, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ey, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ew, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ctl, 1))) return rc;
    {
        dim3 grd((unsigned)((n + 255) / 256), (unsigned)n);
        rthc_expand_square<T, S><<<grd, 256, 0, ctx->stream>>>(d_cond, n, d_W);
        RTHC_LAUNCH_CHECK(ctx);
        std::vector<int32_t> ones((size_t)n, 1);
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_size, ones.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    int64_t nn = n;
    void *args[] = {(void *)&d_W, (void *)&nn, (void *)&method, (void *)&d_size, (void *)&d_chain, (void *)&d_slots,
                    (void *)&d_ex, (void *)&d_ey, (void *)&d_ew, (void *)&d_ctl};
    RTHC_CUDA(ctx, cudaLaunchCooperativeKernel((void *)rthc_nn_chain_kernel<T>, dim3(nblocks), dim3(SEQ_THREADS), args,
                                               0, ctx->stream));
    ctx->launches++;
    ex.resize((size_t)m); ey.resize((size_t)m); ew.resize((size_t)m);
    ChainCtl ctl;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ex.data(), d_ex, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ey.data(), d_ey, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ew.data(), d_ew, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(&ctl, d_ctl, sizeof(ChainCtl), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->row_scans = ctl.scans;
    rthc_dev_free(ctx, &d_W); rthc_dev_free(ctx, &d_size); rthc_dev_free(ctx, &d_chain); rthc_dev_free(ctx, &d_slots);
    rthc_dev_free(ctx, &d_ex); rthc_dev_free(ctx, &d_ey); rthc_dev_free(ctx, &d_ew); rthc_dev_free(ctx, &d_ctl);
    return 0;
}

int rthc_nn_chain(rthc_ctx *ctx, int method, std::vector<int32_t> &ex, std::vector<int32_t> &ey,
                  std::vector<double> &ew)
{
    if (ctx->d_cond) {
        if (method == RTHC_COMPLETE) return run_nn_chain<uint16_t, uint16_t>(ctx, ctx->d_cond, method, ex, ey, ew);
        return run_nn_chain<double, uint16_t>(ctx, ctx->d_cond, method, ex, ey, ew);
    }
    if (ctx->d_cond_f64) return run_nn_chain<double, double>(ctx, ctx->d_cond_f64, method, ex, ey, ew);
    return rthc_fail(ctx, RTHC_E_STATE, "no distance matrix on the device");
}
