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

#include <cstdio>
#include <cstdlib>

#include <algorithm>
#include <cmath>
#include <type_traits>

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

// ---------------------------------------------------------------------------------------------
// Prim in scipy order inside ONE thread-block cluster (uint16 matrices)
// ---------------------------------------------------------------------------------------------
// The loop above is bound by the grid barrier (3.3 us per step whatever the CTA count).  A step only
// needs one row of the matrix (<= 0.5 MB), which a 16-CTA cluster streams in well under a microsecond,
// and the cluster has a hardware barrier and distributed shared memory: every CTA reduces its slice to one
// (weight, index) word, stores it into the slot table of EVERY CTA of the cluster (st.shared::cluster),
// one barrier.cluster later each warp reduces the <= 16 slots of its own CTA.  No global memory, no grid
// barrier, no slot re-read through L2.  Keys live in registers (vertex v = rank*1024 + tid + q*C*1024).
// SQUARE: rows come from a mirrored n x n copy (contiguous); otherwise from the condensed vector, where
// the part v < x of a row is a strided column.
// condensed upper triangle -> full n x n mirror, 64 x 64 tiles through shared memory so that both the
// (i, j) and the (j, i) halves are written with coalesced rows (a direct gather of the lower half reads
// one 32-byte sector per element)
__global__ void __launch_bounds__(256) rthc_mirror_square_u16(const uint16_t *__restrict__ cond, int64_t n, int64_t pitch,
                                                              uint16_t *__restrict__ W)
{
    const int ti = blockIdx.y, tj = blockIdx.x;
    if (tj < ti) return;
    __shared__ uint16_t tile[64][66];
    const int64_t i0 = (int64_t)ti * 64, j0 = (int64_t)tj * 64;
    const int c = threadIdx.x & 63, r4 = threadIdx.x >> 6;      // 4 rows per pass
    for (int r = r4; r < 64; r += 4) {
        const int64_t i = i0 + r, j = j0 + c;
        uint16_t v = 0;
        if (i < n && j < n && j > i) v = cond[n * i - (i * (i + 1)) / 2 + (j - i - 1)];
        tile[r][c] = v;
    }
    __syncthreads();
    for (int r = r4; r < 64; r += 4) {
        const int64_t i = i0 + r, j = j0 + c;
        if (i < n && j < n) {
            if (ti != tj) W[i * pitch + j] = tile[r][c];
            else W[i * pitch + j] = c > r ? tile[r][c] : (c < r ? tile[c][r] : (uint16_t)0);
        }
        if (ti != tj) {                                          // transposed block: row j0 + r, column i0 + c
            const int64_t jj = j0 + r, ii = i0 + c;
            if (jj < n && ii < n) W[jj * pitch + ii] = tile[c][r];
        }
    }
}

#define PC_THREADS 1024
#define PC_MAX_CLUSTER 16

template <int MAXQ, bool SQUARE>
__global__ void __launch_bounds__(PC_THREADS, 1)
rthc_prim_cluster_kernel(const uint16_t *__restrict__ mat, int64_t n, int32_t *ex, int32_t *ey, double *ew)
{
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned C = cluster.num_blocks(), rank = cluster.block_rank();
    __shared__ unsigned long long slot[2][PC_MAX_CLUSTER];      // written by every CTA of the cluster
    __shared__ unsigned int sw[PC_THREADS / 32], si[PC_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t stride = (int64_t)C * PC_THREADS;
    const int64_t v0 = (int64_t)rank * PC_THREADS + threadIdx.x;
    uint32_t key[MAXQ];
    uint32_t gone = 0;                                          // bit q: vertex q is in the tree (or beyond n)
#pragma unroll
    for (int q = 0; q < MAXQ; ++q) {
        key[q] = 0xFFFFFFFFu;
        if (v0 + q * stride >= n) gone |= 1u << q;
    }
    int64_t x = 0;
    for (int64_t k = 0; k < n - 1; ++k) {
        const int64_t xbase = SQUARE ? x * n : n * x - (x * (x + 1)) / 2;
        uint32_t d[MAXQ];
#pragma unroll
        for (int q = 0; q < MAXQ; ++q) {                        // all loads of the step in flight together
            const int64_t v = v0 + q * stride;
            if (v == x) gone |= 1u << q;
            d[q] = 0xFFFFFFFFu;
            if (!((gone >> q) & 1u)) {
                const int64_t idx = SQUARE ? xbase + v
                                           : (v > x ? xbase + (v - x - 1) : n * v - (v * (v + 1)) / 2 + (x - v - 1));
                d[q] = __ldg(mat + idx);
            }
        }
        uint32_t bw = 0xFFFFFFFFu, bi = 0xFFFFFFFFu;
#pragma unroll
        for (int q = 0; q < MAXQ; ++q) {
            if ((gone >> q) & 1u) continue;
            key[q] = min(key[q], d[q]);
            if (key[q] < bw) { bw = key[q]; bi = (uint32_t)(v0 + q * stride); }     // ascending v: lowest index of the minimum
        }
        // CTA minimum: two REDUX per warp (weight, then lowest index among the lanes that hold it)
        uint32_t m = __reduce_min_sync(0xffffffffu, bw);
        uint32_t i = __reduce_min_sync(0xffffffffu, bw == m ? bi : 0xFFFFFFFFu);
        if (lane == 0) { sw[warp] = m; si[warp] = i; }
        __syncthreads();
        const int par = (int)(k & 1);
        if (warp == 0) {
            bw = sw[lane]; bi = si[lane];
            m = __reduce_min_sync(0xffffffffu, bw);
            i = __reduce_min_sync(0xffffffffu, bw == m ? bi : 0xFFFFFFFFu);
            if ((unsigned)lane < C)                              // lane l publishes to CTA l
                *cluster.map_shared_rank(&slot[par][rank], lane) = ((unsigned long long)m << 32) | i;
        }
        cluster.sync();
        const unsigned long long sv = (unsigned)lane < C ? slot[par][lane] : ~0ull;
        bw = (uint32_t)(sv >> 32); bi = (uint32_t)sv;
        m = __reduce_min_sync(0xffffffffu, bw);
        i = __reduce_min_sync(0xffffffffu, bw == m ? bi : 0xFFFFFFFFu);
        if (rank == 0 && threadIdx.x == 0) { ex[k] = (int32_t)x; ey[k] = (int32_t)i; ew[k] = (double)m; }
        x = i;
    }
}

// Same step, rows from the mirrored copy with a pitch that is a multiple of 8: a thread owns 8 CONTIGUOUS
// vertices per chunk (one 16-byte load), keys stay packed 2 x 16 bit (VIMNMX.U16x2), vertices already in the
// tree are held at 0xFFFF by an OR mask and skipped when the index is resolved -- about 30 instructions per
// thread and step instead of about 150, which matters because 32 warps share the four schedulers of an SM.
template <int NCH>
__global__ void __launch_bounds__(PC_THREADS, 1)
rthc_prim_cluster_vec_kernel(const uint16_t *__restrict__ sq, int64_t n, int64_t pitch, int32_t *ex, int32_t *ey, double *ew)
{
    cg::cluster_group cluster = cg::this_cluster();
    const unsigned C = cluster.num_blocks(), rank = cluster.block_rank();
    __shared__ unsigned long long slot[2][PC_MAX_CLUSTER];
    __shared__ unsigned int sw[PC_THREADS / 32], si[PC_THREADS / 32];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t key[NCH][4], gm[NCH][4];
    int64_t vbase[NCH];
#pragma unroll
    for (int c = 0; c < NCH; ++c) {
        vbase[c] = (((int64_t)c * C + rank) * PC_THREADS + threadIdx.x) * 8;
#pragma unroll
        for (int w = 0; w < 4; ++w) {
            key[c][w] = 0xFFFFFFFFu;
            const int64_t v = vbase[c] + 2 * w;
            gm[c][w] = (v >= n ? 0x0000FFFFu : 0u) | (v + 1 >= n ? 0xFFFF0000u : 0u);
        }
    }
    int64_t x = 0;
    for (int64_t k = 0; k < n - 1; ++k) {
        const uint16_t *row = sq + x * pitch;
        uint32_t bw = 0xFFFFFFFFu;
        int bc = 0;
#pragma unroll
        for (int c = 0; c < NCH; ++c) {
            const int64_t off = x - vbase[c];
            if (off >= 0 && off < 8) {                          // x joins the tree; 64-bit shifts instead of a register
                const unsigned long long bits = 0xFFFFull << (16 * (int)(off & 3));      // index (no local memory)
                unsigned long long lo = ((unsigned long long)gm[c][1] << 32) | gm[c][0];
                unsigned long long hi = ((unsigned long long)gm[c][3] << 32) | gm[c][2];
                if (off < 4) lo |= bits; else hi |= bits;
                gm[c][0] = (uint32_t)lo; gm[c][1] = (uint32_t)(lo >> 32);
                gm[c][2] = (uint32_t)hi; gm[c][3] = (uint32_t)(hi >> 32);
            }
            if ((gm[c][0] & gm[c][1] & gm[c][2] & gm[c][3]) == 0xFFFFFFFFu) continue;      // nothing left here
            const uint4 d = __ldg(reinterpret_cast<const uint4 *>(row + vbase[c]));
            key[c][0] = __vminu2(key[c][0], d.x) | gm[c][0];
            key[c][1] = __vminu2(key[c][1], d.y) | gm[c][1];
            key[c][2] = __vminu2(key[c][2], d.z) | gm[c][2];
            key[c][3] = __vminu2(key[c][3], d.w) | gm[c][3];
            const uint32_t m2 = __vminu2(__vminu2(key[c][0], key[c][1]), __vminu2(key[c][2], key[c][3]));
            const uint32_t m = min(m2 & 0xFFFFu, m2 >> 16);
            if (m < bw) { bw = m; bc = c; }
        }
        uint32_t m = __reduce_min_sync(0xffffffffu, bw);
        uint32_t bi = 0xFFFFFFFFu;
        if (bw == m && m != 0xFFFFFFFFu) {                      // few lanes: resolve the lowest alive index that holds m
#pragma unroll
            for (int c = 0; c < NCH; ++c) {
                if (c != bc) continue;
#pragma unroll
                for (int e = 7; e >= 0; --e) {
                    const uint32_t kv = (key[c][e >> 1] >> (16 * (e & 1))) & 0xFFFFu;
                    const bool alive = !((gm[c][e >> 1] >> (16 * (e & 1))) & 1u);
                    if (alive && kv == m) bi = (uint32_t)(vbase[c] + e);
                }
            }
        }
        uint32_t i = __reduce_min_sync(0xffffffffu, bi);
        if (lane == 0) { sw[warp] = m; si[warp] = i; }
        __syncthreads();
        const int par = (int)(k & 1);
        if (warp == 0) {
            bw = sw[lane]; bi = si[lane];
            m = __reduce_min_sync(0xffffffffu, bw);
            i = __reduce_min_sync(0xffffffffu, bw == m ? bi : 0xFFFFFFFFu);
            if ((unsigned)lane < C)
                *cluster.map_shared_rank(&slot[par][rank], lane) = ((unsigned long long)m << 32) | i;
        }
        cluster.sync();
        const unsigned long long sv = (unsigned)lane < C ? slot[par][lane] : ~0ull;
        bw = (uint32_t)(sv >> 32); bi = (uint32_t)sv;
        m = __reduce_min_sync(0xffffffffu, bw);
        i = __reduce_min_sync(0xffffffffu, bw == m ? bi : 0xFFFFFFFFu);
        if (rank == 0 && threadIdx.x == 0) { ex[k] = (int32_t)x; ey[k] = (int32_t)i; ew[k] = (double)m; }
        x = i;
    }
}

template <int NCH>
static cudaError_t launch_prim_cluster_vec(int C, cudaStream_t stream, const uint16_t *sq, int64_t n, int64_t pitch,
                                           int32_t *ex, int32_t *ey, double *ew)
{
    auto kern = rthc_prim_cluster_vec_kernel<NCH>;
    cudaError_t e;
    if (C > 8 && (e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1)) != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)C);
    cfg.blockDim = dim3(PC_THREADS);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int clusters = 0;
    if ((e = cudaOccupancyMaxActiveClusters(&clusters, kern, &cfg)) != cudaSuccess) return e;
    if (clusters < 1) return cudaErrorInvalidConfiguration;
    return cudaLaunchKernelEx(&cfg, kern, sq, n, pitch, ex, ey, ew);
}

template <int MAXQ, bool SQUARE>
static cudaError_t launch_prim_cluster(int C, cudaStream_t stream, const uint16_t *mat, int64_t n, int32_t *ex, int32_t *ey,
                                       double *ew)
{
    auto kern = rthc_prim_cluster_kernel<MAXQ, SQUARE>;
    cudaError_t e;
    if (C > 8 && (e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1)) != cudaSuccess) return e;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)C);
    cfg.blockDim = dim3(PC_THREADS);
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = (unsigned)C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr; cfg.numAttrs = 1;
    int clusters = 0;
    if ((e = cudaOccupancyMaxActiveClusters(&clusters, kern, &cfg)) != cudaSuccess) return e;
    if (clusters < 1) return cudaErrorInvalidConfiguration;
    return cudaLaunchKernelEx(&cfg, kern, mat, n, ex, ey, ew);
}

// returns 1 when the cluster kernel is not applicable (size, RTHC_PRIM_CLUSTER=0, no cluster launch): caller falls
// back to the cooperative kernel; 0 on success; < 0 on error
static int run_prim_cluster(rthc_ctx *ctx, const uint16_t *d_cond, int32_t *d_ex, int32_t *d_ey, double *d_ew)
{
    static const int mode = getenv("RTHC_PRIM_CLUSTER") ? atoi(getenv("RTHC_PRIM_CLUSTER")) : 1;   // 0 off, 1 auto, 2 no square copy
    if (mode == 0) return 1;
    const int64_t n = ctx->n, pitch = (n + 7) / 8 * 8;
    uint16_t *d_sq = nullptr;
    if (mode == 1) {                                             // mirrored copy when 2 n^2 bytes are free (plus slack)
        size_t free_b = 0, total_b = 0;
        cudaMemGetInfo(&free_b, &total_b);
        // a cold pool allocation costs about 7 ms per GB, which at C4 (20 GB) eats what the contiguous rows save in a
        // one-shot run (0.32-0.39 s cold vs 0.35 s from the condensed vector; 0.20 s once the pool is warm)
        static const double mirror_gb = getenv("RTHC_PRIM_MIRROR_GB") ? atof(getenv("RTHC_PRIM_MIRROR_GB")) : 4.0;
        const double need = 2.0 * (double)n * (double)pitch;
        if (need <= mirror_gb * 1e9 && (double)free_b > need + 2e9 && rthc_dev_alloc(ctx, &d_sq, (size_t)(n * pitch)) == 0) {
            const unsigned B = (unsigned)((n + 63) / 64);
            rthc_mirror_square_u16<<<dim3(B, B), 256, 0, ctx->stream>>>(d_cond, n, pitch, d_sq);
            ctx->launches++;
        } else {
            d_sq = nullptr;
        }
    }
    cudaError_t e = cudaErrorInvalidConfiguration;
    for (int C = PC_MAX_CLUSTER; C >= 8 && e != cudaSuccess; C >>= 1) {
        const int64_t per = (n + (int64_t)C * PC_THREADS - 1) / ((int64_t)C * PC_THREADS);
        if (per > 16) break;
        if (d_sq) e = per <= 8 ? launch_prim_cluster_vec<1>(C, ctx->stream, d_sq, n, pitch, d_ex, d_ey, d_ew)
                               : launch_prim_cluster_vec<2>(C, ctx->stream, d_sq, n, pitch, d_ex, d_ey, d_ew);
        else e = per <= 8 ? launch_prim_cluster<8, false>(C, ctx->stream, d_cond, n, d_ex, d_ey, d_ew)
                          : launch_prim_cluster<16, false>(C, ctx->stream, d_cond, n, d_ex, d_ey, d_ew);
        if (e != cudaSuccess) cudaGetLastError();
    }
    if (e == cudaSuccess) ctx->launches++;
    rthc_dev_free(ctx, &d_sq);                                   // stream-ordered: released after the kernel
    return e == cudaSuccess ? 0 : 1;
}

template <typename T>
static int run_prim(rthc_ctx *ctx, const T *d_cond, std::vector<int32_t> &ex, std::vector<int32_t> &ey,
                    std::vector<double> &ew)
{
    const int64_t n = ctx->n, m = n - 1;
    int nblocks = ctx->sm_count;
    int per_thread = (int)((n + (int64_t)nblocks * SEQ_THREADS - 1) / ((int64_t)nblocks * SEQ_THREADS));
    MinSlot *d_slots = nullptr;
    int32_t *d_ex = nullptr, *d_ey = nullptr;
    double *d_ew = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_slots, (size_t)2 * nblocks))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ex, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ey, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ew, (size_t)m))) return rc;
    int64_t nn = n;
    int clustered = 1;
    if constexpr (std::is_same<T, uint16_t>::value) clustered = run_prim_cluster(ctx, d_cond, d_ex, d_ey, d_ew);
    if (clustered < 0) return clustered;
    if (clustered == 1) {
        if (per_thread > 8)
            return rthc_fail(ctx, RTHC_E_RANGE, "Prim-exact supports up to %lld samples", (long long)nblocks * SEQ_THREADS * 8);
        void *args[] = {(void *)&d_cond, (void *)&nn, (void *)&per_thread, (void *)&d_slots, (void *)&d_ex, (void *)&d_ey,
                        (void *)&d_ew};
        RTHC_CUDA(ctx, cudaLaunchCooperativeKernel((void *)rthc_prim_kernel<T>, dim3(nblocks), dim3(SEQ_THREADS), args, 0,
                                                   ctx->stream));
        ctx->launches++;
    }
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
    const int64_t i = blockIdx.x;                              // rows on grid.x: no 65,535 limit
    const int64_t j = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
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
    int64_t scans;      // full row scans (one grid barrier each)
    int64_t fused;      // scans of the next chain tip that rode on a merge's barrier
    int64_t hits;       // chain steps answered by the nearest-neighbour cache (no barrier)
};

// nearest neighbour of every row: lexicographic minimum of (D[r, i], i) over i != r  (one CTA per row)
template <typename T>
__global__ void __launch_bounds__(SEQ_THREADS) rthc_nn_init(const T *__restrict__ W, int64_t n, MinSlot *__restrict__ nn)
{
    __shared__ MinSlot sh[SEQ_THREADS / 32];
    const int64_t r = blockIdx.x;
    const T *row = W + r * n;
    unsigned long long bw = ~0ull;
    unsigned int bi = ~0u;
    for (int64_t i = threadIdx.x; i < n; i += SEQ_THREADS) {
        if (i == r) continue;
        const unsigned long long d = order_bits(row[i]);
        if (slot_less(d, (unsigned int)i, bw, bi)) { bw = d; bi = (unsigned int)i; }
    }
    block_min(bw, bi, sh);
    if (threadIdx.x == 0) { nn[r].w = bw; nn[r].i = bi; }
}

// The merge loop.  scipy's chain step asks for the nearest neighbour of the chain tip x: the previous chain
// element if it ties the minimum, else the lowest index of the minimum of row x.  A grid-wide row scan costs one
// grid barrier (about 2.4 us) however short the row is, so the kernel keeps, for every live row r, the exact
// lexicographic minimum nn[r] = min over live i != r of (D[r, i], i) and only scans when that entry is unknown:
//   * a merge (a dies, b = a u b) already visits every live row i to write D[i, b]; the same thread folds the new
//     entry into the cache -- nn[i] = min(nn[i], (D[i, b], b)) is exact because every other entry of row i is
//     unchanged -- and drops it (index = ~0) when the cached neighbour was a or b;
//   * the candidates (D[b, i], i) of the new row b are reduced over the grid on the merge's own barrier;
//   * so is a scan of the row of p, the chain element that becomes the tip after the merge (its neighbour was a
//     or b by construction): row p only changes in column b, whose new value the thread that owns i == p has.
// A chain step whose tip has a cached neighbour reads it (warp 0 of every CTA, broadcast through shared memory)
// and costs no barrier.  The answer is the same (value, lowest index, previous-wins) triple as the scan, so Z
// stays byte-identical to scipy.  nn[] is double-buffered by merge parity: a merge writes the other copy for every
// live row, so a CTA still walking the chain never reads an entry that a faster CTA has already updated.
// Every nn[r] is written by ONE thread (the owner of index r in the strided loops), reads of scalar state go
// through warp 0 + shared memory (148 x 16 warps loading one address serialise in L2), size[] of the last merge
// is patched from registers until the next barrier has published it.
template <typename T>
__global__ void __launch_bounds__(SEQ_THREADS, 1)
rthc_nn_chain_kernel(T *W, int64_t n, int method, int32_t *size, int32_t *chains, MinSlot *slots, MinSlot *nn2,
                     int32_t *ex, int32_t *ey, double *ew, ChainCtl *ctl)
{
    cg::grid_group grid = cg::this_grid();
    __shared__ MinSlot sh[SEQ_THREADS / 32];
    __shared__ MinSlot bcs[2];                                   // broadcast buffers, alternating
    __shared__ unsigned long long bcd[2];
    __shared__ int32_t bci[2][4];
    const int nblocks = gridDim.x;
    const int64_t gtid = (int64_t)blockIdx.x * SEQ_THREADS + threadIdx.x;
    const int64_t gthreads = (int64_t)nblocks * SEQ_THREADS;
    const bool leader = blockIdx.x == 0 && threadIdx.x == 0;
    const int lane = threadIdx.x & 31;
    const bool w0 = threadIdx.x < 32;
    int32_t *chain = chains + (int64_t)blockIdx.x * n;          // this CTA's copy, touched by its thread 0 only

    // replicated scalar state (identical in every thread of the grid)
    int64_t len = 0, first_live = 0, step = 0, scans = 0, fused = 0, hits = 0;
    int32_t x = -1, prev = -1;
    int32_t dead = -1, grown = -1, grown_size = 0;              // last merge, until the next barrier publishes size[]
    unsigned long long gw = ~0ull; unsigned int gi = ~0u;        // nn of `grown`
    bool have_tip = false;                                       // nn of the tip is in (tw, ti, tdprev)
    unsigned long long tw = ~0ull, tdprev = ~0ull; unsigned int ti = ~0u;
    int q = 0, cur = 0;                                          // broadcast buffer, current nn copy

#define SIZE_OF(i_) ((i_) == dead ? 0 : ((i_) == grown ? grown_size : __ldcg(&size[(i_)])))

    for (int64_t k = 0; k < n - 1;) {
        MinSlot *nn = nn2 + (int64_t)cur * n;
        if (len == 0) {
            if (w0) {                                            // lowest live index, 32 candidates per trip
                int64_t f = first_live;
                for (;;) {
                    const int64_t i = f + lane;
                    const bool live = i < n && SIZE_OF(i) != 0;
                    const unsigned m = __ballot_sync(0xffffffffu, live);
                    if (m) { f += __ffs(m) - 1; break; }
                    f += 32;
                }
                if (lane == 0) { bci[q][0] = (int32_t)f; chain[0] = (int32_t)f; }
            }
            __syncthreads();
            first_live = bci[q][0]; q ^= 1;
            x = (int32_t)first_live; prev = -1; len = 1; have_tip = false;
        }
        // ---- nearest neighbour of the tip x -----------------------------------------------------
        unsigned long long cw, dprev;
        unsigned int ci;
        if (have_tip) {
            cw = tw; ci = ti; dprev = tdprev;
            have_tip = false;
        } else {
            if (threadIdx.x < 3) {                               // three scalars, three lanes: the loads overlap
                if (threadIdx.x == 0) bcs[q].w = x == grown ? gw : __ldcg(&nn[x].w);
                else if (threadIdx.x == 1) bcs[q].i = x == grown ? gi : __ldcg(&nn[x].i);
                else bcd[q] = len > 1 ? order_bits(__ldcg(&W[(int64_t)x * n + prev])) : ~0ull;
            }
            __syncthreads();
            cw = bcs[q].w; ci = bcs[q].i; dprev = bcd[q]; q ^= 1;
            if (ci != ~0u) ++hits;
        }
        if (ci == ~0u) {
            // ---- unknown: scan row x over the grid (lowest index of the minimum over live i != x) --
            unsigned long long bw = ~0ull;
            unsigned int bi = ~0u;
            const T *row = W + (int64_t)x * n;
            for (int64_t i = gtid; i < n; i += gthreads) {
                if (i == x) continue;
                const int32_t szi = __ldcg(&size[i]);            // liveness and the element together: one latency
                const unsigned long long d = order_bits(__ldcg(&row[i]));
                if ((i == dead ? 0 : (i == grown ? grown_size : szi)) == 0) continue;
                if (slot_less(d, (unsigned int)i, bw, bi)) { bw = d; bi = (unsigned int)i; }
            }
            block_min(bw, bi, sh);
            MinSlot *cs = slots + (step & 1) * nblocks;
            if (threadIdx.x == 0) { cs[blockIdx.x].w = bw; cs[blockIdx.x].i = bi; }
            grid.sync();
            grid_min(cs, nblocks, cw, ci, sh);
            ++step; ++scans;
            dead = -1; grown = -1;
            if (gtid == x % gthreads) { nn[x].w = cw; nn[x].i = ci; }
        }
        int32_t y = (int32_t)ci;
        unsigned long long dmin = cw;
        if (len > 1 && dprev <= dmin) { y = prev; dmin = dprev; }  // previous chain element wins ties
        if (!(len > 1 && y == prev)) {
            // ---- extend the chain ---------------------------------------------------------------
            if (threadIdx.x == 0) chain[len] = y;
            prev = x; x = y; ++len;
            continue;
        }
        // ---- merge x and y (reciprocal nearest neighbours) ---------------------------------------
        const int32_t a = x < y ? x : y, b = x < y ? y : x;       // a dies, b survives (scipy: x < y swap)
        if (threadIdx.x == 0) {
            bci[q][0] = SIZE_OF(a); bci[q][1] = SIZE_OF(b);
            const int32_t p0 = len >= 3 ? chain[len - 3] : -1, pp0 = len >= 4 ? chain[len - 4] : -1;
            bci[q][2] = p0; bci[q][3] = pp0;
            bcd[q] = pp0 >= 0 ? order_bits(__ldcg(&W[(int64_t)p0 * n + pp0])) : ~0ull;
        }
        __syncthreads();
        const int32_t na = bci[q][0], nb = bci[q][1], p = bci[q][2], pp = bci[q][3];
        const unsigned long long dpp = bcd[q];
        q ^= 1;
        if (leader) { ex[k] = a; ey[k] = b; ew[k] = from_bits(dmin, T()); }
        const T *ra = W + (int64_t)a * n;
        T *rb = W + (int64_t)b * n;
        const T *rp = W + (int64_t)(p >= 0 ? p : 0) * n;
        MinSlot *nxt = nn2 + (int64_t)(cur ^ 1) * n;
        const double dab = from_bits(dmin, T());
        unsigned long long bwb = ~0ull, bwp = ~0ull;
        unsigned int bib = ~0u, bip = ~0u;
        for (int64_t i = gtid; i < n; i += gthreads) {
            if (i == a || i == b) continue;
            // every load of the element first, liveness included (a dead row costs a few wasted loads; testing size[i]
            // first would put a second memory latency in front of every live one, and the stores below may alias the
            // loads as far as the compiler knows)
            const int32_t szi = __ldcg(&size[i]);
            const T da = __ldcg(&ra[i]), db = __ldcg(&rb[i]);
            unsigned long long ow = __ldcg(&nn[i].w);
            unsigned int oi = __ldcg(&nn[i].i);
            const T dp = p >= 0 ? __ldcg(&rp[i]) : T(0);
            const int32_t ni = i == dead ? 0 : (i == grown ? grown_size : szi);
            if (ni == 0) continue;
            const T v = lance_williams<T>(method, da, db, dab, na, nb, ni);
            rb[i] = v;                       // row b (contiguous)
            W[i * n + b] = v;                // column b (symmetric copy)
            const unsigned long long ob = order_bits(v);
            if (slot_less(ob, (unsigned int)i, bwb, bib)) { bwb = ob; bib = (unsigned int)i; }
            if (i == p) {                    // row p: its only changed entry; the rest is scanned below
                if (slot_less(ob, (unsigned int)b, bwp, bip)) { bwp = ob; bip = (unsigned int)b; }
            } else {
                if (oi == (unsigned int)a || oi == (unsigned int)b) oi = ~0u;
                else if (oi != ~0u && slot_less(ob, (unsigned int)b, ow, oi)) { ow = ob; oi = (unsigned int)b; }
                __stcg(&nxt[i].w, ow); __stcg(&nxt[i].i, oi);
                if (p >= 0) {
                    const unsigned long long d = order_bits(dp);
                    if (slot_less(d, (unsigned int)i, bwp, bip)) { bwp = d; bip = (unsigned int)i; }
                }
            }
        }
        block_min(bwb, bib, sh);
        MinSlot *sb = slots + (2 + (step & 1)) * nblocks, *sp = slots + (4 + (step & 1)) * nblocks;
        if (threadIdx.x == 0) { sb[blockIdx.x].w = bwb; sb[blockIdx.x].i = bib; }
        if (p >= 0) {
            block_min(bwp, bip, sh);
            if (threadIdx.x == 0) { sp[blockIdx.x].w = bwp; sp[blockIdx.x].i = bip; }
        }
        grid.sync();                          // row/column b and the other nn copy are visible
        ++step;
        cur ^= 1;
        grid_min(sb, nblocks, gw, gi, sh);
        if (gtid == b % gthreads) { __stcg(&nxt[b].w, gw); __stcg(&nxt[b].i, gi); }
        if (p >= 0) {
            grid_min(sp, nblocks, tw, ti, sh);
            tdprev = dpp; have_tip = true; ++fused;
            if (gtid == p % gthreads) { __stcg(&nxt[p].w, tw); __stcg(&nxt[p].i, ti); }
        }
        if (leader) { size[a] = 0; size[b] = na + nb; }
        dead = a; grown = b; grown_size = na + nb;
        ++k;
        len -= 2;
        if (len > 0) { x = p; prev = pp; }
    }
#undef SIZE_OF
    if (leader) { ctl->scans = scans; ctl->fused = fused; ctl->hits = hits; }
}

template <typename T, typename S>
static int run_nn_chain(rthc_ctx *ctx, const S *d_cond, int method, std::vector<int32_t> &ex,
                        std::vector<int32_t> &ey, std::vector<double> &ew)
{
    const int64_t n = ctx->n, m = n - 1;
    T *d_W = nullptr;
    int32_t *d_size = nullptr, *d_chain = nullptr, *d_ex = nullptr, *d_ey = nullptr;
    double *d_ew = nullptr;
    MinSlot *d_slots = nullptr, *d_nn = nullptr;
    ChainCtl *d_ctl = nullptr;
    int nblocks = ctx->sm_count;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_W, (size_t)(n * n)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_size, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_chain, (size_t)n * (size_t)nblocks))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_slots, (size_t)6 * nblocks))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_nn, (size_t)2 * (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ex, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ey, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ew, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ctl, 1))) return rc;
    {
        dim3 grd((unsigned)n, (unsigned)((n + 255) / 256));
        rthc_expand_square<T, S><<<grd, 256, 0, ctx->stream>>>(d_cond, n, d_W);
        RTHC_LAUNCH_CHECK(ctx);
        rthc_nn_init<T><<<(unsigned)n, SEQ_THREADS, 0, ctx->stream>>>(d_W, n, d_nn);
        RTHC_LAUNCH_CHECK(ctx);
        std::vector<int32_t> ones((size_t)n, 1);
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_size, ones.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    int64_t nn = n;
    void *args[] = {(void *)&d_W, (void *)&nn, (void *)&method, (void *)&d_size, (void *)&d_chain, (void *)&d_slots,
                    (void *)&d_nn, (void *)&d_ex, (void *)&d_ey, (void *)&d_ew, (void *)&d_ctl};
    RTHC_CUDA(ctx, cudaLaunchCooperativeKernel((void *)rthc_nn_chain_kernel<T>, dim3(nblocks), dim3(SEQ_THREADS), args, 0,
                                               ctx->stream));
    ctx->launches++;
    ex.resize((size_t)m); ey.resize((size_t)m); ew.resize((size_t)m);
    ChainCtl ctl;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ex.data(), d_ex, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ey.data(), d_ey, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ew.data(), d_ew, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(&ctl, d_ctl, sizeof(ChainCtl), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->row_scans = ctl.scans + ctl.fused;
    ctx->chain_full_scans = ctl.scans + m;          // grid barriers: one per scan of its own, one per merge (info[9])
    ctx->chain_cache_hits = ctl.hits;
    rthc_dev_free(ctx, &d_W); rthc_dev_free(ctx, &d_size); rthc_dev_free(ctx, &d_chain); rthc_dev_free(ctx, &d_slots);
    rthc_dev_free(ctx, &d_nn); rthc_dev_free(ctx, &d_ex); rthc_dev_free(ctx, &d_ey); rthc_dev_free(ctx, &d_ew);
    rthc_dev_free(ctx, &d_ctl);
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
