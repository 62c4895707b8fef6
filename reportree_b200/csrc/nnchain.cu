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
__global__ void rthc_expand_square(const S *__restrict__ cond, int64_t n, int64_t pitch, T *__restrict__ W)
{
    const int64_t i = blockIdx.x;                              // rows on grid.x: no 65,535 limit
    const int64_t j = (int64_t)blockIdx.y * blockDim.x + threadIdx.x;
    if (j >= pitch) return;
    T v = 0;
    if (i != j && j < n) {
        const int64_t a = i < j ? i : j, b = i < j ? j : i;
        v = (T)cond[n * a - (a * (a + 1)) / 2 + (b - a - 1)];
    }
    W[i * pitch + j] = v;
}

// all-lanes lexicographic minimum of (w, i) over a warp
__device__ __forceinline__ void warp_min_all(unsigned long long &w, unsigned int &i)
{
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        unsigned long long ow = __shfl_xor_sync(0xffffffffu, w, s);
        unsigned int oi = __shfl_xor_sync(0xffffffffu, i, s);
        if (slot_less(ow, oi, w, i)) { w = ow; i = oi; }
    }
}

// control block in global memory (written by block 0, thread 0 only)
struct ChainCtl {
    int64_t scans;      // row scans that had a barrier of their own (cache entry unknown when the chain needed it)
    int64_t fused;      // scans of the next chain tip that rode on a merge's barrier
    int64_t hits;       // chain steps answered by the cache
    int64_t clk[8];     // SM cycles of CTA 0, thread 0 per section: walk, scan loads + reduce, scan barrier + combine,
                        // merge pre-barrier, merge pass, merge reductions, merge barrier, merge combine
};

// The two lexicographically smallest (value bits, index) entries of a row.  (~0, ~0) = none.
struct __align__(16) Top2 {
    unsigned long long w1, w2;
    unsigned int i1, i2, pad0, pad1;
    __device__ __forceinline__ void clear() { w1 = w2 = ~0ull; i1 = i2 = ~0u; }
    __device__ __forceinline__ void insert(unsigned long long w, unsigned int i)
    {
        if (slot_less(w, i, w1, i1)) { w2 = w1; i2 = i1; w1 = w; i1 = i; }
        else if (slot_less(w, i, w2, i2)) { w2 = w; i2 = i; }
    }
};
__device__ __forceinline__ void warp_top2(Top2 &t)                // all lanes end with the warp's top-2
{
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) {
        const unsigned long long ow1 = __shfl_xor_sync(0xffffffffu, t.w1, s), ow2 = __shfl_xor_sync(0xffffffffu, t.w2, s);
        const unsigned int oi1 = __shfl_xor_sync(0xffffffffu, t.i1, s), oi2 = __shfl_xor_sync(0xffffffffu, t.i2, s);
        t.insert(ow1, oi1);
        t.insert(ow2, oi2);
    }
}
// block-wide top-2; result valid in warp 0 (all its lanes)
__device__ __forceinline__ void block_top2(Top2 &t, Top2 *sh)
{
    warp_top2(t);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) sh[warp] = t;
    __syncthreads();
    if (warp == 0) {
        if (lane < SEQ_THREADS / 32) t = sh[lane]; else t.clear();
        warp_top2(t);
    }
    __syncthreads();
}

// cache entry of a row: E1 = its nearest neighbour (exact) or i1 == ~0: unknown; E2 = the runner-up when `act`,
// otherwise only a bound: every entry of the row other than E1 (and E2 when act) is lexicographically greater
// than (w2, i2)
struct __align__(16) NNEntry {
    unsigned long long w1, w2;
    unsigned int i1, i2, act, pad;
};
__device__ __forceinline__ void nn_store(NNEntry *p, unsigned long long w1, unsigned int i1, unsigned long long w2, unsigned int i2,
                                         unsigned int act)
{
    __stcg(reinterpret_cast<uint4 *>(p), make_uint4((uint32_t)w1, (uint32_t)(w1 >> 32), (uint32_t)w2, (uint32_t)(w2 >> 32)));
    __stcg(reinterpret_cast<uint4 *>(p) + 1, make_uint4(i1, i2, act, 0u));
}

#define NNC_MAX_CLUSTER 16
#define NNC_SCAN 1
#define NNC_MERGE 2
struct ChainBc {                // the decision of warp 0, broadcast through shared memory
    unsigned long long dmin;
    int32_t what, x, a, b, na, nb, p;
};

// The merge loop: scipy's nn_chain inside ONE thread-block cluster.
//
// A chain step asks for the nearest neighbour of the chain tip x: the previous chain element if it ties the minimum,
// else the lowest index of the minimum of row x.  The loop is a chain of dependent steps, so what it costs is
// latency, and the grid-wide version paid 5.5 us per step for a cooperative grid barrier and the L2 round trips
// around it (round 1: 0.3 % of the HBM roofline).  Here
//   * the loop runs in one cluster of 16 CTAs (8 where 16 cannot be scheduled): barrier.cluster instead of a grid
//     barrier, and the per-CTA results of a reduction are written straight into the shared memory of every CTA of
//     the cluster (st.shared::cluster), so nothing is re-read through L2 after a barrier.  A row is <= 0.2 MB:
//     16 SMs stream it in about a microsecond, more SMs would only wait longer at the barrier;
//   * every row keeps a cache entry: its exact nearest neighbour E1 and the runner-up E2 (or a bound below every
//     other entry).  A chain step whose tip has a known E1 is a look-up by warp 0 of every CTA (no barrier; the
//     distance to the previous element and the sizes of the chain elements ride along in the CTA's private chain
//     arrays, so no matrix element is read).  A merge (a dies, b = a u b) visits every live row i to write D[i, b];
//     the same thread updates the entry of row i from E1, E2 / the bound and the new entry N = (D[i, b], b): entries
//     of a and b leave the candidate set, N joins it, the smallest candidate is the new E1 if it does not exceed the
//     bound (everything unknown lies above the bound), the next one the new E2 under the same condition; otherwise
//     E1 becomes unknown and the row is scanned over the cluster when (if) the chain reaches it;
//   * the merge's barrier also carries the top-2 of the new row b and the scan of p, the chain element that becomes
//     the tip after the merge (its neighbour was a or b by construction): row p only changes in column b, whose new
//     value the thread that owns i == p has, so the scan reads nothing that the phase writes.
// A look-up returns the same (value, lowest index, previous-wins) triple as scipy's scan, so Z stays byte-identical
// (tests/test_gpu_exact_linkage.py, the C3 / C5 goldens).  Cache entries and size[] of rows touched after the last
// barrier are carried in the registers of warp 0 (`grown`, `known`); a barrier before every merge pass makes the
// leader's earlier stores visible to the pass.  W is pitched so that every row starts on a 16-byte boundary.
template <typename T>
__global__ void __launch_bounds__(SEQ_THREADS, 1)
rthc_nn_chain_kernel(T *W, int64_t n, int64_t pitch, int method, int32_t *size, int32_t *chains, unsigned long long *chains_d,
                     int32_t *chains_s, NNEntry *nn, int32_t *ex, int32_t *ey, double *ew, ChainCtl *ctl)
{
    cg::cluster_group cluster = cg::this_cluster();
    const int C = (int)cluster.num_blocks(), rank = (int)cluster.block_rank();
    extern __shared__ uint32_t s_live[];                         // bit j: cluster j is alive
    __shared__ Top2 sh[SEQ_THREADS / 32];
    __shared__ Top2 slotA[2][NNC_MAX_CLUSTER], slotB[2][NNC_MAX_CLUSTER];      // written by every CTA of the cluster
    __shared__ ChainBc bc[2];                                    // alternating: one __syncthreads per decision
    const int tid = threadIdx.x, lane = tid & 31;
    const bool w0 = tid < 32;
    const int64_t TT = (int64_t)C * SEQ_THREADS, gt = (int64_t)rank * SEQ_THREADS + tid;
    const bool leader = rank == 0 && tid == 0;
    // this CTA's copy of the chain, touched by its thread 0 only: element, D[element, previous element], size
    int32_t *chain = chains + (int64_t)rank * n;
    unsigned long long *chain_d = chains_d + (int64_t)rank * n;
    int32_t *chain_s = chains_s + (int64_t)rank * n;

    const int64_t live_words = (pitch + 31) / 32;
    for (int64_t w = tid; w < live_words; w += SEQ_THREADS) {
        const int64_t j = w * 32;
        s_live[w] = j + 32 <= n ? 0xFFFFFFFFu : (j >= n ? 0u : ((1u << (int)(n - j)) - 1u));
    }
    for (int64_t i = gt; i < n; i += TT) nn_store(&nn[i], ~0ull, ~0u, ~0ull, ~0u, 0u);      // nothing known yet
    __syncthreads();
    cluster.sync();

    // walk state: warp 0 only (uniform over its lanes)
    int64_t len = 0, first_live = 0, hits = 0, scans = 0, fused = 0;
    int32_t x = -1, prev = -1, sx = 0, sprev = 0;
    unsigned long long dprev = ~0ull;                            // D[x, prev]
    bool need_sx = false;
    // rows whose entry / size was produced after the last barrier (not yet visible in global memory)
    int32_t dead = -1, grown = -1, grown_size = 0, known = -1;
    unsigned long long gw1 = ~0ull, kw1 = ~0ull;
    unsigned int gi1 = ~0u, ki1 = ~0u;
    int32_t nx = -1, nprev = -1, nsx = 0, nsprev = 0;            // the chain below the tip, prefetched
    unsigned long long ndprev = ~0ull;
    // all threads
    int64_t step = 0;
    int q = 0;
    long long clk[8] = {0, 0, 0, 0, 0, 0, 0, 0}, t0 = clock64();
#define NNC_LAP(slot_) do { const long long t1_ = clock64(); clk[slot_] += t1_ - t0; t0 = t1_; } while (0)

    for (int64_t k = 0; k < n - 1;) {
        if (w0) {
            for (;;) {
                if (len == 0) {                                  // lowest live index from the bitmap
                    int64_t w = first_live >> 5;
                    for (;;) {
                        const int64_t ww = w + lane;
                        const uint32_t bits = ww < live_words ? s_live[ww] : 0u;
                        const unsigned m = __ballot_sync(0xffffffffu, bits != 0u);
                        if (m) {
                            const int src = __ffs(m) - 1;
                            const uint32_t b0 = __shfl_sync(0xffffffffu, bits, src);
                            first_live = (w + src) * 32 + (__ffs(b0) - 1);
                            break;
                        }
                        w += 32;
                    }
                    x = (int32_t)first_live; prev = -1; len = 1; dprev = ~0ull; need_sx = true;
                    if (lane == 0) chain[0] = x;
                }
                // look-up: independent loads in separate lanes (entry, size, and the chain below the tip in case
                // this step ends in a merge)
                // (the chain arrays are private to lane 0, which wrote them: program order keeps them coherent)
                unsigned long long v = 0, c3d = ~0ull;
                int32_t c3 = -1, c3s = 0, c4 = -1, c4s = 0;
                if (lane == 0) {
                    v = x == grown ? gw1 : (x == known ? kw1 : __ldcg(&nn[x].w1));
                    if (len >= 3) { c3 = chain[len - 3]; c3s = chain_s[len - 3]; }
                    if (len >= 4) { c4 = chain[len - 4]; c4s = chain_s[len - 4]; c3d = chain_d[len - 3]; }
                } else if (lane == 1) {
                    v = x == grown ? (unsigned long long)gi1 : (x == known ? (unsigned long long)ki1 : (unsigned long long)__ldcg(&nn[x].i1));
                } else if (lane == 2 && need_sx) {
                    v = (unsigned long long)(x == dead ? 0 : (x == grown ? grown_size : __ldcg(&size[x])));
                }
                const unsigned long long cw = __shfl_sync(0xffffffffu, v, 0);
                const unsigned int ci = (unsigned int)__shfl_sync(0xffffffffu, v, 1);
                if (need_sx) {
                    sx = (int32_t)__shfl_sync(0xffffffffu, v, 2);
                    if (lane == 0) chain_s[len - 1] = sx;
                    need_sx = false;
                }
                nx = __shfl_sync(0xffffffffu, c3, 0); nsx = __shfl_sync(0xffffffffu, c3s, 0);
                nprev = __shfl_sync(0xffffffffu, c4, 0); nsprev = __shfl_sync(0xffffffffu, c4s, 0);
                ndprev = __shfl_sync(0xffffffffu, c3d, 0);
                RTHC_DEV_ASSERT(x >= 0 && x < n && ((s_live[x >> 5] >> (x & 31)) & 1u));
                if (ci == ~0u) {                                 // unknown: scan row x over the cluster
                    if (lane == 0) { bc[q].what = NNC_SCAN; bc[q].x = x; }
                    break;
                }
                ++hits;
                RTHC_DEV_ASSERT(ci < (unsigned int)n && ci != (unsigned int)x && ((s_live[ci >> 5] >> (ci & 31)) & 1u));
                RTHC_DEV_ASSERT(sx > 0 && (len == 1 || sprev > 0));
                int32_t y = (int32_t)ci;
                unsigned long long dmin = cw;
                if (len > 1 && dprev <= dmin) { y = prev; dmin = dprev; }      // previous chain element wins ties
                if (len > 1 && y == prev) {
                    // ---- reciprocal nearest neighbours: the merge decision -------------------------------
                    if (lane == 0) {
                        ChainBc &o = bc[q];
                        const bool xs = x < y;                   // a dies, b survives (scipy: x < y swap)
                        o.what = NNC_MERGE;
                        o.a = xs ? x : y; o.b = xs ? y : x; o.na = xs ? sx : sprev; o.nb = xs ? sprev : sx;
                        o.dmin = dmin; o.p = nx;
                    }
                    break;
                }
                // ---- extend the chain -----------------------------------------------------------------
                RTHC_DEV_ASSERT(len < n);
                if (lane == 0) { chain[len] = y; chain_d[len] = dmin; }
                prev = x; sprev = sx; x = y; dprev = dmin; ++len; need_sx = true;
            }
        }
        __syncthreads();
        NNC_LAP(0);
        const int what = bc[q].what;
        const int par = (int)(step & 1);
        if (what == NNC_SCAN) {
            // ---- scan row x: the two smallest (D[x, i], i) over live i != x ---------------------------
            const int32_t xr = bc[q].x;
            q ^= 1;
            const T *row = W + (int64_t)xr * pitch;
            Top2 t;
            t.clear();
            for (int64_t i0 = gt; i0 < n; i0 += TT * 8) {        // up to 8 elements of this thread in flight
                T d[8];
                bool ok[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) {
                    const int64_t i = i0 + u * TT;
                    ok[u] = i < n && i != xr && ((s_live[i >> 5] >> (i & 31)) & 1u);
                    if (ok[u]) d[u] = __ldcg(&row[i]);
                }
#pragma unroll
                for (int u = 0; u < 8; ++u)
                    if (ok[u]) t.insert(order_bits(d[u]), (unsigned int)(i0 + u * TT));
            }
            block_top2(t, sh);
            if (w0 && lane < C) *cluster.map_shared_rank(&slotA[par][rank], lane) = t;      // lane l publishes to CTA l
            NNC_LAP(1);
            cluster.sync();
            ++step;
            if (w0) {
                if (lane < C) t = slotA[par][lane]; else t.clear();
                warp_top2(t);
                known = xr; kw1 = t.w1; ki1 = t.i1;
                if (leader) nn_store(&nn[xr], t.w1, t.i1, t.w2, t.i2, t.i2 != ~0u);
                dead = -1;                                       // the barrier published the last merge's sizes
                ++scans;
            }
            NNC_LAP(2);
            continue;
        }
        // ---- merge a and b ----------------------------------------------------------------------------
        const int32_t a = bc[q].a, b = bc[q].b, na = bc[q].na, nb = bc[q].nb, p = bc[q].p;
        const unsigned long long dminb = bc[q].dmin;
        q ^= 1;
        if (tid == 0) s_live[a >> 5] &= ~(1u << (a & 31));
        if (leader) { ex[k] = a; ey[k] = b; ew[k] = from_bits(dminb, T()); }
        cluster.sync();                       // every CTA has left the walk; the leader's earlier stores are visible
        NNC_LAP(3);
        const T *ra = W + (int64_t)a * pitch;
        T *rb = W + (int64_t)b * pitch;
        const T *rp = W + (int64_t)(p >= 0 ? p : 0) * pitch;
        const double dab = from_bits(dminb, T());
        Top2 tb, tp;
        tb.clear(); tp.clear();
        constexpr int PB = sizeof(T) == 8 ? 2 : 3;               // elements of this thread in flight (register budget)
        for (int64_t i0 = gt; i0 < n; i0 += TT * PB) {
            T da[PB], db[PB], dp[PB];
            uint4 elo[PB], ehi[PB];
            int32_t szi[PB];
            bool ok[PB];
#pragma unroll
            for (int u = 0; u < PB; ++u) {
                const int64_t i = i0 + u * TT;
                ok[u] = i < n && i != b && ((s_live[i >> 5] >> (i & 31)) & 1u);
                if (ok[u]) {
                    da[u] = __ldcg(&ra[i]); db[u] = __ldcg(&rb[i]);
                    elo[u] = __ldcg(reinterpret_cast<const uint4 *>(&nn[i])); ehi[u] = __ldcg(reinterpret_cast<const uint4 *>(&nn[i]) + 1);
                    szi[u] = __ldcg(&size[i]);
                    if (p >= 0) dp[u] = __ldcg(&rp[i]);
                }
            }
#pragma unroll
            for (int u = 0; u < PB; ++u) {
                if (!ok[u]) continue;
                const int64_t i = i0 + u * TT;
                const T v = lance_williams<T>(method, da[u], db[u], dab, na, nb, szi[u]);
                rb[i] = v;                       // row b (contiguous)
                W[i * pitch + b] = v;            // column b (symmetric copy)
                const unsigned long long ob = order_bits(v);
                tb.insert(ob, (unsigned int)i);
                if (i == p) { tp.insert(ob, (unsigned int)b); continue; }       // row p is rescanned as a whole
                if (p >= 0) tp.insert(order_bits(dp[u]), (unsigned int)i);
                const unsigned long long w1 = ((unsigned long long)elo[u].y << 32) | elo[u].x;
                const unsigned long long w2 = ((unsigned long long)elo[u].w << 32) | elo[u].z;
                const unsigned int i1 = ehi[u].x, i2 = ehi[u].y, act = ehi[u].z;
                if (i1 == ~0u) continue;                                         // unknown stays unknown
                // candidates: N = (ob, b), E1 unless it was a or b, E2 if known and not a or b; bound = (w2, i2)
                const bool h1 = i1 == (unsigned int)a || i1 == (unsigned int)b;
                const bool k2 = act && !(i2 == (unsigned int)a || i2 == (unsigned int)b);
                Top2 c;
                c.clear();
                c.insert(ob, (unsigned int)b);
                if (!h1) c.insert(w1, i1);
                if (k2) c.insert(w2, i2);                                        // only the two smallest matter
                if (!slot_less(w2, i2, c.w1, c.i1)) {                            // c1 <= bound: exact
                    const bool second = c.i2 != ~0u && !slot_less(w2, i2, c.w2, c.i2);
                    nn_store(&nn[i], c.w1, c.i1, second ? c.w2 : w2, second ? c.i2 : i2, second ? 1u : 0u);
                } else {
                    nn_store(&nn[i], ~0ull, ~0u, ~0ull, ~0u, 0u);
                }
            }
        }
        NNC_LAP(4);
        block_top2(tb, sh);
        if (w0 && lane < C) *cluster.map_shared_rank(&slotA[par][rank], lane) = tb;
        if (p >= 0) {
            block_top2(tp, sh);
            if (w0 && lane < C) *cluster.map_shared_rank(&slotB[par][rank], lane) = tp;
        }
        NNC_LAP(5);
        cluster.sync();                       // row / column b, the cache entries and the slots are visible
        NNC_LAP(6);
        ++step;
        if (w0) {
            Top2 t;
            if (lane < C) t = slotA[par][lane]; else t.clear();
            warp_top2(t);
            gw1 = t.w1; gi1 = t.i1;
            if (leader) { nn_store(&nn[b], t.w1, t.i1, t.w2, t.i2, t.i2 != ~0u); size[a] = 0; size[b] = na + nb; }
            dead = a; grown = b; grown_size = na + nb;
            known = -1;
            if (p >= 0) {
                if (lane < C) t = slotB[par][lane]; else t.clear();
                warp_top2(t);
                known = p; kw1 = t.w1; ki1 = t.i1;
                if (leader) nn_store(&nn[p], t.w1, t.i1, t.w2, t.i2, t.i2 != ~0u);
                ++fused;
            }
            len -= 2;
            if (len > 0) { x = nx; prev = nprev; sx = nsx; sprev = nsprev; dprev = ndprev; need_sx = false; }
        }
        NNC_LAP(7);
        ++k;
    }
#undef NNC_LAP
    if (leader) {
        ctl->scans = scans; ctl->fused = fused; ctl->hits = hits;
        for (int c = 0; c < 8; ++c) ctl->clk[c] = clk[c];
    }
}

template <typename T, typename S>
static int run_nn_chain(rthc_ctx *ctx, const S *d_cond, int method, std::vector<int32_t> &ex,
                        std::vector<int32_t> &ey, std::vector<double> &ew)
{
    const int64_t n = ctx->n, m = n - 1;
    const int64_t per16 = 16 / (int64_t)sizeof(T), pitch = (n + per16 - 1) / per16 * per16;
    T *d_W = nullptr;
    int32_t *d_size = nullptr, *d_chain = nullptr, *d_chain_s = nullptr, *d_ex = nullptr, *d_ey = nullptr;
    unsigned long long *d_chain_d = nullptr;
    double *d_ew = nullptr;
    NNEntry *d_nn = nullptr;
    ChainCtl *d_ctl = nullptr;
    int rc;
    const size_t smem = sizeof(uint32_t) * (size_t)((pitch + 31) / 32 + 1);
    if (smem > 160u * 1024u) return rthc_fail(ctx, RTHC_E_RANGE, "merge loop: %lld samples exceed the liveness bitmap", (long long)n);
    if ((rc = rthc_dev_alloc(ctx, &d_W, (size_t)(n * pitch)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_size, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_chain, (size_t)n * NNC_MAX_CLUSTER))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_chain_s, (size_t)n * NNC_MAX_CLUSTER))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_chain_d, (size_t)n * NNC_MAX_CLUSTER))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_nn, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ex, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ey, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ew, (size_t)m))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_ctl, 1))) return rc;
    {
        dim3 grd((unsigned)n, (unsigned)((pitch + 255) / 256));
        rthc_expand_square<T, S><<<grd, 256, 0, ctx->stream>>>(d_cond, n, pitch, d_W);
        RTHC_LAUNCH_CHECK(ctx);
        std::vector<int32_t> ones((size_t)n, 1);
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_size, ones.data(), sizeof(int32_t) * (size_t)n, cudaMemcpyHostToDevice, ctx->stream));
        RTHC_CUDA(ctx, cudaMemsetAsync(d_ctl, 0, sizeof(ChainCtl), ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    }
    auto kern = rthc_nn_chain_kernel<T>;
    RTHC_CUDA(ctx, cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    cudaError_t ce = cudaErrorInvalidConfiguration;
    for (int C = NNC_MAX_CLUSTER; C >= 1 && ce != cudaSuccess; C >>= 1) {       // one cluster of 16, 8, ... CTAs
        if (C > 8 && cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1) != cudaSuccess) {
            cudaGetLastError();
            continue;
        }
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned)C);
        cfg.blockDim = dim3(SEQ_THREADS);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = ctx->stream;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = (unsigned)C; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr; cfg.numAttrs = 1;
        int clusters = 0;
        ce = cudaOccupancyMaxActiveClusters(&clusters, kern, &cfg);
        if (ce == cudaSuccess && clusters < 1) ce = cudaErrorInvalidConfiguration;
        if (ce == cudaSuccess)
            ce = cudaLaunchKernelEx(&cfg, kern, d_W, n, pitch, method, d_size, d_chain, d_chain_d, d_chain_s, d_nn, d_ex, d_ey, d_ew, d_ctl);
        if (ce != cudaSuccess) cudaGetLastError();
    }
    if (ce != cudaSuccess) return rthc_fail(ctx, RTHC_E_CUDA, "merge loop: no thread-block cluster could be launched (%s)", cudaGetErrorString(ce));
    ctx->launches++;
    ex.resize((size_t)m); ey.resize((size_t)m); ew.resize((size_t)m);
    ChainCtl ctl;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ex.data(), d_ex, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ey.data(), d_ey, sizeof(int32_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ew.data(), d_ew, sizeof(double) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(&ctl, d_ctl, sizeof(ChainCtl), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (getenv("RTHC_NNC_TIMING")) {                 // section cycles of CTA 0 / thread 0
        static const char *name[8] = {"walk", "scan: loads + CTA top-2", "scan: barrier + combine", "merge: pre-barrier",
                                      "merge: pass", "merge: CTA top-2 x2", "merge: barrier", "merge: combine"};
        long long tot = 0;
        for (int c = 0; c < 8; ++c) tot += ctl.clk[c];
        for (int c = 0; c < 8; ++c)
            fprintf(stderr, "nn_chain %-26s %8.1f Mcycles %5.1f %%\n", name[c], ctl.clk[c] * 1e-6, 100.0 * ctl.clk[c] / (double)(tot ? tot : 1));
        fprintf(stderr, "nn_chain scans %lld fused %lld merges %lld look-ups %lld\n", (long long)ctl.scans, (long long)ctl.fused,
                (long long)m, (long long)ctl.hits);
    }
    ctx->row_scans = ctl.scans + ctl.fused;         // row scans: own barrier + fused with a merge
    ctx->chain_full_scans = ctl.scans + 2 * m;      // cluster barriers (info[9]): one per scan, two per merge
    ctx->chain_cache_hits = ctl.hits;
    rthc_dev_free(ctx, &d_W); rthc_dev_free(ctx, &d_size); rthc_dev_free(ctx, &d_chain);
    rthc_dev_free(ctx, &d_chain_s); rthc_dev_free(ctx, &d_chain_d); rthc_dev_free(ctx, &d_nn);
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
