// distance.cu -- pairwise allelic distance on sm_100a.
//
// Replaces the O(n^2 L) loop of the external `cgmlst-dists` binary and ReporTree's
// 2,000-column chunk/sum protocol around it (scripts/partitioning_HC.py:489-516, call at
// :502):   d(i,j) = #{ l : a_il != a_jl  and  a_il != 0  and  a_jl != 0 }.
//
// Bit-sliced formulation (planes built by recode.cu): for a group of 32 loci with b bit
// planes P_k and a valid plane V,
//     mism = OR_k (Pa_k ^ Pb_k)          one LOP3 per plane per 32 loci per pair
//     d   += popc(mism & Va & Vb)        one LOP3 + POPC + IADD per group per pair
// i.e. (b + 1)/32 ALU lane-ops per allele comparison instead of >= 2 for a direct
// compare-and-accumulate.  No tensor cores: integer compare + popcount only.
//
// Tiling: one CTA computes a 128 x 128 sample tile of the upper triangle with 256
// threads, each holding an 8 x 8 block of pairs in registers (64 accumulators + 64
// running mismatch words).  The plane rows of the 128 + 128 samples of one locus group
// form one pipeline stage; a stage is filled by two TMA bulk copies (cp.async.bulk, one per
// sample block, planes are stored tile-blocked so each is one contiguous run) completing on
// an mbarrier, RTHC_STAGES deep, so loads of groups g+1..g+3 overlap the LOP3 stream of
// group g; operand words of plane k+1 are fetched from shared memory while plane k is combined.  Results go straight to the condensed uint16
// upper triangle (scipy pdist order) -- the square matrix is never materialised.
#include <cstdlib>
#include <cstring>

#include "dist_common.cuh"

#define DIST_THREADS 256

// Plane layout in HBM is blocked by 64 samples: planes[((s / 64) * plane_rows + row) * 64 + s % 64],
// so the np plane rows of one locus group of one 64-sample block are one contiguous run of
// np * 256 bytes = ONE bulk copy.  A stage in shared memory holds 2 + NJ such blocks
// ([max_np][64] words each): the two halves of the 128-sample I block, then NJ J blocks.
//
// NJ = 2: 128 x 128 tile, 8 x 8 pairs per thread, 1 CTA (8 warps) per SM.
// NJ = 1: 128 x  64 tile, 8 x 4 pairs per thread, <= 128 registers, 2 CTAs (16 warps) per SM:
//         the second CTA's LOP3 stream covers the POPC tail and barrier waits of the first.
template <int NJ> struct PlaneOps {
    uint32_t a[8], b[4 * NJ];
};
template <int NJ>
__device__ __forceinline__ PlaneOps<NJ> load_ops(const uint32_t *st, int blk_words, int ty, int tx, int k)
{
    PlaneOps<NJ> o;
#ifdef RTHC_EXPERIMENT_NOLDS
    {
#pragma unroll
        for (int q = 0; q < 8; ++q) o.a[q] = (uint32_t)(k * 7 + ty + q);
#pragma unroll
        for (int q = 0; q < 4 * NJ; ++q) o.b[q] = (uint32_t)(k * 13 + tx + q) ^ st[0];
        return o;
    }
#endif
    const uint4 a0 = *reinterpret_cast<const uint4 *>(st + k * RTHC_BLK + 4 * ty);
    const uint4 a1 = *reinterpret_cast<const uint4 *>(st + blk_words + k * RTHC_BLK + 4 * ty);
    o.a[0] = a0.x; o.a[1] = a0.y; o.a[2] = a0.z; o.a[3] = a0.w;
    o.a[4] = a1.x; o.a[5] = a1.y; o.a[6] = a1.z; o.a[7] = a1.w;
#pragma unroll
    for (int h = 0; h < NJ; ++h) {
        const uint4 b0 = *reinterpret_cast<const uint4 *>(st + (2 + h) * blk_words + k * RTHC_BLK + 4 * tx);
        o.b[4 * h + 0] = b0.x; o.b[4 * h + 1] = b0.y; o.b[4 * h + 2] = b0.z; o.b[4 * h + 3] = b0.w;
    }
    return o;
}

template <int NJ>
__global__ void __launch_bounds__(DIST_THREADS, NJ == 1 ? 2 : 1)
rthc_distance_bitsliced(const uint32_t *__restrict__ planes, const int32_t *__restrict__ grp_off,
                        const int32_t *__restrict__ grp_np, int G, int max_np, int plane_rows,
                        const int2 *__restrict__ tiles, const int64_t *__restrict__ row_base,
                        int64_t n, uint16_t *__restrict__ cond)
{
    constexpr int NB = 2 + NJ;          // 64-sample blocks per stage
    constexpr int TJ = 4 * NJ;          // pairs per thread along j
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[RTHC_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[RTHC_STAGES];

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int2 tile = tiles[blockIdx.x];          // x: 128-sample row block, y: first 64-sample column block
    const int64_t i0 = (int64_t)tile.x * RTHC_TILE, j0 = (int64_t)tile.y * RTHC_BLK;
    const int blk_words = max_np * RTHC_BLK;
    const uint32_t stage_words = (uint32_t)blk_words * NB;
    uint32_t *stage_base = reinterpret_cast<uint32_t *>(smem_raw);
    // group table (offset << 8 | np) behind the stages
    uint32_t *gtab = stage_base + (size_t)RTHC_STAGES * stage_words;
    for (int g = tid; g < G; g += DIST_THREADS) gtab[g] = ((uint32_t)grp_off[g] << 8) | (uint32_t)grp_np[g];

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < RTHC_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], DIST_THREADS / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const size_t blk_stride = (size_t)plane_rows * RTHC_BLK;      // words per 64-sample block in HBM
    auto issue = [&](int g) {   // thread 0 only: one bulk copy per 64-sample block
        const int s = g % RTHC_STAGES;
        const uint32_t e = gtab[g];
        const uint32_t np = e & 0xFFu, off = e >> 8;
        const uint32_t bytes = np * RTHC_BLK * 4u;
        uint32_t *dst = stage_base + (size_t)s * stage_words;
        mbar_expect_tx(&full_bar[s], NB * bytes);
        tma_bulk_g2s(dst, planes + (size_t)(2 * tile.x) * blk_stride + (size_t)off * RTHC_BLK, bytes, &full_bar[s]);
        tma_bulk_g2s(dst + blk_words, planes + (size_t)(2 * tile.x + 1) * blk_stride + (size_t)off * RTHC_BLK,
                     bytes, &full_bar[s]);
#pragma unroll
        for (int h = 0; h < NJ; ++h)
            tma_bulk_g2s(dst + (2 + h) * blk_words,
                         planes + (size_t)(tile.y + h) * blk_stride + (size_t)off * RTHC_BLK, bytes, &full_bar[s]);
    };

    if (tid == 0) {
        const int pre = G < RTHC_STAGES ? G : RTHC_STAGES;
        for (int g = 0; g < pre; ++g) issue(g);
    }

    // Accumulators are packed two 16-bit counters per register (distances <= L <= 65535 cannot
    // carry across): pair (a, b) lives in accp[a][b % 4], low half for b < 4, high half otherwise
    // (NJ = 1 uses the low half only).  mm[][] holds the masked mismatch words of the PREVIOUS
    // group.  Their POPCs run on the XU pipe (16 lanes/clk/SM), so they are issued one slice (two
    // rows) per peeled plane block to overlap the LOP3 stream (ALU pipe) of the current group, and
    // each slice's counts pc[] are folded into accp one block LATER (IMAD on the FMA pipe), so no
    // instruction ever waits on a POPC that was just issued.
    uint32_t accp[8][4];
    uint32_t mm[8][TJ];
    uint32_t pc[2 * TJ];
#pragma unroll
    for (int a = 0; a < 8; ++a) {
#pragma unroll
        for (int b = 0; b < 4; ++b) accp[a][b] = 0;
#pragma unroll
        for (int b = 0; b < TJ; ++b) mm[a][b] = 0;
    }
#pragma unroll
    for (int q = 0; q < 2 * TJ; ++q) pc[q] = 0;

#ifdef RTHC_EXPERIMENT_NOPOPC
#define RTHC_POPC(x) (x)
#else
#define RTHC_POPC(x) ((uint32_t)__popc(x))
#endif
#define RTHC_POPC_SLICE(sl)                                                        \
    _Pragma("unroll") for (int r_ = 0; r_ < 2; ++r_)                               \
        _Pragma("unroll") for (int b_ = 0; b_ < TJ; ++b_)                          \
            pc[r_ * TJ + b_] = RTHC_POPC(mm[2 * (sl) + r_][b_]);
#define RTHC_ACC_SLICE(sl)                                                         \
    _Pragma("unroll") for (int r_ = 0; r_ < 2; ++r_)                               \
        _Pragma("unroll") for (int b_ = 0; b_ < TJ; ++b_)                          \
            accp[2 * (sl) + r_][b_ & 3] += (b_ < 4) ? pc[r_ * TJ + b_] : pc[r_ * TJ + b_] * 65536u;
#define RTHC_COMBINE(op_, P_)                                                      \
    _Pragma("unroll") for (int a_ = 0; a_ < 8; ++a_)                               \
        _Pragma("unroll") for (int b_ = 0; b_ < TJ; ++b_)                          \
            m[a_][b_] op_ (P_).a[a_] ^ (P_).b[b_];

    PlaneOps<NJ> cur;
    if (G > 0) {
        mbar_wait(&full_bar[0], 0);
        cur = load_ops<NJ>(stage_base, blk_words, ty, tx, 0);
    }
    for (int g = 0; g < G; ++g) {
        const int s = g % RTHC_STAGES;
        // refill the stage released one iteration ago (keeps thread 0 off the critical path)
        if (tid == 0 && g >= 1 && g - 1 + RTHC_STAGES < G) {
            const int gp = g - 1;
            mbar_wait(&empty_bar[gp % RTHC_STAGES], (uint32_t)(gp / RTHC_STAGES) & 1u);
            issue(gp + RTHC_STAGES);
        }
        __syncwarp();
        const int np = (int)(gtab[g] & 0xFFu);      // >= RTHC_MIN_BITS + 1
        const uint32_t *st = stage_base + (size_t)s * stage_words;

        uint32_t m[8][TJ];
        // block 0: bit plane 0 (XOR); fold slice 3 (issued in the previous group), issue slice 0
        PlaneOps<NJ> nxt = load_ops<NJ>(st, blk_words, ty, tx, 1);
        RTHC_ACC_SLICE(3)
        RTHC_POPC_SLICE(0)
        RTHC_COMBINE(=, cur)
        cur = nxt;
        // block 1: bit plane 1; fold slice 0, issue slice 1
        nxt = load_ops<NJ>(st, blk_words, ty, tx, 2);
        RTHC_ACC_SLICE(0)
        RTHC_POPC_SLICE(1)
        RTHC_COMBINE(|=, cur)
        cur = nxt;
        // middle bit planes 2 .. np-3, two per iteration with ping-pong operand registers
        int k = 2;
        for (; k + 1 < np - 2; k += 2) {
            nxt = load_ops<NJ>(st, blk_words, ty, tx, k + 1);
            RTHC_COMBINE(|=, cur)
            cur = load_ops<NJ>(st, blk_words, ty, tx, k + 2);
            RTHC_COMBINE(|=, nxt)
        }
        if (k < np - 2) {
            nxt = load_ops<NJ>(st, blk_words, ty, tx, k + 1);
            RTHC_COMBINE(|=, cur)
            cur = nxt;
        }
        // block 2: last bit plane (np-2); fold slice 1, issue slice 2.  The barrier of the next
        // stage is polled here, a whole plane block before its result is needed.
        nxt = load_ops<NJ>(st, blk_words, ty, tx, np - 1);
        const int s1 = (g + 1) % RTHC_STAGES;
        const uint32_t par1 = (uint32_t)((g + 1) / RTHC_STAGES) & 1u;
        uint32_t ready = 1;
        if (g + 1 < G) ready = mbar_try_wait(&full_bar[s1], par1);
        RTHC_ACC_SLICE(1)
        RTHC_POPC_SLICE(2)
        RTHC_COMBINE(|=, cur)
        cur = nxt;
        // block 3: valid plane; fold slice 2, issue slice 3 (rows 6..7) before mm is overwritten;
        // plane 0 of group g+1 is fetched first so that its latency hides here
        if (g + 1 < G) {
            if (!ready) mbar_wait(&full_bar[s1], par1);
            nxt = load_ops<NJ>(stage_base + (size_t)s1 * stage_words, blk_words, ty, tx, 0);
        }
        RTHC_ACC_SLICE(2)
        RTHC_POPC_SLICE(3)
#pragma unroll
        for (int a = 0; a < 8; ++a)
#pragma unroll
            for (int b = 0; b < TJ; ++b) mm[a][b] = m[a][b] & cur.a[a] & cur.b[b];
        cur = nxt;
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty_bar[s]);
    }
    // drain: slice 3 of the second-to-last group is pending in pc[], mm holds the last group
    RTHC_ACC_SLICE(3)
#pragma unroll
    for (int sl = 0; sl < 4; ++sl) {
        RTHC_POPC_SLICE(sl)
        RTHC_ACC_SLICE(sl)
    }
#undef RTHC_POPC_SLICE
#undef RTHC_ACC_SLICE
#undef RTHC_COMBINE

    // epilogue: condensed upper triangle, scipy order: row_base[i] + (j - i - 1)
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int64_t i = i0 + (a < 4 ? 4 * ty + a : 64 + 4 * ty + (a - 4));
        if (i >= n) continue;
        const int64_t base = row_base[i];
        if (base < 0) continue;
#pragma unroll
        for (int b = 0; b < TJ; ++b) {
            const int64_t j = j0 + (b < 4 ? 4 * tx + b : 64 + 4 * tx + (b - 4));
            const uint32_t d = b < 4 ? (accp[a][b] & 0xFFFFu) : (accp[a][b & 3] >> 16);
            if (j > i && j < n) cond[base + (j - i - 1)] = (uint16_t)d;
        }
    }
}

// Independent scalar kernel on the uint16 codes: one warp per requested pair.
__global__ void rthc_direct_pairs_kernel(const uint16_t *__restrict__ codes, int64_t L, int64_t n_pad,
                                         const int32_t *__restrict__ pi, const int32_t *__restrict__ pj,
                                         int64_t npairs, int32_t *__restrict__ out)
{
    const int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= npairs) return;
    const int64_t i = pi[p], j = pj[p];
    int d = 0;
    for (int64_t l = lane; l < L; l += 32) {
        uint16_t a = codes[l * n_pad + i], b = codes[l * n_pad + j];
        d += (a != b) & (a != 0) & (b != 0);
    }
#pragma unroll
    for (int s = 16; s > 0; s >>= 1) d += __shfl_xor_sync(0xffffffffu, d, s);
    if (lane == 0) out[p] = d;
}

int rthc_run_direct_pairs(rthc_ctx *ctx, const int32_t *d_pi, const int32_t *d_pj, int64_t np,
                          int32_t *d_out)
{
    if (np == 0) return 0;
    const int threads = 256;
    const int64_t blocks = (np * 32 + threads - 1) / threads;
    rthc_direct_pairs_kernel<<<(unsigned)blocks, threads, 0, ctx->stream>>>(ctx->d_codes, ctx->L, ctx->n_pad,
                                                                            d_pi, d_pj, np, d_out);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

template <int NJ> static int launch_distance(rthc_ctx *ctx, int max_np)
{
    const size_t smem = (size_t)RTHC_STAGES * (size_t)max_np * (2 + NJ) * RTHC_BLK * sizeof(uint32_t) +
                        (size_t)(ctx->G + 1) * sizeof(uint32_t);
    if (smem > (NJ == 1 ? 100u : 200u) * 1024u)
        return rthc_fail(ctx, RTHC_E_RANGE, "plane stage does not fit shared memory (max_np=%d)", max_np);
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_distance_bitsliced<NJ>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)smem));
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    rthc_distance_bitsliced<NJ><<<(unsigned)ctx->n_tiles, DIST_THREADS, smem, ctx->stream>>>(
        ctx->d_planes, ctx->d_grp_off, ctx->d_grp_np, ctx->G, max_np, ctx->plane_rows, ctx->d_tiles,
        ctx->d_row_base, ctx->n, ctx->d_cond);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    return 0;
}

int rthc_run_distance(rthc_ctx *ctx)
{
    if (ctx->n_tiles == 0) return 0;
    const int max_np = ctx->max_np > 0 ? ctx->max_np : 1;
    // default: warp-specialised persistent kernel (distance_ws.cu); RTHC_DIST_KERNEL=v4 selects the
    // older one-tile-per-CTA kernel of this file (kept for A/B measurements)
    static const bool use_v4 = getenv("RTHC_DIST_KERNEL") && !strcmp(getenv("RTHC_DIST_KERNEL"), "v4");
    int rc;
    if (use_v4 || ctx->dist_nj == 1) {
        rc = ctx->dist_nj == 1 ? launch_distance<1>(ctx, max_np) : launch_distance<2>(ctx, max_np);
    } else {
        RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
        rc = rthc_launch_distance_ws(ctx, max_np);
        if (!rc) RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    }
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[2] = ms;
    return 0;
}
