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
// form one pipeline stage; stages are filled by TMA bulk copies (cp.async.bulk, 512 B
// per plane row) completing on mbarriers, RTHC_STAGES deep, so loads of groups g+1..g+3
// overlap the LOP3 stream of group g.  Results go straight to the condensed uint16
// upper triangle (scipy pdist order) -- the square matrix is never materialised.
#include "rthc_internal.cuh"

#define RTHC_STAGES 4
#define DIST_THREADS 256

__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    uint32_t done = 0;
    const uint32_t addr = smem_u32(bar);
    while (!done) {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
    }
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                             uint64_t *bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
            smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// stage layout in shared memory: [max_np][128] words of the I block, then the same for J
__global__ void __launch_bounds__(DIST_THREADS, 1)
rthc_distance_bitsliced(const uint32_t *__restrict__ planes, const int32_t *__restrict__ grp_off,
                        const int32_t *__restrict__ grp_np, int G, int max_np, int64_t n_pad,
                        const int2 *__restrict__ tiles, const int64_t *__restrict__ row_base,
                        int64_t n, uint16_t *__restrict__ cond)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[RTHC_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[RTHC_STAGES];

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const int2 tile = tiles[blockIdx.x];
    const int64_t i0 = (int64_t)tile.x * RTHC_TILE, j0 = (int64_t)tile.y * RTHC_TILE;
    const uint32_t stage_words = (uint32_t)max_np * 2u * RTHC_TILE;
    uint32_t *stage_base = reinterpret_cast<uint32_t *>(smem_raw);

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < RTHC_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], DIST_THREADS / 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    auto issue = [&](int g) {   // called by thread 0 only
        const int s = g % RTHC_STAGES;
        const int np = grp_np[g];
        const int64_t off = grp_off[g];
        uint32_t *dstA = stage_base + (size_t)s * stage_words;
        uint32_t *dstB = dstA + (size_t)max_np * RTHC_TILE;
        mbar_expect_tx(&full_bar[s], (uint32_t)np * 2u * RTHC_TILE * 4u);
        for (int k = 0; k < np; ++k) {
            const uint32_t *row = planes + (off + k) * n_pad;
            tma_bulk_g2s(dstA + k * RTHC_TILE, row + i0, RTHC_TILE * 4, &full_bar[s]);
            tma_bulk_g2s(dstB + k * RTHC_TILE, row + j0, RTHC_TILE * 4, &full_bar[s]);
        }
    };

    if (tid == 0) {
        const int pre = G < RTHC_STAGES ? G : RTHC_STAGES;
        for (int g = 0; g < pre; ++g) issue(g);
    }

    uint32_t acc[8][8];
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) acc[a][b] = 0;

    for (int g = 0; g < G; ++g) {
        const int s = g % RTHC_STAGES;
        const uint32_t parity = (uint32_t)(g / RTHC_STAGES) & 1u;
        // refill the stage released one iteration ago (keeps thread 0 off the critical path)
        if (tid == 0 && g >= 1 && g - 1 + RTHC_STAGES < G) {
            const int gp = g - 1;
            mbar_wait(&empty_bar[gp % RTHC_STAGES], (uint32_t)(gp / RTHC_STAGES) & 1u);
            issue(gp + RTHC_STAGES);
        }
        __syncwarp();
        mbar_wait(&full_bar[s], parity);

        const int np = grp_np[g];
        const uint32_t *sA = stage_base + (size_t)s * stage_words;
        const uint32_t *sB = sA + (size_t)max_np * RTHC_TILE;
        const uint4 *pa = reinterpret_cast<const uint4 *>(sA) + ty;   // samples 4ty.. and 64+4ty..
        const uint4 *pb = reinterpret_cast<const uint4 *>(sB) + tx;

        uint32_t m[8][8];
        {
            uint4 a0 = pa[0], a1 = pa[16], b0 = pb[0], b1 = pb[16];
            const uint32_t av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const uint32_t bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int a = 0; a < 8; ++a)
#pragma unroll
                for (int b = 0; b < 8; ++b) m[a][b] = av[a] ^ bv[b];
        }
        for (int k = 1; k < np - 1; ++k) {
            const uint4 *qa = pa + k * (RTHC_TILE / 4), *qb = pb + k * (RTHC_TILE / 4);
            uint4 a0 = qa[0], a1 = qa[16], b0 = qb[0], b1 = qb[16];
            const uint32_t av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const uint32_t bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int a = 0; a < 8; ++a)
#pragma unroll
                for (int b = 0; b < 8; ++b) m[a][b] |= av[a] ^ bv[b];
        }
        {
            const int k = np - 1;   // valid plane
            const uint4 *qa = pa + k * (RTHC_TILE / 4), *qb = pb + k * (RTHC_TILE / 4);
            uint4 a0 = qa[0], a1 = qa[16], b0 = qb[0], b1 = qb[16];
            const uint32_t av[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const uint32_t bv[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int a = 0; a < 8; ++a)
#pragma unroll
                for (int b = 0; b < 8; ++b) acc[a][b] += __popc(m[a][b] & av[a] & bv[b]);
        }
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty_bar[s]);
    }

    // epilogue: condensed upper triangle, scipy order: row_base[i] + (j - i - 1)
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        const int64_t i = i0 + (a < 4 ? 4 * ty + a : 64 + 4 * ty + (a - 4));
        if (i >= n) continue;
        const int64_t base = row_base[i];
        if (base < 0) continue;
#pragma unroll
        for (int b = 0; b < 8; ++b) {
            const int64_t j = j0 + (b < 4 ? 4 * tx + b : 64 + 4 * tx + (b - 4));
            if (j > i && j < n) cond[base + (j - i - 1)] = (uint16_t)acc[a][b];
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

int rthc_run_distance(rthc_ctx *ctx)
{
    if (ctx->n_tiles == 0) return 0;
    const int max_np = ctx->max_np > 0 ? ctx->max_np : 1;
    const size_t smem = (size_t)RTHC_STAGES * (size_t)max_np * 2 * RTHC_TILE * sizeof(uint32_t);
    if (smem > 200 * 1024)
        return rthc_fail(ctx, RTHC_E_RANGE, "plane stage does not fit shared memory (max_np=%d)", max_np);
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_distance_bitsliced, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)smem));
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    rthc_distance_bitsliced<<<(unsigned)ctx->n_tiles, DIST_THREADS, smem, ctx->stream>>>(
        ctx->d_planes, ctx->d_grp_off, ctx->d_grp_np, ctx->G, max_np, ctx->n_pad, ctx->d_tiles,
        ctx->d_row_base, ctx->n, ctx->d_cond);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[2] = ms;
    return 0;
}
