// distance.cu -- pairwise allelic distance on sm_100a: formulation, stage driver, check kernel.
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
// The kernel is rthc_distance_ws (distance_ws.cu): persistent, warp-specialised, 128 x 128 sample
// tile per CTA iteration, 8 x 8 pairs per thread, TMA bulk copies + mbarrier ring.  The earlier
// one-tile-per-CTA kernel (v4, 72-74 % of the LOP3 peak) and its 128 x 64 variant were removed after
// their measurements went into DESIGN.md section 7.  Results go straight to the condensed uint16 upper
// triangle (scipy pdist order) -- the square matrix is never materialised.
#include "dist_common.cuh"

// Independent scalar kernel on the re-coded alleles (uint16, or uint32 when a locus has more than 65,535
// alleles): one warp per requested pair.  Not on the product path -- the parity tests use it as a second
// opinion on full-size matrices where the CPU reference only covers a sample.
// It reads the dense ranks (ctx->d_codes), not the shared codes the plane packer gets from collide.cu.
template <typename CT>
__global__ void rthc_direct_pairs_kernel(const CT *__restrict__ codes, int64_t L, int64_t n_pad,
                                         const int32_t *__restrict__ pi, const int32_t *__restrict__ pj,
                                         int64_t npairs, int32_t *__restrict__ out)
{
    const int64_t p = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int lane = threadIdx.x & 31;
    if (p >= npairs) return;
    const int64_t i = pi[p], j = pj[p];
    int d = 0;
    for (int64_t l = lane; l < L; l += 32) {
        const CT a = codes[l * n_pad + i], b = codes[l * n_pad + j];
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
    if (ctx->d_codes32)
        rthc_direct_pairs_kernel<uint32_t><<<(unsigned)blocks, threads, 0, ctx->stream>>>(ctx->d_codes32, ctx->L, ctx->n_pad,
                                                                                          d_pi, d_pj, np, d_out);
    else
        rthc_direct_pairs_kernel<uint16_t><<<(unsigned)blocks, threads, 0, ctx->stream>>>(ctx->d_codes, ctx->L, ctx->n_pad,
                                                                                          d_pi, d_pj, np, d_out);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

int rthc_run_distance(rthc_ctx *ctx)
{
    if (ctx->n_tiles == 0) return 0;
    const int max_np = ctx->max_np > 0 ? ctx->max_np : 1;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    int rc = rthc_launch_distance_ws(ctx, max_np);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    // shared codes: the plane kernel saw two alleles of one code as equal; add those mismatches back (collide.cu)
    if ((rc = rthc_launch_escape_correct(ctx))) return rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev_aux, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[2] = ms;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev1, ctx->ev_aux));
    ctx->t_ms[7] = ms;
    return 0;
}
