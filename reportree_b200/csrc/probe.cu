// probe.cu -- integer-ALU peak microbenchmark (the roofline denominator of the distance kernel).
// MEASURED_PEAKS.json only carries HBM and bf16 numbers; SURVEY.md section 8d asks for the int32
// lane-op rate of the box to be measured live.  Independent LOP3 chains, 16 per thread, at
// full occupancy: the same instruction the bit-sliced inner loop issues.
#include "rthc_internal.cuh"

__global__ void __launch_bounds__(256) rthc_lop3_stream(uint32_t *out, uint32_t x, uint32_t y, int iters)
{
    uint32_t a[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) a[k] = threadIdx.x * 16u + k + blockIdx.x;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k)
            asm volatile("lop3.b32 %0, %0, %1, %2, 0xF6;" : "+r"(a[k]) : "r"(x), "r"(y));   // a | (x ^ y)
    }
    uint32_t s = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) s ^= a[k];
    if (s == 0x12345u) out[0] = s;
}

extern "C" int rthc_alu_peak(rthc_ctx *ctx, double *lane_ops_per_s)
{
    if (!ctx || !lane_ops_per_s) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t *d_out = nullptr;
    int rc = rthc_dev_alloc(ctx, &d_out, 1);
    if (rc) return rc;
    const int iters = 4096, blocks = ctx->sm_count * 8, threads = 256;
    double best = 0;
    for (int rep = 0; rep < 5; ++rep) {
        RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
        rthc_lop3_stream<<<blocks, threads, 0, ctx->stream>>>(d_out, 0x0F0F0F0Fu + rep, 0x33333333u, iters);
        RTHC_LAUNCH_CHECK(ctx);
        RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        float ms = 0;
        RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        double rate = (double)blocks * threads * (double)iters * 16.0 / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    rthc_dev_free(ctx, &d_out);
    *lane_ops_per_s = best;
    return 0;
}

// Same probe with the operand pattern of the distance kernel's inner loop: 64 running words
// m[a][b] |= x[a] ^ y[b] (three distinct register sources per LOP3, 8 x 8 register block).
// This is the ceiling the register file lets that loop reach.
__global__ void __launch_bounds__(256) rthc_lop3_block_stream(uint32_t *out, uint32_t seed, int iters)
{
    uint32_t m[8][8], x[8], y[8];
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        x[a] = threadIdx.x * 8u + a + seed;
        y[a] = blockIdx.x * 8u + a;
#pragma unroll
        for (int b = 0; b < 8; ++b) m[a][b] = 0;
    }
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int a = 0; a < 8; ++a)
#pragma unroll
            for (int b = 0; b < 8; ++b)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xF6;" : "+r"(m[a][b]) : "r"(x[a]), "r"(y[b]));
#pragma unroll
        for (int a = 0; a < 8; ++a) {                      // 16 more lane-ops: keep operands changing
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[a]) : "r"(y[a]), "r"(seed));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[a]) : "r"(x[(a + 1) & 7]), "r"(seed));
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) s ^= m[a][b];
    if (s == 0x12345u) out[0] = s;
}

extern "C" int rthc_alu_peak_block(rthc_ctx *ctx, int warps_per_sm, double *lane_ops_per_s)
{
    if (!ctx || !lane_ops_per_s) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    uint32_t *d_out = nullptr;
    int rc = rthc_dev_alloc(ctx, &d_out, 1);
    if (rc) return rc;
    const int iters = 2048, threads = 256;
    // occupancy is set by the launch width: warps_per_sm / 8 resident CTAs per SM, several waves
    const int blocks = ctx->sm_count * (warps_per_sm / 8 > 0 ? warps_per_sm / 8 : 1);
    double best = 0;
    for (int rep = 0; rep < 4; ++rep) {
        RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
        rthc_lop3_block_stream<<<blocks, threads, 0, ctx->stream>>>(d_out, 0x0F0F0F0Fu + rep, iters);
        RTHC_LAUNCH_CHECK(ctx);
        RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        float ms = 0;
        RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
        double rate = (double)blocks * threads * (double)iters * 80.0 / (ms * 1e-3);
        if (rep > 0 && rate > best) best = rate;
    }
    rthc_dev_free(ctx, &d_out);
    *lane_ops_per_s = best;
    return 0;
}
