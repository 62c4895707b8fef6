// dist_probes.cu -- two micro-experiments behind DESIGN.md section 7 (round 2), NOT part of the product library.
//
// (A) "second pipe": the distance kernel saturates the ALU pipe (LOP3, 64 lanes/clk/SM) while a third of the issue
//     slots stay idle.  Could a slice of the loci go through packed-half compares on the FMA pipe instead?
//     Codes as fp16 bit patterns (missing = NaN), set.ne.f16x2 gives 1.0 / 0.0 per locus for 2 loci per
//     instruction (ordered `ne`: false when either side is NaN, i.e. the reference's missing rule for free) and an
//     HADD2 accumulates: 2 FMA-pipe instructions per 2 loci per pair.  The probe interleaves K such pairs of
//     instructions with every 64 LOP3 of the 8 x 8 register-block pattern, at the real kernel's occupancy
//     (8 warps / SM), and reports cycles per iteration: if the K pairs are free the idea has headroom.
// (B) "escape coding": fewer bit planes if rare alleles share an escape code, corrected afterwards by +1 on every
//     pair of samples that both hold an escaped (and different) allele at a locus.  At C4 that is about
//     200^2 / 2 pairs x 3,000 loci = 6e7 scattered uint16 increments into the 10 GB condensed matrix.  The probe
//     times exactly that access pattern (RED.ADD on the containing 32-bit word).
//
// build: nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o _build/dist_probes dist_probes.cu
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e_), __LINE__); exit(1); } } while (0)

template <int K>
__global__ void __launch_bounds__(256, 1) mix_kernel(uint32_t *out, uint32_t seed, int iters)
{
    uint32_t m[8][8], x[8], y[8];
    uint32_t hacc[8];                                   // f16x2 accumulators of the half path
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        x[a] = threadIdx.x * 8u + a + seed;
        y[a] = blockIdx.x * 8u + a;
        hacc[a] = 0;
#pragma unroll
        for (int b = 0; b < 8; ++b) m[a][b] = 0;
    }
    uint32_t hb = 0x3C004200u + seed, haq[K / 8 + 1];                              // fp16 pairs near (1.0, 2.0) / (1.0, 3.0)
#pragma unroll
    for (int q = 0; q < K / 8 + 1; ++q) haq[q] = 0x3C004000u + (threadIdx.x & 7) + 16 * q;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int a = 0; a < 8; ++a) {
#pragma unroll
            for (int b = 0; b < 8; ++b)
                asm volatile("lop3.b32 %0, %0, %1, %2, 0xF6;" : "+r"(m[a][b]) : "r"(x[a]), "r"(y[b]));
            // K / 8 compare + accumulate pairs behind every LOP3 row
#pragma unroll
            for (int q = 0; q < K / 8; ++q) {
                uint32_t ne;
                asm volatile("set.ne.f16x2.f16x2 %0, %1, %2;" : "=r"(ne) : "r"(haq[q]), "r"(hb));
                asm volatile("add.rn.f16x2 %0, %0, %1;" : "+r"(hacc[(a + q) & 7]) : "r"(ne));
            }
        }
#pragma unroll
        for (int a = 0; a < 8; ++a) {                      // keep the LOP3 operands changing (16 more lane-ops)
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(x[a]) : "r"(y[a]), "r"(seed));
            asm volatile("lop3.b32 %0, %0, %1, %2, 0x96;" : "+r"(y[a]) : "r"(x[(a + 1) & 7]), "r"(seed));
        }
    }
    uint32_t s = 0;
#pragma unroll
    for (int a = 0; a < 8; ++a) {
        s ^= hacc[a];
#pragma unroll
        for (int b = 0; b < 8; ++b) s ^= m[a][b];
    }
    if (s == 0x12345u) out[0] = s;
}

template <int K> static void run_mix(uint32_t *d_out, int sms, double clk_ghz)
{
    const int iters = 4096;
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        CK(cudaEventRecord(e0));
        mix_kernel<K><<<sms, 256>>>(d_out, 0x0F0F0F0Fu + rep, iters);
        CK(cudaEventRecord(e1));
        CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0 && ms < best) best = ms;
    }
    // per SM and iteration: 8 warps x 80 LOP3 (+ K set + K add)
    const double cyc = best * 1e-3 * clk_ghz * 1e9 / iters;
    const double lop3_lane_rate = 8.0 * 32 * 80 / cyc;             // lanes per clock per SM (pipe peak: 64)
    printf("{\"probe\": \"pipe_mix\", \"half_pairs_per_64_lop3\": %d, \"ms\": %.3f, \"cycles_per_iter\": %.1f, "
           "\"lop3_lanes_per_clk_per_sm\": %.2f, \"half_instr_lanes_per_clk_per_sm\": %.2f}\n",
           K, best, cyc, lop3_lane_rate, 8.0 * 32 * 2 * K / cyc);
}

__global__ void scatter_kernel(uint32_t *mat, uint64_t words, uint64_t m, uint32_t seed)
{
    const uint64_t t = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x, T = (uint64_t)gridDim.x * blockDim.x;
    for (uint64_t k = t; k < m; k += T) {
        uint64_t h = (k + seed) * 0x9E3779B97F4A7C15ull;          // pseudo-random entry of the condensed matrix
        h ^= h >> 29; h *= 0xBF58476D1CE4E5B9ull; h ^= h >> 32;
        const uint64_t e = h % (2 * words);
        atomicAdd(&mat[e >> 1], (e & 1) ? 65536u : 1u);            // result unused: RED.ADD
    }
}

int main()
{
    cudaDeviceProp p;
    CK(cudaGetDeviceProperties(&p, 0));
    int clk_khz = 0;
    CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
    const double ghz = clk_khz * 1e-6;
    uint32_t *d_out;
    CK(cudaMalloc(&d_out, 4));
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_ghz_nominal\": %.3f}\n", p.name, p.multiProcessorCount, ghz);
    run_mix<0>(d_out, p.multiProcessorCount, ghz);
    run_mix<8>(d_out, p.multiProcessorCount, ghz);
    run_mix<16>(d_out, p.multiProcessorCount, ghz);
    run_mix<24>(d_out, p.multiProcessorCount, ghz);
    run_mix<32>(d_out, p.multiProcessorCount, ghz);
    run_mix<64>(d_out, p.multiProcessorCount, ghz);

    const uint64_t words = 4999950000ull / 2;                      // the C4 condensed matrix as 32-bit words (10 GB)
    uint32_t *mat;
    CK(cudaMalloc(&mat, words * 4));
    CK(cudaMemset(mat, 0, words * 4));
    cudaEvent_t e0, e1;
    CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    for (uint64_t m : {uint64_t(6e7), uint64_t(2.4e8)}) {
        float best = 1e30f;
        for (int rep = 0; rep < 3; ++rep) {
            CK(cudaEventRecord(e0));
            scatter_kernel<<<p.multiProcessorCount * 16, 256>>>(mat, words, m, 17u + rep);
            CK(cudaEventRecord(e1));
            CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1));
            if (ms < best) best = ms;
        }
        printf("{\"probe\": \"escape_correction_scatter\", \"increments\": %llu, \"matrix_gb\": %.1f, \"ms\": %.3f, "
               "\"increments_per_s\": %.3e}\n", (unsigned long long)m, words * 4 / 1e9, best, m / (best * 1e-3));
    }
    return 0;
}
