// distance_ws.cu -- pairwise allelic distance, warp-specialised persistent kernel (v5).
//
// Same arithmetic and data layout as distance.cu (see there for the bit-sliced formulation
// of cgmlst-dists' rule, scripts/partitioning_HC.py:489-516):
//     mism = OR_k (Pa_k ^ Pb_k);   d += popc(mism & Va & Vb)     per group of 32 loci
// What changes is the schedule, after measuring the v4 kernel at 72-74 % of the LOP3 peak with
// the pure 8 x 8 LOP3 pattern reaching 98 % (rthc_alu_peak_block):
//   * one PRODUCER warp issues every TMA bulk copy; the 8 consumer warps carry no load logic,
//     no divergent branch and no elected-lane code (setmaxnreg moves the producer's registers
//     to the consumers);
//   * CTAs are persistent (one per SM) and the stage ring runs across tile boundaries, so the
//     load pipeline never drains between tiles;
//   * the group body is fully unrolled per plane count NP (template), which lets the 64 POPCs
//     of the previous group (XU pipe, 8 issue cycles each) be spread evenly over all NP plane
//     blocks instead of 16-wide bursts in 4 blocks; groups are sorted by NP, so a run of equal
//     NP is one loop around one body.
#include "dist_common.cuh"

#define WS_CONSUMERS 256
#define WS_THREADS 384            // 8 consumer warps + 1 producer warpgroup (setmaxnreg works per warpgroup; one lane issues)
#define WS_NP_MIN 4               // RTHC_MIN_BITS + valid plane
#define WS_NP_MAX_T 12            // plane counts with a dedicated unrolled body; above: generic

// ptxas hoists every POPC of a straight-line body to its top (XU results are "long latency"),
// which serialises the warp on the XU pipe for hundreds of cycles.  The POPC input is therefore
// routed through one IMAD (FMA pipe, idle here) that adds zero * <a word produced by the LOP3 row
// the POPC should follow>: a true dependency pins the POPC behind that row at no ALU-pipe cost.
// `zero` is a run-time 0 the compiler cannot see through.
__device__ __forceinline__ uint32_t ws_popc_after(uint32_t x, uint32_t after, uint32_t zero)
{
    return (uint32_t)__popc(zero * after + x);
}

struct WsOps {
    uint32_t a[8], b[8];
};

__device__ __forceinline__ uint4 ws_lds128(uint32_t addr)
{
    uint4 v;
    asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
    return v;
}
// st_a / st_b: 32-bit shared addresses of this thread's words in plane 0 of the stage's I block
// (+ 16 ty bytes) and J block (+ 16 tx bytes); blk_bytes = bytes of one 64-sample block slot
__device__ __forceinline__ WsOps ws_load(uint32_t st_a, uint32_t st_b, uint32_t blk_bytes, int k)
{
    WsOps o;
    const uint4 a0 = ws_lds128(st_a + k * (RTHC_BLK * 4));
    const uint4 a1 = ws_lds128(st_a + blk_bytes + k * (RTHC_BLK * 4));
    const uint4 b0 = ws_lds128(st_b + k * (RTHC_BLK * 4));
    const uint4 b1 = ws_lds128(st_b + blk_bytes + k * (RTHC_BLK * 4));
    o.a[0] = a0.x; o.a[1] = a0.y; o.a[2] = a0.z; o.a[3] = a0.w;
    o.a[4] = a1.x; o.a[5] = a1.y; o.a[6] = a1.z; o.a[7] = a1.w;
    o.b[0] = b0.x; o.b[1] = b0.y; o.b[2] = b0.z; o.b[3] = b0.w;
    o.b[4] = b1.x; o.b[5] = b1.y; o.b[6] = b1.z; o.b[7] = b1.w;
    return o;
}

// per-thread state that lives across groups of one tile
struct WsAcc {
    uint32_t accp[8][4];   // 64 distances, two 16-bit counters per register (pair (a,b): accp[a][b&3], high half for b >= 4)
    uint32_t mm[8][8];     // masked mismatch words of the previous group, POPCs pending
};

__device__ __forceinline__ void ws_fold(WsAcc &A, int q, uint32_t v)
{
    const int a = q >> 3, b = q & 7;
    A.accp[a][b & 3] += (b < 4) ? v : v * 65536u;
}

__device__ __forceinline__ void ws_wait(uint32_t bar_addr, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WS_WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra WS_WAIT_%=;\n"
        "}\n" ::"r"(bar_addr), "r"(parity)
        : "memory");
}
__device__ __forceinline__ void ws_arrive(uint32_t bar_addr)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar_addr) : "memory");
}

// The stage ring: `it` counts groups consumed by this CTA since kernel start.  All addresses
// are 32-bit shared-window addresses.
struct WsRing {
    uint32_t base_a, base_b;       // this thread's operand words in plane 0 of stage 0 (I block / J block)
    uint32_t full_bar, empty_bar;  // barrier arrays
    uint32_t stage_bytes, blk_bytes;
    uint32_t it;                   // next group to be consumed
    uint32_t total;                // groups this CTA will consume in all
};

// One group of 32 loci with NP planes (NP-1 bit planes + valid plane), fully unrolled.
// On entry `cur` holds the operands of plane 0 (fetched by the previous group); on exit it
// holds plane 0 of the next group of the ring (if any).
// POPC schedule: the pending words mm[0..6][*] of the previous group are POPC'd and folded
// inside the NP-1 bit-plane blocks, evenly; row 7 is POPC'd at the start of the last block
// (before mm is overwritten) and folded at its end.
template <int NP>
__device__ __forceinline__ void ws_group(WsRing &R, WsAcc &A, WsOps &cur, const uint32_t zero)
{
    const uint32_t s = R.it % RTHC_STAGES;
    const uint32_t st_a = R.base_a + s * R.stage_bytes, st_b = R.base_b + s * R.stage_bytes;
    uint32_t m[8][8];
    constexpr int NB = NP - 1;                 // bit-plane blocks
#pragma unroll
    for (int k = 0; k < NB; ++k) {
        const WsOps nxt = ws_load(st_a, st_b, R.blk_bytes, k + 1);
        const int q0 = (56 * k) / NB, q1 = (56 * (k + 1)) / NB;     // POPCs of this block (folded at compile time)
        const int nq = q1 - q0;
        uint32_t pcs[(56 + NB - 1) / NB + 1];
#pragma unroll
        for (int a = 0; a < 8; ++a) {
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                if (k == 0) m[a][b] = cur.a[a] ^ cur.b[b];
                else m[a][b] |= cur.a[a] ^ cur.b[b];
            }
            const int r0 = (nq * a) / 8, r1 = (nq * (a + 1)) / 8;   // this row's share of the block's POPCs
#pragma unroll
            for (int r = r0; r < r1; ++r) pcs[r] = ws_popc_after(A.mm[(q0 + r) >> 3][(q0 + r) & 7], m[a][7], zero);
        }
#pragma unroll
        for (int r = 0; r < nq; ++r) ws_fold(A, q0 + r, pcs[r]);
        cur = nxt;
    }
    // last block: valid plane (cur).  Plane 0 of the next group is fetched first so that its
    // shared-memory latency hides behind this block's 64 LOP3s.
    // (The producer appends one padding group to the ring, so there always is a next stage.)
    const uint32_t s1 = (R.it + 1) % RTHC_STAGES;
    ws_wait(R.full_bar + 8 * s1, ((R.it + 1) / RTHC_STAGES) & 1u);
    const WsOps nxt = ws_load(R.base_a + s1 * R.stage_bytes, R.base_b + s1 * R.stage_bytes, R.blk_bytes, 0);
    // old row 7 is POPC'd one word per new row 0..6 (two behind row 6), then overwritten last
    uint32_t p7[8];
#pragma unroll
    for (int a = 0; a < 7; ++a) {
#pragma unroll
        for (int b = 0; b < 8; ++b) A.mm[a][b] = m[a][b] & cur.a[a] & cur.b[b];
        p7[a] = ws_popc_after(A.mm[7][a], A.mm[a][7], zero);
        if (a == 6) p7[7] = ws_popc_after(A.mm[7][7], A.mm[6][6], zero);
    }
#pragma unroll
    for (int b = 0; b < 8; ++b) A.mm[7][b] = m[7][b] & cur.a[7] & cur.b[b];
#pragma unroll
    for (int b = 0; b < 8; ++b) ws_fold(A, 56 + b, p7[b]);
    ws_arrive(R.empty_bar + 8 * s);            // every consumer thread arrives (count = 256)
    R.it += 1;
    RTHC_DEV_ASSERT(R.it <= R.total);
    cur = nxt;
}

// Any plane count (loci with more than 2,047 alleles): runtime plane loop, POPCs not overlapped.
__device__ __forceinline__ void ws_group_generic(WsRing &R, WsAcc &A, WsOps &cur, int np)
{
    const uint32_t s = R.it % RTHC_STAGES;
    const uint32_t st_a = R.base_a + s * R.stage_bytes, st_b = R.base_b + s * R.stage_bytes;
    uint32_t m[8][8];
#pragma unroll
    for (int q = 0; q < 64; ++q) ws_fold(A, q, (uint32_t)__popc(A.mm[q >> 3][q & 7]));
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) m[a][b] = cur.a[a] ^ cur.b[b];
    for (int k = 1; k < np - 1; ++k) {
        cur = ws_load(st_a, st_b, R.blk_bytes, k);
#pragma unroll
        for (int a = 0; a < 8; ++a)
#pragma unroll
            for (int b = 0; b < 8; ++b) m[a][b] |= cur.a[a] ^ cur.b[b];
    }
    cur = ws_load(st_a, st_b, R.blk_bytes, np - 1);
#pragma unroll
    for (int a = 0; a < 8; ++a)
#pragma unroll
        for (int b = 0; b < 8; ++b) A.mm[a][b] = m[a][b] & cur.a[a] & cur.b[b];
    ws_arrive(R.empty_bar + 8 * s);
    R.it += 1;
    const uint32_t s1 = R.it % RTHC_STAGES;
    ws_wait(R.full_bar + 8 * s1, (R.it / RTHC_STAGES) & 1u);
    cur = ws_load(R.base_a + s1 * R.stage_bytes, R.base_b + s1 * R.stage_bytes, R.blk_bytes, 0);
}

template <int NP>
__device__ __forceinline__ void ws_run(WsRing &R, WsAcc &A, WsOps &cur, int count, const uint32_t zero)
{
#pragma unroll 1
    for (int c = 0; c < count; ++c) ws_group<NP>(R, A, cur, zero);
}

__global__ void __launch_bounds__(WS_THREADS, 1)
rthc_distance_ws(const uint32_t *__restrict__ planes, const int64_t *__restrict__ grp_base,
                 const int32_t *__restrict__ grp_np, int G, int max_np, int64_t blk_stride_rows,
                 const int2 *__restrict__ tiles, int n_tiles, const int64_t *__restrict__ row_base,
                 int64_t n, uint16_t *__restrict__ cond)
{
    extern __shared__ __align__(128) unsigned char smem_raw[];
    __shared__ __align__(8) uint64_t full_bar[RTHC_STAGES];
    __shared__ __align__(8) uint64_t empty_bar[RTHC_STAGES];

    const int tid = threadIdx.x;
    const int blk_words = max_np * RTHC_BLK;
    const uint32_t stage_words = (uint32_t)blk_words * 4;
    uint32_t *stage_base = reinterpret_cast<uint32_t *>(smem_raw);
    uint64_t *gtab = reinterpret_cast<uint64_t *>(stage_base + (size_t)RTHC_STAGES * stage_words);     // (row offset << 8) | np per group
    for (int g = tid; g < G; g += WS_THREADS) gtab[g] = ((uint64_t)grp_base[g] << 8) | (uint64_t)grp_np[g];
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < RTHC_STAGES; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], WS_CONSUMERS);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    const int my_tiles = ((int)blockIdx.x < n_tiles) ? (n_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

    if (tid >= WS_CONSUMERS) {
        // ------------------------------ producer warp ------------------------------------------
        asm volatile("setmaxnreg.dec.sync.aligned.u32 24;");
        if (tid == WS_CONSUMERS) {
            const size_t blk_stride = (size_t)blk_stride_rows * RTHC_BLK; // words per 64-sample block in HBM
            uint32_t it = 0;
            // one padding group (group 0 of the last tile again) closes the ring: consumers fetch
            // plane 0 of "the next group" unconditionally
            for (int t = 0; t < my_tiles + (my_tiles > 0 && G > 0 ? 1 : 0); ++t) {
                const bool pad = t == my_tiles;
                const int2 tile = tiles[blockIdx.x + (size_t)(pad ? my_tiles - 1 : t) * gridDim.x];
                const uint32_t *srcs[4] = {planes + (size_t)(2 * tile.x) * blk_stride,
                                           planes + (size_t)(2 * tile.x + 1) * blk_stride,
                                           planes + (size_t)tile.y * blk_stride,
                                           planes + (size_t)(tile.y + 1) * blk_stride};
                for (int g = 0; g < (pad ? 1 : G); ++g, ++it) {
                    const uint32_t s = it % RTHC_STAGES;
                    if (it >= RTHC_STAGES) mbar_wait(&empty_bar[s], ((it / RTHC_STAGES) - 1u) & 1u);
                    const uint64_t e = gtab[g];
                    const uint32_t np = (uint32_t)(e & 0xFFu);
                    const size_t off = (size_t)(e >> 8);
                    const uint32_t bytes = np * RTHC_BLK * 4u;
                    uint32_t *dst = stage_base + (size_t)s * stage_words;
                    mbar_expect_tx(&full_bar[s], 4 * bytes);
#pragma unroll
                    for (int h = 0; h < 4; ++h)
                        tma_bulk_g2s(dst + h * blk_words, srcs[h] + off * RTHC_BLK, bytes, &full_bar[s]);
                }
            }
        }
        return;
    }
    // --------------------------------- consumer warps -------------------------------------------
    asm volatile("setmaxnreg.inc.sync.aligned.u32 240;");
    const int tx = tid & 15, ty = tid >> 4;
    const uint32_t zero = (uint32_t)((uint64_t)n >> 40);          // n < 2^24: always 0, opaque to the compiler
    WsRing R;
    R.blk_bytes = (uint32_t)blk_words * 4u;
    R.stage_bytes = stage_words * 4u;
    R.base_a = smem_u32(stage_base) + 16u * (uint32_t)ty;
    R.base_b = smem_u32(stage_base) + 2u * R.blk_bytes + 16u * (uint32_t)tx;
    R.full_bar = smem_u32(full_bar); R.empty_bar = smem_u32(empty_bar);
    R.it = 0; R.total = (uint32_t)my_tiles * (uint32_t)G;
    WsOps cur;
    if (R.total > 0) {
        ws_wait(R.full_bar, 0);
        cur = ws_load(R.base_a, R.base_b, R.blk_bytes, 0);
    }
    for (int t = 0; t < my_tiles; ++t) {
        const int2 tile = tiles[blockIdx.x + (size_t)t * gridDim.x];
        WsAcc A;
#pragma unroll
        for (int a = 0; a < 8; ++a) {
#pragma unroll
            for (int b = 0; b < 4; ++b) A.accp[a][b] = 0;
#pragma unroll
            for (int b = 0; b < 8; ++b) A.mm[a][b] = 0;
        }
        int g = 0;
        while (g < G) {
            const int np = (int)(gtab[g] & 0xFFu);
            int ge = g + 1;
            while (ge < G && (int)(gtab[ge] & 0xFFu) == np) ++ge;
            const int count = ge - g;
            switch (np) {
            case 4: ws_run<4>(R, A, cur, count, zero); break;
            case 5: ws_run<5>(R, A, cur, count, zero); break;
            case 6: ws_run<6>(R, A, cur, count, zero); break;
            case 7: ws_run<7>(R, A, cur, count, zero); break;
            case 8: ws_run<8>(R, A, cur, count, zero); break;
            case 9: ws_run<9>(R, A, cur, count, zero); break;
            case 10: ws_run<10>(R, A, cur, count, zero); break;
            case 11: ws_run<11>(R, A, cur, count, zero); break;
            case 12: ws_run<12>(R, A, cur, count, zero); break;
            default:
                for (int c = 0; c < count; ++c) ws_group_generic(R, A, cur, np);
            }
            g = ge;
        }
        // drain: mm holds the last group
#pragma unroll
        for (int q = 0; q < 64; ++q) ws_fold(A, q, (uint32_t)__popc(A.mm[q >> 3][q & 7]));

        // epilogue: condensed upper triangle, scipy order: row_base[i] + (j - i - 1)
        const int64_t i0 = (int64_t)tile.x * RTHC_TILE, j0 = (int64_t)tile.y * RTHC_BLK;
#pragma unroll
        for (int a = 0; a < 8; ++a) {
            const int64_t i = i0 + (a < 4 ? 4 * ty + a : 64 + 4 * ty + (a - 4));
            if (i >= n) continue;
            const int64_t base = row_base[i];
            if (base < 0) continue;
#pragma unroll
            for (int b = 0; b < 8; ++b) {
                const int64_t j = j0 + (b < 4 ? 4 * tx + b : 64 + 4 * tx + (b - 4));
                const uint32_t d = b < 4 ? (A.accp[a][b] & 0xFFFFu) : (A.accp[a][b & 3] >> 16);
                RTHC_DEV_ASSERT(d <= 0xFFFFu && (int)d <= 32 * G);
                if (j > i && j < n) cond[base + (j - i - 1)] = (uint16_t)d;
            }
        }
    }
}

int rthc_launch_distance_ws(rthc_ctx *ctx, int max_np)
{
    const size_t smem = (size_t)RTHC_STAGES * (size_t)max_np * 4 * RTHC_BLK * sizeof(uint32_t) +
                        (size_t)(ctx->G + 1) * sizeof(uint64_t);
    if (smem > 200u * 1024u)
        return rthc_fail(ctx, RTHC_E_RANGE, "plane stage does not fit shared memory (max_np=%d)", max_np);
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_distance_ws, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t grid = ctx->n_tiles < ctx->sm_count ? ctx->n_tiles : ctx->sm_count;
    rthc_distance_ws<<<(unsigned)grid, WS_THREADS, smem, ctx->stream>>>(
        ctx->d_planes, ctx->d_grp_base, ctx->d_grp_np, ctx->G, max_np, ctx->blk_stride_rows, ctx->d_tiles,
        (int)ctx->n_tiles, ctx->d_row_base, ctx->n, ctx->d_cond);
    RTHC_LAUNCH_CHECK(ctx);
    return 0;
}
