// collide.cu -- fewer bit planes per locus by letting rare alleles SHARE codes, with an exact fix-up.
//
// The plane kernel (distance_ws.cu) pays one LOP3 per pair, plane and 32 loci, and the plane count of a locus is the
// bit width of its number of distinct alleles -- which in cg/wgMLST data is dominated by rare alleles (C4: 5..400
// founder alleles + about 200 singletons per locus => 9..10 bits).  If two alleles of a locus share a code, the plane
// kernel sees "equal" where the reference (cgmlst-dists' rule, scripts/partitioning_HC.py:489-516) says "different",
// for exactly the pairs of samples that hold two DIFFERENT alleles of the same code.  Those pairs are few when the
// sharing alleles are rare, and they are known: every cell whose code is shared is listed, and rthc_collide_fixup adds
// the lost mismatch back with one scattered uint16 increment per pair (2.1e10 increments/s on the 10 GB matrix of C4,
// profiles/r02_dist_probes_a.log).  Distances stay bit-exact; intermediate values are never above the true ones, so
// nothing can overflow.  Round 2's singleton-escape code was the special case "all singletons share one code".
//
// Plan per locus (one CTA, everything in shared memory): alleles sorted by count, descending.  With b bits there are
// C = 2^b - 1 codes: the C most frequent alleles keep a code of their own, the remaining alleles are poured into the k
// least frequent of those codes up to a common level T (water filling: least sum of squared code loads, i.e. fewest
// colliding pairs), k the smallest number for which the level stays below the next code's count.  The bit width is
// the one that minimises  b * pairs_per_bit + colliding pairs  (pairs_per_bit = what one plane of one locus costs,
// expressed in fix-up increments), subject to at most RTHC_ESC_CAP listed cells per locus.
#include <algorithm>
#include <cstdlib>

#include "rthc_internal.cuh"

#define COL_THREADS 512
#define COL_PROBE 4096                    // samples counted before the two most frequent codes are singled out
#define COL_CLAMP ((1u << 18) - 1u)       // sort key: min(count, COL_CLAMP) << 14 | (16383 - code)

__device__ __forceinline__ int col_bit_width(int a) { return a <= 0 ? 0 : 32 - __clz(a); }

// block-wide sum of a 64-bit value (COL_THREADS threads); every thread gets the result
__device__ __forceinline__ unsigned long long col_block_sum(unsigned long long v, unsigned long long *s_red)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    unsigned long long t = 0;
#pragma unroll
    for (int w = 0; w < COL_THREADS / 32; ++w) t += s_red[w];
    return t;
}

// codes: [L][n_pad] dense ranks from recode.cu (1..A, 0 = missing); pcodes: what the plane packer reads.
// nall_io[l]: A on entry, the number of codes in use afterwards.  N2 = power of two >= max A + 1 (shared arrays).
__global__ void __launch_bounds__(COL_THREADS)
rthc_collide(const uint16_t *__restrict__ codes, uint16_t *__restrict__ pcodes, int64_t n, int64_t n_pad,
             const int32_t *__restrict__ nall_in, int32_t *__restrict__ nall_out, int32_t *__restrict__ cnt_out,
             uint64_t *__restrict__ list, int N2, unsigned long long pairs_per_bit)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *hist = reinterpret_cast<uint32_t *>(smem_raw);          // [N2] count per dense code; later the code map
    uint32_t *keys = hist + N2;                                        // [N2] sort keys
    uint32_t *psort = keys + N2;                                       // [N2 + 1] prefix sums of the sorted counts
    __shared__ unsigned long long s_red[COL_THREADS / 32];
    __shared__ int s_k, s_best_b, s_best_k, s_cnt;
    __shared__ unsigned long long s_best_cost, s_top;
    const int tid = threadIdx.x;
    const int64_t l = blockIdx.x;
    const uint16_t *col = codes + l * n_pad;
    uint16_t *pcol = pcodes + l * n_pad;
    const int A = nall_in[l];
    const int full = max(col_bit_width(A), RTHC_MIN_BITS);
    if (A < 2 || full <= RTHC_MIN_BITS || list == nullptr) {           // nothing to save: codes pass through
        for (int64_t s = 2 * tid; s < n_pad; s += 2 * COL_THREADS)
            *reinterpret_cast<uint32_t *>(pcol + s) = *reinterpret_cast<const uint32_t *>(col + s);
        if (tid == 0) { nall_out[l] = A; cnt_out[l] = 0; }
        return;
    }
    for (int w = tid; w < N2; w += COL_THREADS) hist[w] = 0;
    if (tid == 0) s_cnt = 0;
    __syncthreads();
    // ---- pass 1: count per code.  Lanes of a warp that hold the same code would serialise on one shared counter
    // (the most frequent allele of a cgMLST locus covers 10..90 % of the samples), and __match_any_sync costs more than
    // the rest of the loop.  So: the first COL_PROBE samples are counted with match_any, the two most frequent codes
    // so far are then counted by ballot in registers, everything else goes straight to its counter. -------------------
    const int64_t probe_end = min((int64_t)COL_PROBE, n_pad);
    for (int64_t s = 2 * tid; s < probe_end; s += 2 * COL_THREADS) {    // n_pad % 128 == 0: a warp is in or out as a whole
        const uint32_t w = *reinterpret_cast<const uint32_t *>(col + s);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t c = h ? (w >> 16) : (w & 0xFFFFu);
            const uint32_t m = __match_any_sync(0xffffffffu, c);
            if ((tid & 31) == __ffs(m) - 1) atomicAdd(&hist[c], (uint32_t)__popc(m));
        }
    }
    __syncthreads();
    if (probe_end < n_pad) {
        uint32_t top[2];
#pragma unroll
        for (int round = 0; round < 2; ++round) {
            if (tid == 0) s_top = 0ull;
            __syncthreads();
            unsigned long long best = 0;
            for (int c = 1 + tid; c <= A; c += COL_THREADS)
                if (round == 0 || (uint32_t)c != top[0]) best = max(best, ((unsigned long long)hist[c] << 32) | (unsigned)c);
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) best = max(best, __shfl_xor_sync(0xffffffffu, best, o));
            if ((tid & 31) == 0 && best) atomicMax(&s_top, best);
            __syncthreads();
            top[round] = (uint32_t)(s_top & 0xFFFFFFFFull);            // 0 (the missing code) if there is none
            __syncthreads();
        }
        const uint32_t t1 = top[0], t2 = top[1];
        uint32_t n1 = 0, n2 = 0;                                       // lane 0 of every warp
        // 8 samples (one 16-byte load) per lane and trip; the trip count is uniform per warp (ballots), lanes past the
        // end carry a code that matches nothing
        for (int64_t base = probe_end + (int64_t)(tid >> 5) * 256; base < n_pad; base += (COL_THREADS / 32) * 256) {
            const int64_t s = base + 8 * (tid & 31);
            const bool in = s < n_pad;                                  // n_pad % 8 == 0: all 8 or none
            uint4 w4 = make_uint4(0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu, 0xFFFFFFFFu);
            if (in) w4 = *reinterpret_cast<const uint4 *>(col + s);
            const uint32_t ws[4] = {w4.x, w4.y, w4.z, w4.w};
#pragma unroll
            for (int h = 0; h < 8; ++h) {
                const uint32_t c = (h & 1) ? (ws[h >> 1] >> 16) : (ws[h >> 1] & 0xFFFFu);
                const bool is1 = c == t1, is2 = c == t2;
                n1 += __popc(__ballot_sync(0xffffffffu, is1));
                n2 += __popc(__ballot_sync(0xffffffffu, is2));
                if (in && !is1 && !is2) atomicAdd(&hist[c], 1u);
            }
        }
        if ((tid & 31) == 0) { if (n1) atomicAdd(&hist[t1], n1); if (n2 && t2 != t1) atomicAdd(&hist[t2], n2); }
    }
    __syncthreads();
    // ---- sort the alleles by count, descending (bitonic, keys padded with 0) ---------------------------------
    for (int r = tid; r < N2; r += COL_THREADS) {
        const int code = r + 1;
        keys[r] = code <= A ? ((min(hist[code], COL_CLAMP) << 14) | (uint32_t)(16383 - code)) : 0u;
    }
    __syncthreads();
    for (int kk = 2; kk <= N2; kk <<= 1)
        for (int j = kk >> 1; j > 0; j >>= 1) {
            for (int i = tid; i < N2; i += COL_THREADS) {
                const int p = i ^ j;
                if (p > i) {
                    const uint32_t a = keys[i], b = keys[p];
                    const bool desc = (i & kk) == 0;
                    if (desc ? a < b : a > b) { keys[i] = b; keys[p] = a; }
                }
            }
            __syncthreads();
        }
    // ---- psort[r] = number of cells held by the r most frequent alleles (exact counts, not the clamped keys) ----
    {
        const int per = (N2 + COL_THREADS - 1) / COL_THREADS;
        const int r0 = min(tid * per, N2), r1 = min(r0 + per, N2);
        uint32_t sum = 0;
        for (int r = r0; r < r1; ++r) sum += r < A ? hist[16383 - (int)(keys[r] & 16383u)] : 0u;
        // exclusive scan of the per-thread sums
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if ((tid & 31) >= o) inc += t;
        }
        __shared__ uint32_t s_wsum[COL_THREADS / 32];
        if ((tid & 31) == 31) s_wsum[tid >> 5] = inc;
        __syncthreads();
        uint32_t before = inc - sum;
        for (int w = 0; w < (tid >> 5); ++w) before += s_wsum[w];
        for (int r = r0; r < r1; ++r) {
            psort[r] = before;
            before += r < A ? hist[16383 - (int)(keys[r] & 16383u)] : 0u;
        }
        if (r1 == N2 && r0 < N2) psort[N2] = before;
    }
    __syncthreads();
#define COL_CNT(r) (psort[(r) + 1] - psort[(r)])
    // ---- choose the bit width ---------------------------------------------------------------------------------
    if (tid == 0) { s_best_b = full; s_best_k = 0; s_best_cost = (unsigned long long)full * pairs_per_bit; }
    __syncthreads();
    for (int b = RTHC_MIN_BITS; b < full; ++b) {
        const int C = (1 << b) - 1;                                   // < A because b < bit_width(A)
        const uint32_t S = psort[A] - psort[C];                       // cells of the alleles without a code of their own
        if (tid == 0) {
            // smallest k in [1, C] with  S + (cells of the k least frequent coded alleles) <= k * count of the next one
            int lo = 1, hi = C;
            while (lo < hi) {
                const int k = (lo + hi) >> 1;
                const unsigned long long tot = (unsigned long long)S + (psort[C] - psort[C - k]);
                if (tot <= (unsigned long long)k * COL_CNT(C - k - 1)) hi = k; else lo = k + 1;
            }
            s_k = lo;
        }
        __syncthreads();
        const int k = s_k;
        const unsigned long long cells = (unsigned long long)S + (psort[C] - psort[C - k]);
        unsigned long long sq = 0;
        if (cells <= RTHC_ESC_CAP)
            for (int r = C - k + tid; r < A; r += COL_THREADS) { const unsigned long long c = COL_CNT(r); sq += c * c; }
        sq = col_block_sum(sq, s_red);
        // (k * count of the most frequent receiving code <= cells: the level is not below any receiving code; always
        // true for the smallest k of a sorted list -- re-checked because counts >= 2^18 are sorted as equal)
        if (tid == 0 && cells <= RTHC_ESC_CAP && (unsigned long long)k * COL_CNT(C - k) <= cells) {
            // every code of the k receiving ones ends near the level cells / k:  sum of load^2 >= cells^2 / k
            const unsigned long long ld2 = (cells * cells + (unsigned long long)k - 1) / (unsigned long long)k;
            const unsigned long long pairs = ld2 > sq ? (ld2 - sq) / 2 : 0;
            const unsigned long long cost = (unsigned long long)b * pairs_per_bit + pairs;
            if (cost < s_best_cost) { s_best_cost = cost; s_best_b = b; s_best_k = k; }
        }
        __syncthreads();
    }
    const int b = s_best_b, k = s_best_k;
    __syncthreads();                                                  // hist is dead from here on: it becomes the map
    uint32_t *map = hist;                                             // [code] -> new code | merged << 15
    if (b == full) {
        for (int c = tid; c <= A; c += COL_THREADS) map[c] = (uint32_t)c;
    } else {
        const int C = (1 << b) - 1;
        const uint32_t S = psort[A] - psort[C];
        const unsigned long long cells = (unsigned long long)S + (psort[C] - psort[C - k]);
        const unsigned long long T = (cells + (unsigned long long)k - 1) / (unsigned long long)k;     // the level
        if (tid == 0) map[0] = 0;
        __syncthreads();
        for (int r = tid; r < A; r += COL_THREADS) {
            const int code = 16383 - (int)(keys[r] & 16383u);
            uint32_t m;
            if (r < C) {
                m = (uint32_t)(r + 1) | (r >= C - k ? 0x8000u : 0u);
            } else {
                // receiving code j (0 = least frequent coded allele, rank C - 1 - j) has room T - count; the alleles
                // are poured in order: this one starts at cell `pos` of the stream and lands in the code whose room
                // covers that position:  Q(j) <= pos < Q(j + 1),  Q(j) = j T - (cells of the j least frequent coded)
                const unsigned long long pos = psort[r] - psort[C];
                int lo = 0, hi = k - 1;
                while (lo < hi) {
                    const int j = (lo + hi + 1) >> 1;
                    const unsigned long long Q = (unsigned long long)j * T - (psort[C] - psort[C - j]);
                    if (Q <= pos) lo = j; else hi = j - 1;
                }
                m = (uint32_t)(C - lo) | 0x8000u;
            }
            map[code] = m;
        }
    }
    __syncthreads();
    // ---- pass 2: re-map the column, list the cells whose code is shared -------------------------------------------
    uint64_t *mylist = list + l * RTHC_ESC_CAP;
    for (int64_t s = 8 * tid; s < n_pad; s += 8 * COL_THREADS) {       // n_pad % 8 == 0; 16 bytes in, 16 bytes out
        const uint4 w4 = *reinterpret_cast<const uint4 *>(col + s);
        const uint32_t ws[4] = {w4.x, w4.y, w4.z, w4.w};
        uint32_t os[4] = {0, 0, 0, 0};
#pragma unroll
        for (int h = 0; h < 8; ++h) {
            const uint32_t c = (h & 1) ? (ws[h >> 1] >> 16) : (ws[h >> 1] & 0xFFFFu);
            const uint32_t m = map[c];
            if (m & 0x8000u) {
                const int pos = atomicAdd(&s_cnt, 1);
                RTHC_DEV_ASSERT(pos < RTHC_ESC_CAP);
                if (pos < RTHC_ESC_CAP)
                    mylist[pos] = (uint64_t)(s + h) | ((uint64_t)(m & 0x7FFFu) << 24) | ((uint64_t)c << 40);
            }
            os[h >> 1] |= (m & 0x7FFFu) << (16 * (h & 1));
        }
        *reinterpret_cast<uint4 *>(pcol + s) = make_uint4(os[0], os[1], os[2], os[3]);
    }
    __syncthreads();
    if (tid == 0) {
        nall_out[l] = b == full ? A : (1 << b) - 1;
        cnt_out[l] = s_cnt;
    }
#undef COL_CNT
}

// d(i, j) += 1 for every pair of listed cells of a locus that share a code but hold different alleles (one CTA per
// locus).  The cells are first grouped by code (counting sort in shared memory: the shared codes of a locus are one
// contiguous range of at most RTHC_ESC_CAP values), so that only cells of the same code are paired.  Only rows this
// context computed are touched (tile-row ownership as in rthc_setup_geometry); the uint16 entry is incremented through
// the aligned 32-bit word that holds it (no carry: true distances fit 16 bits).
#define FIX_THREADS 256
#define FIX_SMEM (2 * RTHC_ESC_CAP * 8 + (2 * RTHC_ESC_CAP + 2) * 4)
__global__ void __launch_bounds__(FIX_THREADS) rthc_collide_fixup(const int32_t *__restrict__ cnt, const uint64_t *__restrict__ list,
                                                                  const int64_t *__restrict__ row_base, int64_t n_tile_rows, int world,
                                                                  int rank, uint16_t *cond, unsigned long long *__restrict__ pairs_done)
{
    extern __shared__ __align__(16) unsigned char fix_raw[];
    uint64_t *raw = reinterpret_cast<uint64_t *>(fix_raw);            // as listed
    uint64_t *srt = raw + RTHC_ESC_CAP;                                // grouped by code
    uint32_t *start = reinterpret_cast<uint32_t *>(srt + RTHC_ESC_CAP);   // [CAP + 1] first cell of every code
    uint32_t *fill = start + RTHC_ESC_CAP + 1;                         // [CAP] cells placed so far
    __shared__ uint32_t s_min, s_wsum[FIX_THREADS / 32];
    const int64_t l = blockIdx.x;
    const int S = cnt[l];
    if (S < 2) return;
    const int tid = threadIdx.x, lane = tid & 31;
    if (tid == 0) s_min = 0xFFFFFFFFu;
    for (int k = tid; k <= RTHC_ESC_CAP; k += FIX_THREADS) { start[k] = 0; if (k < RTHC_ESC_CAP) fill[k] = 0; }
    __syncthreads();
    uint32_t mn = 0xFFFFFFFFu;
    for (int k = tid; k < S; k += FIX_THREADS) {
        const uint64_t e = list[l * RTHC_ESC_CAP + k];
        raw[k] = e;
        mn = min(mn, (uint32_t)(e >> 24) & 0x7FFFu);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
    if (lane == 0) atomicMin(&s_min, mn);
    __syncthreads();
    const uint32_t base = s_min;
    for (int k = tid; k < S; k += FIX_THREADS) {
        const uint32_t bin = ((uint32_t)(raw[k] >> 24) & 0x7FFFu) - base;
        RTHC_DEV_ASSERT(bin < RTHC_ESC_CAP);
        atomicAdd(&start[bin + 1], 1u);
    }
    __syncthreads();
    {   // inclusive scan of start[1 .. CAP] (counts) => start[b] = first cell of code base + b
        constexpr int PER = RTHC_ESC_CAP / FIX_THREADS;
        uint32_t v[PER], sum = 0;
#pragma unroll
        for (int u = 0; u < PER; ++u) { v[u] = start[1 + tid * PER + u]; sum += v[u]; }
        uint32_t inc = sum;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
            if (lane >= o) inc += t;
        }
        if (lane == 31) s_wsum[tid >> 5] = inc;
        __syncthreads();
        uint32_t run = inc - sum;
        for (int w = 0; w < (tid >> 5); ++w) run += s_wsum[w];
#pragma unroll
        for (int u = 0; u < PER; ++u) { run += v[u]; start[1 + tid * PER + u] = run; }
    }
    __syncthreads();
    for (int k = tid; k < S; k += FIX_THREADS) {
        const uint64_t e = raw[k];
        const uint32_t bin = ((uint32_t)(e >> 24) & 0x7FFFu) - base;
        srt[start[bin] + atomicAdd(&fill[bin], 1u)] = e;
    }
    __syncthreads();
    unsigned int done = 0;
    for (int p = tid; p < S; p += FIX_THREADS) {
        const uint64_t ep = srt[p];
        const uint32_t bin = ((uint32_t)(ep >> 24) & 0x7FFFu) - base;
        const int end = (int)start[bin + 1];
        const int64_t sp = (int64_t)(ep & 0xFFFFFFu);
        // one rank: every cell pairs with the cells behind it in its code's run.  Several ranks: a pair belongs to the
        // rank that owns the row of its SMALLER sample, so a cell whose row is someone else's is skipped as a whole and
        // an owned one walks its whole run for the larger samples -- the walk shards with the rows instead of being
        // repeated on every rank.
        int q0 = p + 1;
        if (world > 1) {
            const int64_t r = sp / RTHC_TILE;
            if ((int)(min(r, n_tile_rows - 1 - r) % world) != rank) continue;
            q0 = (int)start[bin];
        }
        for (int q = q0; q < end; ++q) {
            const uint64_t eq = srt[q];
            if (((ep ^ eq) >> 40) == 0) continue;                      // the same allele: the planes were right
            const int64_t sq = (int64_t)(eq & 0xFFFFFFu);
            if (world > 1 && sq < sp) continue;                        // counted from the other side
            const int64_t i = min(sp, sq), j = max(sp, sq);
            uint16_t *e = cond + row_base[i] + (j - i - 1);
            atomicAdd(reinterpret_cast<unsigned int *>(reinterpret_cast<uintptr_t>(e) & ~(uintptr_t)3),
                      (reinterpret_cast<uintptr_t>(e) & 2) ? 65536u : 1u);
            ++done;
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) done += __shfl_xor_sync(0xffffffffu, done, o);
    if (lane == 0 && done) atomicAdd(pairs_done, (unsigned long long)done);
}

int rthc_launch_escape_correct(rthc_ctx *ctx)
{
    if (!ctx->d_esc_cnt || !ctx->d_esc_list || ctx->escape_cells == 0) return 0;
    RTHC_CUDA(ctx, cudaMemsetAsync(ctx->d_esc_pairs, 0, sizeof(unsigned long long), ctx->stream));
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_collide_fixup, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)FIX_SMEM));
    rthc_collide_fixup<<<(unsigned)ctx->esc_L, FIX_THREADS, FIX_SMEM, ctx->stream>>>(ctx->d_esc_cnt, ctx->d_esc_list, ctx->d_row_base, ctx->n_tiles_edge,
                                                                ctx->world, ctx->rank, ctx->d_cond, ctx->d_esc_pairs);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemcpyAsync(&ctx->escape_pairs_u, ctx->d_esc_pairs, sizeof(unsigned long long), cudaMemcpyDeviceToHost,
                                   ctx->stream));
    return 0;
}

// Host side of the plan: called by rthc_recode_and_pack after the dense ranking of the fast path.  d_nall holds the
// allele counts on entry and the code counts afterwards; ctx->d_pcodes is what the plane packer reads.
int rthc_collide_codes(rthc_ctx *ctx, int32_t *d_nall, int max_alleles, std::vector<int32_t> &nall)
{
    const int64_t n = ctx->n, L = ctx->L, n_pad = ctx->n_pad;
    const bool off = getenv("RTHC_NO_ESCAPE") != nullptr;                   // test hook: plain dense ranking
    // lists of RTHC_ESC_CAP cells per locus only while that stays small (cg/wgMLST widths; a SNP matrix with millions of
    // columns has a handful of alleles per column and nothing to share)
    const bool on = !off && L <= 65536 && max_alleles >= (1 << RTHC_MIN_BITS);
    ctx->escape_loci = ctx->escape_cells = 0; ctx->escape_pairs_u = 0;
    ctx->esc_L = L;
    if (!on) { ctx->d_pcodes = ctx->d_codes; return 0; }
    int rc;
    int32_t *d_nall2 = nullptr;
    ctx->d_pcodes = nullptr;                                              // was the alias of d_codes: not to be freed
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_pcodes, (size_t)(L * n_pad)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_esc_cnt, (size_t)L))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_esc_list, (size_t)L * RTHC_ESC_CAP))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_esc_pairs, 1))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_nall2, (size_t)L))) return rc;
    int N2 = 8;
    while (N2 < max_alleles + 1) N2 <<= 1;
    const size_t smem = (size_t)(3 * N2 + 1) * sizeof(uint32_t);
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_collide, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    // one plane of one locus costs P / 32 LOP3 lane-ops; one fix-up increment costs about as much as 640 of them
    // (C4: 1.6e13 lane-ops/s against 2.5e10 increments/s)  =>  pairs_per_bit = P / 20,480; 24,576 leaves a margin
    const unsigned long long P = (unsigned long long)n * (unsigned long long)(n - 1) / 2;
    unsigned long long div = 24576ull;
    if (const char *e = getenv("RTHC_COLLIDE_DIV")) {                        // test hook: 1 = share codes wherever the cap allows
        const long long v = atoll(e);
        if (v > 0) div = (unsigned long long)v;
    }
    const unsigned long long ppb = std::max<unsigned long long>(P / div, 1ull);
    rthc_collide<<<(unsigned)L, COL_THREADS, smem, ctx->stream>>>(ctx->d_codes, ctx->d_pcodes, n, n_pad, d_nall, d_nall2,
                                                                  ctx->d_esc_cnt, ctx->d_esc_list, N2, ppb);
    RTHC_LAUNCH_CHECK(ctx);
    std::vector<int32_t> cnt((size_t)L);
    RTHC_CUDA(ctx, cudaMemcpyAsync(nall.data(), d_nall2, sizeof(int32_t) * (size_t)L, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(cnt.data(), ctx->d_esc_cnt, sizeof(int32_t) * (size_t)L, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    for (int64_t l = 0; l < L; ++l)
        if (cnt[(size_t)l] > 0) { ctx->escape_loci++; ctx->escape_cells += cnt[(size_t)l]; }
    rthc_dev_free(ctx, &d_nall2);
    return 0;
}
