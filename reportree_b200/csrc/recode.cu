// recode.cu -- device-side allele re-coder.
//
// Replaces the per-cell text parse + implicit allele identity of `cgmlst-dists`
// (call site scripts/partitioning_HC.py:502; contract :293-298) for the distance
// kernel: every locus column is dense-ranked on the device into uint16 codes with 0
// reserved for missing (equality-preserving, so distances are bit-exact), then 32 loci
// at a time are bit-sliced into "plane" words:
//
//   plane row off_g + k      : bit l = bit k of code(sample s, locus perm[32 g + l])   k < bits_g
//   plane row off_g + bits_g : bit l = (code != 0)                                  (valid plane)
// stored tile-blocked, planes[(s / 64) * plane_rows + row][s % 64], so that the rows of one
// group for one 128-sample block are one contiguous run (one TMA bulk copy per stage).
//
// Loci are permuted by bit width (the distance is a sum over loci, so order is free),
// which keeps every group at the minimum number of planes; loci with <= 1 distinct
// called allele can never produce a mismatch and are dropped.
#include <algorithm>
#include <numeric>

#include "rthc_internal.cuh"

#define RANK_THREADS 1024
#define BITMAP_BITS (1u << 20)
#define BITMAP_WORDS (BITMAP_BITS / 32)

// ---- 1. transpose n x L (row-major int32 / uint16) -> L x n_pad int32 (general path only) ----
template <typename IT>
__global__ void rthc_transpose_i32(const IT *__restrict__ in, int32_t *__restrict__ out,
                                   int64_t n, int64_t L, int64_t n_pad, int64_t Ls)
{
    __shared__ int32_t tile[32][33];                       // Ls: row stride of `in` (L, or more when `in` is a column range)
    int64_t l0 = (int64_t)blockIdx.x * 32, s0 = (int64_t)blockIdx.y * 32;
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        int64_t s = s0 + r, l = l0 + threadIdx.x;
        tile[r][threadIdx.x] = (s < n && l < L) ? (int32_t)in[s * Ls + l] : 0;
    }
    __syncthreads();
    for (int r = threadIdx.y; r < 32; r += blockDim.y) {
        int64_t l = l0 + r, s = s0 + threadIdx.x;
        if (l < L && s < n_pad) out[l * n_pad + s] = tile[threadIdx.x][r];
    }
}

// ---- 1'. fast path: 32 consecutive loci per CTA, straight from the row-major input ------------------
// Real cg/wgMLST allele ids are small integers, so the common case needs neither the transposed int32 copy nor
// the 2^20-bit bitmap: a CTA owns 32 neighbouring loci, a warp reads one 128-byte (int32) or 64-byte (uint16)
// row segment per load (lane = locus), marks the ids in shared-memory bitmaps of 2^14 bits, ranks them by
// popcount prefix, re-reads its columns and writes the uint16 codes transposed ([locus][sample], 128-byte runs)
// through a shared tile.  2 reads of the input + 1 write of the codes instead of transpose (read + write) +
// 2 reads.  Any id >= 2^14 in the matrix sets status bit 1 and the driver repeats with the general path below.
//
// Fewer bit planes than the dense rank needs: collide.cu (rthc_collide) re-maps the codes of a locus afterwards.
#define R32_BITS 16384
#define R32_WORDS (R32_BITS / 32)
#define R32_STRIDE (R32_WORDS + 1)       // lane * stride + word: lanes marking the same id hit 32 different banks
#define R32_CHUNK 512                    // samples per transposed tile
#define R32_TSTRIDE (R32_CHUNK + 2)      // uint16 row stride of the tile: 257 words, conflict-free both ways
#define R32_SMEM (32 * R32_STRIDE * 4 + 32 * R32_STRIDE * 2 + 32 * R32_TSTRIDE * 2)

__device__ __forceinline__ int dev_bit_width(int a)   // codes 1..a need this many bits
{
    return a <= 0 ? 0 : 32 - __clz(a);
}

// Two kernels, both with grid (32-locus blocks, row splits), so that a matrix with few loci per device (a locus
// range of the sharded re-coder: 12 blocks per rank for C4 on 8 GPUs) still fills the machine:
//   rthc_rank32_mark  marks the ids of its rows in shared bitmaps and ORs the non-empty words into the
//                     global bitmap gbm[locus][512 words] (zeroed by the caller);
//   rthc_rank32_code  loads the 32 finished bitmaps, builds the popcount prefix tables and codes its rows.
template <typename IT>
__global__ void __launch_bounds__(RANK_THREADS, 1)
rthc_rank32_mark(const IT *__restrict__ in, int64_t n, int64_t L, int64_t Ls, int64_t rows_per_split,
                 uint32_t *__restrict__ gbm, int32_t *__restrict__ status)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *bm = reinterpret_cast<uint32_t *>(smem_raw);                 // seen
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t l0 = (int64_t)blockIdx.x * 32, l = l0 + lane;
    const bool lv = l < L;
    const int64_t r_begin = (int64_t)blockIdx.y * rows_per_split, r_end = min(n, r_begin + rows_per_split);
    for (int w = tid; w < 32 * R32_STRIDE; w += RANK_THREADS) bm[w] = 0;
    __syncthreads();
    // ---- mark: 16 row segments in flight per warp; a bit that is already set is not set again --------
    bool big = false;
    const volatile uint32_t *vbm = bm;
    uint32_t *mybm = bm + lane * R32_STRIDE;
    for (int64_t s0 = r_begin + warp; s0 < r_end; s0 += 32 * 16) {
        int32_t v[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int64_t r = s0 + 32 * u;
            v[u] = (lv && r < r_end) ? (int32_t)__ldg(in + r * Ls + l) : 0;
        }
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            if (v[u] > 0) {
                if (v[u] >= R32_BITS) big = true;
                else {
                    const uint32_t bit = 1u << (v[u] & 31);
                    const int w = lane * R32_STRIDE + (v[u] >> 5);
                    if (!(vbm[w] & bit)) atomicOr(&mybm[v[u] >> 5], bit);
                }
            }
        }
    }
    if (__syncthreads_or(big)) {                   // the whole matrix goes through the general path
        if (tid == 0) atomicOr(status, 1);
        return;
    }
    for (int w = tid; w < 32 * R32_WORDS; w += RANK_THREADS) {
        const int loc = w / R32_WORDS, k = w % R32_WORDS;
        const uint32_t x = bm[loc * R32_STRIDE + k];
        if (x && l0 + loc < L) atomicOr(&gbm[(l0 + loc) * R32_WORDS + k], x);
    }
}

template <typename IT>
__global__ void __launch_bounds__(RANK_THREADS, 1)
rthc_rank32_code(const IT *__restrict__ in, int64_t n, int64_t L, int64_t Ls, int64_t n_pad, int64_t rows_per_split,
                 const uint32_t *__restrict__ gbm, uint16_t *__restrict__ codes, int32_t *__restrict__ n_alleles,
                 const int32_t *__restrict__ status)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *bm = reinterpret_cast<uint32_t *>(smem_raw);                 // seen
    uint16_t *pre = reinterpret_cast<uint16_t *>(bm + 32 * R32_STRIDE);
    uint16_t *tile = pre + 32 * R32_STRIDE;
    __shared__ uint32_t s_tot[32];
    if (*status & 1) return;                       // an id >= 2^14 somewhere: the general path redoes everything
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t l0 = (int64_t)blockIdx.x * 32, l = l0 + lane;
    const bool lv = l < L;
    for (int w = tid; w < 32 * R32_WORDS; w += RANK_THREADS) {
        const int loc = w / R32_WORDS, k = w % R32_WORDS;
        bm[loc * R32_STRIDE + k] = l0 + loc < L ? gbm[(l0 + loc) * R32_WORDS + k] : 0u;
    }
    __syncthreads();
    const uint32_t *mybm = bm + lane * R32_STRIDE;
    // ---- rank tables: warp w scans the bitmap of locus l0 + w ----
    {
        const uint32_t *b = bm + warp * R32_STRIDE;
        uint16_t *p = pre + warp * R32_STRIDE;
        uint32_t running = 0;
        for (int k = 0; k < R32_WORDS / 32; ++k) {
            const int idx = k * 32 + lane;
            const uint32_t c = __popc(b[idx]);
            uint32_t inc = c;
#pragma unroll
            for (int o = 1; o < 32; o <<= 1) {
                const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
                if (lane >= o) inc += t;
            }
            p[idx] = (uint16_t)(running + inc - c);
            running += __shfl_sync(0xffffffffu, inc, 31);
        }
        if (lane == 0) {
            s_tot[warp] = running;                               // number of codes of the locus
            if (blockIdx.y == 0 && l0 + warp < L) n_alleles[l0 + warp] = (int32_t)running;
        }
    }
    __syncthreads();
    // ---- code + transposed write, 512 samples per tile (16 row segments in flight per warp) ------------
    const uint16_t *mypre = pre + lane * R32_STRIDE;
    const int64_t c_begin = (int64_t)blockIdx.y * rows_per_split, c_end = min(n_pad, c_begin + rows_per_split);
    for (int64_t c0 = c_begin; c0 < c_end; c0 += R32_CHUNK) {      // rows_per_split % R32_CHUNK == 0
        int32_t v[16];
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            const int64_t r = c0 + warp + 32 * u;
            v[u] = (lv && r < n) ? (int32_t)__ldg(in + r * Ls + l) : 0;
        }
#pragma unroll
        for (int u = 0; u < 16; ++u) {
            uint16_t code = 0;
            if (v[u] > 0) {
                const uint32_t w = (uint32_t)v[u] >> 5, bit = 1u << (v[u] & 31);
                RTHC_DEV_ASSERT(mybm[w] & bit);                              // the id was marked in pass 1
                code = (uint16_t)(1u + mypre[w] + __popc(mybm[w] & (bit - 1u)));
                RTHC_DEV_ASSERT(code >= 1 && code <= s_tot[lane]);
            }
            tile[lane * R32_TSTRIDE + warp + 32 * u] = code;
        }
        __syncthreads();
        if (l0 + warp < L) {                       // warp w writes locus row l0 + w: 2 samples per lane and store
            const uint32_t *trow = reinterpret_cast<const uint32_t *>(tile + warp * R32_TSTRIDE);
            uint32_t *dst = reinterpret_cast<uint32_t *>(codes + (l0 + warp) * n_pad + c0);
#pragma unroll
            for (int k = 0; k < R32_CHUNK / 64; ++k) {
                const int64_t sidx = c0 + 64 * k + 2 * lane;
                if (sidx < n_pad) dst[32 * k + lane] = trow[32 * k + lane];
            }
        }
        __syncthreads();
    }
}

// ---- 2. per-locus dense rank through a shared-memory bitmap ---------------------------
// One CTA per locus.  Allele ids < 2^20 are marked in a 128 KB bitmap; the rank of an id
// is 1 + (number of set bits below it).  status[0] |= 1 if an id >= 2^20 was seen in the
// fast path, |= 2 if a locus has more than 65,535 distinct called alleles and the codes are
// uint16 (CT): the driver then repeats the ranking with uint32 codes (the reference has no such
// limit, scripts/partitioning_HC.py:293-298: any integer is an allele).
template <typename CT>
__global__ void __launch_bounds__(RANK_THREADS, 1)
rthc_rank_columns(const int32_t *__restrict__ colmajor, CT *__restrict__ codes,
                  int32_t *__restrict__ n_alleles, int32_t *__restrict__ status, int64_t n,
                  int64_t n_pad)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *bm = reinterpret_cast<uint32_t *>(smem_raw);
    uint16_t *pre = reinterpret_cast<uint16_t *>(smem_raw + BITMAP_WORDS * 4);
    __shared__ uint32_t warp_tot[32];
    __shared__ uint32_t warp_off[33];

    const int col = (int)blockIdx.x;
    const int32_t *src = colmajor + (int64_t)col * n_pad;
    CT *dst = codes + (int64_t)col * n_pad;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    __shared__ int32_t s_vmax;
    for (int w = tid; w < (int)BITMAP_WORDS; w += RANK_THREADS) bm[w] = 0;
    if (tid == 0) s_vmax = 0;
    __syncthreads();
    // mark: four independent loads in flight per thread; a bit that is already set (the common alleles, after
    // the first few rows) is not set again, which takes the same-word atomics of a skewed locus off the path
    bool big = false;
    int32_t vmax = 0;
    const volatile uint32_t *vbm = bm;
    for (int64_t r0 = tid; r0 < n; r0 += 4 * RANK_THREADS) {
        int32_t v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t r = r0 + (int64_t)u * RANK_THREADS;
            v[u] = r < n ? __ldg(src + r) : 0;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            if (v[u] > 0) {
                if ((uint32_t)v[u] >= BITMAP_BITS) big = true;
                else {
                    vmax = max(vmax, v[u]);
                    const uint32_t bit = 1u << (v[u] & 31);
                    if (!(vbm[v[u] >> 5] & bit)) atomicOr(&bm[v[u] >> 5], bit);
                }
            }
        }
    }
    if (big) atomicOr(status, 1);
    vmax = __reduce_max_sync(0xffffffffu, vmax);
    if (lane == 0 && vmax > 0) atomicMax(&s_vmax, vmax);
    __syncthreads();
    // warp w owns words [1024 w, 1024 (w+1)); conflict-free row-by-row scan, cut off behind the largest id seen
    const int used_words = (s_vmax >> 5) + 1;
    uint32_t running = 0;
    for (int k = 0; k < 32; ++k) {
        if (warp * 1024 + k * 32 >= used_words) break;             // uniform per warp
        int idx = warp * 1024 + k * 32 + lane;
        uint32_t c = __popc(bm[idx]);
        uint32_t inc = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        pre[idx] = (uint16_t)(running + inc - c);
        running += __shfl_sync(0xffffffffu, inc, 31);
    }
    if (lane == 0) warp_tot[warp] = running;
    __syncthreads();
    if (warp == 0) {
        uint32_t v = warp_tot[lane], inc = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            uint32_t t = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += t;
        }
        warp_off[lane] = inc - v;
        if (lane == 31) warp_off[32] = inc;
    }
    __syncthreads();
    const uint32_t total = warp_off[32];
    if (tid == 0) {
        n_alleles[col] = (int32_t)total;
        if (sizeof(CT) == 2 && total > 65535u) atomicOr(status, 2);
    }
    for (int64_t r0 = tid; r0 < n_pad; r0 += 4 * RANK_THREADS) {
        int32_t v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t r = r0 + (int64_t)u * RANK_THREADS;
            v[u] = r < n ? __ldg(src + r) : 0;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int64_t r = r0 + (int64_t)u * RANK_THREADS;
            if (r >= n_pad) break;
            CT code = 0;
            if (v[u] > 0 && (uint32_t)v[u] < BITMAP_BITS) {
                uint32_t w = (uint32_t)v[u] >> 5;
                uint32_t below = __popc(bm[w] & ((1u << (v[u] & 31)) - 1u));
                code = (CT)(1u + warp_off[w >> 10] + pre[w] + below);
            }
            dst[r] = code;
        }
    }
}

// ---- 2b. slow path for loci holding allele ids >= 2^20 --------------------------------
// One CTA per such locus, run one locus at a time against a 2^31-bit bitmap in HBM
// (256 MB scratch).  Ranks are assigned by scanning only the 2,048-bit windows that a
// level-1 shared bitmap marks as occupied.
template <typename CT>
__global__ void __launch_bounds__(RANK_THREADS, 1)
rthc_rank_column_wide(const int32_t *__restrict__ colmajor, CT *__restrict__ codes,
                      int32_t *__restrict__ n_alleles, int32_t *__restrict__ status, int64_t n,
                      int64_t n_pad, int col, uint32_t *__restrict__ gbm,
                      uint32_t *__restrict__ win_prefix)
{
    // level-1: bit b set <=> some id in [2048 b, 2048 (b+1)) occurs; 2^20 bits = 128 KB
    extern __shared__ __align__(16) unsigned char smem_raw[];
    uint32_t *l1 = reinterpret_cast<uint32_t *>(smem_raw);
    __shared__ uint32_t s_total;
    const int32_t *src = colmajor + (int64_t)col * n_pad;
    CT *dst = codes + (int64_t)col * n_pad;
    const int tid = threadIdx.x;

    for (int w = tid; w < (int)BITMAP_WORDS; w += RANK_THREADS) l1[w] = 0;
    if (tid == 0) s_total = 0;
    __syncthreads();
    for (int64_t r = tid; r < n; r += RANK_THREADS) {
        int32_t v = src[r];
        if (v > 0) {
            atomicOr(&gbm[(uint32_t)v >> 5], 1u << (v & 31));
            uint32_t b = (uint32_t)v >> 11;
            atomicOr(&l1[b >> 5], 1u << (b & 31));
        }
    }
    __threadfence();
    __syncthreads();
    // sequential over occupied windows (at most n of them), one warp scans a 64-word window
    // win_prefix[b] = number of set bits in all windows before window b (only for occupied b)
    if (tid < 32) {
        uint32_t run = 0;
        for (int w = 0; w < (int)BITMAP_WORDS; ++w) {
            uint32_t bits = l1[w];
            while (bits) {
                int bpos = __ffs(bits) - 1;
                bits &= bits - 1;
                uint32_t b = (uint32_t)w * 32u + (uint32_t)bpos;
                uint32_t c = __popc(gbm[b * 64u + tid]) + __popc(gbm[b * 64u + 32u + tid]);
#pragma unroll
                for (int d = 16; d > 0; d >>= 1) c += __shfl_xor_sync(0xffffffffu, c, d);
                if (tid == 0) win_prefix[b] = run;
                run += c;
            }
        }
        if (tid == 0) s_total = run;
    }
    __threadfence();
    __syncthreads();
    const uint32_t total = s_total;
    if (tid == 0) {
        n_alleles[col] = (int32_t)total;
        if (sizeof(CT) == 2 && total > 65535u) atomicOr(status, 2);
    }
    for (int64_t r = tid; r < n_pad; r += RANK_THREADS) {
        CT code = 0;
        if (r < n) {
            int32_t v = src[r];
            if (v > 0) {
                uint32_t b = (uint32_t)v >> 11, w = (uint32_t)v >> 5;
                uint32_t below = win_prefix[b];
                for (uint32_t q = b * 64u; q < w; ++q) below += __popc(gbm[q]);
                below += __popc(gbm[w] & ((1u << (v & 31)) - 1u));
                code = (CT)(1u + below);
            }
        }
        dst[r] = code;
    }
    // leave the HBM bitmap clean for the next locus
    __syncthreads();
    for (int64_t r = tid; r < n; r += RANK_THREADS) {
        int32_t v = src[r];
        if (v > 0) gbm[(uint32_t)v >> 5] = 0;
    }
}

// ---- 3. bit-slice 32 loci into plane words --------------------------------------------
// grid.y = group; the plane count of a group is uniform per CTA, so the bit loop is compiled per width
template <int BITS, typename CT>
__device__ __forceinline__ void pack_group(const CT *__restrict__ codes, const int32_t *__restrict__ perm, int g,
                                           int64_t off, uint32_t *__restrict__ planes, int64_t n_pad, int plane_rows, int64_t s)
{
    uint32_t pl[BITS];
#pragma unroll
    for (int k = 0; k < BITS; ++k) pl[k] = 0;
    uint32_t valid = 0;
#pragma unroll 8
    for (int l = 0; l < 32; ++l) {
        int col = perm[g * 32 + l];
        uint32_t c = col < 0 ? 0u : (uint32_t)codes[(int64_t)col * n_pad + s];
        valid |= (c != 0u ? 1u : 0u) << l;
#pragma unroll
        for (int k = 0; k < BITS; ++k) pl[k] |= ((c >> k) & 1u) << l;
    }
    // tile-blocked layout: [s / 64][plane row][s % 64]
    const int64_t blk = (s / RTHC_BLK) * (int64_t)plane_rows;
    const int64_t lane_s = s % RTHC_BLK;
#pragma unroll
    for (int k = 0; k < BITS; ++k) planes[(blk + off + k) * RTHC_BLK + lane_s] = pl[k];
    planes[(blk + off + BITS) * RTHC_BLK + lane_s] = valid;
}

template <typename CT>
__global__ void rthc_pack_planes(const CT *__restrict__ codes, const int32_t *__restrict__ perm,
                                 const int32_t *__restrict__ grp_off, const int32_t *__restrict__ grp_np,
                                 uint32_t *__restrict__ planes, int64_t n_pad, int plane_rows)
{
    const int g = blockIdx.y;
    const int64_t s = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (s >= n_pad) return;
    const int bits = grp_np[g] - 1;
    const int64_t off = grp_off[g];
#define RTHC_PACK_CASE(B) case B: pack_group<B, CT>(codes, perm, g, off, planes, n_pad, plane_rows, s); break;
    switch (bits) {
        RTHC_PACK_CASE(3) RTHC_PACK_CASE(4) RTHC_PACK_CASE(5) RTHC_PACK_CASE(6) RTHC_PACK_CASE(7) RTHC_PACK_CASE(8)
        RTHC_PACK_CASE(9) RTHC_PACK_CASE(10) RTHC_PACK_CASE(11) RTHC_PACK_CASE(12) RTHC_PACK_CASE(13)
        RTHC_PACK_CASE(14) RTHC_PACK_CASE(15) RTHC_PACK_CASE(16)
    default:
        if constexpr (sizeof(CT) == 4) {             // 17..24 bits: n < 2^24 samples bounds the alleles of a locus
            switch (bits) {
                RTHC_PACK_CASE(17) RTHC_PACK_CASE(18) RTHC_PACK_CASE(19) RTHC_PACK_CASE(20) RTHC_PACK_CASE(21)
                RTHC_PACK_CASE(22) RTHC_PACK_CASE(23)
            default: pack_group<24, CT>(codes, perm, g, off, planes, n_pad, plane_rows, s); break;
            }
        }
        break;
    }
#undef RTHC_PACK_CASE
}

// ---- host driver -----------------------------------------------------------------------
// dense-ranks every locus column into `codes` (uint16 or uint32); fills nall[] and status
template <typename CT>
static int rank_all(rthc_ctx *ctx, const int32_t *d_colmajor, CT *d_codes, int32_t *d_nall, int32_t *d_status,
                    std::vector<int32_t> &nall, int32_t &status)
{
    const int64_t n = ctx->n, L = ctx->L, n_pad = ctx->n_pad;
    int rc;
    RTHC_CUDA(ctx, cudaMemsetAsync(d_status, 0, sizeof(int32_t), ctx->stream));
    const size_t rank_smem = BITMAP_WORDS * 4 + BITMAP_WORDS * 2;
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_rank_columns<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)rank_smem));
    rthc_rank_columns<CT><<<(unsigned)L, RANK_THREADS, rank_smem, ctx->stream>>>(d_colmajor, d_codes, d_nall, d_status, n, n_pad);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemcpyAsync(&status, d_status, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(nall.data(), d_nall, sizeof(int32_t) * (size_t)L, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (!(status & 1)) return 0;
    // some loci hold allele ids >= 2^20 (the rare path): find them and redo those through the HBM bitmap
    std::vector<int32_t> colmax((size_t)L, 0), tmp((size_t)n_pad);
    for (int64_t l = 0; l < L; ++l) {
        RTHC_CUDA(ctx, cudaMemcpy(tmp.data(), d_colmajor + l * n_pad, sizeof(int32_t) * (size_t)n, cudaMemcpyDeviceToHost));
        int32_t mx = 0;
        for (int64_t r = 0; r < n; ++r) mx = std::max(mx, tmp[(size_t)r]);
        colmax[(size_t)l] = mx;
    }
    uint32_t *d_gbm = nullptr, *d_winpre = nullptr;
    if ((rc = rthc_dev_alloc(ctx, &d_gbm, (size_t)1 << 26))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_winpre, (size_t)1 << 20))) return rc;
    RTHC_CUDA(ctx, cudaMemsetAsync(d_gbm, 0, sizeof(uint32_t) << 26, ctx->stream));
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_rank_column_wide<CT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                        (int)(BITMAP_WORDS * 4)));
    for (int64_t l = 0; l < L; ++l) {
        if ((uint32_t)colmax[(size_t)l] < BITMAP_BITS) continue;
        rthc_rank_column_wide<CT><<<1, RANK_THREADS, BITMAP_WORDS * 4, ctx->stream>>>(d_colmajor, d_codes, d_nall, d_status, n,
                                                                                      n_pad, (int)l, d_gbm, d_winpre);
        RTHC_LAUNCH_CHECK(ctx);
    }
    RTHC_CUDA(ctx, cudaMemcpyAsync(&status, d_status, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(nall.data(), d_nall, sizeof(int32_t) * (size_t)L, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_gbm);
    rthc_dev_free(ctx, &d_winpre);
    return 0;
}

static inline int bit_width(int a)   // codes 1..a need this many bits
{
    int b = 0;
    while ((1 << b) <= a) ++b;
    return b;
}

template <typename IT>
static int rank_fast(rthc_ctx *ctx, const IT *d_rowmajor, int32_t *d_nall, int32_t *d_status, std::vector<int32_t> &nall,
                     int32_t &status)
{
    const int64_t n = ctx->n, L = ctx->L, n_pad = ctx->n_pad, Ls = ctx->L_stride ? ctx->L_stride : L;
    RTHC_CUDA(ctx, cudaMemsetAsync(d_status, 0, sizeof(int32_t), ctx->stream));
    const unsigned nblocks = (unsigned)((L + 31) / 32);
    // row splits: enough CTAs for about four waves of the code kernel (one CTA per SM), rows per split a multiple of
    // the 512-sample tile
    int64_t splits = std::max<int64_t>(1, (4 * (int64_t)ctx->sm_count + nblocks - 1) / nblocks);
    int64_t rows_per_split = ((n_pad + splits - 1) / splits + R32_CHUNK - 1) / R32_CHUNK * R32_CHUNK;
    splits = (n_pad + rows_per_split - 1) / rows_per_split;
    // the global bitmaps stay with the context (grow only): a 6 MB block that comes and goes between the GB-sized ones
    // made the pool split and re-grow its large blocks for several calls
    int rc;
    if (ctx->gbm_cap < (size_t)L * R32_WORDS) {
        ctx->gbm_cap = 0;
        if ((rc = rthc_dev_alloc(ctx, &ctx->d_gbm, (size_t)L * R32_WORDS))) return rc;
        ctx->gbm_cap = (size_t)L * R32_WORDS;
    }
    uint32_t *d_gbm = ctx->d_gbm;
    RTHC_CUDA(ctx, cudaMemsetAsync(d_gbm, 0, sizeof(uint32_t) * (size_t)L * R32_WORDS, ctx->stream));
    const int mark_smem = 32 * R32_STRIDE * 4;
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_rank32_mark<IT>, cudaFuncAttributeMaxDynamicSharedMemorySize, mark_smem));
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_rank32_code<IT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)R32_SMEM));
    const dim3 grid(nblocks, (unsigned)splits);
    rthc_rank32_mark<IT><<<grid, RANK_THREADS, mark_smem, ctx->stream>>>(d_rowmajor, n, L, Ls, rows_per_split, d_gbm, d_status);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_rank32_code<IT><<<grid, RANK_THREADS, R32_SMEM, ctx->stream>>>(d_rowmajor, n, L, Ls, n_pad, rows_per_split, d_gbm,
                                                                       ctx->d_codes, d_nall, d_status);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemcpyAsync(&status, d_status, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(nall.data(), d_nall, sizeof(int32_t) * (size_t)L, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

// general path: transposed int32 copy, 2^20-bit bitmap per locus, HBM bitmap for ids >= 2^20, uint32 codes when a
// locus has more than 65,535 alleles
template <typename IT>
static int rank_general(rthc_ctx *ctx, const IT *d_rowmajor, int32_t *d_nall, int32_t *d_status, std::vector<int32_t> &nall,
                        int32_t &status)
{
    const int64_t n = ctx->n, L = ctx->L, n_pad = ctx->n_pad;
    int32_t *d_colmajor = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_colmajor, (size_t)(L * n_pad)))) return rc;
    {
        dim3 blk(32, 8), grd((unsigned)((L + 31) / 32), (unsigned)((n_pad + 31) / 32));
        rthc_transpose_i32<IT><<<grd, blk, 0, ctx->stream>>>(d_rowmajor, d_colmajor, n, L, n_pad, ctx->L_stride ? ctx->L_stride : L);
        RTHC_LAUNCH_CHECK(ctx);
    }
    if ((rc = rank_all<uint16_t>(ctx, d_colmajor, ctx->d_codes, d_nall, d_status, nall, status))) return rc;
    if (status & 2) {
        // a locus with more than 65,535 distinct called alleles (needs n > 65,535): the codes of this matrix are
        // kept as uint32 instead and the wide loci get 17..24 bit planes (generic group body of the distance kernel)
        rthc_dev_free(ctx, &ctx->d_codes);
        if ((rc = rthc_dev_alloc(ctx, &ctx->d_codes32, (size_t)(L * n_pad)))) return rc;
        if ((rc = rank_all<uint32_t>(ctx, d_colmajor, ctx->d_codes32, d_nall, d_status, nall, status))) return rc;
    }
    rthc_dev_free(ctx, &d_colmajor);
    return 0;
}

// in_u16: the caller's matrix is uint16 (0 = missing) instead of int32 (<= 0 = missing)
int rthc_recode_and_pack(rthc_ctx *ctx, const void *d_rowmajor, bool in_u16)
{
    const int64_t L = ctx->L, n_pad = ctx->n_pad;
    int32_t *d_nall = nullptr, *d_status = nullptr, *d_perm = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_nall, (size_t)L))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_status, 1))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_codes, (size_t)(L * n_pad)))) return rc;
    std::vector<int32_t> nall((size_t)L);
    int32_t status = 0;
    rc = in_u16 ? rank_fast<uint16_t>(ctx, (const uint16_t *)d_rowmajor, d_nall, d_status, nall, status)
                : rank_fast<int32_t>(ctx, (const int32_t *)d_rowmajor, d_nall, d_status, nall, status);
    if (rc) return rc;
    ctx->recode_general = (status & 1) != 0;
    if (status & 1) {
        rc = in_u16 ? rank_general<uint16_t>(ctx, (const uint16_t *)d_rowmajor, d_nall, d_status, nall, status)
                    : rank_general<int32_t>(ctx, (const int32_t *)d_rowmajor, d_nall, d_status, nall, status);
        if (rc) return rc;
    }

    int max_all = 0;                                                   // distinct alleles of the widest locus
    for (int64_t l = 0; l < L; ++l) max_all = std::max(max_all, (int)nall[(size_t)l]);
    // fast path only: rare alleles of a locus may share codes (fewer bit planes, exact fix-up after the plane kernel);
    // nall[] becomes the number of codes in use
    ctx->escape_loci = ctx->escape_cells = 0; ctx->escape_pairs_u = 0;
    ctx->d_pcodes = ctx->d_codes;
    if (!ctx->recode_general && (rc = rthc_collide_codes(ctx, d_nall, max_all, nall))) return rc;

    // order loci by bit width (widest first), drop loci that cannot mismatch
    std::vector<int32_t> kept;
    for (int64_t l = 0; l < L; ++l)
        if (nall[(size_t)l] >= 2) kept.push_back((int32_t)l);
    std::stable_sort(kept.begin(), kept.end(), [&](int32_t a, int32_t b) {
        return bit_width(nall[(size_t)a]) > bit_width(nall[(size_t)b]);
    });
    if (kept.size() > 65535) {
        rthc_dev_free(ctx, &d_nall); rthc_dev_free(ctx, &d_status);
        return rthc_fail(ctx, RTHC_E_RANGE, "%lld of the %lld loci have two or more called alleles; distances are "
                         "stored as uint16, so at most 65,535 such loci are supported",
                         (long long)kept.size(), (long long)L);
    }
    ctx->max_alleles = max_all;                                        // distinct alleles, not codes
    ctx->L_kept = (int64_t)kept.size();
    const int G = (int)((kept.size() + 31) / 32);
    std::vector<int32_t> perm((size_t)G * 32, -1), goff((size_t)std::max(G, 1)), gnp((size_t)std::max(G, 1));
    int rows = 0, max_np = 0;
    for (int g = 0; g < G; ++g) {
        int bits = RTHC_MIN_BITS;      // the distance kernel peels 3 bit planes + the valid plane per group
        for (int l = 0; l < 32; ++l) {
            size_t idx = (size_t)g * 32 + (size_t)l;
            if (idx < kept.size()) {
                perm[idx] = kept[idx];
                bits = std::max(bits, bit_width(nall[(size_t)kept[idx]]));
            }
        }
        goff[(size_t)g] = rows;
        gnp[(size_t)g] = bits + 1;
        rows += bits + 1;
        max_np = std::max(max_np, bits + 1);
    }
    ctx->G = G;
    ctx->plane_rows = rows;
    ctx->blk_stride_rows = rows;
    ctx->max_np = max_np;
    ctx->h_grp_np.assign(gnp.begin(), gnp.begin() + G);
    std::vector<int64_t> gbase((size_t)std::max(G, 1));
    for (int g = 0; g < G; ++g) gbase[(size_t)g] = goff[(size_t)g];

    if ((rc = rthc_dev_alloc(ctx, &ctx->d_planes, (size_t)std::max(rows, 1) * (size_t)n_pad))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_grp_off, (size_t)std::max(G, 1)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_grp_base, (size_t)std::max(G, 1)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_grp_np, (size_t)std::max(G, 1)))) return rc;
    if (G > 0) {
        RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_grp_base, gbase.data(), (size_t)G * sizeof(int64_t), cudaMemcpyHostToDevice,
                                       ctx->stream));
        if ((rc = rthc_dev_alloc(ctx, &d_perm, perm.size()))) return rc;
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_perm, perm.data(), perm.size() * sizeof(int32_t),
                                       cudaMemcpyHostToDevice, ctx->stream));
        RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_grp_off, goff.data(), (size_t)G * sizeof(int32_t),
                                       cudaMemcpyHostToDevice, ctx->stream));
        RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_grp_np, gnp.data(), (size_t)G * sizeof(int32_t),
                                       cudaMemcpyHostToDevice, ctx->stream));
        dim3 grd((unsigned)((n_pad + 255) / 256), (unsigned)G);
        if (ctx->d_codes32)
            rthc_pack_planes<uint32_t><<<grd, 256, 0, ctx->stream>>>(ctx->d_codes32, d_perm, ctx->d_grp_off, ctx->d_grp_np,
                                                                     ctx->d_planes, n_pad, rows);
        else
            rthc_pack_planes<uint16_t><<<grd, 256, 0, ctx->stream>>>(ctx->d_pcodes, d_perm, ctx->d_grp_off, ctx->d_grp_np,
                                                                     ctx->d_planes, n_pad, rows);
        RTHC_LAUNCH_CHECK(ctx);
    }
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_perm);
    rthc_dev_free(ctx, &d_nall);
    rthc_dev_free(ctx, &d_status);
    return 0;
}
