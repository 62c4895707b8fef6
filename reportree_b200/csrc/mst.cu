// mst.cu -- single linkage on the device-resident condensed matrix.
//
// Replaces scipy's `_hierarchy.mst_single_linkage` behind
// linkage(y, "single") (scripts/partitioning_HC.py:197; scipy _hierarchy_impl.py:960-961).
//
// (1) Boruvka minimum spanning forest over the condensed uint16 rows a context owns,
//     under the strict total order key = (w, i, j) packed in 64 bits, so the forest is
//     unique and identical for 1/2/4/8-way shards.  Each round is one streaming pass
//     over the owned part of the matrix (HBM-bound: 2 bytes per pair per round).
// (2) A bitonic sort of the <= 8 (n-1) packed edges.
// (3) Device Prim in scipy's visiting order (first strictly-smaller key wins) for the
//     mode that reproduces scipy's Z byte for byte.
#include <algorithm>
#include <cstdlib>

#include "rthc_internal.cuh"

__global__ void rthc_iota_i32(int32_t *a, int64_t n)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) a[v] = (int32_t)v;
}
__global__ void rthc_fill_u64(uint64_t *a, int64_t n, uint64_t val)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) a[v] = val;
}
// Round state on the device, so that several Boruvka rounds can be queued without a host round trip each:
// [0] edges emitted so far, [1] the same at the end of the previous round, [2] done (spanning tree complete, or a
// round added nothing: the forest of a disconnected edge set is complete), [3] rounds executed.  Every kernel of a
// round returns at once when done is set.
#define BST_COUNT 0
#define BST_PREV 1
#define BST_DONE 2
#define BST_ROUNDS 3
__global__ void rthc_boruvka_round_end(unsigned int *state, int64_t n)
{
    if (state[BST_DONE]) return;
    state[BST_ROUNDS] += 1;
    if (state[BST_COUNT] == state[BST_PREV] || (int64_t)state[BST_COUNT] >= n - 1) state[BST_DONE] = 1;
    state[BST_PREV] = state[BST_COUNT];
}
__global__ void rthc_fill_u64_round(uint64_t *a, int64_t n, uint64_t val, const unsigned int *state)
{
    if (state[BST_DONE]) return;
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) a[v] = val;
}

// TILE PASS (row and column candidates in ONE sweep over the matrix).  A CTA owns TS_ROWS
// consecutive rows x TS_COLS columns of the square view; thread t owns the 8 strided columns
// c0 + t + 256 k, so for a fixed row a warp reads 64 contiguous bytes of the condensed row and
// keeps 8 loads in flight.  comp[] of the thread's columns and the running column minima live
// in registers, comp[] / base of the rows in shared memory -- no gathers in the inner loop.
// Candidates are compared as 32-bit local keys (w << 16 | local index): for a fixed row the
// order (w, j) and for a fixed column the order (w, i) are exactly the global (w, i, j) order.
// Row minima: warp REDUX + one shared atomic per warp per row; per tile at most one filtered
// 64-bit atomicMin per row and per column.
#define TS_ROWS 64
#define TS_COLS 2048
#define TS_KEY_INF 0xFFFFFFFFu

// Same-address pressure on best[] is what limits the late rounds (one giant component: every
// column of every tile holds a candidate for the same entry), so candidates are merged first
// inside the warp (match_any groups + REDUX), then inside the CTA (direct-mapped shared table),
// and the filter reads best[] through L2 (ld.global.cg; an L1-stale +inf would let every
// candidate through to the atomic unit).
#define TS_TAB 512
__device__ __forceinline__ void best_push(uint64_t *best, int32_t c, uint64_t key)
{
    if (key < __ldcg(reinterpret_cast<const unsigned long long *>(&best[c])))
        atomicMin((unsigned long long *)&best[c], (unsigned long long)key);
}
// must be called by all 32 lanes of a warp
__device__ __forceinline__ void tile_push(bool has, int32_t c, uint64_t key, int32_t *tab_c,
                                          unsigned long long *tab_k, uint64_t *best)
{
    const unsigned act = __ballot_sync(0xffffffffu, has);
    if (!has) return;
    // Only the component of the first active lane is merged inside the warp: that is the giant
    // component of the late rounds (the same-address case), and every lane of that group passes the
    // SAME mask, so the reductions are single instructions.  (Reducing every match_any group made the
    // compiler loop over the distinct masks: 32 iterations per call in the first round.)
    const unsigned grp = __match_any_sync(act, c);
    const unsigned lead = __shfl_sync(act, grp, __ffs(act) - 1);
    if (grp == lead) {
        const uint32_t hi = (uint32_t)(key >> 32), lo = (uint32_t)key;
        const uint32_t mh = __reduce_min_sync(lead, hi);
        const uint32_t ml = __reduce_min_sync(lead, hi == mh ? lo : 0xFFFFFFFFu);
        if (hi != mh || lo != ml) return;                  // keys are distinct: one lane of the group survives
    }
    const uint32_t slot = ((uint32_t)c * 2654435761u) >> 23;
    const int32_t old = atomicCAS(&tab_c[slot], -1, c);
    if (old == -1 || old == c) atomicMin(&tab_k[slot], (unsigned long long)key);
    else best_push(best, c, key);
}

template <bool CHECK>
__device__ __forceinline__ void scan_tile_body(const uint16_t *__restrict__ cond, const int64_t *s_off,
                                               const int32_t *s_ci, uint32_t *s_rowmin, int64_t r0,
                                               int64_t c0, int64_t n, const int32_t (&cj)[8],
                                               uint32_t (&colmin)[8])
{
    const int tid = threadIdx.x;
    const int64_t jt = c0 + tid;
#pragma unroll 2
    for (int r = 0; r < TS_ROWS; ++r) {
        const int64_t off = s_off[r];               // row_base[i] - i - 1, or < -n if the row is absent
        if (off < -n) continue;                     // uniform
        const int32_t ci = s_ci[r];
        const int64_t i = r0 + r;
        const uint16_t *row = cond + off + jt;
        uint32_t w[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int64_t j = jt + 256 * k;
            w[k] = 0;
            if (!CHECK || (j > i && j < n)) w[k] = row[256 * k];
        }
        uint32_t rowmin = TS_KEY_INF;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            const int64_t j = jt + 256 * k;
            if ((!CHECK || (j > i && j < n)) && cj[k] != ci) {
                const uint32_t hi = w[k] << 16;
                colmin[k] = min(colmin[k], hi | (uint32_t)r);
                rowmin = min(rowmin, hi | (uint32_t)(tid + 256 * k));
            }
        }
        rowmin = __reduce_min_sync(0xffffffffu, rowmin);
        if ((tid & 31) == 0 && rowmin != TS_KEY_INF) atomicMin(&s_rowmin[r], rowmin);
    }
}

// Interior tiles (no diagonal, no right edge): thread t owns the 8 CONTIGUOUS columns c0 + 8 t + e and
// reads them with 16-byte loads.  Row i's segment starts at element off_i + c0 of the condensed vector,
// which is 16-byte aligned only when s = (off_i + c0) mod 8 is 0; the thread therefore loads the two
// aligned chunks that cover its columns and realigns by s (uniform per row) with funnel shifts.
// Keys are built by PRMT straight from the packed words: (w << 16) | index in one instruction.
struct TileRow {          // one row's two aligned 16-byte chunks in flight
    uint4 lo, hi;
    int s;                // misalignment of the row segment in elements (0..7), -1: row absent
};
// s_seg[r] = element index of the aligned chunk that holds column c0 of row r, shifted left by 3, plus the
// misalignment (0..7) in the low bits; -1: row absent.  One 64-bit shared load per row.
__device__ __forceinline__ TileRow tile_row_load(const uint16_t *__restrict__ cond, const int64_t *s_seg, int r, int tid)
{
    TileRow R;
    const int64_t seg = s_seg[r];
    if (seg < 0) { R.s = -1; R.lo = make_uint4(0, 0, 0, 0); R.hi = R.lo; return R; }     // uniform
    R.s = (int)(seg & 7);
    const uint4 *p = reinterpret_cast<const uint4 *>(cond + (seg & ~(int64_t)7)) + tid;
    R.lo = __ldg(p);
    R.hi = __ldg(p + 1);
    return R;
}

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel)
{
    uint32_t d;
    asm("prmt.b32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(sel));
    return d;
}

// one row of an interior tile: thread t holds columns c0 + 8 t + e
__device__ __forceinline__ void tile_row_scan(const TileRow &R, int r, int32_t ci, const int32_t (&cj)[8], uint32_t (&colmin)[8],
                                              uint32_t jl0, uint32_t *s_rowmin8)
{
    if (R.s < 0) return;                            // uniform
    const uint4 lo = R.lo, hi = R.hi;
    uint32_t W[4];
    switch (R.s) {                                  // uniform
    case 0: W[0] = lo.x; W[1] = lo.y; W[2] = lo.z; W[3] = lo.w; break;
    case 1: W[0] = __funnelshift_r(lo.x, lo.y, 16); W[1] = __funnelshift_r(lo.y, lo.z, 16);
            W[2] = __funnelshift_r(lo.z, lo.w, 16); W[3] = __funnelshift_r(lo.w, hi.x, 16); break;
    case 2: W[0] = lo.y; W[1] = lo.z; W[2] = lo.w; W[3] = hi.x; break;
    case 3: W[0] = __funnelshift_r(lo.y, lo.z, 16); W[1] = __funnelshift_r(lo.z, lo.w, 16);
            W[2] = __funnelshift_r(lo.w, hi.x, 16); W[3] = __funnelshift_r(hi.x, hi.y, 16); break;
    case 4: W[0] = lo.z; W[1] = lo.w; W[2] = hi.x; W[3] = hi.y; break;
    case 5: W[0] = __funnelshift_r(lo.z, lo.w, 16); W[1] = __funnelshift_r(lo.w, hi.x, 16);
            W[2] = __funnelshift_r(hi.x, hi.y, 16); W[3] = __funnelshift_r(hi.y, hi.z, 16); break;
    case 6: W[0] = lo.w; W[1] = hi.x; W[2] = hi.y; W[3] = hi.z; break;
    default: W[0] = __funnelshift_r(lo.w, hi.x, 16); W[1] = __funnelshift_r(hi.x, hi.y, 16);
             W[2] = __funnelshift_r(hi.y, hi.z, 16); W[3] = __funnelshift_r(hi.z, hi.w, 16); break;
    }
    // keys by PRMT straight from the packed words: column key (w << 16) | r; row key (w << 16) | idx(e)
    // with idx = e for e < 4 and 0x100 + (e - 4) above, both bytes taken from ONE constant register
    // (0x03020100: byte k = k), so that no per-column index registers are needed
    const uint32_t IDX = 0x03020100u;
    uint32_t rk = TS_KEY_INF;
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        if (cj[e] != ci) {
            const uint32_t wsel = (e & 1) ? 0x3200u : 0x1000u;
            colmin[e] = min(colmin[e], prmt(W[e >> 1], (uint32_t)r, wsel | 0x54u));
            rk = min(rk, prmt(W[e >> 1], IDX, wsel | ((e < 4 ? 4u : 5u) << 4) | (4u + (e & 3))));
        }
    }
    if (rk != TS_KEY_INF) rk = (rk & 0xFFFF0000u) | (jl0 + (rk & 3u) + ((rk >> 6) & 4u));
    rk = __reduce_min_sync(0xffffffffu, rk);
    if ((threadIdx.x & 31) == 0) s_rowmin8[r * 8 + (threadIdx.x >> 5)] = rk;      // plain store: one slot per (row, warp)
}

__device__ __forceinline__ void scan_tile_body_vec(const uint16_t *__restrict__ cond, const int64_t *s_seg,
                                                   const int32_t *s_ci, uint32_t *s_rowmin8,
                                                   const int32_t (&cj)[8], uint32_t (&colmin)[8])
{
    const int tid = threadIdx.x;
    const uint32_t jl0 = 8u * (uint32_t)tid;
    TileRow A = tile_row_load(cond, s_seg, 0, tid);
#pragma unroll 1
    for (int r = 0; r < TS_ROWS; r += 2) {          // two rows per trip: A / B swap roles without register copies
        const TileRow B = tile_row_load(cond, s_seg, r + 1, tid);
        tile_row_scan(A, r, s_ci[r], cj, colmin, jl0, s_rowmin8);
        if (r + 2 < TS_ROWS) A = tile_row_load(cond, s_seg, r + 2, tid);
        tile_row_scan(B, r + 1, s_ci[r + 1], cj, colmin, jl0, s_rowmin8);
    }
}

__global__ void __launch_bounds__(256)
rthc_boruvka_scan_tile(const uint16_t *__restrict__ cond, const int64_t *__restrict__ row_base, int64_t n,
                       const int32_t *__restrict__ comp, uint64_t *best, const int32_t *__restrict__ bands,
                       const unsigned int *__restrict__ state)
{
    if (state[BST_DONE]) return;                            // uniform: a queued round that is not needed
    // bands: the 64-row bands that hold rows of this shard (a sharded context owns 1/world of them)
    const int64_t r0 = (int64_t)bands[blockIdx.y] * TS_ROWS, c0 = (int64_t)blockIdx.x * TS_COLS;
    if (c0 + TS_COLS - 1 <= r0) return;                     // no column j > i in this tile
    __shared__ int64_t s_off[TS_ROWS];
    __shared__ int32_t s_ci[TS_ROWS];
    __shared__ uint32_t s_rowmin[TS_ROWS];
    __shared__ int64_t s_seg[TS_ROWS];            // interior tiles: aligned segment start | misalignment
    __shared__ uint32_t s_rowmin8[TS_ROWS * 8];   // interior tiles: row minimum of every warp
    __shared__ int32_t tab_c[TS_TAB];
    __shared__ unsigned long long tab_k[TS_TAB];
    const int tid = threadIdx.x;
    for (int q = tid; q < TS_TAB; q += 256) { tab_c[q] = -1; tab_k[q] = RTHC_KEY_INF; }
    if (tid < TS_ROWS) {
        const int64_t i = r0 + tid;
        int64_t off = -n - 1;
        int32_t ci = -1;
        if (i < n - 1) {
            const int64_t base = row_base[i];
            if (base >= 0) { off = base - i - 1; ci = comp[i]; }
        }
        s_off[tid] = off; s_ci[tid] = ci; s_rowmin[tid] = TS_KEY_INF;
        s_seg[tid] = off < -n ? (int64_t)-1 : off + c0;        // (off + c0): low 3 bits = misalignment of the segment
    }
    // edge tiles (diagonal / right border): strided columns c0 + t + 256 k with bounds checks;
    // interior tiles: contiguous columns c0 + 8 t + k, 16-byte loads
    const bool edge = (c0 <= r0 + TS_ROWS - 1) || (c0 + TS_COLS > n);
    const int cmul = edge ? 1 : 8, kmul = edge ? 256 : 1;
    int32_t cj[8];
    uint32_t colmin[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const int64_t j = c0 + cmul * tid + kmul * k;
        cj[k] = j < n ? comp[j] : -2;
        colmin[k] = TS_KEY_INF;
    }
    __syncthreads();
    if (edge) scan_tile_body<true>(cond, s_off, s_ci, s_rowmin, r0, c0, n, cj, colmin);
    else scan_tile_body_vec(cond, s_seg, s_ci, s_rowmin8, cj, colmin);
    __syncthreads();
    if (!edge && tid < TS_ROWS) {
        uint32_t m = TS_KEY_INF;
#pragma unroll
        for (int w = 0; w < 8; ++w) m = min(m, s_rowmin8[tid * 8 + w]);
        s_rowmin[tid] = s_seg[tid] < 0 ? TS_KEY_INF : m;       // absent rows were never written
    }
    // candidates of this tile -> at most one global operation per distinct component
    {
        bool has = false;
        int32_t c = -1;
        uint64_t key = RTHC_KEY_INF;
        if (tid < TS_ROWS) {                                   // warps 0 and 1, convergent
            const uint32_t rm = s_rowmin[tid];
            if (rm != TS_KEY_INF) {
                has = true;
                c = s_ci[tid];
                key = rthc_pack_key(rm >> 16, (uint32_t)(r0 + tid), (uint32_t)(c0 + (rm & 0xFFFFu)));
            }
            tile_push(has, c, key, tab_c, tab_k, best);
        }
    }
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        const uint32_t cm = colmin[k];
        const bool has = cm != TS_KEY_INF;
        const uint64_t key = rthc_pack_key(cm >> 16, (uint32_t)(r0 + (cm & 0xFFFFu)), (uint32_t)(c0 + cmul * tid + kmul * k));
        tile_push(has, cj[k], key, tab_c, tab_k, best);
    }
    __syncthreads();
    for (int slot = tid; slot < TS_TAB; slot += 256) {
        const int32_t c = tab_c[slot];
        if (c != -1) best_push(best, c, tab_k[slot]);
    }
}

// same round step over an explicit edge list (merge of the ranks' forests)
__global__ void rthc_boruvka_scan_edges(const uint64_t *__restrict__ keys, int64_t m,
                                        const int32_t *__restrict__ comp, uint64_t *best,
                                        const unsigned int *__restrict__ state)
{
    int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= m || state[BST_DONE]) return;
    const uint64_t key = keys[e];
    if (key == RTHC_KEY_INF) return;                // padding of a fixed-size gather
    const int32_t ci = comp[rthc_key_i(key)], cj = comp[rthc_key_j(key)];
    if (ci == cj) return;
    if (key < best[ci]) atomicMin((unsigned long long *)&best[ci], (unsigned long long)key);
    if (key < best[cj]) atomicMin((unsigned long long *)&best[cj], (unsigned long long)key);
}

// every component root hooks along its lightest edge; a mutual pair keeps the smaller id as root
__global__ void rthc_boruvka_hook(const uint64_t *__restrict__ best, const int32_t *__restrict__ comp,
                                  int32_t *__restrict__ parent, int64_t n, uint64_t *__restrict__ out_keys,
                                  unsigned int *out_count)
{
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n || out_count[BST_DONE]) return;
    if (comp[c] != (int32_t)c) { parent[c] = comp[c]; return; }
    const uint64_t key = best[c];
    if (key == RTHC_KEY_INF) { parent[c] = (int32_t)c; return; }
    RTHC_DEV_ASSERT(rthc_key_i(key) < rthc_key_j(key) && (int64_t)rthc_key_j(key) < n);
    const int32_t ci = comp[rthc_key_i(key)], cj = comp[rthc_key_j(key)];
    RTHC_DEV_ASSERT((ci == (int32_t)c) != (cj == (int32_t)c));       // the lightest edge LEAVES the component
    const int32_t other = (ci == (int32_t)c) ? cj : ci;
    const bool mutual = best[other] == key;
    if (mutual && (int32_t)c < other) { parent[c] = (int32_t)c; return; }
    parent[c] = other;
    out_keys[atomicAdd(out_count, 1u)] = key;
}

__global__ void rthc_pointer_jump(int32_t *parent, int64_t n, const unsigned int *__restrict__ state)
{
    int64_t c = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (c >= n || state[BST_DONE]) return;
    const int32_t p = parent[c];
    const int32_t pp = parent[p];
    if (pp != p) parent[c] = pp;
}
// comp[v] = root of v in the hooked forest.  Every thread chases its own chain (read-only, so no
// ordering issues); chains are short because each round at least halves the components and a
// hooked root points straight at a previous root.
__global__ void rthc_relabel(int32_t *comp, const int32_t *__restrict__ parent, int64_t n, const unsigned int *__restrict__ state)
{
    int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v >= n || state[BST_DONE]) return;
    int32_t r = parent[v];
    for (;;) {
        const int32_t p = parent[r];
        if (p == r) break;
        r = p;
    }
    comp[v] = r;
}

__global__ void rthc_bitonic_step(uint64_t *a, uint32_t j, uint32_t k, uint32_t N)
{
    uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N) return;
    uint32_t ixj = i ^ j;
    if (ixj > i) {
        uint64_t x = a[i], y = a[ixj];
        bool asc = (i & k) == 0;
        if ((x > y) == asc) { a[i] = y; a[ixj] = x; }
    }
}

// every bitonic step whose partner distance j is below BS_CHUNK stays inside one aligned chunk of BS_CHUNK keys:
// one CTA runs all of them for k = k_first .. k_last in shared memory (one pair per thread per step)
#define BS_CHUNK 2048u
__global__ void __launch_bounds__(BS_CHUNK / 2) rthc_bitonic_local(uint64_t *a, uint32_t N, uint32_t k_first, uint32_t k_last)
{
    __shared__ uint64_t s[BS_CHUNK];
    const uint32_t base = blockIdx.x * BS_CHUNK, tid = threadIdx.x;
    for (uint32_t t = tid; t < BS_CHUNK; t += BS_CHUNK / 2) s[t] = base + t < N ? a[base + t] : RTHC_KEY_INF;
    __syncthreads();
    for (uint32_t k = k_first; k <= k_last; k <<= 1) {
        for (uint32_t j = min(k >> 1, BS_CHUNK >> 1); j > 0; j >>= 1) {
            const uint32_t lo = ((tid & ~(j - 1)) << 1) | (tid & (j - 1)), hi = lo | j;
            const bool asc = ((base + lo) & k) == 0;
            const uint64_t x = s[lo], y = s[hi];
            if ((x > y) == asc) { s[lo] = y; s[hi] = x; }
            __syncthreads();
        }
    }
    for (uint32_t t = tid; t < BS_CHUNK; t += BS_CHUNK / 2)
        if (base + t < N) a[base + t] = s[t];
}

// ascending sort of m packed keys: returns a new pool buffer padded to a power of two with +inf keys
static int sort_keys_device(rthc_ctx *ctx, const uint64_t *d_keys, int64_t m, uint64_t **d_sorted_out)
{
    uint32_t N = 1;
    while ((int64_t)N < m) N <<= 1;
    uint64_t *d_pad = nullptr;
    int rc = rthc_dev_alloc(ctx, &d_pad, (size_t)N);
    if (rc) return rc;
    if ((int64_t)N > m) {
        rthc_fill_u64<<<(unsigned)((N - m + 255) / 256), 256, 0, ctx->stream>>>(d_pad + m, (int64_t)N - m, RTHC_KEY_INF);
        RTHC_LAUNCH_CHECK(ctx);
    }
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_pad, d_keys, sizeof(uint64_t) * (size_t)m, cudaMemcpyDeviceToDevice,
                                   ctx->stream));
    const unsigned nchunks = (N + BS_CHUNK - 1) / BS_CHUNK;
    rthc_bitonic_local<<<nchunks, BS_CHUNK / 2, 0, ctx->stream>>>(d_pad, N, 2, std::min(N, BS_CHUNK));
    RTHC_LAUNCH_CHECK(ctx);
    for (uint32_t k = 2 * BS_CHUNK; k <= N && k != 0; k <<= 1) {
        for (uint32_t j = k >> 1; j >= BS_CHUNK; j >>= 1) {
            rthc_bitonic_step<<<(N + 255) / 256, 256, 0, ctx->stream>>>(d_pad, j, k, N);
            RTHC_LAUNCH_CHECK(ctx);
        }
        rthc_bitonic_local<<<nchunks, BS_CHUNK / 2, 0, ctx->stream>>>(d_pad, N, k, k);
        RTHC_LAUNCH_CHECK(ctx);
    }
    *d_sorted_out = d_pad;
    return 0;
}

// ---- contraction after the first round ------------------------------------------------------------------------------
// Round 1 (every sample hooks to its nearest neighbour) leaves a few thousand components where there were n samples
// (C4: 100,000 -> about 3,000), yet every further round would stream the whole matrix again just to find the lightest
// edge between those few components.  Instead ONE more pass reduces the matrix to the lightest edge of every PAIR of
// components (K x K keys under the same (w, i, j) order), and the remaining rounds scan that edge list (tens of MB
// instead of 10 GB each).  Borůvka's later rounds only ever pick, for a merged component, the lightest edge to another
// one, and that edge is the lightest edge of some pair of round-1 components -- so the forest is the same, edge for edge.
//   rows are grouped by component (counting sort) so that a CTA keeps the K candidate keys of "its" component in
//   shared memory and writes them to the K x K table once per run of rows.
#define PM_ROWS 64
#define PM_KMAX 8192

// dense[r] = rank of root r among the roots (single CTA), *k_out = number of roots
__global__ void __launch_bounds__(1024) rthc_pm_dense_roots(const int32_t *__restrict__ comp, int64_t n, int32_t *__restrict__ dense,
                                                            int32_t *__restrict__ k_out)
{
    __shared__ uint32_t s_w[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int64_t per = (n + 1023) / 1024, v0 = min(n, (int64_t)tid * per), v1 = min(n, v0 + per);
    uint32_t cnt = 0;
    for (int64_t v = v0; v < v1; ++v) cnt += comp[v] == (int32_t)v;
    uint32_t inc = cnt;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    uint32_t run = inc - cnt;
    for (int w = 0; w < warp; ++w) run += s_w[w];
    for (int64_t v = v0; v < v1; ++v)
        if (comp[v] == (int32_t)v) dense[v] = (int32_t)run++;
    if (tid == 1023) *k_out = (int32_t)run;
}
__global__ void rthc_pm_comp16(const int32_t *__restrict__ comp, const int32_t *__restrict__ dense, int64_t n,
                               uint16_t *__restrict__ comp16)
{
    const int64_t v = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (v < n) comp16[v] = (uint16_t)dense[comp[v]];
}
// rows with storage (owned, i < n - 1) counted per component, then placed (two launches around a scan)
__global__ void rthc_pm_count(const int32_t *__restrict__ rows, int64_t n_rows, int64_t n, const uint16_t *__restrict__ comp16,
                              uint32_t *__restrict__ hist)
{
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n_rows && rows[q] < n - 1) atomicAdd(&hist[comp16[rows[q]]], 1u);
}
__global__ void __launch_bounds__(1024) rthc_pm_scan(uint32_t *hist, int K, uint32_t *total)      // exclusive, in place, K <= PM_KMAX
{
    __shared__ uint32_t s_w[32];
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    constexpr int PER = PM_KMAX / 1024;
    uint32_t v[PER], sum = 0;
#pragma unroll
    for (int u = 0; u < PER; ++u) { const int e = tid * PER + u; v[u] = e < K ? hist[e] : 0u; sum += v[u]; }
    uint32_t inc = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, inc, o);
        if (lane >= o) inc += t;
    }
    if (lane == 31) s_w[warp] = inc;
    __syncthreads();
    uint32_t run = inc - sum;
    for (int w = 0; w < warp; ++w) run += s_w[w];
#pragma unroll
    for (int u = 0; u < PER; ++u) { const int e = tid * PER + u; if (e < K) hist[e] = run; run += v[u]; }
    if (tid == 1023) *total = run;
}
__global__ void rthc_pm_place(const int32_t *__restrict__ rows, int64_t n_rows, int64_t n, const uint16_t *__restrict__ comp16,
                              uint32_t *__restrict__ cursor, int32_t *__restrict__ order)
{
    const int64_t q = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (q < n_rows && rows[q] < n - 1) order[atomicAdd(&cursor[comp16[rows[q]]], 1u)] = rows[q];
}

// lightest edge from the component of a run of rows to every other component.  A CTA takes PM_ROWS grouped rows; for
// every run of rows of one component it first reduces over the ROWS in registers (per column the smallest
// (w, row) -- the column-minimum half of the tile scan, no shared memory in the inner loop), then merges the columns
// of equal component through the K-entry table in shared memory, and writes the table to M[ci][*] once per run.
__global__ void __launch_bounds__(256)
rthc_pm_pair_min(const uint16_t *__restrict__ cond, const int64_t *__restrict__ row_base, int64_t n,
                 const uint16_t *__restrict__ comp16, const int32_t *__restrict__ order, const uint32_t *__restrict__ n_rows_p,
                 int K, unsigned long long *__restrict__ M)
{
    extern __shared__ unsigned long long pm_tab[];
    __shared__ int32_t s_row[PM_ROWS];
    __shared__ int64_t s_base[PM_ROWS];
    const int tid = threadIdx.x;
    const int64_t n_rows = *n_rows_p;
    const int64_t q0 = (int64_t)blockIdx.x * PM_ROWS, q1 = min(n_rows, q0 + PM_ROWS);
    if (q0 >= n_rows) return;
    for (int e = tid; e < K; e += 256) pm_tab[e] = RTHC_KEY_INF;
    int64_t q = q0;
    while (q < q1) {                                           // one run of rows of the same component (uniform)
        const int ci = comp16[order[q]];
        int64_t qe = q + 1;
        while (qe < q1 && (int)comp16[order[qe]] == ci) ++qe;
        const int nr = (int)(qe - q);
        __syncthreads();                                       // table reset / previous run's rows no longer in use
        if (tid < nr) {                                        // rows in ascending order: ties in w go to the smaller row
            const int32_t my = order[q + tid];
            int rank = 0;
            for (int t = 0; t < nr; ++t) rank += order[q + t] < my;
            s_row[rank] = my;
            s_base[rank] = row_base[my] - my - 1;              // (cond + base)[j] = d(my, j) for j > my
        }
        __syncthreads();
        const int64_t i_lo = s_row[0];
        int r_full = 0;                                        // rows entirely left of the current column block (rows ascend)
        for (int64_t jb = i_lo + 1; jb < n; jb += 8 * 256) {
            uint32_t colmin[8];
#pragma unroll
            for (int e = 0; e < 8; ++e) colmin[e] = 0xFFFFFFFFu;
            const int64_t j0 = jb + tid;
            const bool whole = jb + 8 * 256 <= n;              // no column of the block is past the end
            while (r_full < nr && s_row[r_full] < jb) ++r_full;
            int r = 0;
            if (whole) {                                       // rows left of the block: plain streaming, 16 loads in flight
#pragma unroll 2
                for (; r < r_full; ++r) {
                    const uint16_t *row = cond + s_base[r] + j0;
                    uint32_t w[8];
#pragma unroll
                    for (int e = 0; e < 8; ++e) w[e] = row[256 * e];
#pragma unroll
                    for (int e = 0; e < 8; ++e) colmin[e] = min(colmin[e], (w[e] << 16) | (uint32_t)r);
                }
            }
            for (; r < nr; ++r) {                              // rows that start inside the block (or the last block): checked
                const int64_t i = s_row[r];
                if (i >= jb + 8 * 256 - 1) break;              // this row and the following ones start right of the block
                const uint16_t *row = cond + s_base[r] + j0;
#pragma unroll
                for (int e = 0; e < 8; ++e) {
                    const int64_t j = j0 + 256 * e;
                    if (j > i && j < n) colmin[e] = min(colmin[e], ((uint32_t)row[256 * e] << 16) | (uint32_t)r);
                }
            }
#pragma unroll
            for (int e = 0; e < 8; ++e) {
                if (colmin[e] == 0xFFFFFFFFu) continue;
                const int64_t j = j0 + 256 * e;
                const int cj = comp16[j];
                if (cj == ci) continue;
                const unsigned long long key = rthc_pack_key(colmin[e] >> 16, (uint32_t)s_row[colmin[e] & 0xFFFFu], (uint32_t)j);
                if (key < pm_tab[cj]) atomicMin(&pm_tab[cj], key);
            }
        }
        __syncthreads();
        for (int e = tid; e < K; e += 256) {
            const unsigned long long k = pm_tab[e];
            if (k != RTHC_KEY_INF) { atomicMin(&M[(size_t)ci * K + e], k); pm_tab[e] = RTHC_KEY_INF; }
        }
        q = qe;
    }
}
// M is filled from the side of the row's component only: the lightest edge of the pair {a, b} is the smaller of
// M[a][b] and M[b][a]; pairs without an edge (a shard sees only its rows) are skipped
__global__ void rthc_pm_extract(const unsigned long long *__restrict__ M, int K, unsigned long long *__restrict__ edges,
                                unsigned long long *__restrict__ n_edges)
{
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (int64_t)K * K) return;
    const int a = (int)(t / K), b = (int)(t % K);
    if (a >= b) return;
    const unsigned long long k = min(M[t], M[(size_t)b * K + a]);
    if (k != RTHC_KEY_INF) edges[atomicAdd(n_edges, 1ull)] = k;
}

// after round 1: returns 0 and sets *d_edges / *m_edges (inside the context's scratch block) when the contraction applies,
// 1 when it does not
static int pair_min_contract(rthc_ctx *ctx, int64_t n, const int32_t *d_comp, int64_t comps, uint64_t **d_edges, int64_t *m_edges)
{
    const int64_t n_owned = (int64_t)ctx->h_owned_rows.size();
    if (comps < 2 || comps > PM_KMAX || n_owned == 0 || !ctx->d_cond) return 1;
    const int K = (int)comps;
    // one scratch block that stays with the context (grow only): blocks of tens of MB that come and go between the
    // GB-sized ones make the stream-ordered pool split and re-grow its large blocks (6 ms per call measured here)
    auto up = [](size_t b) { return (b + 255) / 256 * 256; };
    const size_t o_dense = 0, o_comp16 = o_dense + up(4 * (size_t)n), o_hist = o_comp16 + up(2 * (size_t)n),
                 o_cursor = o_hist + up(4 * ((size_t)K + 1)), o_order = o_cursor + up(4 * ((size_t)K + 1)),
                 o_k = o_order + up(4 * (size_t)n_owned), o_ne = o_k + 256, o_M = o_ne + 256,
                 o_e = o_M + up(8 * (size_t)K * K), total = o_e + up(8 * ((size_t)K * (K - 1) / 2 + 1));
    if (ctx->pm_cap < total) {
        ctx->pm_cap = 0;
        if (rthc_dev_alloc(ctx, &ctx->d_pm, total + total / 4)) { ctx->err.clear(); return 1; }     // no room: plain rounds
        ctx->pm_cap = total + total / 4;
    }
    unsigned char *base = ctx->d_pm;
    int32_t *d_dense = reinterpret_cast<int32_t *>(base + o_dense), *d_k = reinterpret_cast<int32_t *>(base + o_k);
    int32_t *d_order = reinterpret_cast<int32_t *>(base + o_order);
    uint16_t *d_comp16 = reinterpret_cast<uint16_t *>(base + o_comp16);
    uint32_t *d_hist = reinterpret_cast<uint32_t *>(base + o_hist), *d_cursor = reinterpret_cast<uint32_t *>(base + o_cursor);
    unsigned long long *d_M = reinterpret_cast<unsigned long long *>(base + o_M);
    unsigned long long *d_e = reinterpret_cast<unsigned long long *>(base + o_e);
    unsigned long long *d_ne = reinterpret_cast<unsigned long long *>(base + o_ne);
    const unsigned nb = (unsigned)((n + 255) / 256), nbo = (unsigned)((n_owned + 255) / 256);
    rthc_pm_dense_roots<<<1, 1024, 0, ctx->stream>>>(d_comp, n, d_dense, d_k);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_pm_comp16<<<nb, 256, 0, ctx->stream>>>(d_comp, d_dense, n, d_comp16);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemsetAsync(d_hist, 0, sizeof(uint32_t) * ((size_t)K + 1), ctx->stream));
    rthc_pm_count<<<nbo, 256, 0, ctx->stream>>>(ctx->d_owned_rows, n_owned, n, d_comp16, d_hist);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_pm_scan<<<1, 1024, 0, ctx->stream>>>(d_hist, K, d_hist + K);          // hist[K] = rows in all
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_cursor, d_hist, sizeof(uint32_t) * ((size_t)K + 1), cudaMemcpyDeviceToDevice, ctx->stream));
    rthc_pm_place<<<nbo, 256, 0, ctx->stream>>>(ctx->d_owned_rows, n_owned, n, d_comp16, d_cursor, d_order);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemsetAsync(d_M, 0xFF, sizeof(unsigned long long) * (size_t)K * K, ctx->stream));
    RTHC_CUDA(ctx, cudaMemsetAsync(d_ne, 0, sizeof(unsigned long long), ctx->stream));
    const size_t smem = sizeof(unsigned long long) * (size_t)K;
    RTHC_CUDA(ctx, cudaFuncSetAttribute(rthc_pm_pair_min, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    rthc_pm_pair_min<<<(unsigned)((n_owned + PM_ROWS - 1) / PM_ROWS), 256, smem, ctx->stream>>>(
        ctx->d_cond, ctx->d_row_base, n, d_comp16, d_order, d_hist + K, K, d_M);
    RTHC_LAUNCH_CHECK(ctx);
    rthc_pm_extract<<<(unsigned)(((int64_t)K * K + 255) / 256), 256, 0, ctx->stream>>>(d_M, K, d_e, d_ne);
    RTHC_LAUNCH_CHECK(ctx);
    unsigned long long ne = 0;
    int32_t k_dev = 0;
    RTHC_CUDA(ctx, cudaMemcpyAsync(&ne, d_ne, sizeof(ne), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(&k_dev, d_k, sizeof(k_dev), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (k_dev != K) return rthc_fail(ctx, RTHC_E_STATE, "contraction: %d components on the device, %d expected", (int)k_dev, K);
    *d_edges = reinterpret_cast<uint64_t *>(d_e);
    *m_edges = (int64_t)ne;
    return 0;
}

// shared round loop: `scan` launches the lightest-edge pass for the current components.  The forest stays on the
// device: *d_out (caller frees) holds *m_out keys, ascending when `sorted`.
template <typename ScanFn>
static int boruvka_rounds(rthc_ctx *ctx, int64_t n, ScanFn scan, bool sorted, int *hint, uint64_t **d_out, int64_t *m_out,
                          int64_t full_count = -1, bool contract = false)
{
    // contract: after the first round the matrix may be reduced to the lightest edge per pair of components
    // (pair_min_contract), which the remaining rounds scan instead of calling `scan`
    // full_count: edges of the complete forest when the caller knows it (n - 1 for a connected edge set; fewer for a
    // shard whose first rows belong to other ranks: their samples are isolated here).  Without it a disconnected
    // edge set needs one more round -- one more pass over the matrix -- just to see that nothing is added.
    const int64_t stop_n = full_count >= 0 ? full_count + 1 : n;
    int32_t *d_comp = nullptr, *d_parent = nullptr;
    uint64_t *d_best = nullptr, *d_keys = nullptr;
    unsigned int *d_state = nullptr;
    int rc;
    *d_out = nullptr; *m_out = 0;
    if ((rc = rthc_dev_alloc(ctx, &d_comp, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_parent, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_best, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_keys, (size_t)n))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_state, 4))) return rc;
    const unsigned nb = (unsigned)((n + 255) / 256);
    rthc_iota_i32<<<nb, 256, 0, ctx->stream>>>(d_comp, n);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemsetAsync(d_state, 0, 4 * sizeof(unsigned int), ctx->stream));
    // Rounds are queued without waiting for each other: as many as the previous call of this kind needed (a queued
    // round that turns out not to be needed costs nine empty launches), then one by one until the device says done.
    unsigned int st[4] = {0, 0, 0, 0};
    int batch = contract ? 1 : std::max(1, *hint);
    bool contraction_tried = !contract;
    uint64_t *d_pair = nullptr;
    int64_t m_pair = 0;
    while (!st[BST_DONE] && (int64_t)st[BST_COUNT] < stop_n - 1) {
        for (int r = 0; r < batch; ++r) {
            rthc_fill_u64_round<<<nb, 256, 0, ctx->stream>>>(d_best, n, RTHC_KEY_INF, d_state);
            RTHC_LAUNCH_CHECK(ctx);
            if (d_pair) {
                if (m_pair > 0) {
                    rthc_boruvka_scan_edges<<<(unsigned)((m_pair + 255) / 256), 256, 0, ctx->stream>>>(d_pair, m_pair, d_comp, d_best, d_state);
                    RTHC_LAUNCH_CHECK(ctx);
                }
            } else if ((rc = scan(d_comp, d_best, d_state))) return rc;
            rthc_boruvka_hook<<<nb, 256, 0, ctx->stream>>>(d_best, d_comp, d_parent, n, d_keys, d_state);
            RTHC_LAUNCH_CHECK(ctx);
            rthc_boruvka_round_end<<<1, 1, 0, ctx->stream>>>(d_state, stop_n);
            RTHC_LAUNCH_CHECK(ctx);
            for (int it = 0; it < 4; ++it) {      // 16x path compression bounds the chase below on adversarial chains
                rthc_pointer_jump<<<nb, 256, 0, ctx->stream>>>(d_parent, n, d_state);
                RTHC_LAUNCH_CHECK(ctx);
            }
            rthc_relabel<<<nb, 256, 0, ctx->stream>>>(d_comp, d_parent, n, d_state);
            RTHC_LAUNCH_CHECK(ctx);
        }
        RTHC_CUDA(ctx, cudaMemcpyAsync(st, d_state, sizeof(st), cudaMemcpyDeviceToHost, ctx->stream));
        RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        batch = 1;
        if (!contraction_tried && !st[BST_DONE] && (int64_t)st[BST_COUNT] < stop_n - 1) {
            contraction_tried = true;
            rc = pair_min_contract(ctx, n, d_comp, n - (int64_t)st[BST_COUNT], &d_pair, &m_pair);
            if (rc < 0) return rc;
            batch = std::max(1, *hint - 1);                   // the rounds the previous call needed after its first
        }
    }
    const unsigned int total = st[BST_COUNT];
    ctx->boruvka_rounds = (int)st[BST_ROUNDS];
    *hint = (int)st[BST_ROUNDS];
    rthc_dev_free(ctx, &d_comp); rthc_dev_free(ctx, &d_parent); rthc_dev_free(ctx, &d_best);
    rthc_dev_free(ctx, &d_state);
    if (sorted && total > 0) {
        uint64_t *d_sorted = nullptr;
        if ((rc = sort_keys_device(ctx, d_keys, (int64_t)total, &d_sorted))) return rc;
        rthc_dev_free(ctx, &d_keys);
        d_keys = d_sorted;
    }
    *d_out = d_keys;
    *m_out = (int64_t)total;
    return 0;
}

int rthc_keys_to_host(rthc_ctx *ctx, const uint64_t *d_keys, int64_t m, const uint64_t **h_keys)
{
    uint64_t *h = (uint64_t *)rthc_pinned(ctx, 0, sizeof(uint64_t) * (size_t)std::max<int64_t>(m, 1));
    if (!h) return RTHC_E_NOMEM;
    if (m > 0)
        RTHC_CUDA(ctx, cudaMemcpyAsync(h, d_keys, sizeof(uint64_t) * (size_t)m, cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    *h_keys = h;
    return 0;
}

int rthc_boruvka_owned(rthc_ctx *ctx, bool sorted, uint64_t **d_keys, int64_t *m)
{
    const int64_t n = ctx->n;
    const int64_t n_owned = (int64_t)ctx->h_owned_rows.size();
    std::vector<int32_t> bands;
    for (int32_t i : ctx->h_owned_rows)
        if (bands.empty() || bands.back() != i / TS_ROWS) bands.push_back(i / TS_ROWS);      // rows are ascending
    int32_t *d_bands = nullptr;
    int rcb = rthc_dev_alloc(ctx, &d_bands, bands.size());
    if (rcb) return rcb;
    if (!bands.empty())
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_bands, bands.data(), sizeof(int32_t) * bands.size(), cudaMemcpyHostToDevice, ctx->stream));
    auto scan = [&](int32_t *d_comp, uint64_t *d_best, const unsigned int *d_state) -> int {
        if (n_owned > 0) {
            dim3 grd((unsigned)((n + TS_COLS - 1) / TS_COLS), (unsigned)bands.size());
            rthc_boruvka_scan_tile<<<grd, 256, 0, ctx->stream>>>(ctx->d_cond, ctx->d_row_base, n, d_comp, d_best, d_bands, d_state);
            RTHC_LAUNCH_CHECK(ctx);
        }
        return 0;
    };
    // the owned rows i reach every sample j > i, so the shard's forest spans exactly the samples from its first row on
    const int64_t full_count = n_owned > 0 ? n - 1 - (int64_t)ctx->h_owned_rows.front() : 0;
    // The contraction pays when a pass over the owned rows is long against its own fixed costs (eight small launches,
    // one more host round trip, and K table entries written per run of rows, which does not shrink with the shard):
    // C4 measured 16.4 -> 7.8 ms on one GPU, 6.8 -> 4.6 on two, 4.5 -> 3.4 on four, but 2.5 -> 3.3 ms on eight
    // (6.2e8 owned pairs).  RTHC_PAIRMIN_MIN_N (test hook) replaces the rule by "n at least this".
    const char *env = getenv("RTHC_PAIRMIN_MIN_N");
    const bool contract = env ? n >= atoll(env) : ctx->owned_pairs >= 1000000000ll;
    const int rc = boruvka_rounds(ctx, n, scan, sorted, &ctx->boruvka_hint_owned, d_keys, m, full_count, contract);
    rthc_dev_free(ctx, &d_bands);
    return rc;
}

// d_in: m_in packed edges on the device, any order; +inf entries (padding of a fixed-size gather) are skipped
int rthc_boruvka_edge_list(rthc_ctx *ctx, const uint64_t *d_in, int64_t m_in, int64_t n, uint64_t **d_keys, int64_t *m)
{
    auto scan = [&](int32_t *d_comp, uint64_t *d_best, const unsigned int *d_state) -> int {
        if (m_in > 0) {
            rthc_boruvka_scan_edges<<<(unsigned)((m_in + 255) / 256), 256, 0, ctx->stream>>>(d_in, m_in, d_comp, d_best, d_state);
            RTHC_LAUNCH_CHECK(ctx);
        }
        return 0;
    };
    return boruvka_rounds(ctx, n, scan, true, &ctx->boruvka_hint_edges, d_keys, m);
}
