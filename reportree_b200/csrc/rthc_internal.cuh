// rthc_internal.cuh -- context layout and helpers shared by the kernels of libreportree_hc.so
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string>
#include <vector>

#include "../../include/rthc.h"

#define RTHC_TILE 128            // samples per tile-row (ownership granularity) of the distance kernel
#define RTHC_BLK 64              // samples per plane block in HBM / per bulk copy
#define RTHC_MIN_BITS 3          // every locus group carries at least 3 bit planes (zero-padded)
#define RTHC_ESC_CAP 4096        // shared codes (collide.cu): at most this many cells of a locus hold a shared code and are listed
#define RTHC_KEY_INF 0xFFFFFFFFFFFFFFFFull
#define RTHC_COND_PAD 64         // spare uint16 entries behind d_cond: 16-byte row loads may run 15 entries past a segment

// (w, i, j) packed as w<<48 | i<<24 | j  (i < j < 2^24, w < 2^16): one total order
__host__ __device__ inline uint64_t rthc_pack_key(uint32_t w, uint32_t i, uint32_t j)
{
    return ((uint64_t)w << 48) | ((uint64_t)i << 24) | (uint64_t)j;
}
__host__ __device__ inline uint32_t rthc_key_w(uint64_t k) { return (uint32_t)(k >> 48); }
__host__ __device__ inline uint32_t rthc_key_i(uint64_t k) { return (uint32_t)((k >> 24) & 0xFFFFFFu); }
__host__ __device__ inline uint32_t rthc_key_j(uint64_t k) { return (uint32_t)(k & 0xFFFFFFu); }

struct rthc_ctx {
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;            // label download of one threshold piece under the kernels of the next (cut.cu)
    cudaEvent_t ev_piece[4] = {nullptr, nullptr, nullptr, nullptr};
    cudaMemPool_t pool = nullptr;   // stream-ordered allocation pool of this context
    std::string err;
    int sm_count = 148;

    // shard
    int rank = 0, world = 1;
    bool full_storage = false;     // sharded tiles, but d_cond is the FULL condensed vector (single-process multi-device)

    // problem
    int64_t n = 0, L = 0, n_pad = 0;
    int64_t n_tiles_edge = 0;      // ceil(n / 128)
    int64_t L_kept = 0;
    int G = 0;                     // groups of 32 loci
    int plane_rows = 0;            // sum over groups of (bits + 1)
    int max_np = 0;                // max planes per group (bits + valid)
    int max_alleles = 0;

    // device buffers (distance half)
    uint16_t *d_codes = nullptr;   // [L][n_pad] dense-ranked codes, column-major per locus
    // shared codes (collide.cu): what the plane packer reads (d_codes itself when no locus shares codes), per locus the
    // number of listed cells, the cells (sample | code << 24 | dense rank << 40), pairs incremented by the last fix-up
    uint16_t *d_pcodes = nullptr;
    int32_t *d_esc_cnt = nullptr;
    uint64_t *d_esc_list = nullptr;
    unsigned long long *d_esc_pairs = nullptr;
    unsigned long long escape_pairs_u = 0;
    int64_t escape_loci = 0, escape_cells = 0;
    int64_t geom_n = -1; int geom_rank = -1, geom_world = -1; bool geom_full = false;      // what d_tiles / d_row_base were built for
    bool recode_general = false;   // last re-code left the fast path (an allele id >= 2^14 somewhere)
    uint32_t *d_codes32 = nullptr; // same, used INSTEAD of d_codes when a locus has more than 65,535 called alleles
    uint32_t *d_planes = nullptr;  // [n_pad / 64][blk_stride_rows][64]: plane row `base` of a group, 64-sample block b, lives at row b * blk_stride_rows + base
    int32_t *d_grp_off = nullptr;  // [G] first plane row of a group inside its block (what the packer writes)
    int64_t *d_grp_base = nullptr; // [G] the same as the distance kernel wants it: row offset inside block 0 of the WHOLE buffer
    int32_t *d_grp_np = nullptr;   // [G]
    int64_t blk_stride_rows = 0;   // plane_rows for a matrix re-coded here; rows_max of the widest rank for gathered planes
    // planes re-coded by locus ranges on several ranks and gathered (rthc_recode_shard / rthc_export_shard / rthc_import_planes):
    // d_planes, d_esc_cnt and d_esc_list then point into the CALLER's gathered buffers and are not freed here
    unsigned char *d_pm = nullptr; size_t pm_cap = 0;   // scratch of the Boruvka contraction (mst.cu), kept across calls
    uint32_t *d_gbm = nullptr; size_t gbm_cap = 0;     // allele-id bitmaps of the fast re-coder path (recode.cu), kept across calls
    bool planes_borrowed = false;
    int64_t L_stride = 0;          // row stride of the caller's matrix while a column range is re-coded (0: L)
    int64_t L_global = 0, esc_L = 0;      // loci of the whole matrix; entries of d_esc_cnt
    std::vector<int32_t> h_grp_np; // plane counts of the groups built by the last re-code (header of rthc_recode_shard)
    int2 *d_tiles = nullptr;       // tile list (I, J), J >= I, owned tile rows only
    int64_t n_tiles = 0;
    int64_t *d_row_base = nullptr; // [n_pad] offset of (i, i+1) in d_cond, -1 if row not owned
    std::vector<int64_t> h_row_base;
    std::vector<int32_t> h_owned_rows;
    int32_t *d_owned_rows = nullptr;
    int64_t owned_pairs = 0;
    uint16_t *d_cond = nullptr;    // condensed uint16 distances of the owned rows
    int64_t cond_cap = 0;          // allocated entries of d_cond (kept across calls of equal size)
    bool have_profiles = false, have_dist = false;

    // externally supplied condensed matrix (-d_mx)
    double *d_cond_f64 = nullptr;
    bool ext_integer = false;

    // clustering state
    int single_algo = RTHC_SL_BORUVKA;
    int64_t m_edges = 0;           // n - 1 after a successful linkage
    int32_t *d_tpar = nullptr;     // [2n-1] dendrogram parent of every node (leaves 0..n-1, merge k = n+k), root: INT_MAX
    int32_t *d_event = nullptr;    // [2n-1] position of every node in scipy's fcluster numbering walk
    int32_t *d_up = nullptr;       // [levels][2n-1] binary-lifting table over d_tpar (built by the first cut)
    int up_levels = 0;
    std::vector<double> h_heights;
    std::vector<int32_t> h_uf, h_cid, h_nsize;     // dendrogram bookkeeping scratch (linkage.cu), kept across calls
    std::vector<int32_t> h_ex, h_ey;               // merges unpacked from packed keys
    std::vector<double> h_ew;
    void *h_pin[3] = {nullptr, nullptr, nullptr};  // page-locked staging: 0 keys (D2H), 1 tpar, 2 event (H2D)
    size_t h_pin_cap[3] = {0, 0, 0};
    bool have_linkage = false;
    int boruvka_rounds = 0;
    int boruvka_hint_owned = 1, boruvka_hint_edges = 1;   // rounds the last call of each kind needed: queued without a host round trip
    int64_t row_scans = 0;         // NN-chain: row scans executed (full + fused with a merge)
    int64_t chain_full_scans = 0;  // ... of which paid their own grid barrier
    int64_t chain_cache_hits = 0;  // chain steps answered by the nearest-neighbour cache

    // timings
    double t_ms[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    int64_t launches = 0;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr, ev_aux = nullptr;
    cudaEvent_t ev_mark[2] = {nullptr, nullptr};   // rthc_mark(): caller-placed events on the library stream
};

// Checked build (python -m reportree_b200.build --checked -> build/lib_checked.so, used by scripts/gpu_checked.sh):
// device-side invariants compiled in; a violation prints its location and traps, which the host sees as a CUDA
// error.  compute-sanitizer is not available on the GPU pool, so these asserts + parity against scipy on small and
// ragged shapes are the memory / ordering evidence.
#ifdef RTHC_CHECKED
#include <cstdio>
#define RTHC_DEV_ASSERT(cond)                                                                              \
    do {                                                                                                   \
        if (!(cond)) {                                                                                     \
            printf("RTHC_DEV_ASSERT failed %s:%d: %s (block %d, thread %d)\n", __FILE__, __LINE__, #cond,   \
                   (int)blockIdx.x, (int)threadIdx.x);                                                     \
            __trap();                                                                                      \
        }                                                                                                  \
    } while (0)
#else
#define RTHC_DEV_ASSERT(cond) ((void)0)
#endif

extern std::string g_rthc_create_err;

int rthc_fail(rthc_ctx *ctx, int code, const char *fmt, ...);

#define RTHC_CUDA(ctx, expr)                                                              \
    do {                                                                                  \
        cudaError_t _e = (expr);                                                          \
        if (_e != cudaSuccess)                                                            \
            return rthc_fail((ctx), RTHC_E_CUDA, "%s failed at %s:%d: %s", #expr, __FILE__, \
                             __LINE__, cudaGetErrorString(_e));                           \
    } while (0)

#define RTHC_LAUNCH_CHECK(ctx)                                                            \
    do {                                                                                  \
        (ctx)->launches++;                                                                \
        RTHC_CUDA((ctx), cudaGetLastError());                                             \
    } while (0)

// Device memory comes from a per-context stream-ordered pool (cudaMallocFromPoolAsync) whose
// release threshold is unlimited: after the first pass over a workload no call reaches the
// driver's allocator again (cudaMalloc / cudaFree of GB-sized buffers cost tens of ms and
// synchronise the device).  Everything is allocated, used and freed on ctx->stream.
template <typename T> inline void rthc_dev_free(rthc_ctx *ctx, T **p)
{
    if (*p) { cudaFreeAsync((void *)*p, ctx->stream); *p = nullptr; }
}
template <typename T> inline int rthc_dev_alloc(rthc_ctx *ctx, T **p, size_t count)
{
    rthc_dev_free(ctx, p);
    if (count == 0) count = 1;
    cudaError_t e = cudaMallocFromPoolAsync((void **)p, count * sizeof(T), ctx->pool, ctx->stream);
    if (e != cudaSuccess) {                 // give cached blocks back to the driver and retry once
        cudaGetLastError();
        cudaStreamSynchronize(ctx->stream);
        cudaMemPoolTrimTo(ctx->pool, 0);
        e = cudaMallocFromPoolAsync((void **)p, count * sizeof(T), ctx->pool, ctx->stream);
    }
    if (e != cudaSuccess) {
        *p = nullptr;
        cudaGetLastError();
        return rthc_fail(ctx, RTHC_E_NOMEM, "device allocation of %zu bytes failed: %s", count * sizeof(T),
                         cudaGetErrorString(e));
    }
    return 0;
}

// context state helpers (api.cu)
void rthc_free_distance_state(rthc_ctx *ctx);
void rthc_free_geometry(rthc_ctx *ctx);
void rthc_free_linkage_state(rthc_ctx *ctx);
int rthc_setup_geometry(rthc_ctx *ctx, int64_t n);

// stage entry points implemented in the other translation units
int rthc_recode_and_pack(rthc_ctx *ctx, const void *d_rowmajor, bool in_u16);   // recode.cu
int rthc_run_distance(rthc_ctx *ctx);                                 // distance.cu
int rthc_launch_distance_ws(rthc_ctx *ctx, int max_np);                // distance_ws.cu
int rthc_launch_escape_correct(rthc_ctx *ctx);                           // collide.cu
int rthc_collide_codes(rthc_ctx *ctx, int32_t *d_nall, int max_alleles, std::vector<int32_t> &nall);   // collide.cu
int rthc_run_direct_pairs(rthc_ctx *ctx, const int32_t *d_pi, const int32_t *d_pj, int64_t np,
                          int32_t *d_out);                            // distance.cu
// Boruvka results stay on the device: *d_keys (pool allocation, caller frees with rthc_dev_free) holds *m packed
// edges, ascending when `sorted`; rthc_keys_to_host copies them into the page-locked slot 0 and synchronises
int rthc_boruvka_owned(rthc_ctx *ctx, bool sorted, uint64_t **d_keys, int64_t *m);              // mst.cu
int rthc_boruvka_edge_list(rthc_ctx *ctx, const uint64_t *d_in, int64_t m_in, int64_t n, uint64_t **d_keys,
                           int64_t *m);                                                         // mst.cu
int rthc_keys_to_host(rthc_ctx *ctx, const uint64_t *d_keys, int64_t m, const uint64_t **h_keys); // mst.cu
void *rthc_pinned(rthc_ctx *ctx, int slot, size_t bytes);                                       // api.cu
int rthc_prim_exact(rthc_ctx *ctx, std::vector<int32_t> &ex, std::vector<int32_t> &ey,
                    std::vector<double> &ew);                         // mst.cu
int rthc_nn_chain(rthc_ctx *ctx, int method, std::vector<int32_t> &ex, std::vector<int32_t> &ey,
                  std::vector<double> &ew);                           // nnchain.cu
int rthc_run_cut(rthc_ctx *ctx, const double *t, int T, int32_t *labels_out);   // cut.cu
