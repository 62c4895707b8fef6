// api.cu -- the extern "C" surface declared in include/rthc.h (context, geometry, copies).
#include <stdarg.h>

#include <algorithm>
#include <cmath>
#include <cstring>
#include <cstdlib>

#include "rthc_internal.cuh"

std::string g_rthc_create_err;

int rthc_fail(rthc_ctx *ctx, int code, const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf;
    else g_rthc_create_err = buf;
    return code;
}

extern "C" int rthc_version(void) { return 100; }

extern "C" const char *rthc_last_error(const rthc_ctx *ctx)
{
    return ctx ? ctx->err.c_str() : g_rthc_create_err.c_str();
}

extern "C" int rthc_prewarm(int device)
{
    if (cudaSetDevice(device) != cudaSuccess || cudaFree(nullptr) != cudaSuccess) {
        cudaGetLastError();
        return RTHC_E_CUDA;                      // not fatal here: rthc_create reports the reason
    }
    return RTHC_OK;
}

extern "C" int rthc_create(rthc_ctx **out, int device)
{
    if (!out) return rthc_fail(nullptr, RTHC_E_ARG, "rthc_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return rthc_fail(nullptr, RTHC_E_CUDA, "no usable CUDA device (%s); this library has no CPU fallback",
                         e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count)
        return rthc_fail(nullptr, RTHC_E_ARG, "device %d out of range (0..%d)", device, count - 1);
    rthc_ctx *ctx = new rthc_ctx();
    ctx->device = device;
    if ((e = cudaSetDevice(device)) != cudaSuccess || (e = cudaStreamCreate(&ctx->stream)) != cudaSuccess ||
        (e = cudaEventCreate(&ctx->ev0)) != cudaSuccess || (e = cudaEventCreate(&ctx->ev1)) != cudaSuccess ||
        (e = cudaEventCreate(&ctx->ev_aux)) != cudaSuccess) {
        rthc_fail(nullptr, RTHC_E_CUDA, "context setup failed: %s", cudaGetErrorString(e));
        delete ctx;
        return RTHC_E_CUDA;
    }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) == cudaSuccess) {
        ctx->sm_count = prop.multiProcessorCount;
        if (prop.major < 10) {
            rthc_fail(nullptr, RTHC_E_CUDA, "device %d is sm_%d%d; this library is built for sm_100a only",
                      device, prop.major, prop.minor);
            delete ctx;
            return RTHC_E_CUDA;
        }
    }
    {
        cudaMemPoolProps props = {};
        props.allocType = cudaMemAllocationTypePinned;
        props.handleTypes = cudaMemHandleTypeNone;
        props.location.type = cudaMemLocationTypeDevice;
        props.location.id = device;
        unsigned long long keep = ~0ull;
        if ((e = cudaMemPoolCreate(&ctx->pool, &props)) != cudaSuccess ||
            (e = cudaMemPoolSetAttribute(ctx->pool, cudaMemPoolAttrReleaseThreshold, &keep)) != cudaSuccess) {
            rthc_fail(nullptr, RTHC_E_CUDA, "memory pool setup failed: %s", cudaGetErrorString(e));
            delete ctx;
            return RTHC_E_CUDA;
        }
    }
    if ((e = cudaEventCreate(&ctx->ev_mark[0])) != cudaSuccess || (e = cudaEventCreate(&ctx->ev_mark[1])) != cudaSuccess) {
        rthc_fail(nullptr, RTHC_E_CUDA, "context setup failed: %s", cudaGetErrorString(e));
        delete ctx;
        return RTHC_E_CUDA;
    }
    *out = ctx;
    return 0;
}

extern "C" int rthc_mark(rthc_ctx *ctx, int slot)
{
    if (!ctx || slot < 0 || slot > 1) return rthc_fail(ctx, RTHC_E_ARG, "rthc_mark: bad argument");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev_mark[slot], ctx->stream));
    return 0;
}

extern "C" int rthc_mark_elapsed_ms(rthc_ctx *ctx, double *ms)
{
    if (!ctx || !ms) return rthc_fail(ctx, RTHC_E_ARG, "rthc_mark_elapsed_ms: NULL argument");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    RTHC_CUDA(ctx, cudaEventSynchronize(ctx->ev_mark[1]));
    float f = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&f, ctx->ev_mark[0], ctx->ev_mark[1]));
    *ms = f;
    return 0;
}

void rthc_free_distance_state(rthc_ctx *ctx)
{
    if (ctx->planes_borrowed) {                                             // the caller's gathered buffers (rthc_import_planes)
        ctx->d_planes = nullptr; ctx->d_esc_cnt = nullptr; ctx->d_esc_list = nullptr;
        ctx->planes_borrowed = false;
    }
    ctx->L_stride = 0;
    if (ctx->d_pcodes == ctx->d_codes) ctx->d_pcodes = nullptr;              // an alias when no locus shares codes
    rthc_dev_free(ctx, &ctx->d_pcodes);
    rthc_dev_free(ctx, &ctx->d_codes); rthc_dev_free(ctx, &ctx->d_codes32); rthc_dev_free(ctx, &ctx->d_planes);
    rthc_dev_free(ctx, &ctx->d_esc_cnt); rthc_dev_free(ctx, &ctx->d_esc_list); rthc_dev_free(ctx, &ctx->d_esc_pairs);
    rthc_dev_free(ctx, &ctx->d_grp_off); rthc_dev_free(ctx, &ctx->d_grp_np); rthc_dev_free(ctx, &ctx->d_grp_base);
    rthc_dev_free(ctx, &ctx->d_cond_f64);
    ctx->have_profiles = ctx->have_dist = false;
}
// tile list, row offsets and owned rows depend on (n, rank, world, storage mode) only: they survive a new matrix of
// the same shape (the tile list of C4 has 306,153 entries: 1.2 ms of host work + copies per call otherwise)
void rthc_free_geometry(rthc_ctx *ctx)
{
    rthc_dev_free(ctx, &ctx->d_tiles); rthc_dev_free(ctx, &ctx->d_row_base);
    rthc_dev_free(ctx, &ctx->d_owned_rows);
    ctx->geom_n = -1;
}
void rthc_free_linkage_state(rthc_ctx *ctx)
{
    rthc_dev_free(ctx, &ctx->d_tpar); rthc_dev_free(ctx, &ctx->d_event); rthc_dev_free(ctx, &ctx->d_up);
    ctx->up_levels = 0;
    ctx->have_linkage = false;
}

extern "C" void rthc_destroy(rthc_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    rthc_free_distance_state(ctx);
    rthc_free_geometry(ctx);
    rthc_dev_free(ctx, &ctx->d_cond);
    rthc_dev_free(ctx, &ctx->d_gbm);
    rthc_dev_free(ctx, &ctx->d_pm);
    rthc_free_linkage_state(ctx);
    if (ctx->ev0) cudaEventDestroy(ctx->ev0);
    if (ctx->ev1) cudaEventDestroy(ctx->ev1);
    if (ctx->ev_aux) cudaEventDestroy(ctx->ev_aux);
    for (int k = 0; k < 3; ++k)
        if (ctx->h_pin[k]) cudaFreeHost(ctx->h_pin[k]);
    for (int k = 0; k < 2; ++k)
        if (ctx->ev_mark[k]) cudaEventDestroy(ctx->ev_mark[k]);
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    if (ctx->copy_stream) { cudaStreamSynchronize(ctx->copy_stream); cudaStreamDestroy(ctx->copy_stream); }
    for (int k = 0; k < 4; ++k)
        if (ctx->ev_piece[k]) cudaEventDestroy(ctx->ev_piece[k]);
    if (ctx->pool) cudaMemPoolDestroy(ctx->pool);
    if (ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

// page-locked staging slot of at least `bytes` (grown geometrically, contents not preserved); nullptr on failure
void *rthc_pinned(rthc_ctx *ctx, int slot, size_t bytes)
{
    if (bytes <= ctx->h_pin_cap[slot] && ctx->h_pin[slot]) return ctx->h_pin[slot];
    if (ctx->h_pin[slot]) { cudaFreeHost(ctx->h_pin[slot]); ctx->h_pin[slot] = nullptr; ctx->h_pin_cap[slot] = 0; }
    const size_t cap = std::max<size_t>(bytes + bytes / 2, 4096);
    if (cudaMallocHost(&ctx->h_pin[slot], cap) != cudaSuccess) {
        cudaGetLastError();
        ctx->h_pin[slot] = nullptr;
        rthc_fail(ctx, RTHC_E_NOMEM, "page-locked allocation of %zu bytes failed", cap);
        return nullptr;
    }
    ctx->h_pin_cap[slot] = cap;
    return ctx->h_pin[slot];
}

// tile-row r of B is owned by rank (min(r, B-1-r) mod world): rows pair up long + short
static inline int owner_of_tile_row(int64_t r, int64_t B, int world)
{
    int64_t p = std::min(r, B - 1 - r);
    return (int)(p % world);
}

extern "C" int rthc_set_shard(rthc_ctx *ctx, int rank, int world)
{
    if (!ctx) return RTHC_E_ARG;
    if (world < 1 || rank < 0 || rank >= world) return rthc_fail(ctx, RTHC_E_ARG, "bad shard %d/%d", rank, world);
    ctx->rank = rank;
    ctx->world = world;
    ctx->full_storage = false;
    return 0;
}

// ---- single-process multi-device (the CLI is ONE process, SURVEY 8b seam B0) -------------------
extern "C" int rthc_set_shard_full_storage(rthc_ctx *ctx, int rank, int world)
{
    int rc = rthc_set_shard(ctx, rank, world);
    if (rc) return rc;
    ctx->full_storage = true;
    return 0;
}

extern "C" int rthc_import_shard(rthc_ctx *dst, rthc_ctx *src)
{
    if (!dst || !src) return rthc_fail(dst, RTHC_E_ARG, "rthc_import_shard: NULL argument");
    if (!dst->full_storage || !dst->have_dist) return rthc_fail(dst, RTHC_E_STATE, "destination needs full storage and its own distances");
    if (!src->have_dist || !src->d_cond || src->n != dst->n || src->full_storage)
        return rthc_fail(dst, RTHC_E_STATE, "source shard has no distances of the same matrix");
    RTHC_CUDA(dst, cudaSetDevice(dst->device));
    const int64_t n = dst->n, B = dst->n_tiles_edge;
    for (int64_t r = 0; r < B; ++r) {
        if (owner_of_tile_row(r, B, src->world) != src->rank) continue;
        const int64_t i0 = r * RTHC_TILE, i1 = std::min(n, (r + 1) * RTHC_TILE);
        int64_t len = 0;
        for (int64_t i = i0; i < i1; ++i) len += n - 1 - i;
        if (len == 0) continue;
        RTHC_CUDA(dst, cudaMemcpyPeerAsync(dst->d_cond + dst->h_row_base[(size_t)i0], dst->device,
                                           src->d_cond + src->h_row_base[(size_t)i0], src->device,
                                           sizeof(uint16_t) * (size_t)len, dst->stream));
    }
    RTHC_CUDA(dst, cudaStreamSynchronize(dst->stream));
    return 0;
}

extern "C" int rthc_finish_import(rthc_ctx *ctx)
{
    if (!ctx || !ctx->full_storage || !ctx->have_dist) return rthc_fail(ctx, RTHC_E_STATE, "rthc_finish_import: nothing to finish");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    ctx->rank = 0; ctx->world = 1; ctx->full_storage = false;      // from here on: an ordinary full shard
    ctx->geom_n = -1;                                              // the owned-row list below is not the sharded one
    ctx->h_owned_rows.resize((size_t)ctx->n);
    for (int64_t i = 0; i < ctx->n; ++i) ctx->h_owned_rows[(size_t)i] = (int32_t)i;
    int rc = rthc_dev_alloc(ctx, &ctx->d_owned_rows, (size_t)ctx->n);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_owned_rows, ctx->h_owned_rows.data(), sizeof(int32_t) * (size_t)ctx->n,
                                   cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return 0;
}

extern "C" int rthc_set_single_algo(rthc_ctx *ctx, int algo)
{
    if (!ctx) return RTHC_E_ARG;
    if (algo != RTHC_SL_BORUVKA && algo != RTHC_SL_PRIM) return rthc_fail(ctx, RTHC_E_ARG, "bad algo %d", algo);
    ctx->single_algo = algo;
    return 0;
}


// geometry of the owned part of the condensed matrix + the tile list
int rthc_setup_geometry(rthc_ctx *ctx, int64_t n)
{
    if (ctx->geom_n == n && ctx->geom_rank == ctx->rank && ctx->geom_world == ctx->world && ctx->geom_full == ctx->full_storage &&
        ctx->d_tiles && ctx->d_row_base) {
        ctx->n = n;
        return 0;
    }
    rthc_free_geometry(ctx);
    ctx->n = n;
    ctx->n_pad = (n + RTHC_TILE - 1) / RTHC_TILE * RTHC_TILE;
    const int64_t B = ctx->n_pad / RTHC_TILE;
    ctx->n_tiles_edge = B;
    ctx->h_row_base.assign((size_t)ctx->n_pad, -1);
    ctx->h_owned_rows.clear();
    std::vector<int64_t> owned_tile_rows;
    int64_t off = 0;
    for (int64_t r = 0; r < B; ++r) {
        const bool owned = owner_of_tile_row(r, B, ctx->world) == ctx->rank;
        if (owned) owned_tile_rows.push_back(r);
        if (!owned && !ctx->full_storage) continue;
        for (int64_t i = r * RTHC_TILE; i < std::min(n, (r + 1) * RTHC_TILE); ++i) {
            ctx->h_row_base[(size_t)i] = off;          // full storage: the global scipy offset of row i
            off += n - 1 - i;
            if (owned) ctx->h_owned_rows.push_back((int32_t)i);
        }
    }
    ctx->owned_pairs = off;
    // tiles are (128-sample row block I, first 64-sample column block J); supertiles of
    // 16 row blocks x 32 column blocks keep the blocks of concurrently running CTAs in L2
    std::vector<int2> tiles;
    const int64_t S = 16, nj = 2, B64 = 2 * B;     // nj: 64-sample column blocks per tile (128 x 128)
    for (size_t rc0 = 0; rc0 < owned_tile_rows.size(); rc0 += (size_t)S)
        for (int64_t c0 = 0; c0 < B64; c0 += 2 * S)
            for (size_t q = rc0; q < std::min(owned_tile_rows.size(), rc0 + (size_t)S); ++q) {
                const int64_t I = owned_tile_rows[q];
                for (int64_t J = std::max(c0, 2 * I); J < std::min(B64, c0 + 2 * S); J += nj)
                    tiles.push_back(make_int2((int)I, (int)J));
            }
    ctx->n_tiles = (int64_t)tiles.size();
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_row_base, (size_t)ctx->n_pad))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_owned_rows, ctx->h_owned_rows.size()))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_tiles, tiles.size()))) return rc;
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_row_base, ctx->h_row_base.data(), sizeof(int64_t) * (size_t)ctx->n_pad,
                                   cudaMemcpyHostToDevice, ctx->stream));
    if (!ctx->h_owned_rows.empty())
        RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_owned_rows, ctx->h_owned_rows.data(),
                                       sizeof(int32_t) * ctx->h_owned_rows.size(), cudaMemcpyHostToDevice,
                                       ctx->stream));
    if (!tiles.empty())
        RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_tiles, tiles.data(), sizeof(int2) * tiles.size(),
                                       cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->geom_n = n; ctx->geom_rank = ctx->rank; ctx->geom_world = ctx->world; ctx->geom_full = ctx->full_storage;
    return 0;
}

static int set_profiles_common(rthc_ctx *ctx, const void *d_rowmajor, int64_t n, int64_t L, bool in_u16)
{
    int rc;
    ctx->L = L;
    if ((rc = rthc_setup_geometry(ctx, n))) return rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    if ((rc = rthc_recode_and_pack(ctx, d_rowmajor, in_u16))) return rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[1] = ms;
    ctx->have_profiles = true;
    return 0;
}

static int check_shape(rthc_ctx *ctx, int64_t n, int64_t L)
{
    if (n < 2) return rthc_fail(ctx, RTHC_E_ARG, "need at least 2 samples (n=%lld)", (long long)n);
    if (L < 1) return rthc_fail(ctx, RTHC_E_ARG, "need at least 1 locus (L=%lld)", (long long)L);
    if (n >= (1ll << 24)) return rthc_fail(ctx, RTHC_E_RANGE, "n=%lld exceeds 2^24-1 samples", (long long)n);
    // SNP matrices can be far wider than 65,535 columns; what bounds a uint16 distance is the number of
    // loci that can mismatch at all (>= 2 called alleles), checked by the re-coder once it knows them
    if (L >= (1ll << 24)) return rthc_fail(ctx, RTHC_E_RANGE, "L=%lld exceeds 2^24-1 loci", (long long)L);
    return 0;
}

static int set_profiles_host(rthc_ctx *ctx, const void *codes, int64_t n, int64_t L, bool in_u16)
{
    if (!ctx || !codes) return rthc_fail(ctx, RTHC_E_ARG, "rthc_set_profiles: NULL argument");
    int rc = check_shape(ctx, n, L);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    rthc_free_distance_state(ctx);
    rthc_free_linkage_state(ctx);
    const size_t bytes = (size_t)(n * L) * (in_u16 ? sizeof(uint16_t) : sizeof(int32_t));
    unsigned char *d_in = nullptr;
    if ((rc = rthc_dev_alloc(ctx, &d_in, bytes))) return rc;
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_in, codes, bytes, cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[0] = ms;
    rc = set_profiles_common(ctx, d_in, n, L, in_u16);
    rthc_dev_free(ctx, &d_in);
    return rc;
}

static int set_profiles_device(rthc_ctx *ctx, const void *dev_codes, int64_t n, int64_t L, bool in_u16)
{
    if (!ctx || !dev_codes) return rthc_fail(ctx, RTHC_E_ARG, "rthc_set_profiles_*_device: NULL argument");
    int rc = check_shape(ctx, n, L);
    if (rc) return rc;
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    rthc_free_distance_state(ctx);
    rthc_free_linkage_state(ctx);
    ctx->t_ms[0] = 0;
    RTHC_CUDA(ctx, cudaDeviceSynchronize());   // the caller's producer stream is unknown
    return set_profiles_common(ctx, dev_codes, n, L, in_u16);
}

extern "C" int rthc_set_profiles_i32(rthc_ctx *ctx, const int32_t *codes, int64_t n, int64_t L)
{
    return set_profiles_host(ctx, codes, n, L, false);
}
extern "C" int rthc_set_profiles_i32_device(rthc_ctx *ctx, const void *dev_codes, int64_t n, int64_t L)
{
    return set_profiles_device(ctx, dev_codes, n, L, false);
}
extern "C" int rthc_set_profiles_u16(rthc_ctx *ctx, const uint16_t *codes, int64_t n, int64_t L)
{
    return set_profiles_host(ctx, codes, n, L, true);
}
extern "C" int rthc_set_profiles_u16_device(rthc_ctx *ctx, const void *dev_codes, int64_t n, int64_t L)
{
    return set_profiles_device(ctx, dev_codes, n, L, true);
}

// ---- re-coder sharded by locus ranges (everything that does not shard caps the scaling) ----------------
// With several ranks the distance tiles shard but the re-coder would run in full on every rank (C4: 3.6 ms of a
// 39 ms step on 8 GPUs).  The loci are independent until the planes are built, so rank q re-codes and packs only the
// loci [q * chunk, (q + 1) * chunk) -- the SAME code path as a whole matrix, applied to a column range -- and the
// ranks then gather three buffers (planes, listed-cell counts, listed cells) with the caller's collective:
//     rthc_recode_shard   ->  header (groups, plane rows, plane count per group, ...) of this rank
//     (caller: all_gather of the headers, rows_max = widest rank, allocates the gathered buffers)
//     rthc_export_shard   ->  this rank's slot of the gathered buffers filled
//     (caller: all_gather_into_tensor x 3)
//     rthc_import_planes  ->  context ready for rthc_distance, planes = [rank][64-sample block][rows_max][64]
// The groups of a rank never mix with another rank's, so the result is the same set of loci in another order: the
// distances are identical (tests: every cell against the CPU restatement).
static inline int64_t shard_chunk(int64_t L, int world)
{
    return ((L + world - 1) / world + 31) / 32 * 32;
}
extern "C" int64_t rthc_shard_chunk(int64_t L, int world) { return world < 1 ? 0 : shard_chunk(L, world); }

extern "C" int rthc_recode_shard(rthc_ctx *ctx, const void *dev_codes, int in_u16, int64_t n, int64_t L, int32_t *hdr,
                                 int64_t hdr_len)
{
    if (!ctx || !dev_codes || !hdr) return rthc_fail(ctx, RTHC_E_ARG, "rthc_recode_shard: NULL argument");
    int rc = check_shape(ctx, n, L);
    if (rc) return rc;
    const int64_t chunk = shard_chunk(L, ctx->world);
    if (hdr_len < RTHC_SHARD_HDR + chunk / 32) return rthc_fail(ctx, RTHC_E_ARG, "rthc_recode_shard: header too short");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    rthc_free_distance_state(ctx);
    rthc_free_linkage_state(ctx);
    ctx->t_ms[0] = 0;
    RTHC_CUDA(ctx, cudaDeviceSynchronize());   // the caller's producer stream is unknown
    const int64_t l0 = std::min(L, ctx->rank * chunk), l1 = std::min(L, l0 + chunk), Lloc = l1 - l0;
    for (int64_t k = 0; k < hdr_len; ++k) hdr[k] = 0;
    ctx->L_global = L;
    if (Lloc == 0) {                           // more ranks than 32-locus blocks: this one contributes nothing
        ctx->L = 0; ctx->G = 0; ctx->plane_rows = 0; ctx->max_np = 0; ctx->L_kept = 0; ctx->max_alleles = 0;
        ctx->escape_loci = ctx->escape_cells = 0; ctx->h_grp_np.clear(); ctx->t_ms[1] = 0;
        return rthc_setup_geometry(ctx, n);
    }
    ctx->L_stride = L;
    const size_t elem = in_u16 ? sizeof(uint16_t) : sizeof(int32_t);
    rc = set_profiles_common(ctx, (const unsigned char *)dev_codes + (size_t)l0 * elem, n, Lloc, in_u16 != 0);
    ctx->L_stride = 0;
    ctx->have_profiles = false;                // not before rthc_import_planes
    if (rc) return rc;
    hdr[0] = ctx->G; hdr[1] = ctx->plane_rows; hdr[2] = ctx->max_np; hdr[3] = (int32_t)ctx->L_kept;
    hdr[4] = ctx->max_alleles; hdr[5] = (int32_t)ctx->escape_loci;
    hdr[6] = (int32_t)std::min<int64_t>(ctx->escape_cells, INT32_MAX); hdr[7] = (int32_t)Lloc;
    for (int g = 0; g < ctx->G; ++g) hdr[RTHC_SHARD_HDR + g] = ctx->h_grp_np[(size_t)g];
    return 0;
}

extern "C" int rthc_export_shard(rthc_ctx *ctx, void *d_planes_slot, int64_t rows_max, int32_t *d_cnt_slot,
                                 uint64_t *d_list_slot)
{
    if (!ctx || !d_planes_slot || !d_cnt_slot || !d_list_slot) return rthc_fail(ctx, RTHC_E_ARG, "rthc_export_shard: NULL argument");
    if (ctx->L_global <= 0 || ctx->planes_borrowed) return rthc_fail(ctx, RTHC_E_STATE, "rthc_export_shard before rthc_recode_shard");
    if (rows_max < ctx->plane_rows) return rthc_fail(ctx, RTHC_E_ARG, "rthc_export_shard: rows_max below this rank's plane rows");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev0, ctx->stream));
    const int64_t chunk = shard_chunk(ctx->L_global, ctx->world), nblk = ctx->n_pad / RTHC_BLK, Lloc = ctx->L;
    if (ctx->plane_rows > 0)
        RTHC_CUDA(ctx, cudaMemcpy2DAsync(d_planes_slot, (size_t)rows_max * RTHC_BLK * 4, ctx->d_planes,
                                         (size_t)ctx->plane_rows * RTHC_BLK * 4, (size_t)ctx->plane_rows * RTHC_BLK * 4,
                                         (size_t)nblk, cudaMemcpyDeviceToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaMemsetAsync(d_cnt_slot, 0, sizeof(int32_t) * (size_t)chunk, ctx->stream));
    if (ctx->d_esc_cnt && ctx->d_esc_list && ctx->escape_cells > 0 && Lloc > 0) {
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_cnt_slot, ctx->d_esc_cnt, sizeof(int32_t) * (size_t)Lloc, cudaMemcpyDeviceToDevice, ctx->stream));
        RTHC_CUDA(ctx, cudaMemcpyAsync(d_list_slot, ctx->d_esc_list, sizeof(uint64_t) * (size_t)Lloc * RTHC_ESC_CAP,
                                       cudaMemcpyDeviceToDevice, ctx->stream));
    }
    RTHC_CUDA(ctx, cudaEventRecord(ctx->ev1, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));        // the collective runs on the caller's stream
    float ms = 0;
    RTHC_CUDA(ctx, cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1));
    ctx->t_ms[1] += ms;
    return 0;
}

extern "C" int rthc_import_planes(rthc_ctx *ctx, const void *d_planes_all, int64_t rows_max, const int32_t *hdr_all,
                                  int64_t hdr_len, const int32_t *d_cnt_all, const uint64_t *d_list_all)
{
    if (!ctx || !d_planes_all || !hdr_all || !d_cnt_all || !d_list_all)
        return rthc_fail(ctx, RTHC_E_ARG, "rthc_import_planes: NULL argument");
    if (ctx->L_global <= 0 || ctx->planes_borrowed) return rthc_fail(ctx, RTHC_E_STATE, "rthc_import_planes before rthc_recode_shard");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t L = ctx->L_global, chunk = shard_chunk(L, ctx->world), nblk = ctx->n_pad / RTHC_BLK;
    // this rank's own re-code products are in the gathered buffers now
    if (ctx->d_pcodes == ctx->d_codes) ctx->d_pcodes = nullptr;
    rthc_dev_free(ctx, &ctx->d_pcodes); rthc_dev_free(ctx, &ctx->d_codes); rthc_dev_free(ctx, &ctx->d_codes32);
    rthc_dev_free(ctx, &ctx->d_planes); rthc_dev_free(ctx, &ctx->d_esc_cnt); rthc_dev_free(ctx, &ctx->d_esc_list);
    rthc_dev_free(ctx, &ctx->d_grp_off); rthc_dev_free(ctx, &ctx->d_grp_np); rthc_dev_free(ctx, &ctx->d_grp_base);
    std::vector<int64_t> gbase;
    std::vector<int32_t> gnp;
    int64_t rows = 0, kept = 0, cells = 0, loci = 0;
    int max_np = 0, max_all = 0;
    for (int q = 0; q < ctx->world; ++q) {
        const int32_t *h = hdr_all + (size_t)q * (size_t)hdr_len;
        if (h[0] < 0 || RTHC_SHARD_HDR + (int64_t)h[0] > hdr_len || h[1] > rows_max)
            return rthc_fail(ctx, RTHC_E_ARG, "rthc_import_planes: header of rank %d is inconsistent", q);
        int64_t off = 0;
        for (int g = 0; g < h[0]; ++g) {
            const int np = h[RTHC_SHARD_HDR + g];
            if (np < RTHC_MIN_BITS + 1 || np > 25)
                return rthc_fail(ctx, RTHC_E_ARG, "rthc_import_planes: group %d of rank %d claims %d planes", g, q, np);
            max_np = std::max(max_np, np);                   // (not the header's own maximum: it sizes shared memory)
            gbase.push_back((int64_t)q * nblk * rows_max + off);
            gnp.push_back(np);
            off += np;
        }
        if (off != h[1]) return rthc_fail(ctx, RTHC_E_ARG, "rthc_import_planes: plane rows of rank %d do not add up", q);
        rows += h[1]; kept += h[3]; max_all = std::max(max_all, (int)h[4]);
        loci += h[5]; cells += h[6];
    }
    if (kept > 65535)
        return rthc_fail(ctx, RTHC_E_RANGE, "%lld of the %lld loci have two or more called alleles; distances are "
                         "stored as uint16, so at most 65,535 such loci are supported", (long long)kept, (long long)L);
    const int G = (int)gnp.size();
    {   // the distance kernel runs one unrolled body per plane count: keep groups of equal width together, widest first,
        // as the unsharded re-coder does (every change of body is an instruction-cache refill)
        std::vector<int> order((size_t)G);
        for (int g = 0; g < G; ++g) order[(size_t)g] = g;
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return gnp[(size_t)a] > gnp[(size_t)b]; });
        std::vector<int64_t> b2((size_t)G);
        std::vector<int32_t> n2((size_t)G);
        for (int g = 0; g < G; ++g) { b2[(size_t)g] = gbase[(size_t)order[(size_t)g]]; n2[(size_t)g] = gnp[(size_t)order[(size_t)g]]; }
        gbase.swap(b2); gnp.swap(n2);
    }
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_grp_base, (size_t)std::max(G, 1)))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_grp_np, (size_t)std::max(G, 1)))) return rc;
    if (!ctx->d_esc_pairs && (rc = rthc_dev_alloc(ctx, &ctx->d_esc_pairs, 1))) return rc;
    if (G > 0) {
        RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_grp_base, gbase.data(), sizeof(int64_t) * (size_t)G, cudaMemcpyHostToDevice, ctx->stream));
        RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_grp_np, gnp.data(), sizeof(int32_t) * (size_t)G, cudaMemcpyHostToDevice, ctx->stream));
    }
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    ctx->d_planes = const_cast<uint32_t *>(static_cast<const uint32_t *>(d_planes_all));
    ctx->d_esc_cnt = const_cast<int32_t *>(d_cnt_all);
    ctx->d_esc_list = const_cast<uint64_t *>(d_list_all);
    ctx->planes_borrowed = true;
    ctx->blk_stride_rows = rows_max;
    ctx->esc_L = (int64_t)ctx->world * chunk;
    ctx->L = L; ctx->G = G; ctx->plane_rows = (int)rows; ctx->max_np = max_np; ctx->L_kept = kept; ctx->max_alleles = max_all;
    ctx->escape_loci = loci; ctx->escape_cells = cells; ctx->escape_pairs_u = 0;
    ctx->have_profiles = true;
    return 0;
}

extern "C" int rthc_distance(rthc_ctx *ctx)
{
    if (!ctx) return RTHC_E_ARG;
    if (!ctx->have_profiles) return rthc_fail(ctx, RTHC_E_STATE, "rthc_distance before rthc_set_profiles_i32");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    int rc;
    if (!ctx->d_cond || ctx->cond_cap < ctx->owned_pairs) {
        ctx->cond_cap = 0;
        if ((rc = rthc_dev_alloc(ctx, &ctx->d_cond, (size_t)std::max<int64_t>(ctx->owned_pairs, 1) + RTHC_COND_PAD))) return rc;
        ctx->cond_cap = std::max<int64_t>(ctx->owned_pairs, 1);
    }
    if ((rc = rthc_run_distance(ctx))) return rc;
    ctx->have_dist = true;
    return 0;
}

extern "C" int64_t rthc_owned_pairs(const rthc_ctx *ctx) { return ctx ? ctx->owned_pairs : 0; }

extern "C" int rthc_get_owned_condensed_u16(rthc_ctx *ctx, uint16_t *out)
{
    if (!ctx || !out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_dist) return rthc_fail(ctx, RTHC_E_STATE, "no distance matrix yet");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    RTHC_CUDA(ctx, cudaMemcpy(out, ctx->d_cond, sizeof(uint16_t) * (size_t)ctx->owned_pairs,
                              cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int rthc_get_condensed_range_u16(rthc_ctx *ctx, int64_t first, int64_t count, uint16_t *out)
{
    if (!ctx || !out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_dist) return rthc_fail(ctx, RTHC_E_STATE, "no distance matrix yet");
    if (first < 0 || count < 0 || first + count > ctx->owned_pairs) return rthc_fail(ctx, RTHC_E_ARG, "condensed range out of bounds");
    if (count == 0) return 0;
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    RTHC_CUDA(ctx, cudaMemcpy(out, ctx->d_cond + first, sizeof(uint16_t) * (size_t)count, cudaMemcpyDeviceToHost));
    return 0;
}

extern "C" int rthc_get_condensed_u16(rthc_ctx *ctx, uint16_t *out)
{
    if (ctx && ctx->world != 1) return rthc_fail(ctx, RTHC_E_STATE, "full condensed copy needs a full shard");
    return rthc_get_owned_condensed_u16(ctx, out);
}

__global__ void rthc_square_rows(const uint16_t *__restrict__ cond, int64_t n, int64_t row0, int64_t nrows,
                                 uint16_t *__restrict__ out)
{
    const int64_t r = blockIdx.y;
    const int64_t j = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= nrows || j >= n) return;
    const int64_t i = row0 + r;
    uint16_t v = 0;
    if (i != j) {
        const int64_t a = i < j ? i : j, b = i < j ? j : i;
        v = cond[n * a - (a * (a + 1)) / 2 + (b - a - 1)];
    }
    out[r * n + j] = v;
}

extern "C" int rthc_copy_rows_u16(rthc_ctx *ctx, int64_t row0, int64_t nrows, uint16_t *out)
{
    if (!ctx || !out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_dist) return rthc_fail(ctx, RTHC_E_STATE, "no distance matrix yet");
    if (ctx->world != 1) return rthc_fail(ctx, RTHC_E_STATE, "square rows need a full shard");
    if (row0 < 0 || nrows < 0 || row0 + nrows > ctx->n) return rthc_fail(ctx, RTHC_E_ARG, "row range out of bounds");
    if (nrows == 0) return 0;
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    uint16_t *d_out = nullptr;
    int rc = rthc_dev_alloc(ctx, &d_out, (size_t)(nrows * ctx->n));
    if (rc) return rc;
    dim3 grd((unsigned)((ctx->n + 255) / 256), (unsigned)nrows);
    rthc_square_rows<<<grd, 256, 0, ctx->stream>>>(ctx->d_cond, ctx->n, row0, nrows, d_out);
    RTHC_LAUNCH_CHECK(ctx);
    RTHC_CUDA(ctx, cudaMemcpyAsync(out, d_out, sizeof(uint16_t) * (size_t)(nrows * ctx->n),
                                   cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_out);
    return 0;
}

extern "C" int rthc_direct_pairs(rthc_ctx *ctx, const int32_t *pi, const int32_t *pj, int64_t np, int32_t *out)
{
    if (!ctx || !pi || !pj || !out) return rthc_fail(ctx, RTHC_E_ARG, "NULL argument");
    if (!ctx->have_profiles) return rthc_fail(ctx, RTHC_E_STATE, "no profiles loaded");
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    int32_t *d_pi = nullptr, *d_pj = nullptr, *d_out = nullptr;
    int rc;
    if ((rc = rthc_dev_alloc(ctx, &d_pi, (size_t)np))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_pj, (size_t)np))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_out, (size_t)np))) return rc;
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_pi, pi, sizeof(int32_t) * (size_t)np, cudaMemcpyHostToDevice, ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(d_pj, pj, sizeof(int32_t) * (size_t)np, cudaMemcpyHostToDevice, ctx->stream));
    rc = rthc_run_direct_pairs(ctx, d_pi, d_pj, np, d_out);
    if (!rc) {
        cudaError_t e = cudaMemcpy(out, d_out, sizeof(int32_t) * (size_t)np, cudaMemcpyDeviceToHost);
        if (e != cudaSuccess) rc = rthc_fail(ctx, RTHC_E_CUDA, "D2H failed: %s", cudaGetErrorString(e));
    }
    rthc_dev_free(ctx, &d_pi); rthc_dev_free(ctx, &d_pj); rthc_dev_free(ctx, &d_out);
    return rc;
}

extern "C" int rthc_info(const rthc_ctx *ctx, int64_t *info, int n_info)
{
    if (!ctx || !info) return RTHC_E_ARG;
    const int64_t v[14] = {ctx->n, ctx->L, ctx->L_kept, ctx->G, ctx->plane_rows, ctx->n_tiles,
                           ctx->boruvka_rounds, ctx->max_alleles, ctx->row_scans, ctx->chain_full_scans,
                           ctx->chain_cache_hits, ctx->escape_loci, ctx->escape_cells, (int64_t)ctx->escape_pairs_u};
    for (int k = 0; k < n_info && k < 14; ++k) info[k] = v[k];
    return 0;
}

extern "C" int rthc_timings(const rthc_ctx *ctx, double *ms, int n_ms)
{
    if (!ctx || !ms) return RTHC_E_ARG;
    for (int k = 0; k < n_ms && k < 8; ++k) ms[k] = k == 5 ? (double)ctx->launches : ctx->t_ms[k];
    return 0;
}

// ---- `-d_mx` entry: externally supplied condensed distances --------------------------------
__global__ void rthc_classify_f64(const double *__restrict__ y, int64_t P, uint16_t *__restrict__ out, int *flags)
{
    const int64_t p = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (p >= P) return;
    const double v = y[p];
    int f = 0;
    if (!(v >= 0.0) || isinf(v)) f |= 1;                       // negative, NaN or infinite
    else if (v > 65535.0 || v != floor(v)) f |= 2;             // not representable as uint16
    else out[p] = (uint16_t)v;
    if (f) atomicOr(flags, f);
}

extern "C" int rthc_set_condensed_f64(rthc_ctx *ctx, const double *y, int64_t n)
{
    if (!ctx || !y) return rthc_fail(ctx, RTHC_E_ARG, "rthc_set_condensed_f64: NULL argument");
    if (n < 2) return rthc_fail(ctx, RTHC_E_ARG, "need at least 2 samples (n=%lld)", (long long)n);
    if (n >= (1ll << 24)) return rthc_fail(ctx, RTHC_E_RANGE, "n=%lld exceeds 2^24-1 samples", (long long)n);
    RTHC_CUDA(ctx, cudaSetDevice(ctx->device));
    rthc_free_distance_state(ctx);
    rthc_free_linkage_state(ctx);
    const int save_rank = ctx->rank, save_world = ctx->world;
    ctx->rank = 0; ctx->world = 1;                              // an external matrix is always a full shard
    int rc = rthc_setup_geometry(ctx, n);
    ctx->rank = save_rank; ctx->world = save_world;
    if (rc) return rc;
    ctx->world = 1; ctx->rank = 0;
    ctx->L = 0; ctx->G = 0; ctx->plane_rows = 0; ctx->L_kept = 0;
    const int64_t P = n * (n - 1) / 2;
    int *d_flags = nullptr;
    if ((rc = rthc_dev_alloc(ctx, &ctx->d_cond_f64, (size_t)P))) return rc;
    if ((rc = rthc_dev_alloc(ctx, &d_flags, 1))) return rc;
    if (!ctx->d_cond || ctx->cond_cap < P) {
        ctx->cond_cap = 0;
        if ((rc = rthc_dev_alloc(ctx, &ctx->d_cond, (size_t)P + RTHC_COND_PAD))) return rc;
        ctx->cond_cap = P;
    }
    RTHC_CUDA(ctx, cudaMemsetAsync(d_flags, 0, sizeof(int), ctx->stream));
    RTHC_CUDA(ctx, cudaMemcpyAsync(ctx->d_cond_f64, y, sizeof(double) * (size_t)P, cudaMemcpyHostToDevice, ctx->stream));
    rthc_classify_f64<<<(unsigned)((P + 255) / 256), 256, 0, ctx->stream>>>(ctx->d_cond_f64, P, ctx->d_cond, d_flags);
    RTHC_LAUNCH_CHECK(ctx);
    int flags = 0;
    RTHC_CUDA(ctx, cudaMemcpyAsync(&flags, d_flags, sizeof(int), cudaMemcpyDeviceToHost, ctx->stream));
    RTHC_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    rthc_dev_free(ctx, &d_flags);
    if (flags & 1) {
        rthc_dev_free(ctx, &ctx->d_cond_f64);
        return rthc_fail(ctx, RTHC_E_RANGE, "distance matrix holds negative, NaN or infinite entries");
    }
    if (flags & 2) {
        rthc_dev_free(ctx, &ctx->d_cond);                             // non-integer distances: float64 path only
        ctx->cond_cap = 0;
        ctx->ext_integer = false;
    } else {
        rthc_dev_free(ctx, &ctx->d_cond_f64);                         // integer distances: uint16 path
        ctx->ext_integer = true;
    }
    ctx->have_dist = true;
    ctx->have_profiles = false;
    return 0;
}
