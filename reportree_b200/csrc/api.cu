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
    if (ctx->d_pcodes == ctx->d_codes) ctx->d_pcodes = nullptr;              // an alias when no locus shares codes
    rthc_dev_free(ctx, &ctx->d_pcodes);
    rthc_dev_free(ctx, &ctx->d_codes); rthc_dev_free(ctx, &ctx->d_codes32); rthc_dev_free(ctx, &ctx->d_planes);
    rthc_dev_free(ctx, &ctx->d_esc_cnt); rthc_dev_free(ctx, &ctx->d_esc_list); rthc_dev_free(ctx, &ctx->d_esc_pairs);
    rthc_dev_free(ctx, &ctx->d_grp_off); rthc_dev_free(ctx, &ctx->d_grp_np);
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
