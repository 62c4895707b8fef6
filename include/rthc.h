/*
 * rthc.h -- C ABI of libreportree_hc.so: the B200 (sm_100a) replacement for the
 * numeric layer underneath ReporTree's scripts/partitioning_HC.py.
 *
 * The reference has no FFI for this path; the seams these entry points replace are
 * (citations relative to the reference tree):
 *
 *   rthc_set_profiles_i32 + rthc_distance
 *        <- the `cgmlst-dists` subprocess and the 2,000-column chunk/sum loop,
 *           scripts/partitioning_HC.py:489-516 (call at :502); same seam in
 *           scripts/partitioning_grapetree.py:452-483 (call at :466)
 *   rthc_copy_rows_u16 / rthc_get_condensed_u16
 *        <- `<out>_dist_hamming.tsv` write + re-read (:518-520) and
 *           scipy.spatial.distance.squareform in dist_mx() (:173-188)
 *   rthc_set_condensed_f64
 *        <- the `-d_mx` entry (:522-525) feeding dist_mx() (:186)
 *   rthc_linkage
 *        <- scipy.cluster.hierarchy.linkage(y, method) in hcluster() (:197) and
 *           maxdists(Z)[-1] (:198) (Z[:,2] is monotone for the supported methods)
 *   rthc_cut
 *        <- scipy.cluster.hierarchy.fcluster(Z, t, criterion="distance") at
 *           :649, :689, :722, :751
 *
 * Conventions: every function returns 0 on success or a negative RTHC_E_* code;
 * rthc_last_error() gives the message.  No C++ exceptions cross the ABI.  The caller
 * allocates every host buffer; the library owns all device memory inside the context.
 * A context is bound to one CUDA device, is not thread-safe, and every call is
 * synchronous on return.  There is NO CPU fallback: without a usable CUDA device
 * rthc_create() fails.
 */
#ifndef RTHC_H
#define RTHC_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef struct rthc_ctx rthc_ctx;

enum {
    RTHC_OK = 0,
    RTHC_E_CUDA = -1,      /* CUDA runtime error (message has the CUDA string)    */
    RTHC_E_ARG = -2,       /* bad argument                                         */
    RTHC_E_STATE = -3,     /* call order violated (e.g. cut before linkage)        */
    RTHC_E_RANGE = -4,     /* input outside the supported range (see message)      */
    RTHC_E_NOMEM = -5
};

/* linkage methods (names follow scipy / --HC-threshold) */
enum { RTHC_SINGLE = 0, RTHC_COMPLETE = 1, RTHC_AVERAGE = 2, RTHC_WEIGHTED = 3, RTHC_WARD = 4 };

/* single-linkage algorithm selection for rthc_linkage(RTHC_SINGLE):
 *   RTHC_SL_BORUVKA : GPU Boruvka MST under the total order (w, i, j); membership at
 *                     every threshold equals scipy's, fcluster ids may differ on ties.
 *   RTHC_SL_PRIM    : device Prim in scipy's visiting order -> Z identical to scipy's
 *                     (byte-identical cluster_<id> names and Newick).                  */
enum { RTHC_SL_BORUVKA = 0, RTHC_SL_PRIM = 1 };

/* ---- lifetime -------------------------------------------------------------- */
int rthc_create(rthc_ctx **out, int device);
/* Creates the CUDA primary context of `device` and nothing else (0.7 s of driver start-up on a B200 box),
 * so that a host can pay it on a helper thread from its very first instruction -- before its own imports and
 * argument parsing -- and find rthc_create() taking milliseconds later.  Needs no rthc_ctx; safe to skip. */
int rthc_prewarm(int device);
void rthc_destroy(rthc_ctx *ctx);
const char *rthc_last_error(const rthc_ctx *ctx);   /* ctx may be NULL: last create error */
int rthc_version(void);

/* ---- sharding (multi-GPU): this context owns 128-sample tile-rows r with
 *      owner(r) == rank under the balanced pairing (r, B-1-r) round-robin.
 *      Default rank 0 / world 1 = everything.  Call before rthc_set_profiles_i32. */
int rthc_set_shard(rthc_ctx *ctx, int rank, int world);

/* single-process multi-device (the reference spawns ONE partitioning_HC.py process, reportree.py:1796):
 * the context computes only the tiles of shard rank/world but stores them in the FULL condensed
 * vector; rthc_import_shard copies another context's owned rows (any device of the box, peer copy)
 * into it, rthc_finish_import turns it into an ordinary full shard (linkage, cut, row copies).  */
int rthc_set_shard_full_storage(rthc_ctx *ctx, int rank, int world);
int rthc_import_shard(rthc_ctx *dst, rthc_ctx *src);
int rthc_finish_import(rthc_ctx *ctx);

/* ---- distance half ----------------------------------------------------------
 * codes: n x L row-major int32 allele codes, 0 (or negative) = missing.
 * Copies H2D, dense-ranks every locus on the device (uint16, 0 = missing) and packs
 * bit planes.  Loci with <= 1 distinct called allele cannot contribute and are dropped. */
int rthc_set_profiles_i32(rthc_ctx *ctx, const int32_t *codes_rowmajor, int64_t n, int64_t L);
/* same, but the matrix is already in device memory (n x L row-major int32) */
int rthc_set_profiles_i32_device(rthc_ctx *ctx, const void *dev_codes_rowmajor, int64_t n, int64_t L);
/* same two entry points for callers whose allele ids fit 16 bits (0 = missing): half the host-to-device
 * bytes of the int32 form.  The tokenizer that feeds the seam (scripts/partitioning_HC.py:293-298 contract)
 * knows the largest id of the matrix and picks the narrower form when it can.                            */
int rthc_set_profiles_u16(rthc_ctx *ctx, const uint16_t *codes_rowmajor, int64_t n, int64_t L);
int rthc_set_profiles_u16_device(rthc_ctx *ctx, const void *dev_codes_rowmajor, int64_t n, int64_t L);

/* d(i,j) = #{l : a_il != a_jl, a_il != 0, a_jl != 0} for every owned pair, written to
 * the device-resident condensed uint16 matrix (scipy pdist order for a full shard). */
int rthc_distance(rthc_ctx *ctx);

/* number of condensed entries this context owns (n(n-1)/2 for a full shard) */
int64_t rthc_owned_pairs(const rthc_ctx *ctx);
/* full condensed vector (scipy order), full shard only; out has n(n-1)/2 entries */
int rthc_get_condensed_u16(rthc_ctx *ctx, uint16_t *out);
/* owned part of the condensed matrix, rows in increasing order; out has rthc_owned_pairs() */
int rthc_get_owned_condensed_u16(rthc_ctx *ctx, uint16_t *out);
/* entries [first, first+count) of the owned condensed vector (block-wise checks of matrices
 * that should not be copied to the host in one piece; same seam as rthc_get_condensed_u16) */
int rthc_get_condensed_range_u16(rthc_ctx *ctx, int64_t first, int64_t count, uint16_t *out);
/* rows [row0, row0+nrows) of the SQUARE matrix (nrows x n, uint16), full shard only:
 * the block the `_dist_hamming.tsv` writer streams.                                 */
int rthc_copy_rows_u16(rthc_ctx *ctx, int64_t row0, int64_t nrows, uint16_t *out);
/* distances of an arbitrary pair list straight from the uint16 codes (independent
 * scalar kernel; used to cross-check the bit-sliced kernel at sizes no CPU can).    */
int rthc_direct_pairs(rthc_ctx *ctx, const int32_t *pi, const int32_t *pj, int64_t npairs,
                      int32_t *out);

/* `-d_mx` entry: externally supplied condensed distances (scipy order, n(n-1)/2).
 * Any finite non-negative doubles.  Integer-valued matrices <= 65535 take the uint16 paths
 * (Boruvka / integer complete linkage); others run Prim and the merge loop in float64.  */
int rthc_set_condensed_f64(rthc_ctx *ctx, const double *y, int64_t n);

/* ---- clustering half --------------------------------------------------------- */
int rthc_set_single_algo(rthc_ctx *ctx, int algo);
/* Z_out: (n-1) x 4 float64, scipy linkage layout.  Full shard only.
 * method = RTHC_SINGLE .. RTHC_WARD: replaces scipy linkage(y, method) at
 * scripts/partitioning_HC.py:197 (complete / average / weighted / ward = scipy nn_chain,
 * same merge order, same floating-point operation order, so Z is byte-identical).   */
int rthc_linkage(rthc_ctx *ctx, int method, double *Z_out);
/* multi-GPU single linkage: minimum spanning forest of the owned edges.
 * keys_out holds up to n-1 packed edges (w<<48 | i<<24 | j); *m_out = count.       */
int rthc_local_forest(rthc_ctx *ctx, uint64_t *keys_out, int64_t *m_out);
/* finish single linkage from the union of the ranks' forests (m keys, any order).  */
int rthc_linkage_from_forest(rthc_ctx *ctx, const uint64_t *keys, int64_t m, int64_t n,
                             double *Z_out);
/* Device-resident variants for the NCCL exchange ("one fixed-size all_gather of packed edges per
 * rank"): d_keys_out is DEVICE memory of the caller (e.g. a torch tensor) with `capacity` >= n-1 slots; the
 * forest is written unsorted and the remaining slots are set to +inf (0xFFFFFFFFFFFFFFFF), so the buffer can
 * go straight into an all_gather.  Synchronous on return.                                              */
int rthc_local_forest_device(rthc_ctx *ctx, uint64_t *d_keys_out, int64_t capacity, int64_t *m_out);
/* d_keys: m slots of DEVICE memory (the gathered buffers, any order, +inf slots ignored); the caller
 * must have synchronised whatever produced them.                                                       */
int rthc_linkage_from_forest_device(rthc_ctx *ctx, const uint64_t *d_keys, int64_t m, int64_t n,
                                    double *Z_out);
/* ---- re-coder sharded by locus ranges ------------------------------------------------------
 * Replaces, for world > 1, the replicated call of rthc_set_profiles_*_device: the per-locus work of the
 * reference's input stage (the cell parse / allele identity of cgmlst-dists behind
 * scripts/partitioning_HC.py:502, cleaning :293-298) is independent per locus, so rank q of the context's shard
 * (rthc_set_shard) re-codes and bit-slices only the loci [q * chunk, (q + 1) * chunk), chunk =
 * rthc_shard_chunk(L, world) (a multiple of 32), and the caller gathers the ranks' products with its own
 * collective (torch.distributed all_gather_into_tensor over NCCL in reportree_b200/sharded.py):
 *   1. rthc_recode_shard(ctx, dev_codes, in_u16, n, L, hdr, hdr_len): dev_codes = the WHOLE n x L row-major
 *      matrix in device memory (int32, or uint16 when in_u16); hdr (host, hdr_len >= RTHC_SHARD_HDR + chunk / 32
 *      int32) receives [0] groups of 32 loci, [1] plane rows, [2] most planes in a group, [3] loci kept,
 *      [4] most alleles in a locus, [5] loci with shared codes, [6] cells listed for the fix-up, [7] loci of
 *      this rank, [RTHC_SHARD_HDR + g] planes of group g.
 *   2. the caller gathers the headers (host or device), takes rows_max = max over ranks of hdr[1] and provides
 *      three DEVICE buffers: planes uint32[world][n_pad / 64][rows_max][64] (n_pad = n rounded up to 128),
 *      counts int32[world][chunk], cells uint64[world][chunk][4096].
 *   3. rthc_export_shard(ctx, planes slot of this rank, rows_max, counts slot, cells slot): fills this rank's
 *      slots (synchronous on return); the caller runs one all_gather per buffer.
 *   4. rthc_import_planes(ctx, planes, rows_max, hdr_all (host, world x hdr_len), hdr_len, counts, cells):
 *      the context is then in the state rthc_set_profiles_* leaves (ready for rthc_distance).  The three
 *      gathered buffers are BORROWED: they must stay allocated and unchanged until the next
 *      rthc_set_profiles_* / rthc_recode_shard / rthc_destroy of this context.
 * Distances are identical to the unsharded path (the loci are the same, their grouping differs);
 * rthc_direct_pairs is not available in this state.                                                     */
#define RTHC_SHARD_HDR 8
int64_t rthc_shard_chunk(int64_t L, int world);
int rthc_recode_shard(rthc_ctx *ctx, const void *dev_codes, int in_u16, int64_t n, int64_t L, int32_t *hdr,
                      int64_t hdr_len);
int rthc_export_shard(rthc_ctx *ctx, void *d_planes_slot, int64_t rows_max, int32_t *d_cnt_slot,
                      uint64_t *d_list_slot);
int rthc_import_planes(rthc_ctx *ctx, const void *d_planes_all, int64_t rows_max, const int32_t *hdr_all,
                       int64_t hdr_len, const int32_t *d_cnt_all, const uint64_t *d_list_all);
/* labels_out: T x n int32, row k = fcluster(Z, t[k], "distance") of the last linkage. */
int rthc_cut(rthc_ctx *ctx, const double *t, int T, int32_t *labels_out);

/* ---- zoom-in on the resident matrix ----------------------------- */
/* Replaces get_closest_samples (reportree.py:591-617), which re-reads one column of
 * <out>_dist_hamming.tsv per sample of interest: every sample j with
 * d(sample, j) <= d_sorted[k], where d_sorted is the ascending row of `sample` including its own 0
 * (the reference's `max_dist = distances[int(n)]`).  out_idx / out_dist need room for n entries
 * (ties can select any number), filled in no particular order; *out_count = how many.           */
int rthc_closest(rthc_ctx *ctx, int64_t sample, int64_t k, int64_t *out_idx, uint16_t *out_dist,
                 int64_t *out_count, int *out_max_dist);
/* rthc_closest without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops,
 * harnesses that forbid it).  Same arguments, results and error codes as rthc_closest; the scratch (row copy,
 * selection lists, 65,536-bin histogram) lives in one file-scope `static __device__` array of the library
 * (about 3.3 MiB, held by every GPU that loads the library until exit), so n is limited to
 * RTHC_NOALLOC_MAX_N samples (RTHC_E_RANGE beyond).  The call makes no cudaMalloc / cudaFree / pool
 * allocation.  All calls on a device share the array: calls that overlap on the GPU race -- serialise them. */
#define RTHC_NOALLOC_MAX_N 262144
int rthc_closest_noalloc(rthc_ctx *ctx, int64_t sample, int64_t k, int64_t *out_idx, uint16_t *out_dist,
                         int64_t *out_count, int *out_max_dist);
/* Replaces the distance half of a zoom-in re-run (reportree.py:1965-2055 re-invokes
 * partitioning_HC.py with --loci <out>_loci_used.txt on the subset, cgmlst-dists included):
 * `child` (same device, created by rthc_create) receives the m x m sub-matrix of `parent`'s resident
 * distances for the samples idx[0..m) (distinct) and is then ready for rthc_linkage / rthc_cut /
 * rthc_copy_rows_u16.                                                                            */
int rthc_subset_from(rthc_ctx *child, rthc_ctx *parent, const int64_t *idx, int64_t m);

/* ---- cluster stability over neighbouring thresholds -------------- */
/* Replaces getContTable / getContTableTotals / getMismatchMatrix (scripts/ComparingPartitions/
 * metrics.py:23-91) for the adjacent-column comparisons of metrics4stability
 * (scripts/ComparingPartitions/comparing_partitions_v2.py:74-113).
 * codes: T x n int32 (host), row t = cluster id of every sample in partition column t, any
 * numbering with 0 <= id < n.  For every pair p = (column p, column p+1) the non-empty cells of
 * the contingency table are returned in the iteration order of the reference's dict of dicts
 * getContTable(col[p+1], col[p]): rows = clusters of column p+1 by first appearance, cells of a row
 * by first appearance.  cell_off: T entries (cell_off[0] = 0; pair p owns [cell_off[p], cell_off[p+1]));
 * cell_row / cell_col / cell_count: room for (T-1) * n entries each.                            */
int rthc_contingency_adjacent(rthc_ctx *ctx, const int32_t *codes, int T, int64_t n, int64_t *cell_off,
                              int32_t *cell_row, int32_t *cell_col, int32_t *cell_count);
/* Host finish of one table (no device work): the two row sums of getAdjustedWallace
 * (metrics.py:147-165: rSumW1, rSumW2) accumulated in the reference's order with libm pow where the
 * reference writes `** 2.0` / `** 3.0`, from the cells of one pair as returned above.            */
int rthc_adjusted_wallace_sums(const int32_t *cell_row, const int32_t *cell_count, int64_t m, double *sum_w1,
                               double *sum_w2);

/* ---- introspection ----------------------------------------------------------- */
/* info[0]=n, [1]=L, [2]=loci kept, [3]=groups of 32 loci, [4]=plane rows,
 * [5]=tiles launched, [6]=Boruvka rounds of the last linkage, [7]=max alleles/locus,
 * [8]=row scans of the last merge loop (own barrier + fused with a merge),
 * [9]=grid barriers of the last merge loop, [10]=chain steps answered by the nearest-neighbour
 * cache (no scan), [11]=loci re-coded with an escape code for their singleton alleles,
 * [12]=cells holding it, [13]=pairs fixed up after the plane kernel (rthc_escape_correct)   */
int rthc_info(const rthc_ctx *ctx, int64_t *info, int n_info);
/* device-side durations (CUDA events, ms) of the last call of each stage:
 * [0]=h2d, [1]=recode+pack, [2]=distance kernel, [3]=mst/merge loop, [4]=cut,
 * [5]=number of kernel launches since create, [6]=rthc_contingency_adjacent        */
int rthc_timings(const rthc_ctx *ctx, double *ms, int n_ms);
/* caller-placed CUDA events on the library's own stream (slot 0 / 1): bench.py brackets its timed
 * region with them, so the step time is measured on the device, on the stream the kernels run on */
int rthc_mark(rthc_ctx *ctx, int slot);
int rthc_mark_elapsed_ms(rthc_ctx *ctx, double *ms);   /* waits for slot 1, returns t(1) - t(0) */
/* integer-ALU roofline probe: independent LOP3 streams at full occupancy, lane-ops per second */
int rthc_alu_peak(rthc_ctx *ctx, double *lane_ops_per_s);
/* same, with the distance kernel's operand pattern (8 x 8 register block, three distinct register
 * sources per LOP3) at warps_per_sm resident warps: the ceiling of that loop on this chip */
int rthc_alu_peak_block(rthc_ctx *ctx, int warps_per_sm, double *lane_ops_per_s);

#ifdef __cplusplus
}
#endif
#endif /* RTHC_H */
