/*
 * rthio.h -- C ABI of librthc_hostio.so: host-side text ingest / artefact writers of the drop-in
 * partitioning_HC.py (plain C++, no CUDA).  SURVEY.md section 8(f) N2: once the numeric layer runs
 * on the GPU, wall time is the reference's pandas string handling; these entry points replace it
 * with the same string-level semantics (citations relative to the reference tree):
 *
 *   rthio_profile_open      <- pandas.read_table(dtype=str) + the two regex replaces + the
 *                              --missing-code handling, scripts/partitioning_HC.py:362-372,
 *                              conv_nucl (:28-47) and the cell contract of cgmlst-dists (:293-298)
 *   rthio_profile_write     <- DataFrame.to_csv of <out>_missing0.tsv (:372), _subset_matrix.tsv
 *                              (:392), _clean_missing_matrix.tsv (:443), _flt_samples_matrix.tsv (:477)
 *   rthio_square_*          <- dist_df.to_csv(<out>_dist_hamming.tsv) (:517-518)
 *   rthio_write_partitions  <- the naming loop (:651-670), get_cluster_composition (:217-230) and
 *                              the two output files (:779-787)
 *
 * rthio_profile_open returns RTHIO_IRREGULAR for inputs whose pandas parse is not the plain
 * rectangular case (quotes, CR, ragged rows, empty or duplicated header cells, BOM): the caller
 * then takes the pandas path, which stays the specification.
 */
#ifndef RTHIO_H
#define RTHIO_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

typedef struct rthio_profile rthio_profile;
typedef struct rthio_file rthio_file;

enum { RTHIO_OK = 0, RTHIO_IRREGULAR = 1, RTHIO_E_IO = -1, RTHIO_E_ARG = -2, RTHIO_E_NOMEM = -3 };

const char *rthio_last_error(void);
int rthio_version(void);

/* ---- allele / SNP profile --------------------------------------------------------------------
 * missing_code: the --missing-code argument ("0" = nothing to do, "empty" = NaN cells, else a
 * whole-cell string).  n_threads <= 0: all cores. */
int rthio_profile_open(const char *path, const char *missing_code, int n_threads, rthio_profile **out);
void rthio_profile_close(rthio_profile *p);
int64_t rthio_profile_rows(const rthio_profile *p);          /* samples                        */
int64_t rthio_profile_cols(const rthio_profile *p);          /* loci (id column not counted)   */
/* header cell `col` (0 = name of the id column, 1..L = loci) */
int rthio_profile_header(const rthio_profile *p, int64_t col, const char **s, int64_t *len);
/* cleaned sample id of `row`; *is_nan = 1 when pandas would have parsed it as NaN */
int rthio_profile_sample(const rthio_profile *p, int64_t row, const char **s, int64_t *len, int *is_nan);
/* every name in two calls' worth of data instead of n + L calls: header cells joined by '\t', cleaned sample ids
 * joined by '\n' (neither can contain its separator), id_is_nan[row] = 1 where pandas has NaN.  Owned by the profile. */
int rthio_profile_names(const rthio_profile *p, const char **header_blob, int64_t *header_len, const char **ids_blob,
                        int64_t *ids_len, const char **id_is_nan);
/* n x L int32 codes, row-major, owned by the profile:
 *    > 0  allele code (integer value, or 1..52 for a single ASCII letter)
 *      0  the cell is the literal string "0" after cleaning / missing-code replacement
 *     -1  any other cell (NaN, "LNF", "", "00", ...): missing for the distance, but != "0"
 * so the reference's string filters are (code != 0) / (code == 0) and the GPU re-coder, which
 * treats every value <= 0 as missing, takes the matrix as is. */
const int32_t *rthio_profile_codes(const rthio_profile *p);
/* out[r * ncols + c] = codes[rows[r]][cols[c]] (cols are 0-based locus indices) */
int rthio_profile_gather(const rthio_profile *p, const int64_t *rows, int64_t nrows, const int64_t *cols,
                         int64_t ncols, int32_t *out);
/* out[r] = #{c : codes[rows[r]][cols[c]] == value}: the per-sample count behind
 * `isin(["0"]).sum(axis=1)` (scripts/partitioning_HC.py:451) with value = 0, without materialising the sub-matrix */
int rthio_profile_count_eq(const rthio_profile *p, const int64_t *rows, int64_t nrows, const int64_t *cols,
                           int64_t ncols, int32_t value, int64_t *out);
/* TSV of the selected rows / loci with the cleaned cell text, as DataFrame.to_csv writes it */
int rthio_profile_write(const rthio_profile *p, const char *path, const int64_t *rows, int64_t nrows,
                        const int64_t *cols, int64_t ncols);

/* ---- square distance matrix ------------------------------------------------------------------ */
/* names: n strings concatenated in `blob`, string k = blob[offs[k] .. offs[k+1]) */
int rthio_square_open(const char *path, const char *corner, const char *blob, const int64_t *offs, int64_t n,
                      int n_threads, rthio_file **out);
/* rows [row0, row0 + nrows) of the square matrix, uint16, row-major nrows x n */
int rthio_square_append_u16(rthio_file *f, int64_t row0, int64_t nrows, const uint16_t *block);
int rthio_square_close(rthio_file *f);

/* ---- partitions + cluster composition --------------------------------------------------------
 * labels: T x n int32 fcluster ids (1..k per row).  colnames: T partition column names.
 * Writes <path_partitions> (first column "sequence") and <path_composition>. */
int rthio_write_partitions(const char *path_partitions, const char *path_composition, const char *names_blob,
                           const int64_t *names_offs, int64_t n, const int32_t *labels, int64_t T,
                           const char *col_blob, const int64_t *col_offs, int n_threads);

/* ---- synthetic inputs (bench / tests): n x L int32 matrix -> profile TSV with header
 * <corner>\t<col_prefix>00001.. and ids <row_prefix>0000001..                                   */
int rthio_write_int_matrix(const char *path, const char *corner, const char *row_prefix, const char *col_prefix,
                           const int32_t *codes, int64_t n, int64_t L, int n_threads);

#ifdef __cplusplus
}
#endif
#endif /* RTHIO_H */
