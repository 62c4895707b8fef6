/*
 * oracle/cgmlst_dists.c -- TEST INFRASTRUCTURE ONLY (CPU oracle, not the product path).
 *
 * Plain-C restatement of the contract of the external program `cgmlst-dists`
 * (conda package cgmlst-dists_genpat-it 0.4.0, reportree_env.yml:11; bioconda
 * cgmlst-dists 0.4.0, Dockerfile:5) that ReporTree shells out to at
 * scripts/partitioning_HC.py:502 and scripts/partitioning_grapetree.py:466.
 * The program's source is NOT vendored under /root/reference, so this file
 * restates the published behaviour:
 *
 *   - input : TSV, first row = header (corner cell + locus names), first
 *             column = sample id               (scripts/partitioning_HC.py:496-499)
 *   - cells : decimal non-negative integer -> that integer; anything else -> 0
 *             ("All the other values will be considered as missing data",
 *             scripts/partitioning_HC.py:293-298)
 *   - dist  : d(i,j) = #{ l : a_il != a_jl  and  a_il != 0  and  a_jl != 0 }
 *   - output: full square TSV, header = "cgmlst-dists" corner + ids, one row
 *             per sample: id + n integers    (read back at :508-510)
 *
 * PARITY PIN: the distance rule is pinned by the golden pair shipped with the
 * reference, examples/Listeria/output/Lm_flt_samples_matrix.tsv ->
 * examples/Listeria/output/Lm_dist_hamming.tsv (all 861x861 cells equal; see
 * tests/test_oracle_golden.py and tests/golden/make_golden.py).  Behaviour on
 * tokens that are neither decimal integers nor single letters (e.g. "LNF") is
 * "follows the stated contract, upstream behaviour unverified".
 *
 * Build:  gcc -O3 -march=native -shared -fPIC -o liboracle.so ...   (library)
 *         gcc -O3 -march=native -DORACLE_MAIN -o cgmlst-dists ...   (CLI stand-in)
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* token -> allele code; 0 = missing.  Clean non-negative decimal integer only. */
int32_t oracle_parse_cell(const char *tok, size_t len)
{
    if (len == 0 || len > 9) return 0;          /* > 9 digits: out of int32 contract */
    int32_t v = 0;
    for (size_t k = 0; k < len; ++k) {
        unsigned c = (unsigned char)tok[k];
        if (c < '0' || c > '9') return 0;
        v = v * 10 + (int32_t)(c - '0');
    }
    return v;
}

/* Square n x n int32 distance matrix from an n x L row-major int32 code matrix. */
void oracle_dist_square_i32(const int32_t *codes, int64_t n, int64_t L, int32_t *out)
{
    for (int64_t i = 0; i < n; ++i) {
        const int32_t *a = codes + i * L;
        out[i * n + i] = 0;
        for (int64_t j = i + 1; j < n; ++j) {
            const int32_t *b = codes + j * L;
            int32_t d = 0;
            for (int64_t l = 0; l < L; ++l)
                d += (a[l] != b[l]) & (a[l] != 0) & (b[l] != 0);
            out[i * n + j] = d;
            out[j * n + i] = d;
        }
    }
}

/* Condensed (scipy pdist order) uint16 distances for rows [row0, row1) only.
 * out must hold sum_{i in rows}(n-1-i) entries, packed row after row.        */
void oracle_dist_condensed_rows_u16(const int32_t *codes, int64_t n, int64_t L,
                                    int64_t row0, int64_t row1, uint16_t *out)
{
    int64_t p = 0;
    for (int64_t i = row0; i < row1; ++i) {
        const int32_t *a = codes + i * L;
        for (int64_t j = i + 1; j < n; ++j) {
            const int32_t *b = codes + j * L;
            int32_t d = 0;
            for (int64_t l = 0; l < L; ++l)
                d += (a[l] != b[l]) & (a[l] != 0) & (b[l] != 0);
            out[p++] = (uint16_t)d;
        }
    }
}

/* Chunk-sum protocol of scripts/partitioning_HC.py:489-516: distances computed per
 * 2,000-column chunk and added.  Mathematically equal to one pass; kept so that the
 * equivalence is tested rather than assumed.                                     */
void oracle_dist_square_chunked_i32(const int32_t *codes, int64_t n, int64_t L,
                                    int64_t chunk, int32_t *out)
{
    memset(out, 0, sizeof(int32_t) * (size_t)(n * n));
    for (int64_t c0 = 0; c0 < L; c0 += chunk) {
        int64_t c1 = c0 + chunk < L ? c0 + chunk : L;
        for (int64_t i = 0; i < n; ++i)
            for (int64_t j = i + 1; j < n; ++j) {
                const int32_t *a = codes + i * L, *b = codes + j * L;
                int32_t d = 0;
                for (int64_t l = c0; l < c1; ++l)
                    d += (a[l] != b[l]) & (a[l] != 0) & (b[l] != 0);
                out[i * n + j] += d;
                out[j * n + i] += d;
            }
    }
}

#ifdef ORACLE_MAIN
/* CLI stand-in:  cgmlst-dists in.tsv > out.tsv */
static char *read_all(const char *path, size_t *len)
{
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    char *buf = (char *)malloc((size_t)sz + 2);
    if (fread(buf, 1, (size_t)sz, f) != (size_t)sz) { fclose(f); free(buf); return NULL; }
    fclose(f);
    if (sz > 0 && buf[sz - 1] != '\n') buf[sz++] = '\n';
    buf[sz] = 0;
    *len = (size_t)sz;
    return buf;
}

int main(int argc, char **argv)
{
    if (argc < 2) { fprintf(stderr, "usage: cgmlst-dists in.tsv > out.tsv\n"); return 2; }
    size_t len = 0;
    char *buf = read_all(argv[1], &len);
    if (!buf) { fprintf(stderr, "ERROR: cannot read %s\n", argv[1]); return 1; }

    /* count rows / header columns */
    int64_t nlines = 0;
    for (size_t k = 0; k < len; ++k) nlines += buf[k] == '\n';
    char *p = buf, *eol = memchr(p, '\n', len);
    int64_t L = 0;
    for (char *q = p; q < eol; ++q) L += *q == '\t';
    int64_t n = nlines - 1;
    if (n < 1 || L < 1) { fprintf(stderr, "ERROR: empty matrix\n"); return 1; }

    char **ids = (char **)calloc((size_t)n, sizeof(char *));
    int32_t *codes = (int32_t *)calloc((size_t)(n * L), sizeof(int32_t));
    p = eol + 1;
    int64_t r = 0;
    while (r < n && p < buf + len) {
        eol = memchr(p, '\n', (size_t)(buf + len - p));
        char *e = eol;
        if (e > p && e[-1] == '\r') --e;
        if (e == p) { p = eol + 1; continue; }          /* blank line */
        char *t = memchr(p, '\t', (size_t)(e - p));
        if (!t) t = e;
        *t = 0;
        ids[r] = p;
        char *c = t + 1;
        for (int64_t l = 0; l < L && c <= e; ++l) {
            char *ce = memchr(c, '\t', c < e ? (size_t)(e - c) : 0);
            if (!ce) ce = e;
            codes[r * L + l] = oracle_parse_cell(c, (size_t)(ce - c));
            c = ce + 1;
        }
        *e = 0;
        ++r;
        p = eol + 1;
    }
    n = r;
    int32_t *out = (int32_t *)malloc(sizeof(int32_t) * (size_t)(n * n));
    oracle_dist_square_i32(codes, n, L, out);

    fputs("cgmlst-dists", stdout);
    for (int64_t i = 0; i < n; ++i) { fputc('\t', stdout); fputs(ids[i], stdout); }
    fputc('\n', stdout);
    for (int64_t i = 0; i < n; ++i) {
        fputs(ids[i], stdout);
        for (int64_t j = 0; j < n; ++j) fprintf(stdout, "\t%d", out[i * n + j]);
        fputc('\n', stdout);
    }
    return 0;
}
#endif
