/*
 * oracle/hclust.c -- TEST INFRASTRUCTURE ONLY (CPU oracle, not the product path).
 *
 * Plain-C restatement of the scipy.cluster.hierarchy routines that
 * scripts/partitioning_HC.py calls in-process:
 *
 *   linkage(y, "single")                      scripts/partitioning_HC.py:197
 *       -> _hierarchy.mst_single_linkage + label   (scipy _hierarchy_impl.py:960-961)
 *   linkage(y, "complete"|"average"|"weighted"|"ward")
 *       -> _hierarchy.nn_chain + label             (scipy _hierarchy_impl.py:962-963)
 *   maxdists(Z)[-1]                           scripts/partitioning_HC.py:198
 *   fcluster(Z, t, criterion="distance")      scripts/partitioning_HC.py:649,689,722,751
 *       -> _hierarchy.cluster_dist                 (scipy _hierarchy_impl.py:2589-2590)
 *
 * scipy ships these only as a compiled extension (no .pyx in the wheel); the
 * algorithms are restated from scipy's published description (SURVEY.md section 8,
 * rows A7-A9) and PINNED against the live scipy 1.18.1 in this image on
 * tie-heavy integer matrices (tests/test_oracle_golden.py): identical Z and
 * identical fcluster labels.  scipy is importable on the GPU box too, so GPU
 * tests compare with scipy directly as well.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

static inline int64_t cidx(int64_t n, int64_t i, int64_t j)
{
    if (i > j) { int64_t t = i; i = j; j = t; }
    return n * i - (i * (i + 1)) / 2 + (j - i - 1);
}

/* stable merge sort of rows (x, y, d) by d: indices only */
static void stable_order_by_dist(const double *Z3, int64_t m, int64_t *order)
{
    int64_t *tmp = (int64_t *)malloc(sizeof(int64_t) * (size_t)(m > 0 ? m : 1));
    for (int64_t i = 0; i < m; ++i) order[i] = i;
    for (int64_t w = 1; w < m; w *= 2) {
        for (int64_t lo = 0; lo < m; lo += 2 * w) {
            int64_t mid = lo + w < m ? lo + w : m, hi = lo + 2 * w < m ? lo + 2 * w : m;
            int64_t a = lo, b = mid, k = lo;
            while (a < mid && b < hi)
                tmp[k++] = (Z3[3 * order[b] + 2] < Z3[3 * order[a] + 2]) ? order[b++] : order[a++];
            while (a < mid) tmp[k++] = order[a++];
            while (b < hi) tmp[k++] = order[b++];
        }
        memcpy(order, tmp, sizeof(int64_t) * (size_t)m);
    }
    free(tmp);
}

/* scipy `label`: union-find relabel of sorted merges into scipy Z ids. */
static void label_rows(const double *Z3, const int64_t *order, int64_t n, double *Z)
{
    int64_t m = n - 1;
    int64_t *parent = (int64_t *)malloc(sizeof(int64_t) * (size_t)(2 * n));
    int64_t *size = (int64_t *)malloc(sizeof(int64_t) * (size_t)(2 * n));
    for (int64_t i = 0; i < 2 * n; ++i) { parent[i] = -1; size[i] = i < n ? 1 : 0; }
    int64_t next = n;
    for (int64_t k = 0; k < m; ++k) {
        const double *r = Z3 + 3 * order[k];
        int64_t x = (int64_t)r[0], y = (int64_t)r[1];
        int64_t rx = x, ry = y, t;
        while (parent[rx] != -1) rx = parent[rx];
        while (parent[ry] != -1) ry = parent[ry];
        while (parent[x] != -1) { t = parent[x]; parent[x] = rx; x = t; }   /* compress */
        while (parent[y] != -1) { t = parent[y]; parent[y] = ry; y = t; }
        Z[4 * k + 0] = (double)(rx < ry ? rx : ry);
        Z[4 * k + 1] = (double)(rx < ry ? ry : rx);
        Z[4 * k + 2] = r[2];
        size[next] = size[rx] + size[ry];
        Z[4 * k + 3] = (double)size[next];
        parent[rx] = next;
        parent[ry] = next;
        ++next;
    }
    free(parent);
    free(size);
}

/* scipy mst_single_linkage: Prim from vertex 0, first strictly-smaller wins. */
void oracle_linkage_single(const double *y, int64_t n, double *Z)
{
    int64_t m = n - 1;
    double *Z3 = (double *)malloc(sizeof(double) * 3 * (size_t)(m > 0 ? m : 1));
    double *D = (double *)malloc(sizeof(double) * (size_t)n);
    char *merged = (char *)calloc((size_t)n, 1);
    for (int64_t i = 0; i < n; ++i) D[i] = INFINITY;
    int64_t x = 0, yv = 0;
    for (int64_t k = 0; k < m; ++k) {
        double cur = INFINITY;
        merged[x] = 1;
        for (int64_t i = 0; i < n; ++i) {
            if (merged[i]) continue;
            double d = y[cidx(n, x, i)];
            if (D[i] > d) D[i] = d;
            if (D[i] < cur) { yv = i; cur = D[i]; }
        }
        Z3[3 * k] = (double)x; Z3[3 * k + 1] = (double)yv; Z3[3 * k + 2] = cur;
        x = yv;
    }
    int64_t *order = (int64_t *)malloc(sizeof(int64_t) * (size_t)(m > 0 ? m : 1));
    stable_order_by_dist(Z3, m, order);
    label_rows(Z3, order, n, Z);
    free(order); free(merged); free(D); free(Z3);
}

/* Prim-order edge list only (x, y, w), before sorting: used to check the device Prim. */
void oracle_prim_edges(const double *y, int64_t n, double *Z3)
{
    int64_t m = n - 1;
    double *D = (double *)malloc(sizeof(double) * (size_t)n);
    char *merged = (char *)calloc((size_t)n, 1);
    for (int64_t i = 0; i < n; ++i) D[i] = INFINITY;
    int64_t x = 0, yv = 0;
    for (int64_t k = 0; k < m; ++k) {
        double cur = INFINITY;
        merged[x] = 1;
        for (int64_t i = 0; i < n; ++i) {
            if (merged[i]) continue;
            double d = y[cidx(n, x, i)];
            if (D[i] > d) D[i] = d;
            if (D[i] < cur) { yv = i; cur = D[i]; }
        }
        Z3[3 * k] = (double)x; Z3[3 * k + 1] = (double)yv; Z3[3 * k + 2] = cur;
        x = yv;
    }
    free(merged); free(D);
}

/* method: 1 complete, 2 average, 3 weighted, 4 ward (0 single is allowed too: min).
 * Ward: scipy's update takes t = 1/(nx+ny+ni) first and multiplies left to right; any other
 * association rounds differently (pinned on the live scipy, tests/test_oracle_golden.py). */
static inline double lance_williams(int method, double dxi, double dyi, double dxy, int64_t nx, int64_t ny,
                                    int64_t ni)
{
    switch (method) {
    case 0: return dxi < dyi ? dxi : dyi;
    case 1: return dxi > dyi ? dxi : dyi;
    case 2: return ((double)nx * dxi + (double)ny * dyi) / ((double)nx + (double)ny);
    case 4: {
        double t = 1.0 / (double)(nx + ny + ni);
        return sqrt((double)(ni + nx) * t * dxi * dxi + (double)(ni + ny) * t * dyi * dyi
                    - (double)ni * t * dxy * dxy);
    }
    default: return 0.5 * (dxi + dyi);
    }
}

/* scipy nn_chain.  y is copied (scipy works on a copy of the condensed matrix). */
void oracle_linkage_nn_chain(const double *y_in, int64_t n, int method, double *Z)
{
    int64_t m = n - 1, P = n * (n - 1) / 2;
    double *D = (double *)malloc(sizeof(double) * (size_t)(P > 0 ? P : 1));
    memcpy(D, y_in, sizeof(double) * (size_t)P);
    double *Z3 = (double *)malloc(sizeof(double) * 3 * (size_t)(m > 0 ? m : 1));
    int64_t *size = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    int64_t *chain = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    for (int64_t i = 0; i < n; ++i) size[i] = 1;
    int64_t len = 0;
    for (int64_t k = 0; k < m; ++k) {
        int64_t x = 0, yv = 0;
        double cur = 0;
        if (len == 0) {
            len = 1;
            for (int64_t i = 0; i < n; ++i)
                if (size[i] > 0) { chain[0] = i; break; }
        }
        for (;;) {
            x = chain[len - 1];
            if (len > 1) { yv = chain[len - 2]; cur = D[cidx(n, x, yv)]; }
            else cur = INFINITY;
            for (int64_t i = 0; i < n; ++i) {
                if (size[i] == 0 || x == i) continue;
                double d = D[cidx(n, x, i)];
                if (d < cur) { cur = d; yv = i; }
            }
            if (len > 1 && yv == chain[len - 2]) break;
            chain[len++] = yv;
        }
        len -= 2;
        if (x > yv) { int64_t t = x; x = yv; yv = t; }
        int64_t nx = size[x], ny = size[yv];
        Z3[3 * k] = (double)x; Z3[3 * k + 1] = (double)yv; Z3[3 * k + 2] = cur;
        size[x] = 0; size[yv] = nx + ny;
        for (int64_t i = 0; i < n; ++i) {
            if (size[i] == 0 || i == yv) continue;
            D[cidx(n, i, yv)] = lance_williams(method, D[cidx(n, i, x)], D[cidx(n, i, yv)], cur, nx, ny, size[i]);
        }
    }
    int64_t *order = (int64_t *)malloc(sizeof(int64_t) * (size_t)(m > 0 ? m : 1));
    stable_order_by_dist(Z3, m, order);
    label_rows(Z3, order, n, Z);
    free(order); free(chain); free(size); free(Z3); free(D);
}

/* scipy get_max_dist_for_each_cluster (maxdists): MD[k] = max(Z[k,2], MD[children]). */
void oracle_maxdists(const double *Z, int64_t n, double *MD)
{
    for (int64_t k = 0; k < n - 1; ++k) {
        double md = Z[4 * k + 2];
        int64_t l = (int64_t)Z[4 * k], r = (int64_t)Z[4 * k + 1];
        if (l >= n && MD[l - n] > md) md = MD[l - n];
        if (r >= n && MD[r - n] > md) md = MD[r - n];
        MD[k] = md;
    }
}

/* scipy cluster_dist = cluster_monocrit(Z, maxdists(Z), T, cutoff, n):
 * iterative DFS, internal children first (left, then right); leader = first node on
 * the path with MD <= cutoff; leaf children numbered when their parent is popped. */
void oracle_fcluster_distance(const double *Z, int64_t n, double cutoff, int32_t *T)
{
    if (n == 1) { T[0] = 1; return; }
    double *MD = (double *)malloc(sizeof(double) * (size_t)(n - 1));
    oracle_maxdists(Z, n, MD);
    int64_t *stack = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    char *visited = (char *)calloc((size_t)(2 * n), 1);
    int64_t k = 0, leader = -1;
    int32_t ncl = 0;
    stack[0] = 2 * n - 2;
    while (k >= 0) {
        int64_t root = stack[k] - n;
        int64_t lc = (int64_t)Z[4 * root], rc = (int64_t)Z[4 * root + 1];
        if (leader == -1 && MD[root] <= cutoff) { leader = root; ++ncl; }
        if (lc >= n && !visited[lc]) { visited[lc] = 1; stack[++k] = lc; continue; }
        if (rc >= n && !visited[rc]) { visited[rc] = 1; stack[++k] = rc; continue; }
        if (lc < n) { if (leader == -1) ++ncl; T[lc] = ncl; }
        if (rc < n) { if (leader == -1) ++ncl; T[rc] = ncl; }
        if (leader == root) leader = -1;
        --k;
    }
    free(visited); free(stack); free(MD);
}

/* ---- uint16 condensed input (integer allelic distances as the distance kernel stores them) ----
 * Same algorithms, same visiting order and tie rules as the float64 entry points above; only the
 * storage type of the condensed matrix differs, so that configs[3] (100,000 samples: 10 GB instead of
 * 40 GB) and configs[4] (2.3 GB instead of 9 GB) fit a workstation.  Every uint16 is exactly
 * representable as a double, so Z is bit-identical to the float64 path (tests/test_oracle_golden.py
 * checks that on the same tie-heavy matrices that pin the float64 path on scipy). */
void oracle_linkage_single_u16(const uint16_t *y, int64_t n, double *Z)
{
    int64_t m = n - 1;
    double *Z3 = (double *)malloc(sizeof(double) * 3 * (size_t)(m > 0 ? m : 1));
    double *D = (double *)malloc(sizeof(double) * (size_t)n);
    char *merged = (char *)calloc((size_t)n, 1);
    for (int64_t i = 0; i < n; ++i) D[i] = INFINITY;
    int64_t x = 0, yv = 0;
    for (int64_t k = 0; k < m; ++k) {
        double cur = INFINITY;
        merged[x] = 1;
        for (int64_t i = 0; i < n; ++i) {
            if (merged[i]) continue;
            double d = (double)y[cidx(n, x, i)];
            if (D[i] > d) D[i] = d;
            if (D[i] < cur) { yv = i; cur = D[i]; }
        }
        Z3[3 * k] = (double)x; Z3[3 * k + 1] = (double)yv; Z3[3 * k + 2] = cur;
        x = yv;
    }
    int64_t *order = (int64_t *)malloc(sizeof(int64_t) * (size_t)(m > 0 ? m : 1));
    stable_order_by_dist(Z3, m, order);
    label_rows(Z3, order, n, Z);
    free(order); free(merged); free(D); free(Z3);
}

/* complete (method 1) and single-by-chain (0) never leave the integers: the working copy stays uint16.
 * average / weighted / ward need the float64 working copy and go through oracle_linkage_nn_chain. */
void oracle_linkage_nn_chain_u16(const uint16_t *y_in, int64_t n, int method, double *Z)
{
    int64_t m = n - 1, P = n * (n - 1) / 2;
    if (method != 0 && method != 1) {
        double *yf = (double *)malloc(sizeof(double) * (size_t)(P > 0 ? P : 1));
        for (int64_t p = 0; p < P; ++p) yf[p] = (double)y_in[p];
        oracle_linkage_nn_chain(yf, n, method, Z);
        free(yf);
        return;
    }
    uint16_t *D = (uint16_t *)malloc(sizeof(uint16_t) * (size_t)(P > 0 ? P : 1));
    memcpy(D, y_in, sizeof(uint16_t) * (size_t)P);
    double *Z3 = (double *)malloc(sizeof(double) * 3 * (size_t)(m > 0 ? m : 1));
    int64_t *size = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    int64_t *chain = (int64_t *)malloc(sizeof(int64_t) * (size_t)n);
    for (int64_t i = 0; i < n; ++i) size[i] = 1;
    int64_t len = 0;
    for (int64_t k = 0; k < m; ++k) {
        int64_t x = 0, yv = 0;
        int32_t cur = 0;                       /* 0x10000 plays +inf: above every uint16 */
        if (len == 0) {
            len = 1;
            for (int64_t i = 0; i < n; ++i)
                if (size[i] > 0) { chain[0] = i; break; }
        }
        for (;;) {
            x = chain[len - 1];
            if (len > 1) { yv = chain[len - 2]; cur = D[cidx(n, x, yv)]; }
            else cur = 0x10000;
            for (int64_t i = 0; i < n; ++i) {
                if (size[i] == 0 || x == i) continue;
                int32_t d = D[cidx(n, x, i)];
                if (d < cur) { cur = d; yv = i; }
            }
            if (len > 1 && yv == chain[len - 2]) break;
            chain[len++] = yv;
        }
        len -= 2;
        if (x > yv) { int64_t t = x; x = yv; yv = t; }
        int64_t nx = size[x], ny = size[yv];
        Z3[3 * k] = (double)x; Z3[3 * k + 1] = (double)yv; Z3[3 * k + 2] = (double)cur;
        size[x] = 0; size[yv] = nx + ny;
        for (int64_t i = 0; i < n; ++i) {
            if (size[i] == 0 || i == yv) continue;
            uint16_t dxi = D[cidx(n, i, x)], dyi = D[cidx(n, i, yv)];
            D[cidx(n, i, yv)] = method == 1 ? (dxi > dyi ? dxi : dyi) : (dxi < dyi ? dxi : dyi);
        }
    }
    int64_t *order = (int64_t *)malloc(sizeof(int64_t) * (size_t)(m > 0 ? m : 1));
    stable_order_by_dist(Z3, m, order);
    label_rows(Z3, order, n, Z);
    free(order); free(chain); free(size); free(Z3); free(D);
}
