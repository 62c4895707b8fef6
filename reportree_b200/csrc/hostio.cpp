// hostio.cpp -- librthc_hostio.so: text ingest and artefact writers of the drop-in CLI (include/rthio.h).
//
// Plain C++ (no CUDA).  Every function restates, at the string level, what the reference does with
// pandas in scripts/partitioning_HC.py; the pandas code path in reportree_b200/partitioning_HC.py
// stays the specification and the fallback for irregular files (RTHIO_IRREGULAR).
#include <fcntl.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>
#if defined(__x86_64__)
#include <immintrin.h>
#endif

#include <algorithm>
#include <atomic>
#include <chrono>
#include <string>
#include <thread>
#include <unordered_set>
#include <vector>

#include "../../include/rthio.h"

static thread_local std::string g_err;
static int fail(int code, const char *fmt, ...)
{
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    g_err = buf;
    return code;
}
extern "C" const char *rthio_last_error(void) { return g_err.c_str(); }
extern "C" int rthio_version(void) { return 100; }

static int pick_threads(int n_threads)
{
    if (n_threads > 0) return n_threads;
    unsigned hc = std::thread::hardware_concurrency();
    return hc ? (int)hc : 1;
}

// run fn(t, lo, hi) over [0, count) split into contiguous ranges, one per thread
template <typename F> static void parallel_ranges(int64_t count, int n_threads, F fn)
{
    int T = (int)std::min<int64_t>(std::max<int64_t>(count, 1), n_threads);
    if (T <= 1) { fn(0, (int64_t)0, count); return; }
    std::vector<std::thread> th;
    for (int t = 0; t < T; ++t) {
        const int64_t lo = count * t / T, hi = count * (t + 1) / T;
        th.emplace_back([=]() { fn(t, lo, hi); });
    }
    for (auto &x : th) x.join();
}

// ---- pandas' default NA tokens (read_table, keep_default_na=True; they apply with dtype=str too)
static bool is_na_token(const char *s, size_t n)
{
    switch (n) {
    case 0: return true;
    case 2: return !memcmp(s, "NA", 2);
    case 3: return !memcmp(s, "N/A", 3) || !memcmp(s, "NaN", 3) || !memcmp(s, "nan", 3) || !memcmp(s, "n/a", 3) ||
                   !memcmp(s, "#NA", 3);
    case 4: return !memcmp(s, "NULL", 4) || !memcmp(s, "null", 4) || !memcmp(s, "None", 4) || !memcmp(s, "#N/A", 4) ||
                   !memcmp(s, "-NaN", 4) || !memcmp(s, "-nan", 4) || !memcmp(s, "<NA>", 4);
    case 6: return !memcmp(s, "1.#IND", 6);
    case 7: return !memcmp(s, "-1.#IND", 7) || !memcmp(s, "1.#QNAN", 7);
    case 8: return !memcmp(s, "-1.#QNAN", 8) || !memcmp(s, "#N/A N/A", 8);
    default: return false;
    }
}

// The reference's cleaning (scripts/partitioning_HC.py:363-364): one regex pass removing every
// "INF-", then one removing every "*".  Returns the cleaned token (view into `tmp` if changed).
struct Tok {
    const char *s;
    size_t n;
};
static inline Tok clean_token(const char *s, size_t n, std::string &tmp)
{
    const bool has_star = memchr(s, '*', n) != nullptr;
    bool has_inf = false;
    if (n >= 4) {
        const char *p = s, *end = s + n - 3;
        while (p < end && (p = (const char *)memchr(p, 'I', (size_t)(end - p)))) {
            if (p[1] == 'N' && p[2] == 'F' && p[3] == '-') { has_inf = true; break; }
            ++p;
        }
    }
    if (!has_star && !has_inf) return Tok{s, n};
    tmp.clear();
    size_t i = 0;
    while (i < n) {
        if (has_inf && i + 4 <= n && s[i] == 'I' && s[i + 1] == 'N' && s[i + 2] == 'F' && s[i + 3] == '-') { i += 4; continue; }
        tmp.push_back(s[i++]);
    }
    if (has_star) tmp.erase(std::remove(tmp.begin(), tmp.end(), '*'), tmp.end());
    return Tok{tmp.data(), tmp.size()};
}

enum { MISS_NONE = 0, MISS_EMPTY = 1, MISS_STRING = 2 };

struct rthio_profile {
    const char *buf = nullptr;       // mmap of the file
    size_t size = 0;
    int fd = -1;
    int miss_mode = MISS_NONE;
    std::string miss_code;
    int n_threads = 1;
    int64_t n = 0, L = 0;
    std::vector<std::string> header;            // L + 1 cells
    std::vector<size_t> row_begin, row_end;     // data rows (blank lines skipped)
    std::vector<std::string> ids;               // cleaned sample ids
    std::vector<char> id_nan;
    std::vector<char> row_plain;                // 1: the final text of every cell of the row is its raw text
    int32_t *codes = nullptr;                   // n x L, anonymous mapping: pages are first touched (zero-filled) by the
    size_t codes_bytes = 0;                     // parser threads, not by one thread clearing 1.2 GB up front
    std::string ids_blob, header_blob;          // ids joined by '\n' / header cells joined by '\t' (one call instead of n)
};

// final text of a DATA cell (not the id column): cleaning + --missing-code replacement.
// *nan_out = the cell is NaN in the DataFrame (prints as the empty string).
static inline Tok final_cell(const rthio_profile *p, const char *s, size_t n, std::string &tmp, bool *nan_out)
{
    static const char ZERO[] = "0";
    if (is_na_token(s, n)) {
        if (p->miss_mode == MISS_EMPTY) { *nan_out = false; return Tok{ZERO, 1}; }
        *nan_out = true;
        return Tok{s, 0};
    }
    *nan_out = false;
    Tok t = clean_token(s, n, tmp);
    if (p->miss_mode == MISS_STRING && t.n == p->miss_code.size() && !memcmp(t.s, p->miss_code.data(), t.n))
        return Tok{ZERO, 1};
    return t;
}

// conv_nucl + the cgmlst-dists cell contract, see rthio.h
static inline int32_t cell_code(Tok t, bool nan)
{
    if (nan || t.n == 0) return -1;
    const unsigned char c0 = (unsigned char)t.s[0];
    if (t.n == 1) {
        if (c0 >= 'a' && c0 <= 'z') return c0 - 'a' + 1;
        if (c0 >= 'A' && c0 <= 'Z') return c0 - 'A' + 27;
    }
    if (t.n > 9) return -1;
    int32_t v = 0;
    for (size_t i = 0; i < t.n; ++i) {
        const unsigned d = (unsigned char)t.s[i] - '0';
        if (d > 9) return -1;
        v = v * 10 + (int32_t)d;
    }
    if (v == 0) return t.n == 1 ? 0 : -1;      // "0" is the literal missing code; "00" is not
    return v;
}

// ---------------------------------------------------------------------------------------------
// profile ingest
// ---------------------------------------------------------------------------------------------
extern "C" void rthio_profile_close(rthio_profile *p)
{
    if (!p) return;
    if (p->buf && p->size) munmap((void *)p->buf, p->size);
    if (p->codes && p->codes_bytes) munmap((void *)p->codes, p->codes_bytes);
    if (p->fd >= 0) close(p->fd);
    delete p;
}

// ---------------------------------------------------------------------------------------------
// data cells of one row -> int32 codes
// ---------------------------------------------------------------------------------------------
struct RowSink {
    const rthio_profile *p;
    int32_t *dst;
    int64_t L, c;
    bool plain;
    std::string *tmp;
};

// any token: the reference's cleaning + missing-code + conv_nucl + cell contract
static inline void emit_general(RowSink &S, const char *a, size_t len)
{
    if (S.c >= S.L) { S.c = S.L + 1; return; }
    bool nan;
    Tok t = final_cell(S.p, a, len, *S.tmp, &nan);
    if (nan || t.s != a || t.n != len) S.plain = false;
    S.dst[S.c++] = cell_code(t, nan);
}

// 1..8 decimal digits ending at `end`, without a loop: the 8 bytes before `end` are loaded, the bytes in front of
// the token are overwritten with '0', every byte is checked to be a digit and the value comes from three
// multiplications (first character in the low byte).  Needs end - 8 inside the mapping.
static inline bool digits8(const char *end, size_t len, uint32_t *out)
{
    uint64_t v;
    memcpy(&v, end - 8, 8);
    const unsigned sh = 8u * (unsigned)(8 - len);
    const uint64_t low = sh ? (~0ull >> (64 - sh)) : 0ull;
    v = (v & ~low) | (0x3030303030303030ull & low);
    const bool ok = ((v & 0xF0F0F0F0F0F0F0F0ull) == 0x3030303030303030ull) &
                    (((v + 0x0606060606060606ull) & 0xF0F0F0F0F0F0F0F0ull) == 0x3030303030303030ull);
    v -= 0x3030303030303030ull;
    v = v * 10 + (v >> 8);
    v = (((v & 0x000000FF000000FFull) * 0x000F424000000064ull) + (((v >> 16) & 0x000000FF000000FFull) * 0x0000271000000001ull)) >> 32;
    *out = (uint32_t)v;
    return ok;
}

#if defined(__x86_64__)
// The row's cells from `tok` (first data cell) to the line end `e`: tabs are found 32 bytes at a time, so the
// only data-dependent branch per cell is "is it 1..8 digits" (nearly always yes) instead of one unpredictable
// loop exit per token.  Caller guarantees tok - 8 and e + 32 lie inside the mapping.
__attribute__((target("avx2,bmi"))) static void row_cells_avx2(RowSink &S, const char *tok, const char *e)
{
    const __m256i tabs = _mm256_set1_epi8('\t');
    const char *pos = tok;
    while (pos < e) {
        const __m256i blk = _mm256_loadu_si256(reinterpret_cast<const __m256i *>(pos));
        uint32_t m = (uint32_t)_mm256_movemask_epi8(_mm256_cmpeq_epi8(blk, tabs));
        const size_t rem = (size_t)(e - pos);
        if (rem < 32) m &= (1u << rem) - 1u;
        while (m) {
            const char *end = pos + _tzcnt_u32(m);
            m &= m - 1;
            const size_t len = (size_t)(end - tok);
            uint32_t v;
            if (len - 1 < 8 && S.c < S.L && digits8(end, len, &v)) S.dst[S.c++] = v ? (int32_t)v : (len == 1 ? 0 : -1);
            else emit_general(S, tok, len);
            tok = end + 1;
        }
        pos += 32;
    }
    const size_t len = (size_t)(e - tok);
    uint32_t v;
    if (len - 1 < 8 && S.c < S.L && digits8(e, len, &v)) S.dst[S.c++] = v ? (int32_t)v : (len == 1 ? 0 : -1);
    else emit_general(S, tok, len);
}
#endif

// RTHIO_TIMING=1: phase times of the ingest on stderr
struct PhaseClock {
    bool on = getenv("RTHIO_TIMING") != nullptr;
    std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
    void mark(const char *what)
    {
        if (!on) return;
        const auto now = std::chrono::steady_clock::now();
        fprintf(stderr, "[rthio] %-24s %8.3f s\n", what, std::chrono::duration<double>(now - t).count());
        t = now;
    }
};

extern "C" int rthio_profile_open(const char *path, const char *missing_code, int n_threads, rthio_profile **out)
{
    if (!path || !out) return fail(RTHIO_E_ARG, "rthio_profile_open: NULL argument");
    *out = nullptr;
    PhaseClock clk;
    rthio_profile *p = new rthio_profile();
    p->n_threads = pick_threads(n_threads);
    const std::string mc = missing_code ? missing_code : "0";
    if (mc == "0") p->miss_mode = MISS_NONE;
    else if (mc == "empty") p->miss_mode = MISS_EMPTY;
    else { p->miss_mode = MISS_STRING; p->miss_code = mc; }
    p->fd = open(path, O_RDONLY);
    if (p->fd < 0) { rthio_profile_close(p); return fail(RTHIO_E_IO, "cannot open %s", path); }
    struct stat st;
    if (fstat(p->fd, &st) != 0 || st.st_size <= 0) { rthio_profile_close(p); return fail(RTHIO_IRREGULAR, "empty or unreadable file"); }
    p->size = (size_t)st.st_size;
    void *m = mmap(nullptr, p->size, PROT_READ, MAP_PRIVATE, p->fd, 0);
    if (m == MAP_FAILED) { p->size = 0; rthio_profile_close(p); return fail(RTHIO_E_IO, "mmap of %s failed", path); }
    p->buf = (const char *)m;
    madvise(m, p->size, MADV_SEQUENTIAL);
    const char *buf = p->buf;
    const size_t size = p->size;
    // anything pandas' C parser would treat specially -> the caller takes the pandas path
    if (size >= 3 && (unsigned char)buf[0] == 0xEF && (unsigned char)buf[1] == 0xBB && (unsigned char)buf[2] == 0xBF) {
        rthio_profile_close(p); return fail(RTHIO_IRREGULAR, "byte-order mark");
    }
    {
        std::atomic<int> odd(0);
        parallel_ranges((int64_t)size, p->n_threads, [&](int, int64_t lo, int64_t hi) {
            if (memchr(buf + lo, '"', (size_t)(hi - lo)) || memchr(buf + lo, '\r', (size_t)(hi - lo)) ||
                memchr(buf + lo, '\0', (size_t)(hi - lo)))
                odd.store(1);
        });
        if (odd.load()) { rthio_profile_close(p); return fail(RTHIO_IRREGULAR, "quote, CR or NUL in the file"); }
    }
    clk.mark("map + irregular scan");
    // lines (blank lines are skipped like skip_blank_lines=True)
    std::vector<size_t> lb, le;
    {
        std::vector<std::vector<size_t>> nls((size_t)p->n_threads);          // newline offsets, found in parallel
        parallel_ranges((int64_t)size, p->n_threads, [&](int t, int64_t lo, int64_t hi) {
            std::vector<size_t> &v = nls[(size_t)t];
            const char *q = buf + lo, *end = buf + hi;
            while (q < end && (q = (const char *)memchr(q, '\n', (size_t)(end - q)))) { v.push_back((size_t)(q - buf)); ++q; }
        });
        size_t pos = 0;
        for (const std::vector<size_t> &v : nls)
            for (const size_t end : v) {
                if (end > pos) { lb.push_back(pos); le.push_back(end); }
                pos = end + 1;
            }
        if (pos < size) { lb.push_back(pos); le.push_back(size); }            // last line without a newline
    }
    if (lb.size() < 1) { rthio_profile_close(p); return fail(RTHIO_IRREGULAR, "no header line"); }
    clk.mark("line split");
    // header
    {
        std::unordered_set<std::string> seen;
        size_t a = lb[0];
        const size_t e = le[0];
        for (;;) {
            const char *tab = (const char *)memchr(buf + a, '\t', e - a);
            const size_t b = tab ? (size_t)(tab - buf) : e;
            std::string cell(buf + a, b - a);
            if (cell.empty() || !seen.insert(cell).second) {
                rthio_profile_close(p); return fail(RTHIO_IRREGULAR, "empty or duplicated header cell");
            }
            p->header.push_back(cell);
            if (!tab) break;
            a = b + 1;
        }
    }
    if (p->header.size() < 2) { rthio_profile_close(p); return fail(RTHIO_IRREGULAR, "fewer than two columns"); }
    p->L = (int64_t)p->header.size() - 1;
    p->n = (int64_t)lb.size() - 1;
    p->row_begin.assign(lb.begin() + 1, lb.end());
    p->row_end.assign(le.begin() + 1, le.end());
    try {
        p->codes_bytes = std::max<size_t>(sizeof(int32_t) * (size_t)(p->n * p->L), 4096);
        void *cm = mmap(nullptr, p->codes_bytes, PROT_READ | PROT_WRITE, MAP_PRIVATE | MAP_ANONYMOUS, -1, 0);
        if (cm == MAP_FAILED) { p->codes_bytes = 0; throw std::bad_alloc(); }
        p->codes = (int32_t *)cm;
#ifdef MADV_HUGEPAGE
        madvise(cm, p->codes_bytes, MADV_HUGEPAGE);
#endif
        p->ids.resize((size_t)p->n);
        p->id_nan.assign((size_t)p->n, 0);
        p->row_plain.assign((size_t)p->n, 0);
    } catch (...) { rthio_profile_close(p); return fail(RTHIO_E_NOMEM, "out of host memory for %lld x %lld codes", (long long)p->n, (long long)p->L); }
    clk.mark("header + allocations");
    std::atomic<int> ragged(0);
    const int64_t L = p->L;
#if defined(__x86_64__)
    const bool simd = __builtin_cpu_supports("avx2") && __builtin_cpu_supports("bmi") && !getenv("RTHIO_NO_SIMD");
#endif
    parallel_ranges(p->n, p->n_threads, [&](int, int64_t lo, int64_t hi) {
        std::string tmp;
        for (int64_t r = lo; r < hi && !ragged.load(std::memory_order_relaxed); ++r) {
            const char *a = buf + p->row_begin[(size_t)r], *e = buf + p->row_end[(size_t)r];
            int32_t *dst = p->codes + (size_t)(r * L);
            // id cell
            const char *tab = (const char *)memchr(a, '\t', (size_t)(e - a));
            const char *b = tab ? tab : e;
            bool plain = true;
            if (is_na_token(a, (size_t)(b - a))) { p->id_nan[(size_t)r] = 1; plain = false; }
            else { Tok t = clean_token(a, (size_t)(b - a), tmp); p->ids[(size_t)r].assign(t.s, t.n); plain = t.s == a; }
            int64_t c = 0;
#if defined(__x86_64__)
            if (simd && tab && p->miss_mode != MISS_STRING && a - buf >= 8 && e + 32 <= buf + size) {
                RowSink S{p, dst, L, 0, plain, &tmp};
                row_cells_avx2(S, tab + 1, e);
                c = S.c; plain = S.plain;
                tab = nullptr;
            }
#endif
            while (tab) {                        // scalar path: no AVX2, --missing-code <string>, first / last bytes of the file
                a = tab + 1;
                if (c >= L) { c = L + 1; break; }
                tab = (const char *)memchr(a, '\t', (size_t)(e - a));
                b = tab ? tab : e;
                bool nan;
                const size_t len = (size_t)(b - a);
                if (len >= 1 && len <= 9 && p->miss_mode != MISS_STRING) {
                    int32_t v = 0; bool digits = true;
                    for (size_t i = 0; i < len; ++i) { const unsigned d = (unsigned char)a[i] - '0'; if (d > 9) { digits = false; break; } v = v * 10 + (int32_t)d; }
                    if (digits) { dst[c++] = v ? v : (len == 1 ? 0 : -1); continue; }
                }
                Tok t = final_cell(p, a, len, tmp, &nan);
                if (nan || t.s != a || t.n != len) plain = false;
                dst[c++] = cell_code(t, nan);
            }
            if (c != L) ragged.store(1);
            p->row_plain[(size_t)r] = plain ? 1 : 0;
        }
    });
    if (ragged.load()) { rthio_profile_close(p); return fail(RTHIO_IRREGULAR, "a row does not have %lld cells", (long long)(L + 1)); }
    clk.mark("cells");
    for (int64_t r = 0; r < p->n; ++r) {
        if (r) p->ids_blob.push_back('\n');
        p->ids_blob += p->ids[(size_t)r];
    }
    for (size_t c = 0; c < p->header.size(); ++c) {
        if (c) p->header_blob.push_back('\t');
        p->header_blob += p->header[c];
    }
    clk.mark("name blobs");
    *out = p;
    return RTHIO_OK;
}

extern "C" int64_t rthio_profile_rows(const rthio_profile *p) { return p ? p->n : 0; }
extern "C" int64_t rthio_profile_cols(const rthio_profile *p) { return p ? p->L : 0; }
extern "C" const int32_t *rthio_profile_codes(const rthio_profile *p) { return p ? p->codes : nullptr; }

extern "C" int rthio_profile_names(const rthio_profile *p, const char **header_blob, int64_t *header_len, const char **ids_blob,
                                   int64_t *ids_len, const char **id_is_nan)
{
    if (!p || !header_blob || !header_len || !ids_blob || !ids_len || !id_is_nan) return fail(RTHIO_E_ARG, "rthio_profile_names: NULL argument");
    *header_blob = p->header_blob.data(); *header_len = (int64_t)p->header_blob.size();
    *ids_blob = p->ids_blob.data(); *ids_len = (int64_t)p->ids_blob.size();
    *id_is_nan = p->id_nan.data();
    return RTHIO_OK;
}

extern "C" int rthio_profile_count_eq(const rthio_profile *p, const int64_t *rows, int64_t nrows, const int64_t *cols,
                                      int64_t ncols, int32_t value, int64_t *out)
{
    if (!p || (!rows && nrows) || (!cols && ncols) || !out) return fail(RTHIO_E_ARG, "rthio_profile_count_eq: NULL argument");
    for (int64_t r = 0; r < nrows; ++r) if (rows[r] < 0 || rows[r] >= p->n) return fail(RTHIO_E_ARG, "row index out of range");
    for (int64_t c = 0; c < ncols; ++c) if (cols[c] < 0 || cols[c] >= p->L) return fail(RTHIO_E_ARG, "column index out of range");
    const bool all_cols = ncols == p->L;       // ascending distinct columns are not required here, so test the identity
    bool identity = all_cols;
    for (int64_t c = 0; identity && c < ncols; ++c) identity = cols[c] == c;
    parallel_ranges(nrows, p->n_threads, [&](int, int64_t lo, int64_t hi) {
        for (int64_t r = lo; r < hi; ++r) {
            const int32_t *src = p->codes + (size_t)(rows[r] * p->L);
            int64_t k = 0;
            if (identity) for (int64_t c = 0; c < ncols; ++c) k += src[c] == value;
            else for (int64_t c = 0; c < ncols; ++c) k += src[cols[c]] == value;
            out[r] = k;
        }
    });
    return RTHIO_OK;
}

extern "C" int rthio_profile_header(const rthio_profile *p, int64_t col, const char **s, int64_t *len)
{
    if (!p || !s || !len || col < 0 || col > p->L) return fail(RTHIO_E_ARG, "rthio_profile_header: bad argument");
    *s = p->header[(size_t)col].data();
    *len = (int64_t)p->header[(size_t)col].size();
    return RTHIO_OK;
}

extern "C" int rthio_profile_sample(const rthio_profile *p, int64_t row, const char **s, int64_t *len, int *is_nan)
{
    if (!p || !s || !len || !is_nan || row < 0 || row >= p->n) return fail(RTHIO_E_ARG, "rthio_profile_sample: bad argument");
    *s = p->ids[(size_t)row].data();
    *len = (int64_t)p->ids[(size_t)row].size();
    *is_nan = p->id_nan[(size_t)row];
    return RTHIO_OK;
}

extern "C" int rthio_profile_gather(const rthio_profile *p, const int64_t *rows, int64_t nrows, const int64_t *cols,
                                    int64_t ncols, int32_t *out)
{
    if (!p || !rows || !cols || !out) return fail(RTHIO_E_ARG, "rthio_profile_gather: NULL argument");
    for (int64_t r = 0; r < nrows; ++r) if (rows[r] < 0 || rows[r] >= p->n) return fail(RTHIO_E_ARG, "row index out of range");
    for (int64_t c = 0; c < ncols; ++c) if (cols[c] < 0 || cols[c] >= p->L) return fail(RTHIO_E_ARG, "column index out of range");
    parallel_ranges(nrows, p->n_threads, [&](int, int64_t lo, int64_t hi) {
        for (int64_t r = lo; r < hi; ++r) {
            const int32_t *src = p->codes + (size_t)(rows[r] * p->L);
            int32_t *dst = out + (size_t)(r * ncols);
            for (int64_t c = 0; c < ncols; ++c) dst[c] = src[cols[c]];
        }
    });
    return RTHIO_OK;
}

static int write_all(int fd, const char *s, size_t n)
{
    while (n) {
        ssize_t w = write(fd, s, n);
        if (w <= 0) return -1;
        s += w; n -= (size_t)w;
    }
    return 0;
}

extern "C" int rthio_profile_write(const rthio_profile *p, const char *path, const int64_t *rows, int64_t nrows,
                                   const int64_t *cols, int64_t ncols)
{
    if (!p || !path || (!rows && nrows) || (!cols && ncols)) return fail(RTHIO_E_ARG, "rthio_profile_write: NULL argument");
    for (int64_t c = 0; c < ncols; ++c)
        if (cols[c] < 0 || cols[c] >= p->L || (c && cols[c] <= cols[c - 1])) return fail(RTHIO_E_ARG, "columns must be ascending and in range");
    for (int64_t r = 0; r < nrows; ++r) if (rows[r] < 0 || rows[r] >= p->n) return fail(RTHIO_E_ARG, "row index out of range");
    const int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return fail(RTHIO_E_IO, "cannot create %s", path);
    std::string head = p->header[0];
    for (int64_t c = 0; c < ncols; ++c) { head.push_back('\t'); head += p->header[(size_t)cols[c] + 1]; }
    head.push_back('\n');
    int rc = write_all(fd, head.data(), head.size());
    const int T = p->n_threads;
    const int64_t wave = (int64_t)T * 512;
    const bool all_cols = ncols == p->L;
    // two buffer sets: wave w is written by a helper thread while wave w + 1 is formatted
    std::vector<std::string> outb[2] = {std::vector<std::string>((size_t)T), std::vector<std::string>((size_t)T)};
    std::thread writer;
    std::atomic<int> wrc(0);
    int flip = 0;
    for (int64_t w0 = 0; w0 < nrows && !rc && !wrc.load(); w0 += wave, flip ^= 1) {
        const int64_t w1 = std::min(nrows, w0 + wave);
        std::vector<std::string> &set = outb[flip];
        for (auto &s : set) s.clear();
        parallel_ranges(w1 - w0, T, [&](int t, int64_t lo, int64_t hi) {
            std::string &ob = set[(size_t)t];
            std::string tmp;
            for (int64_t q = w0 + lo; q < w0 + hi; ++q) {
                const int64_t r = rows[q];
                const char *a = p->buf + p->row_begin[(size_t)r], *e = p->buf + p->row_end[(size_t)r];
                const bool plain = p->row_plain[(size_t)r] != 0;
                if (plain && all_cols) { ob.append(a, (size_t)(e - a)); ob.push_back('\n'); continue; }
                if (!p->id_nan[(size_t)r]) ob += p->ids[(size_t)r];
                const char *tab = (const char *)memchr(a, '\t', (size_t)(e - a));
                int64_t c = 0, want = 0;
                while (tab && want < ncols) {
                    a = tab + 1;
                    tab = (const char *)memchr(a, '\t', (size_t)(e - a));
                    if (c == cols[want]) {
                        const char *b = tab ? tab : e;
                        ob.push_back('\t');
                        if (plain) ob.append(a, (size_t)(b - a));
                        else {
                            bool nan;
                            Tok t = final_cell(p, a, (size_t)(b - a), tmp, &nan);
                            if (!nan) ob.append(t.s, t.n);
                        }
                        ++want;
                    }
                    ++c;
                }
                ob.push_back('\n');
            }
        });
        if (writer.joinable()) writer.join();
        writer = std::thread([&set, &wrc, fd, T]() {
            for (int t = 0; t < T; ++t)
                if (!set[(size_t)t].empty() && write_all(fd, set[(size_t)t].data(), set[(size_t)t].size())) { wrc.store(1); return; }
        });
    }
    if (writer.joinable()) writer.join();
    if (wrc.load()) rc = -1;
    if (close(fd) != 0) rc = -1;
    return rc ? fail(RTHIO_E_IO, "write to %s failed", path) : RTHIO_OK;
}

// ---------------------------------------------------------------------------------------------
// square distance matrix -> <out>_dist_hamming.tsv
// ---------------------------------------------------------------------------------------------
struct rthio_file {
    int fd = -1;
    int n_threads = 1;
    int64_t n = 0;
    int64_t pos = 0;                            // bytes written so far (file offset of the next row block)
    bool use_map = true;                        // row blocks go through a shared mapping (parallel), else write(2)
    std::vector<std::string> names;
    std::string path;
    std::vector<std::vector<char>> bufs;        // per-thread text buffers, kept across calls
    double t_format = 0, t_copy = 0;            // seconds per phase, printed at close when RTHIO_TIMING=1
};

static inline char *put_u32(char *o, uint32_t v)
{
    char tmp[10];
    int k = 0;
    do { tmp[k++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (k) *o++ = tmp[--k];
    return o;
}

// "\t<decimal>" of every uint16 as one 8-byte word (byte 7 = its length, 2..6): a cell is one table load,
// one unaligned 8-byte store and a pointer bump -- no division, no data-dependent branch.  The bytes
// after the text are overwritten by the next cell.  512 KB, of which distances touch the low end.
struct U16Text {
    uint64_t w[65536];
    U16Text()
    {
        for (uint32_t v = 0; v < 65536; ++v) {
            char b[8] = {'\t', 0, 0, 0, 0, 0, 0, 0};
            char *e = put_u32(b + 1, v);
            b[7] = (char)(e - b);
            memcpy(&w[v], b, 8);
        }
    }
};
static const U16Text &u16_text()
{
    static const U16Text t;
    return t;
}

extern "C" int rthio_square_open(const char *path, const char *corner, const char *blob, const int64_t *offs, int64_t n,
                                 int n_threads, rthio_file **out)
{
    if (!path || !corner || !blob || !offs || !out || n < 0) return fail(RTHIO_E_ARG, "rthio_square_open: bad argument");
    rthio_file *f = new rthio_file();
    f->n_threads = pick_threads(n_threads);
    f->n = n;
    f->path = path;
    f->names.resize((size_t)n);
    std::string head = corner;
    for (int64_t k = 0; k < n; ++k) {
        f->names[(size_t)k].assign(blob + offs[k], (size_t)(offs[k + 1] - offs[k]));
        head.push_back('\t');
        head += f->names[(size_t)k];
    }
    head.push_back('\n');
    f->fd = open(path, O_RDWR | O_CREAT | O_TRUNC, 0644);          // O_RDWR: the row blocks are written through MAP_SHARED
    if (f->fd < 0 || write_all(f->fd, head.data(), head.size())) {
        if (f->fd >= 0) close(f->fd);
        delete f;
        return fail(RTHIO_E_IO, "cannot write %s", path);
    }
    f->pos = (int64_t)head.size();
    f->bufs.resize((size_t)f->n_threads);
    if (const char *e = getenv("RTHIO_SQUARE_MMAP")) f->use_map = atoi(e) != 0;
    *out = f;
    return RTHIO_OK;
}

extern "C" int rthio_square_append_u16(rthio_file *f, int64_t row0, int64_t nrows, const uint16_t *block)
{
    if (!f || !block || row0 < 0 || nrows < 0 || row0 + nrows > f->n) return fail(RTHIO_E_ARG, "rthio_square_append_u16: bad argument");
    const int64_t n = f->n;
    const int T = f->n_threads;
    std::vector<size_t> used((size_t)T, 0);
    const auto c0 = std::chrono::steady_clock::now();
    const uint64_t *text = u16_text().w;
    // Shared mapping: the rows are formatted straight into the file's pages.  A first pass adds up the text length of
    // every row (one table load per cell), the prefix sum places the rows, and every thread then writes its rows where
    // they belong -- no staging buffer, no second copy of 50 GB at 100,000 samples.
    if (f->use_map && nrows > 0) {
        std::vector<size_t> off((size_t)nrows + 1, 0);
        parallel_ranges(nrows, T, [&](int, int64_t lo, int64_t hi) {
            for (int64_t r = lo; r < hi; ++r) {
                const uint16_t *src = block + (size_t)(r * n);
                uint64_t len = 0;
                for (int64_t j = 0; j < n; ++j) len += text[src[j]] >> 56;
                off[(size_t)r + 1] = (size_t)len + f->names[(size_t)(row0 + r)].size() + 1;
            }
        });
        for (int64_t r = 0; r < nrows; ++r) off[(size_t)r + 1] += off[(size_t)r];
        const size_t total = off[(size_t)nrows];
        const int64_t page = (int64_t)sysconf(_SC_PAGESIZE);
        const int64_t end = f->pos + (int64_t)total, map_off = f->pos / page * page;
        void *m = MAP_FAILED;
        if (ftruncate(f->fd, end) == 0)
            m = mmap(nullptr, (size_t)(end - map_off), PROT_READ | PROT_WRITE, MAP_SHARED, f->fd, map_off);
        if (m == MAP_FAILED) {
            f->use_map = false;                 // e.g. a file system without shared writable mappings
        } else {
            char *dst = (char *)m + (f->pos - map_off);
            parallel_ranges(nrows, T, [&](int, int64_t lo, int64_t hi) {
                for (int64_t r = lo; r < hi; ++r) {
                    char *o = dst + off[(size_t)r];
                    const std::string &nm = f->names[(size_t)(row0 + r)];
                    memcpy(o, nm.data(), nm.size());
                    o += nm.size();
                    const uint16_t *src = block + (size_t)(r * n);
                    const int64_t n_wide = n > 4 ? n - 4 : 0;
                    for (int64_t j = 0; j < n_wide; ++j) {
                        const uint64_t w = text[src[j]];
                        memcpy(o, &w, 8);       // up to 6 bytes past the text: they belong to later cells of THIS row
                        o += w >> 56;           // (4 more cells + newline = at least 9 bytes follow)
                    }
                    for (int64_t j = n_wide; j < n; ++j) {     // row tail: exact lengths, the next row may already be there
                        const uint64_t w = text[src[j]];
                        memcpy(o, &w, (size_t)(w >> 56));
                        o += w >> 56;
                    }
                    *o = '\n';
                }
            });
            if (munmap(m, (size_t)(end - map_off)) != 0) return fail(RTHIO_E_IO, "munmap of %s failed", f->path.c_str());
            f->pos = end;
            f->t_format += std::chrono::duration<double>(std::chrono::steady_clock::now() - c0).count();
            return RTHIO_OK;
        }
    }
    parallel_ranges(nrows, T, [&](int t, int64_t lo, int64_t hi) {
        size_t cap = 0;
        for (int64_t r = lo; r < hi; ++r) cap += f->names[(size_t)(row0 + r)].size() + (size_t)n * 6 + 2;
        cap += 8;                                             // the last cell's 8-byte store
        std::vector<char> &ob = f->bufs[(size_t)t];
        if (ob.size() < cap) ob.resize(cap);
        char *o = ob.data();
        for (int64_t r = lo; r < hi; ++r) {
            const std::string &nm = f->names[(size_t)(row0 + r)];
            memcpy(o, nm.data(), nm.size());
            o += nm.size();
            const uint16_t *src = block + (size_t)(r * n);
            for (int64_t j = 0; j < n; ++j) {
                const uint64_t w = text[src[j]];
                memcpy(o, &w, 8);
                o += w >> 56;
            }
            *o++ = '\n';
        }
        used[(size_t)t] = (size_t)(o - ob.data());
    });
    std::vector<size_t> start((size_t)T + 1, 0);
    for (int t = 0; t < T; ++t) start[(size_t)t + 1] = start[(size_t)t] + used[(size_t)t];
    const size_t total = start[(size_t)T];
    if (total == 0) return RTHIO_OK;
    const auto c1 = std::chrono::steady_clock::now();
    f->t_format += std::chrono::duration<double>(c1 - c0).count();
    struct CopyTimer {
        rthio_file *f; std::chrono::steady_clock::time_point t0;
        ~CopyTimer() { f->t_copy += std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); }
    } copy_timer{f, c1};
    // no shared mapping: write(2) from the staging buffers (one memcpy under the inode lock, ~2.5 GB/s)
    if (lseek(f->fd, (off_t)f->pos, SEEK_SET) < 0) return fail(RTHIO_E_IO, "seek in %s failed", f->path.c_str());
    for (int t = 0; t < T; ++t)
        if (used[(size_t)t] && write_all(f->fd, f->bufs[(size_t)t].data(), used[(size_t)t]))
            return fail(RTHIO_E_IO, "write to %s failed", f->path.c_str());
    f->pos += (int64_t)total;
    return RTHIO_OK;
}

extern "C" int rthio_square_close(rthio_file *f)
{
    if (!f) return RTHIO_OK;
    if (const char *e = getenv("RTHIO_TIMING"))
        if (atoi(e)) fprintf(stderr, "[rthio] %s: format %.2f s, copy into the file %.2f s (%s), %.2f GB\n", f->path.c_str(), f->t_format,
                             f->t_copy, f->use_map ? "shared mapping" : "write(2)", (double)f->pos / 1e9);
    int rc = close(f->fd);
    std::string path = f->path;
    delete f;
    return rc ? fail(RTHIO_E_IO, "close of %s failed", path.c_str()) : RTHIO_OK;
}

// ---------------------------------------------------------------------------------------------
// partitions.tsv + clusterComposition.tsv
// ---------------------------------------------------------------------------------------------
// Naming rule (scripts/partitioning_HC.py:651-670): walking the samples in matrix order, a
// cluster of size 1 becomes singleton_<running count>, any other cluster_<fcluster id>.
// code[t][i] > 0: cluster_<code>;  < 0: singleton_<-code>.
static inline char *put_name(char *o, int32_t code)
{
    if (code > 0) { memcpy(o, "cluster_", 8); return put_u32(o + 8, (uint32_t)code); }
    memcpy(o, "singleton_", 10);
    return put_u32(o + 10, (uint32_t)(-(int64_t)code));
}

extern "C" int rthio_write_partitions(const char *path_partitions, const char *path_composition, const char *names_blob,
                                      const int64_t *names_offs, int64_t n, const int32_t *labels, int64_t T,
                                      const char *col_blob, const int64_t *col_offs, int n_threads)
{
    if (!path_partitions || !path_composition || !names_blob || !names_offs || (!labels && T) || !col_blob || !col_offs || n < 1 || T < 0)
        return fail(RTHIO_E_ARG, "rthio_write_partitions: bad argument");
    const int NT = pick_threads(n_threads);
    std::vector<int32_t> code;
    try { code.resize((size_t)(T * n)); } catch (...) { return fail(RTHIO_E_NOMEM, "out of host memory for %lld x %lld names", (long long)T, (long long)n); }
    const int fdc = open(path_composition, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fdc < 0) return fail(RTHIO_E_IO, "cannot create %s", path_composition);
    static const char CHEAD[] = "#partition\tcluster\tcluster_length\tsamples\n";
    int rc = write_all(fdc, CHEAD, sizeof CHEAD - 1);
    std::atomic<int> bad(0);
    // thresholds in waves of NT: names + composition text per threshold in parallel, written in order
    for (int64_t t0 = 0; t0 < T && !rc; t0 += NT) {
        const int64_t t1 = std::min(T, t0 + NT);
        std::vector<std::string> comp((size_t)(t1 - t0));
        parallel_ranges(t1 - t0, NT, [&](int, int64_t lo, int64_t hi) {
            std::vector<int32_t> count, start, fill, first, order;
            for (int64_t t = t0 + lo; t < t0 + hi; ++t) {
                const int32_t *lab = labels + (size_t)(t * n);
                int32_t *cd = code.data() + (size_t)(t * n);
                int32_t k = 0;
                for (int64_t i = 0; i < n; ++i) { if (lab[i] < 1) bad.store(1); k = std::max(k, lab[i]); }
                if (bad.load()) return;
                count.assign((size_t)k + 1, 0);
                for (int64_t i = 0; i < n; ++i) count[(size_t)lab[i]]++;
                int32_t singles = 0;
                for (int64_t i = 0; i < n; ++i) cd[i] = count[(size_t)lab[i]] == 1 ? -(++singles) : lab[i];
                // members of every cluster in row order (counting sort), clusters by first appearance
                start.assign((size_t)k + 2, 0);
                for (int32_t c = 1; c <= k; ++c) start[(size_t)c + 1] = start[(size_t)c] + count[(size_t)c];
                fill.assign(start.begin(), start.end());
                order.resize((size_t)n);
                first.clear();
                for (int64_t i = 0; i < n; ++i) {
                    const int32_t c = lab[i];
                    if (fill[(size_t)c] == start[(size_t)c]) first.push_back(c);
                    order[(size_t)fill[(size_t)c]++] = (int32_t)i;
                }
                std::string &out = comp[(size_t)(t - t0)];
                const std::string part(col_blob + col_offs[t], (size_t)(col_offs[t + 1] - col_offs[t]));
                char num[32];
                for (int32_t c : first) {
                    const int32_t sz = count[(size_t)c];
                    out += part;
                    out.push_back('\t');
                    char *e = put_name(num, cd[order[(size_t)start[(size_t)c]]]);
                    out.append(num, (size_t)(e - num));
                    out.push_back('\t');
                    e = put_u32(num, (uint32_t)sz);
                    out.append(num, (size_t)(e - num));
                    out.push_back('\t');
                    for (int32_t q = 0; q < sz; ++q) {
                        const int32_t i = order[(size_t)(start[(size_t)c] + q)];
                        if (q) out.push_back(',');
                        out.append(names_blob + names_offs[i], (size_t)(names_offs[i + 1] - names_offs[i]));
                    }
                    out.push_back('\n');
                }
            }
        });
        if (bad.load()) { close(fdc); return fail(RTHIO_E_ARG, "cluster ids must be >= 1"); }
        for (auto &s : comp)
            if (!rc && !s.empty()) rc = write_all(fdc, s.data(), s.size());
    }
    if (close(fdc) != 0) rc = -1;
    if (rc) return fail(RTHIO_E_IO, "write to %s failed", path_composition);
    // partitions.tsv: one row per sample, one column per threshold
    const int fdp = open(path_partitions, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fdp < 0) return fail(RTHIO_E_IO, "cannot create %s", path_partitions);
    std::string head = "sequence";
    for (int64_t t = 0; t < T; ++t) { head.push_back('\t'); head.append(col_blob + col_offs[t], (size_t)(col_offs[t + 1] - col_offs[t])); }
    head.push_back('\n');
    rc = write_all(fdp, head.data(), head.size());
    const int64_t wave = std::max<int64_t>(1, (int64_t)NT * std::max<int64_t>(1, (int64_t)(1 << 22) / std::max<int64_t>(T * 16, 1)));
    std::vector<std::vector<char>> outb((size_t)NT);
    std::vector<size_t> used((size_t)NT);
    for (int64_t r0 = 0; r0 < n && !rc; r0 += wave) {
        const int64_t r1 = std::min(n, r0 + wave);
        std::fill(used.begin(), used.end(), 0);
        parallel_ranges(r1 - r0, NT, [&](int th, int64_t lo, int64_t hi) {
            size_t cap = 0;
            for (int64_t i = r0 + lo; i < r0 + hi; ++i) cap += (size_t)(names_offs[i + 1] - names_offs[i]) + (size_t)T * 22 + 2;
            std::vector<char> &ob = outb[(size_t)th];
            if (ob.size() < cap) ob.resize(cap);
            char *o = ob.data();
            for (int64_t i = r0 + lo; i < r0 + hi; ++i) {
                const size_t ln = (size_t)(names_offs[i + 1] - names_offs[i]);
                memcpy(o, names_blob + names_offs[i], ln);
                o += ln;
                for (int64_t t = 0; t < T; ++t) { *o++ = '\t'; o = put_name(o, code[(size_t)(t * n + i)]); }
                *o++ = '\n';
            }
            used[(size_t)th] = (size_t)(o - ob.data());
        });
        for (int th = 0; th < NT && !rc; ++th)
            if (used[(size_t)th]) rc = write_all(fdp, outb[(size_t)th].data(), used[(size_t)th]);
    }
    if (close(fdp) != 0) rc = -1;
    return rc ? fail(RTHIO_E_IO, "write to %s failed", path_partitions) : RTHIO_OK;
}

// ---------------------------------------------------------------------------------------------
// synthetic-input helper: an int32 allele matrix as a profile TSV (bench / test inputs)
// ---------------------------------------------------------------------------------------------
extern "C" int rthio_write_int_matrix(const char *path, const char *corner, const char *row_prefix, const char *col_prefix,
                                      const int32_t *codes, int64_t n, int64_t L, int n_threads)
{
    if (!path || !corner || !row_prefix || !col_prefix || !codes || n < 0 || L < 0) return fail(RTHIO_E_ARG, "rthio_write_int_matrix: bad argument");
    const int T = pick_threads(n_threads);
    const int fd = open(path, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return fail(RTHIO_E_IO, "cannot create %s", path);
    std::string head = corner;
    char num[32];
    for (int64_t c = 0; c < L; ++c) { snprintf(num, sizeof num, "\t%s%05lld", col_prefix, (long long)(c + 1)); head += num; }
    head.push_back('\n');
    int rc = write_all(fd, head.data(), head.size());
    const int64_t wave = (int64_t)T * 256;
    std::vector<std::vector<char>> ob((size_t)T);
    std::vector<size_t> used((size_t)T);
    const size_t plen = strlen(row_prefix);
    for (int64_t r0 = 0; r0 < n && !rc; r0 += wave) {
        const int64_t r1 = std::min(n, r0 + wave);
        std::fill(used.begin(), used.end(), 0);
        parallel_ranges(r1 - r0, T, [&](int t, int64_t lo, int64_t hi) {
            const size_t cap = (size_t)(hi - lo) * (plen + 16 + (size_t)L * 12);
            if (ob[(size_t)t].size() < cap) ob[(size_t)t].resize(cap);
            char *o = ob[(size_t)t].data();
            for (int64_t r = r0 + lo; r < r0 + hi; ++r) {
                memcpy(o, row_prefix, plen); o += plen;
                o += snprintf(o, 16, "%07lld", (long long)(r + 1));
                const int32_t *src = codes + (size_t)(r * L);
                for (int64_t c = 0; c < L; ++c) {
                    *o++ = '\t';
                    const int32_t v = src[c];
                    if (v < 0) { *o++ = '-'; o = put_u32(o, (uint32_t)(-(int64_t)v)); }
                    else o = put_u32(o, (uint32_t)v);
                }
                *o++ = '\n';
            }
            used[(size_t)t] = (size_t)(o - ob[(size_t)t].data());
        });
        for (int t = 0; t < T && !rc; ++t)
            if (used[(size_t)t]) rc = write_all(fd, ob[(size_t)t].data(), used[(size_t)t]);
    }
    if (close(fd) != 0) rc = -1;
    return rc ? fail(RTHIO_E_IO, "write to %s failed", path) : RTHIO_OK;
}
