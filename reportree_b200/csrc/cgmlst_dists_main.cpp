// cgmlst_dists_main.cpp -- `cgmlst-dists`-compatible executable backed by libreportree_hc.so.
//
// Seam B2 of SURVEY.md section 8(b): the reference shells out to the external program
//     cgmlst-dists <in.tsv> > <out.tsv>
// once per 2,000-column chunk (scripts/partitioning_HC.py:502, scripts/partitioning_grapetree.py:466)
// and reads the full square TSV back (:508-510).  Putting this binary first on PATH gives the
// UNMODIFIED reference scripts (HC and GrapeTree paths) the GPU distance kernel.
//
// Contract (scripts/partitioning_HC.py:293-298, :496-499): first row = header, first column = sample
// id; a cell that is a clean non-negative decimal integer is that allele, anything else is missing
// (0); d(i,j) = #{l : a_il != a_jl, a_il != 0, a_jl != 0}; output = corner cell + ids, then one row
// per sample (id + n integers), tab-separated.  No CPU fallback: without a CUDA device it fails.
//
// Environment: REPORTREE_HC_DEVICE=<int> selects the GPU (default 0).
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <algorithm>
#include <string>
#include <thread>
#include <vector>

#include "../../include/rthc.h"

static inline int32_t parse_cell(const char *s, size_t n)
{
    if (n == 0 || n > 9) return 0;
    int32_t v = 0;
    for (size_t k = 0; k < n; ++k) {
        const unsigned d = (unsigned char)s[k] - '0';
        if (d > 9) return 0;
        v = v * 10 + (int32_t)d;
    }
    return v;
}

static inline char *put_u32(char *o, uint32_t v)
{
    char tmp[10];
    int k = 0;
    do { tmp[k++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (k) *o++ = tmp[--k];
    return o;
}

int main(int argc, char **argv)
{
    if (argc < 2 || !strcmp(argv[1], "-h") || !strcmp(argv[1], "--help")) {
        fprintf(stderr, "usage: cgmlst-dists in.tsv > out.tsv   (B200 backend, libreportree_hc.so)\n");
        return argc < 2 ? 2 : 0;
    }
    FILE *f = fopen(argv[1], "rb");
    if (!f) { fprintf(stderr, "ERROR: cannot read %s\n", argv[1]); return 1; }
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    std::vector<char> buf((size_t)sz + 1);
    if (sz > 0 && fread(buf.data(), 1, (size_t)sz, f) != (size_t)sz) { fprintf(stderr, "ERROR: short read on %s\n", argv[1]); return 1; }
    fclose(f);
    size_t len = (size_t)sz;
    if (len == 0 || buf[len - 1] != '\n') buf[len++] = '\n';
    // non-blank lines
    std::vector<size_t> lb, le;
    for (size_t pos = 0; pos < len;) {
        const char *nl = (const char *)memchr(buf.data() + pos, '\n', len - pos);
        size_t end = (size_t)(nl - buf.data());
        size_t e = end;
        if (e > pos && buf[e - 1] == '\r') --e;
        if (e > pos) { lb.push_back(pos); le.push_back(e); }
        pos = end + 1;
    }
    if (lb.size() < 2) { fprintf(stderr, "ERROR: empty matrix\n"); return 1; }
    int64_t L = 0;
    for (size_t k = lb[0]; k < le[0]; ++k) L += buf[k] == '\t';
    const int64_t n = (int64_t)lb.size() - 1;
    if (L < 1) { fprintf(stderr, "ERROR: no locus columns\n"); return 1; }
    std::vector<int32_t> codes((size_t)(n * L), 0);
    std::vector<std::string> ids((size_t)n);
    unsigned hc = std::thread::hardware_concurrency();
    const int T = (int)std::max<int64_t>(1, std::min<int64_t>(hc ? hc : 1, n));
    {
        std::vector<std::thread> th;
        for (int t = 0; t < T; ++t)
            th.emplace_back([&, t]() {
                for (int64_t r = n * t / T; r < n * (t + 1) / T; ++r) {
                    const char *a = buf.data() + lb[(size_t)r + 1], *e = buf.data() + le[(size_t)r + 1];
                    const char *tab = (const char *)memchr(a, '\t', (size_t)(e - a));
                    ids[(size_t)r].assign(a, (size_t)((tab ? tab : e) - a));
                    int64_t c = 0;
                    while (tab && c < L) {                 // short rows: remaining cells stay missing
                        a = tab + 1;
                        tab = (const char *)memchr(a, '\t', (size_t)(e - a));
                        codes[(size_t)(r * L + c++)] = parse_cell(a, (size_t)((tab ? tab : e) - a));
                    }
                }
            });
        for (auto &x : th) x.join();
    }
    rthc_ctx *ctx = nullptr;
    const char *dev = getenv("REPORTREE_HC_DEVICE");
    if (rthc_create(&ctx, dev ? atoi(dev) : 0) != 0) { fprintf(stderr, "ERROR: %s\n", rthc_last_error(nullptr)); return 1; }
    std::string out = "cgmlst-dists";
    for (int64_t i = 0; i < n; ++i) { out.push_back('\t'); out += ids[(size_t)i]; }
    out.push_back('\n');
    if (n == 1) {                                           // a single sample: 1 x 1 matrix, nothing to compute
        out += ids[0] + "\t0\n";
        fwrite(out.data(), 1, out.size(), stdout);
        rthc_destroy(ctx);
        return 0;
    }
    if (rthc_set_profiles_i32(ctx, codes.data(), n, L) != 0 || rthc_distance(ctx) != 0) {
        fprintf(stderr, "ERROR: %s\n", rthc_last_error(ctx));
        rthc_destroy(ctx);
        return 1;
    }
    fwrite(out.data(), 1, out.size(), stdout);
    const int64_t block = std::max<int64_t>(1, std::min<int64_t>(n, (int64_t)(1 << 26) / n));
    std::vector<uint16_t> rows((size_t)(block * n));
    std::vector<std::vector<char>> ob((size_t)T);
    std::vector<size_t> used((size_t)T);
    for (int64_t r0 = 0; r0 < n; r0 += block) {
        const int64_t nr = std::min(block, n - r0);
        if (rthc_copy_rows_u16(ctx, r0, nr, rows.data()) != 0) { fprintf(stderr, "ERROR: %s\n", rthc_last_error(ctx)); return 1; }
        std::vector<std::thread> th;
        for (int t = 0; t < T; ++t)
            th.emplace_back([&, t]() {
                const int64_t lo = nr * t / T, hi = nr * (t + 1) / T;
                size_t cap = 0;
                for (int64_t r = lo; r < hi; ++r) cap += ids[(size_t)(r0 + r)].size() + (size_t)n * 6 + 2;
                if (ob[(size_t)t].size() < cap) ob[(size_t)t].resize(cap);
                char *o = ob[(size_t)t].data();
                for (int64_t r = lo; r < hi; ++r) {
                    const std::string &nm = ids[(size_t)(r0 + r)];
                    memcpy(o, nm.data(), nm.size());
                    o += nm.size();
                    const uint16_t *src = rows.data() + (size_t)(r * n);
                    for (int64_t j = 0; j < n; ++j) { *o++ = '\t'; o = put_u32(o, src[j]); }
                    *o++ = '\n';
                }
                used[(size_t)t] = (size_t)(o - ob[(size_t)t].data());
            });
        for (auto &x : th) x.join();
        for (int t = 0; t < T; ++t)
            if (used[(size_t)t] && fwrite(ob[(size_t)t].data(), 1, used[(size_t)t], stdout) != used[(size_t)t]) {
                fprintf(stderr, "ERROR: write failed\n");
                return 1;
            }
    }
    rthc_destroy(ctx);
    return fflush(stdout) == 0 ? 0 : 1;
}
