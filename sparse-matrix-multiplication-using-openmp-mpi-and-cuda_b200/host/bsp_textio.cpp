// bsp_textio.cpp -- the reference's text format, either side of the hot path.
//
// Reader: "rows cols", "blocks", then per block "brow bcol" and k*k whitespace-separated
// uint64 values (cu:345-383).  The reference parses with ifstream >> (one task per file on
// 16 OpenMP threads, cu:334-339); here a file is read whole and parsed with a hand-rolled
// decimal scanner, split across host threads at token boundaries (two passes: count, parse).
// Writer: cu:595-608 -- "R C", block count, then per tile (in (brow,bcol) order) "i j" and k
// lines of k values separated by single spaces, no trailing space, '\n' line ends.
#include <charconv>
#include <cstdio>
#include <cstdlib>
#include <string>
#include <thread>
#include <vector>

#include "../csrc/bsp_internal.h"

namespace {

inline const char *skip_ws(const char *p, const char *e)
{
    while (p < e && (unsigned char)*p <= ' ') p++;
    return p;
}

inline const char *parse_u64(const char *p, const char *e, uint64_t *v, bool *ok)
{
    p = skip_ws(p, e);
    uint64_t x = 0;
    const char *s = p;
    while (p < e && (unsigned)(*p - '0') <= 9u) { x = x * 10u + (uint64_t)(*p - '0'); p++; }
    if (p == s) *ok = false;
    *v = x;
    return p;
}

inline const char *parse_i32(const char *p, const char *e, int32_t *v, bool *ok)
{
    p = skip_ws(p, e);
    bool neg = false;
    if (p < e && (*p == '-' || *p == '+')) { neg = (*p == '-'); p++; }
    uint64_t x;
    p = parse_u64(p, e, &x, ok);
    *v = neg ? -(int32_t)x : (int32_t)x;
    return p;
}

bool read_file(const char *path, std::string *out)
{
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    out->resize((size_t)sz);
    bool ok = sz == 0 || fread(&(*out)[0], 1, (size_t)sz, f) == (size_t)sz;
    fclose(f);
    return ok;
}

} // namespace

extern "C" int bsp_mat_read_text(bsp_ctx *ctx, const char *path, int32_t k, bsp_mat **out)
{
    if (!ctx || !out || !path) return BSP_ERR_ARG;
    *out = nullptr;
    if (k < 1) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_read_text: k=%d", k);
    std::string buf;
    if (!read_file(path, &buf)) BSP_FAIL(ctx, BSP_ERR_IO, "cannot open %s", path);
    const char *p = buf.data(), *e = p + buf.size();
    bool ok = true;
    int32_t rows = 0, cols = 0, blocks = 0;
    p = parse_i32(p, e, &rows, &ok);
    p = parse_i32(p, e, &cols, &ok);
    p = parse_i32(p, e, &blocks, &ok);
    if (!ok || blocks < 0) BSP_FAIL(ctx, BSP_ERR_IO, "%s: malformed header", path);
    const size_t kk = (size_t)k * k;
    std::vector<int32_t> br((size_t)blocks), bc((size_t)blocks);
    std::vector<uint64_t> v((size_t)blocks * kk);
    // The body is blocks * (2 + k*k) whitespace-separated tokens.  Parse it on several threads:
    // cut the text at whitespace, count the tokens of every piece, prefix-sum the counts so each
    // piece knows the global number of its first token, then parse the pieces independently
    // (token g belongs to block g / (2+k*k), field g % (2+k*k)).
    const size_t per_block = 2 + kk;
    const size_t want_tokens = (size_t)blocks * per_block;
    unsigned nt = std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > 32) nt = 32;
    const size_t body = (size_t)(e - p);
    if (body < (1u << 20)) nt = 1;
    std::vector<const char *> cut(nt + 1);
    cut[0] = p; cut[nt] = e;
    for (unsigned i = 1; i < nt; i++) {
        const char *c = p + body * i / nt;
        while (c < e && (unsigned char)*c > ' ') c++; // move to the end of the token we landed in
        cut[i] = c;
    }
    std::vector<size_t> ntok(nt + 1, 0);
    auto count = [&](unsigned i) {
        size_t n = 0;
        const char *c = cut[i], *ce = cut[i + 1];
        while (c < ce) {
            while (c < ce && (unsigned char)*c <= ' ') c++;
            if (c < ce) { n++; while (c < ce && (unsigned char)*c > ' ') c++; }
        }
        ntok[i + 1] = n;
    };
    auto run_all = [&](auto fn) {
        if (nt == 1) { fn(0u); return; }
        std::vector<std::thread> th;
        for (unsigned i = 0; i < nt; i++) th.emplace_back(fn, i);
        for (auto &t : th) t.join();
    };
    run_all(count);
    for (unsigned i = 0; i < nt; i++) ntok[i + 1] += ntok[i];
    if (ntok[nt] < want_tokens) BSP_FAIL(ctx, BSP_ERR_IO, "%s: truncated block list (%zu of %zu values)", path, ntok[nt], want_tokens);
    std::vector<char> bad(nt, 0);
    auto parse = [&](unsigned i) {
        const char *c = cut[i], *ce = cut[i + 1];
        size_t g = ntok[i];
        bool okk = true;
        while (g < want_tokens) {
            c = skip_ws(c, ce);
            if (c >= ce) break;
            const size_t b = g / per_block, f = g - b * per_block;
            if (f < 2) {
                int32_t x;
                c = parse_i32(c, ce, &x, &okk);
                (f == 0 ? br : bc)[b] = x;
            } else {
                c = parse_u64(c, ce, &v[b * kk + (f - 2)], &okk);
            }
            if (!okk) { bad[i] = 1; return; }
            g++;
        }
    };
    run_all(parse);
    for (unsigned i = 0; i < nt; i++) ok = ok && !bad[i];
    if (!ok) BSP_FAIL(ctx, BSP_ERR_IO, "%s: truncated or malformed block list", path);
    return bsp_mat_from_coo(ctx, rows, cols, k, blocks, br.data(), bc.data(), v.data(), out);
}

extern "C" int bsp_mat_write_text(bsp_ctx *ctx, const bsp_mat *m, const char *path)
{
    if (!ctx || !m || !path) return BSP_ERR_ARG;
    if (m->esize != 8) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_write_text: the reference's text format holds uint64 entries only");
    int32_t rows, cols, k;
    int64_t nnzb;
    bsp_mat_info(m, &rows, &cols, &k, &nnzb);
    const size_t kk = (size_t)k * k;
    std::vector<int32_t> br((size_t)nnzb), bc((size_t)nnzb);
    std::vector<uint64_t> v((size_t)nnzb * kk);
    BSP_TRY(bsp_mat_to_coo(ctx, m, br.data(), bc.data(), v.data()));

    // format in parallel chunks of tiles, then write the chunks in order
    unsigned nt = std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > 32) nt = 32;
    if ((int64_t)nt > nnzb / 1024 + 1) nt = (unsigned)(nnzb / 1024 + 1);
    std::vector<std::string> parts(nt);
    auto work = [&](unsigned ti) {
        const int64_t lo = nnzb * ti / nt, hi = nnzb * (ti + 1) / nt;
        std::string &s = parts[ti];
        s.reserve((size_t)(hi - lo) * (24 + kk * 21));
        char tmp[24];
        for (int64_t t = lo; t < hi; t++) {
            auto r1 = std::to_chars(tmp, tmp + 24, br[t]);
            s.append(tmp, r1.ptr); s.push_back(' ');
            r1 = std::to_chars(tmp, tmp + 24, bc[t]);
            s.append(tmp, r1.ptr); s.push_back('\n');
            const uint64_t *tv = v.data() + (size_t)t * kk;
            for (int r = 0; r < k; r++)
                for (int c = 0; c < k; c++) {
                    auto r2 = std::to_chars(tmp, tmp + 24, (unsigned long long)tv[r * k + c]);
                    s.append(tmp, r2.ptr);
                    s.push_back(c + 1 < k ? ' ' : '\n');
                }
        }
    };
    if (nt == 1) work(0);
    else {
        std::vector<std::thread> th;
        for (unsigned i = 0; i < nt; i++) th.emplace_back(work, i);
        for (auto &t : th) t.join();
    }
    FILE *f = fopen(path, "wb");
    if (!f) BSP_FAIL(ctx, BSP_ERR_IO, "cannot create %s", path);
    fprintf(f, "%d %d\n%lld\n", rows, cols, (long long)nnzb);
    bool ok = true;
    for (auto &s : parts) ok = ok && (s.empty() || fwrite(s.data(), 1, s.size(), f) == s.size());
    ok = (fclose(f) == 0) && ok;
    if (!ok) BSP_FAIL(ctx, BSP_ERR_IO, "short write to %s", path);
    return BSP_OK;
}
