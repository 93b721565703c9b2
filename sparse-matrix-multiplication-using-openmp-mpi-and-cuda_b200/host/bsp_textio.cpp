// bsp_textio.cpp -- the reference's text format, either side of the hot path.
//
// Reader: "rows cols", "blocks", then per block "brow bcol" and k*k whitespace-separated
// uint64 values (cu:345-383).  The reference parses with ifstream >> (one task per file on
// 16 OpenMP threads, cu:334-339); here a file is read whole and parsed with a hand-rolled
// decimal scanner, split across host threads at token boundaries (two passes: count, parse).
// Writer: cu:595-608 -- "R C", block count, then per tile (in (brow,bcol) order) "i j" and k
// lines of k values separated by single spaces, no trailing space, '\n' line ends.
#include <charconv>
#include <condition_variable>
#include <cstdio>
#include <cstdlib>
#include <deque>
#include <mutex>
#include <new>
#include <stdexcept>
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
    while (p < e && (unsigned)(*p - '0') <= 9u) {
        const uint64_t d = (uint64_t)(*p - '0');
        // beyond 2^64-1 the reference's ifstream >> sets failbit (cu:375-376): refuse instead of wrapping
        if (x > (UINT64_MAX - d) / 10u) *ok = false;
        x = x * 10u + d;
        p++;
    }
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
    if (x > (neg ? 2147483648ull : 2147483647ull)) { *ok = false; x = 0; } // does not fit an int
    *v = neg ? (int32_t)(0 - (int64_t)x) : (int32_t)x;
    return p;
}

bool read_file(const char *path, std::string *out)
{
    FILE *f = fopen(path, "rb");
    if (!f) return false;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    bool ok = sz >= 0;
    try { out->resize(ok ? (size_t)sz : 0); } catch (const std::exception &) { ok = false; sz = 0; }
    ok = ok && (sz == 0 || fread(&(*out)[0], 1, (size_t)sz, f) == (size_t)sz);
    fclose(f);
    return ok;
}

// Parses the body of a matrix file: blocks * (2 + k*k) whitespace-separated tokens starting at p.
// Runs on nt threads: cut the text at whitespace, count the tokens of every piece, prefix-sum the
// counts so each piece knows the global number of its first token, then parse the pieces
// independently (token g belongs to block g / (2+k*k), field g % (2+k*k)).
// Returns an empty string on success, else the reason.
std::string parse_body(const char *p, const char *e, int64_t blocks, int32_t k, unsigned nt, int32_t *br, int32_t *bc,
                       uint64_t *v)
{
    const size_t kk = (size_t)k * k;
    const size_t per_block = 2 + kk;
    const size_t want_tokens = (size_t)blocks * per_block;
    const size_t body = (size_t)(e - p);
    if (nt < 1 || body < (1u << 20)) nt = 1;
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
    if (nt > 1) { // a single piece starts at token 0 and needs no count
        run_all(count);
        for (unsigned i = 0; i < nt; i++) ntok[i + 1] += ntok[i];
        if (ntok[nt] < want_tokens)
            return "truncated block list (" + std::to_string(ntok[nt]) + " of " + std::to_string(want_tokens) + " values)";
    }
    std::vector<char> bad(nt, 0);
    std::vector<size_t> reached(nt, 0);
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
        reached[i] = g;
    };
    run_all(parse);
    for (unsigned i = 0; i < nt; i++)
        if (bad[i]) return "truncated or malformed block list";
    if (reached[nt - 1] < want_tokens)
        return "truncated block list (" + std::to_string(reached[nt - 1]) + " of " + std::to_string(want_tokens) + " values)";
    return std::string();
}

// "rows cols" and "blocks" (cu:345-360); returns the first byte of the body or nullptr
const char *parse_header(const char *p, const char *e, int32_t *rows, int32_t *cols, int32_t *blocks)
{
    bool ok = true;
    p = parse_i32(p, e, rows, &ok);
    p = parse_i32(p, e, cols, &ok);
    p = parse_i32(p, e, blocks, &ok);
    return (ok && *blocks >= 0) ? p : nullptr;
}

// A body of `blocks` tiles has blocks*(2+k*k) tokens of at least one digit and one separator each:
// a header that promises more than the file can hold is refused before anything is sized from it.
bool body_can_hold(const char *p, const char *e, int64_t blocks, int32_t k)
{
    const uint64_t tokens = (uint64_t)blocks * (2u + (uint64_t)k * (uint64_t)k);
    return tokens <= ((uint64_t)(e - p) + 1u) / 2u;
}

unsigned host_threads()
{
    unsigned nt = std::thread::hardware_concurrency();
    if (nt < 1) nt = 1;
    if (nt > 32) nt = 32;
    return nt;
}

} // namespace

extern "C" int bsp_mat_read_text(bsp_ctx *ctx, const char *path, int32_t k, bsp_mat **out)
{
    if (!ctx || !out || !path) return BSP_ERR_ARG;
    *out = nullptr;
    if (k < 1) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_read_text: k=%d", k);
    std::string buf;
    if (!read_file(path, &buf)) BSP_FAIL(ctx, BSP_ERR_IO, "cannot open %s", path);
    const char *e = buf.data() + buf.size();
    int32_t rows = 0, cols = 0, blocks = 0;
    const char *p = parse_header(buf.data(), e, &rows, &cols, &blocks);
    if (!p) BSP_FAIL(ctx, BSP_ERR_IO, "%s: malformed header", path);
    if (!body_can_hold(p, e, blocks, k))
        BSP_FAIL(ctx, BSP_ERR_IO, "%s: header announces %d tiles, the file is too short for them", path, blocks);
    const size_t kk = (size_t)k * k;
    try { // nothing may leave through the C boundary
        std::vector<int32_t> br((size_t)blocks), bc((size_t)blocks);
        std::vector<uint64_t> v((size_t)blocks * kk);
        const std::string why = parse_body(p, e, blocks, k, host_threads(), br.data(), bc.data(), v.data());
        if (!why.empty()) BSP_FAIL(ctx, BSP_ERR_IO, "%s: %s", path, why.c_str());
        return bsp_mat_from_coo(ctx, rows, cols, k, blocks, br.data(), bc.data(), v.data(), out);
    } catch (const std::bad_alloc &) {
        BSP_FAIL(ctx, BSP_ERR_NOMEM, "%s: host allocation failed", path);
    } catch (const std::exception &ex) {
        BSP_FAIL(ctx, BSP_ERR_IO, "%s: %s", path, ex.what());
    }
}

// ---- several files at once: parsing overlapped with upload and ingestion ----
// The reference loads its files with one OpenMP task per file into pageable std::map nodes
// (cu:334-339) and uploads nothing until the first product; its report describes pinned buffers
// and asynchronous copies that the code does not have.  Here worker
// threads parse files straight into pinned staging slots while the calling thread uploads the
// finished ones (cudaMemcpyAsync from pinned memory) and builds their CSR-of-blocks on the
// device, so the GPU-side work of file i hides behind the parsing of the files after it.
namespace {

struct stage_slot {
    char *buf = nullptr;
    size_t cap = 0;
    int32_t rows = 0, cols = 0, max_row = -1, max_col = -1;
    int64_t blocks = 0;
    int file = -1;
};

struct ingest_state {
    std::mutex mu;
    std::condition_variable cv;
    std::vector<stage_slot> slots;
    std::deque<int> free_slots, ready_slots;
    int next_file = 0, n_files = 0, workers_alive = 0;
    bool abort = false;
    std::string error;
};

} // namespace

extern "C" int bsp_mat_read_text_many(bsp_ctx *ctx, const char *const *paths, int32_t n, int32_t k, bsp_mat **out)
{
    if (!ctx || !paths || !out || n < 0) return BSP_ERR_ARG;
    for (int i = 0; i < n; i++) out[i] = nullptr;
    if (k < 1) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_read_text_many: k=%d", k);
    if (n == 0) return BSP_OK;
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    const unsigned hw = host_threads();
    const unsigned n_workers = (unsigned)n < hw ? (unsigned)n : hw;
    const unsigned per_file = hw / n_workers > 0 ? hw / n_workers : 1; // few files: threads left over split each file
    const size_t kk = (size_t)k * k;
    ingest_state st;
    st.n_files = n;
    st.slots.resize(n_workers + 2);
    for (int i = 0; i < (int)st.slots.size(); i++) st.free_slots.push_back(i);
    st.workers_alive = (int)n_workers;
    const int device = ctx->device;

    auto worker = [&]() {
        cudaSetDevice(device);
        for (;;) {
            int file, sl;
            {
                std::unique_lock<std::mutex> lk(st.mu);
                if (st.abort || st.next_file >= st.n_files) break;
                file = st.next_file++;
                st.cv.wait(lk, [&] { return st.abort || !st.free_slots.empty(); });
                if (st.abort) break;
                sl = st.free_slots.front();
                st.free_slots.pop_front();
            }
            stage_slot &S = st.slots[sl];
            std::string why, buf;
            int32_t blocks32 = 0;
            const char *p = nullptr;
            if (!read_file(paths[file], &buf)) why = "cannot open";
            else if (!(p = parse_header(buf.data(), buf.data() + buf.size(), &S.rows, &S.cols, &blocks32))) why = "malformed header";
            else if (!body_can_hold(p, buf.data() + buf.size(), blocks32, k)) why = "header announces more tiles than the file can hold";
            if (why.empty()) {
                S.blocks = blocks32;
                const size_t need = (size_t)S.blocks * (kk * 8 + 8) + 64;
                if (need > S.cap) { // values first (8-byte aligned), then the two coordinate arrays
                    if (S.buf) cudaFreeHost(S.buf);
                    S.buf = nullptr; S.cap = 0;
                    const size_t want = need + need / 4;
                    if (cudaMallocHost((void **)&S.buf, want) != cudaSuccess) { cudaGetLastError(); why = "pinned staging buffer allocation failed"; }
                    else S.cap = want;
                }
            }
            if (why.empty()) {
                uint64_t *v = (uint64_t *)S.buf;
                int32_t *br = (int32_t *)(S.buf + (size_t)S.blocks * kk * 8), *bc = br + S.blocks;
                why = parse_body(p, buf.data() + buf.size(), S.blocks, k, per_file, br, bc, v);
                S.max_row = S.max_col = -1;
                for (int64_t i = 0; why.empty() && i < S.blocks; i++) {
                    if (br[i] < 0 || bc[i] < 0) why = "negative block coordinate at entry " + std::to_string(i);
                    if (br[i] > S.max_row) S.max_row = br[i];
                    if (bc[i] > S.max_col) S.max_col = bc[i];
                }
            }
            std::lock_guard<std::mutex> lk(st.mu);
            if (!why.empty()) {
                if (st.error.empty()) st.error = std::string(paths[file]) + ": " + why;
                st.abort = true;
                st.free_slots.push_back(sl);
                st.cv.notify_all();
                break;
            }
            S.file = file;
            st.ready_slots.push_back(sl);
            st.cv.notify_all();
        }
        std::lock_guard<std::mutex> lk(st.mu);
        st.workers_alive--;
        st.cv.notify_all();
    };
    std::vector<std::thread> th;
    for (unsigned i = 0; i < n_workers; i++) th.emplace_back(worker);

    int rc = BSP_OK, done = 0;
    while (done < n) {
        int sl = -1;
        {
            std::unique_lock<std::mutex> lk(st.mu);
            st.cv.wait(lk, [&] { return !st.ready_slots.empty() || st.abort || st.workers_alive == 0; });
            if (!st.ready_slots.empty()) { sl = st.ready_slots.front(); st.ready_slots.pop_front(); }
            else break; // aborted, or the workers are gone without delivering everything
        }
        stage_slot &S = st.slots[sl];
        // upload from the pinned slot and build the matrix; bsp_ingest_coo waits for the device
        // once (tile count after duplicate removal), after which the slot can be parsed into again
        bsp_scratch sc(ctx);
        int32_t *d_r = nullptr, *d_c = nullptr;
        char *d_v = nullptr;
        rc = sc.alloc(&d_r, (size_t)S.blocks);
        if (rc == BSP_OK) rc = sc.alloc(&d_c, (size_t)S.blocks);
        if (rc == BSP_OK) rc = sc.alloc(&d_v, (size_t)S.blocks * kk * 8);
        if (rc == BSP_OK && S.blocks > 0) {
            const char *coords = S.buf + (size_t)S.blocks * kk * 8;
            cudaError_t e1 = cudaMemcpyAsync(d_v, S.buf, (size_t)S.blocks * kk * 8, cudaMemcpyHostToDevice, ctx->stream);
            cudaError_t e2 = cudaMemcpyAsync(d_r, coords, (size_t)S.blocks * 4, cudaMemcpyHostToDevice, ctx->stream);
            cudaError_t e3 = cudaMemcpyAsync(d_c, coords + (size_t)S.blocks * 4, (size_t)S.blocks * 4, cudaMemcpyHostToDevice, ctx->stream);
            if (e1 != cudaSuccess || e2 != cudaSuccess || e3 != cudaSuccess) {
                snprintf(ctx->err, sizeof(ctx->err), "bsp_mat_read_text_many: upload of %s failed", paths[S.file]);
                rc = BSP_ERR_CUDA;
            }
        }
        if (rc == BSP_OK)
            rc = bsp_ingest_coo(ctx, S.rows, S.cols, k, 8, S.blocks, d_r, d_c, d_v, S.max_row, S.max_col, &out[S.file]);
        if (rc == BSP_OK) rc = cudaStreamSynchronize(ctx->stream) == cudaSuccess ? BSP_OK : BSP_ERR_CUDA;
        std::lock_guard<std::mutex> lk(st.mu);
        st.free_slots.push_back(sl);
        if (rc != BSP_OK) st.abort = true;
        st.cv.notify_all();
        if (rc != BSP_OK) break;
        done++;
    }
    {
        std::lock_guard<std::mutex> lk(st.mu);
        if (done < n) st.abort = true;
        st.cv.notify_all();
    }
    for (auto &t : th) t.join();
    for (auto &S : st.slots)
        if (S.buf) cudaFreeHost(S.buf);
    if (rc == BSP_OK && done < n) {
        snprintf(ctx->err, sizeof(ctx->err), "%s", st.error.empty() ? "bsp_mat_read_text_many: reader threads stopped early" : st.error.c_str());
        rc = BSP_ERR_IO;
    }
    if (rc != BSP_OK)
        for (int i = 0; i < n; i++) { bsp_mat_free(ctx, out[i]); out[i] = nullptr; }
    return rc;
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
