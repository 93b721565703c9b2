// bsp_capi.cu -- the extern "C" boundary declared in include/bspgemm.h.
#include <algorithm>
#include <chrono>
#include <cstdlib>
#include <new>
#include <vector>

#include "bsp_internal.h"

namespace {

bool valid_mat(const bsp_mat *m) { return m && m->rowptr; }

int new_empty_like(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t nbr, int32_t nbc, bsp_mat **out,
                   int32_t esize = 8)
{
    bsp_mat *m = new (std::nothrow) bsp_mat();
    if (!m) BSP_FAIL(ctx, BSP_ERR_NOMEM, "host allocation failed");
    m->rows = rows; m->cols = cols; m->k = k; m->nbr = nbr; m->nbc = nbc; m->nnzb = 0; m->device = ctx->device;
    m->esize = esize;
    m->rowptr = nullptr; m->colidx = nullptr; m->vals = nullptr;
    *out = m;
    return BSP_OK;
}

void accumulate(bsp_stats *dst, const bsp_stats &s)
{
    dst->nnzb_a += s.nnzb_a; dst->nnzb_b += s.nnzb_b; dst->nnzb_c += s.nnzb_c; dst->nprod += s.nprod;
    dst->ms_symbolic += s.ms_symbolic; dst->ms_numeric += s.ms_numeric; dst->ms_total += s.ms_total;
    dst->launches += s.launches; dst->exact_refolds += s.exact_refolds;
    dst->numeric_path = s.numeric_path;
}

int spgemm_impl(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t row_lo, int32_t row_hi, bsp_mat **C,
                bsp_stats *stats)
{
    if (!ctx || !C) return BSP_ERR_ARG;
    *C = nullptr;
    if (!valid_mat(A) || !valid_mat(B)) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_spgemm: null matrix");
    if (A->k != B->k) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_spgemm: tile edge differs (%d vs %d)", A->k, B->k);
    if (A->esize != B->esize) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_spgemm: entry types differ (uint64 vs float32)");
    if (A->device != ctx->device || B->device != ctx->device)
        BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_spgemm: matrices live on another device than the context");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    const int64_t launches0 = ctx->launches;
    bsp_stats st;
    memset(&st, 0, sizeof(st));
    st.nnzb_a = A->nnzb; st.nnzb_b = B->nnzb;

    bsp_mat *c = nullptr;
    BSP_TRY(new_empty_like(ctx, A->rows, B->cols, A->k, A->nbr, B->nbc, &c, A->esize)); // cu:280-282
    BSP_CUDA(ctx, cudaEventRecord(ctx->ev[0], ctx->stream));
    bsp_trace_mark(ctx, "start");
    bsp_symbolic_out sym;
    memset(&sym, 0, sizeof(sym));
    sym.pair_budget = ctx->pair_budget;
    int rc = bsp_symbolic(ctx, A, B, row_lo, row_hi, &sym);
    if (rc != BSP_OK) { delete c; return rc; }
    if (sym.n_split > 0) {
        // The pair lists (8 B per tile product) of this product do not fit the budget: compute it in
        // row panels whose lists do and put the panels together.  Rows are independent and every
        // output element keeps its accumulation order, so the result is the same bit for bit.
        delete c;
        std::vector<bsp_mat *> parts((size_t)sym.n_split, nullptr);
        for (int p = 0; p < sym.n_split && rc == BSP_OK; p++) {
            bsp_stats ps;
            rc = spgemm_impl(ctx, A, B, sym.split[p], sym.split[p + 1], &parts[(size_t)p], &ps);
            if (rc == BSP_OK) { const int64_t na = st.nnzb_a, nb = st.nnzb_b; accumulate(&st, ps); st.nnzb_a = na; st.nnzb_b = nb; }
        }
        free(sym.split);
        if (rc == BSP_OK) rc = bsp_mat_concat_panels(ctx, parts.data(), (int32_t)parts.size(), C);
        for (auto *m : parts) bsp_mat_free(ctx, m);
        if (rc != BSP_OK) { bsp_mat_free(ctx, *C); *C = nullptr; return rc; }
        st.launches = ctx->launches - launches0;
        if (stats) *stats = st;
        return BSP_OK;
    }
    c->rowptr = sym.c_rowptr; // from here on bsp_mat_free(c) releases the pattern arrays
    c->colidx = sym.c_colidx;
    c->nnzb = sym.nnzb_c;
    rc = BSP_OK;
    if (cudaEventRecord(ctx->ev[1], ctx->stream) != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "bsp_spgemm: cudaEventRecord failed: %s", cudaGetErrorString(cudaGetLastError()));
        rc = BSP_ERR_CUDA;
    }
    if (rc == BSP_OK) rc = bsp_alloc(ctx, (char **)&c->vals, (size_t)sym.nnzb_c * A->k * A->k * A->esize);
    bsp_trace_mark(ctx, "alloc_c");
    if (rc == BSP_OK) {
        if (A->esize == 8) rc = bsp_numeric(ctx, A, B, &sym, c->vals, &st.exact_refolds, &st.numeric_path);
        else rc = bsp_numeric_f32(ctx, A, B, &sym, (float *)c->vals, &st.numeric_path);
    }
    bsp_trace_mark(ctx, "numeric");
    if (rc == BSP_OK && cudaEventRecord(ctx->ev[2], ctx->stream) != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "bsp_spgemm: cudaEventRecord failed: %s", cudaGetErrorString(cudaGetLastError()));
        rc = BSP_ERR_CUDA;
    }
    bsp_release(ctx, sym.pair_ptr);
    bsp_release(ctx, sym.pairs);
    if (rc != BSP_OK) { bsp_mat_free(ctx, c); return rc; }
    *C = c;
    BSP_CUDA(ctx, cudaEventSynchronize(ctx->ev[2]));
    float ms01 = 0, ms12 = 0;
    BSP_CUDA(ctx, cudaEventElapsedTime(&ms01, ctx->ev[0], ctx->ev[1]));
    BSP_CUDA(ctx, cudaEventElapsedTime(&ms12, ctx->ev[1], ctx->ev[2]));
    if (st.numeric_path != 0 && A->esize == 8) st.exact_refolds = (int64_t) * (unsigned long long *)((char *)ctx->pinned + 64);
    st.nnzb_c = sym.nnzb_c; st.nprod = sym.nprod;
    st.ms_symbolic = ms01; st.ms_numeric = ms12; st.ms_total = ms01 + ms12;
    st.launches = ctx->launches - launches0;
    bsp_trace_dump(ctx);
    if (stats) *stats = st;
    return BSP_OK;
}

int clone_mat(bsp_ctx *ctx, const bsp_mat *s, bsp_mat **out)
{
    bsp_mat *m = nullptr;
    BSP_TRY(new_empty_like(ctx, s->rows, s->cols, s->k, s->nbr, s->nbc, &m, s->esize));
    *out = m;
    m->nnzb = s->nnzb;
    const size_t tb = (size_t)s->k * s->k * s->esize; // bytes per tile
    BSP_TRY(bsp_alloc(ctx, &m->rowptr, (size_t)s->nbr + 1));
    BSP_TRY(bsp_alloc(ctx, &m->colidx, (size_t)s->nnzb));
    BSP_TRY(bsp_alloc(ctx, (char **)&m->vals, (size_t)s->nnzb * s->k * s->k * s->esize));
    // cudaMemcpyDefault: the source may live on a peer device (a4 gathers partial products)
    BSP_CUDA(ctx, cudaMemcpyAsync(m->rowptr, s->rowptr, ((size_t)s->nbr + 1) * 4, cudaMemcpyDefault, ctx->stream));
    if (s->nnzb) {
        BSP_CUDA(ctx, cudaMemcpyAsync(m->colidx, s->colidx, (size_t)s->nnzb * 4, cudaMemcpyDefault, ctx->stream));
        BSP_CUDA(ctx, cudaMemcpyAsync(m->vals, s->vals, (size_t)s->nnzb * tb, cudaMemcpyDefault, ctx->stream));
    }
    return BSP_OK;
}

// One tree level of a chain as ONE product (see bsp_matrix.cu, "batching of independent
// products"): worthwhile when the level has several small products, whose cost is launch and
// synchronisation latency (about 0.2 ms each), not work.
int batched_level(bsp_ctx *ctx, const std::vector<bsp_mat *> &as, const std::vector<bsp_mat *> &bs,
                  std::vector<bsp_mat *> *cs, bsp_stats *st)
{
    const int m = (int)as.size();
    const int k = as[0]->k, esize = as[0]->esize;
    int64_t na = 0, nb = 0;
    int32_t ra = 0, rb = 0, nbc_a = 0, nbc_b = 0;
    std::vector<int32_t> shifts(m), limits(m);
    for (int i = 0; i < m; i++) {
        shifts[i] = rb; limits[i] = bs[i]->nbr;
        na += as[i]->nnzb; nb += bs[i]->nnzb; ra += as[i]->nbr; rb += bs[i]->nbr;
        nbc_b = bs[i]->nbc > nbc_b ? bs[i]->nbc : nbc_b;
    }
    nbc_a = rb;
    bsp_mat *A = nullptr, *B = nullptr, *C = nullptr;
    int rc = new_empty_like(ctx, as[0]->rows, as[0]->cols, k, ra, nbc_a, &A, esize);
    if (rc == BSP_OK) rc = new_empty_like(ctx, bs[0]->rows, bs[0]->cols, k, rb, nbc_b, &B, esize);
    auto fill = [&](bsp_mat *M, int64_t nnzb) {
        M->nnzb = nnzb;
        int r = bsp_alloc(ctx, &M->rowptr, (size_t)M->nbr + 1);
        if (r == BSP_OK) r = bsp_alloc(ctx, &M->colidx, (size_t)nnzb);
        if (r == BSP_OK) r = bsp_alloc(ctx, (char **)&M->vals, (size_t)nnzb * k * k * esize);
        return r;
    };
    if (rc == BSP_OK) rc = fill(A, na);
    if (rc == BSP_OK) rc = fill(B, nb);
    if (rc == BSP_OK) rc = bsp_stack_rows_device(ctx, as.data(), m, shifts.data(), limits.data(), 0x7ffffff0, A);
    if (rc == BSP_OK) rc = bsp_stack_rows_device(ctx, bs.data(), m, nullptr, nullptr, 0, B);
    if (rc == BSP_OK) rc = spgemm_impl(ctx, A, B, 0, INT32_MAX, &C, st);
    bsp_mat_free(ctx, A);
    bsp_mat_free(ctx, B);
    if (rc != BSP_OK) { bsp_mat_free(ctx, C); return rc; }
    // row pointers of the stacked product at the block boundaries -> tile ranges of the parts
    std::vector<int32_t> bounds(m + 1), brow(m + 1);
    {
        int32_t row = 0;
        for (int i = 0; i <= m; i++) {
            brow[i] = row;
            if (i < m) row += as[i]->nbr;
        }
        rc = bsp_gather_i32_to_host(ctx, C->rowptr, brow.data(), m + 1, bounds.data());
    }
    std::vector<int64_t> toff(m);
    for (int i = 0; i < m && rc == BSP_OK; i++) {
        bsp_mat *P = nullptr;
        rc = new_empty_like(ctx, as[i]->rows, bs[i]->cols, k, as[i]->nbr, bs[i]->nbc, &P, esize);
        if (rc == BSP_OK) rc = fill(P, bounds[i + 1] - bounds[i]);
        if (rc != BSP_OK) { bsp_mat_free(ctx, P); break; }
        toff[i] = bounds[i];
        cs->push_back(P);
    }
    if (rc == BSP_OK) rc = bsp_split_rows_device(ctx, C, m, brow.data(), toff.data(), cs->data());
    bsp_mat_free(ctx, C);
    if (rc != BSP_OK) {
        if (rc == BSP_ERR_CUDA) snprintf(ctx->err, sizeof(ctx->err), "batched chain level: CUDA copy failed");
        for (auto *p : *cs) bsp_mat_free(ctx, p);
        cs->clear();
    }
    return rc;
}

} // namespace

extern "C" {

const char *bsp_version(void) { return "bspgemm 0.1 (sm_100a)"; }

int bsp_ctx_create(int device, void *stream, bsp_ctx **out)
{
    if (!out) return BSP_ERR_ARG;
    *out = nullptr;
    bsp_ctx *ctx = new (std::nothrow) bsp_ctx();
    if (!ctx) return BSP_ERR_NOMEM;
    memset(ctx, 0, sizeof(*ctx));
    ctx->device = device;
    *out = ctx; // returned even on failure so that bsp_last_error can be read; destroy it either way
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || ndev == 0) {
        snprintf(ctx->err, sizeof(ctx->err), "no CUDA device available (%s); this library has no CPU fallback",
                 e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
        cudaGetLastError();
        return BSP_ERR_CUDA;
    }
    if (device < 0 || device >= ndev) BSP_FAIL(ctx, BSP_ERR_ARG, "device %d out of range (0..%d)", device, ndev - 1);
    BSP_CUDA(ctx, cudaSetDevice(device));
    cudaDeviceProp prop;
    BSP_CUDA(ctx, cudaGetDeviceProperties(&prop, device));
    ctx->sm_count = prop.multiProcessorCount;
    if (stream) {
        ctx->stream = (cudaStream_t)stream;
        ctx->owns_stream = false;
    } else {
        BSP_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        ctx->owns_stream = true;
    }
    for (int i = 0; i < 4; i++) BSP_CUDA(ctx, cudaEventCreate(&ctx->ev[i]));
    BSP_CUDA(ctx, cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking)); // downloads overlapped with compute
    ctx->pinned_bytes = 4096;
    BSP_CUDA(ctx, cudaMallocHost(&ctx->pinned, ctx->pinned_bytes));
    memset(ctx->pinned, 0, ctx->pinned_bytes);
    BSP_CUDA(ctx, cudaHostGetDevicePointer(&ctx->pinned_dev, ctx->pinned, 0));
    BSP_CUDA(ctx, cudaMalloc(&ctx->d_counters, 256));
    BSP_CUDA(ctx, cudaMemset(ctx->d_counters, 0, 256));
    {
        const char *fg = getenv("BSP_FORCE_GENERIC");
        ctx->force_generic = fg && fg[0] == '1';
        const char *tr = getenv("BSP_TRACE");
        ctx->trace = tr && tr[0] == '1';
        const char *nb = getenv("BSP_NO_BATCH");
        ctx->no_batch = nb && nb[0] == '1';
        // pair lists above this many entries (8 B each) make a product run in row panels
        ctx->pair_budget = (int64_t)1 << 30;
        if (const char *pb = getenv("BSP_PAIR_BUDGET")) {
            const long long v = atoll(pb);
            if (v > 0) ctx->pair_budget = v;
        }
    }
    // keep freed blocks in the stream-ordered pool: a chain allocates and frees every product
    cudaMemPool_t pool;
    BSP_CUDA(ctx, cudaDeviceGetDefaultMemPool(&pool, device));
    uint64_t thr = UINT64_MAX;
    BSP_CUDA(ctx, cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &thr));
    return BSP_OK;
}

void bsp_ctx_destroy(bsp_ctx *ctx)
{
    if (!ctx) return;
    if (ctx->stream) cudaStreamSynchronize(ctx->stream);
    for (int i = 0; i < 4; i++)
        if (ctx->ev[i]) cudaEventDestroy(ctx->ev[i]);
    for (int i = 0; i < 32; i++)
        if (ctx->trace_ev[i]) cudaEventDestroy(ctx->trace_ev[i]);
    if (ctx->pinned) cudaFreeHost(ctx->pinned);
    for (int i = 0; i < 2; i++) if (ctx->f32_pack[i]) cudaFree(ctx->f32_pack[i]);
    for (int i = 0; i < 2; i++) {
        if (ctx->io_buf[i]) cudaFreeHost(ctx->io_buf[i]);
        if (ctx->io_ev[i]) cudaEventDestroy(ctx->io_ev[i]);
    }
    if (ctx->d_counters) cudaFree(ctx->d_counters);
    if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    if (ctx->owns_stream && ctx->stream) cudaStreamDestroy(ctx->stream);
    delete ctx;
}

const char *bsp_last_error(const bsp_ctx *ctx) { return ctx ? ctx->err : "null context"; }

int bsp_ctx_sync(bsp_ctx *ctx)
{
    if (!ctx) return BSP_ERR_ARG;
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BSP_OK;
}

int64_t bsp_ctx_launch_count(const bsp_ctx *ctx) { return ctx ? ctx->launches : 0; }

int bsp_mat_from_device_coo(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int64_t nnzb, const int32_t *d_brow,
                            const int32_t *d_bcol, const uint64_t *d_vals, bsp_mat **out)
{
    if (!ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (k < 1 || nnzb < 0) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_from_device_coo: k=%d nnzb=%lld", k, (long long)nnzb);
    if (nnzb > 0x7fffffffLL) BSP_FAIL(ctx, BSP_ERR_LIMIT, "more than 2^31-1 tiles");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    // coordinate range (device arrays): copy back and scan on the host -- ingest is not the hot path
    int32_t max_row = -1, max_col = -1;
    if (nnzb > 0) {
        std::vector<int32_t> hr((size_t)nnzb), hc((size_t)nnzb);
        BSP_CUDA(ctx, cudaMemcpyAsync(hr.data(), d_brow, (size_t)nnzb * 4, cudaMemcpyDeviceToHost, ctx->stream));
        BSP_CUDA(ctx, cudaMemcpyAsync(hc.data(), d_bcol, (size_t)nnzb * 4, cudaMemcpyDeviceToHost, ctx->stream));
        BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        for (int64_t i = 0; i < nnzb; i++) {
            if (hr[i] < 0 || hc[i] < 0) BSP_FAIL(ctx, BSP_ERR_ARG, "negative block coordinate at entry %lld", (long long)i);
            if (hr[i] > max_row) max_row = hr[i];
            if (hc[i] > max_col) max_col = hc[i];
        }
    }
    BSP_TRY(bsp_ingest_coo(ctx, rows, cols, k, 8, nnzb, d_brow, d_bcol, d_vals, max_row, max_col, out));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // "copied, not adopted": the caller may reuse its arrays on return
    return BSP_OK;
}

static int from_coo_impl(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t esize, int64_t nnzb,
                         const int32_t *brow, const int32_t *bcol, const void *vals, bsp_mat **out)
{
    if (!ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (k < 1 || nnzb < 0 || (nnzb > 0 && (!brow || !bcol || !vals)))
        BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_from_coo: bad arguments (k=%d nnzb=%lld)", k, (long long)nnzb);
    if (nnzb > 0x7fffffffLL) BSP_FAIL(ctx, BSP_ERR_LIMIT, "more than 2^31-1 tiles");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    int32_t max_row = -1, max_col = -1;
    for (int64_t i = 0; i < nnzb; i++) {
        if (brow[i] < 0 || bcol[i] < 0) BSP_FAIL(ctx, BSP_ERR_ARG, "negative block coordinate at entry %lld", (long long)i);
        if (brow[i] > max_row) max_row = brow[i];
        if (bcol[i] > max_col) max_col = bcol[i];
    }
    const size_t tb = (size_t)k * k * esize;
    int32_t *d_r = nullptr, *d_c = nullptr;
    char *d_v = nullptr;
    BSP_TRY(bsp_alloc(ctx, &d_r, (size_t)nnzb));
    BSP_TRY(bsp_alloc(ctx, &d_c, (size_t)nnzb));
    BSP_TRY(bsp_alloc(ctx, &d_v, (size_t)nnzb * tb));
    if (nnzb) {
        BSP_CUDA(ctx, cudaMemcpyAsync(d_r, brow, (size_t)nnzb * 4, cudaMemcpyHostToDevice, ctx->stream));
        BSP_CUDA(ctx, cudaMemcpyAsync(d_c, bcol, (size_t)nnzb * 4, cudaMemcpyHostToDevice, ctx->stream));
        BSP_CUDA(ctx, cudaMemcpyAsync(d_v, vals, (size_t)nnzb * tb, cudaMemcpyHostToDevice, ctx->stream));
    }
    int rc = bsp_ingest_coo(ctx, rows, cols, k, esize, nnzb, d_r, d_c, d_v, max_row, max_col, out);
    bsp_release(ctx, d_r); bsp_release(ctx, d_c); bsp_release(ctx, d_v);
    if (rc == BSP_OK) BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // host buffers may be reused on return
    return rc;
}

int bsp_mat_from_coo(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int64_t nnzb, const int32_t *brow,
                     const int32_t *bcol, const uint64_t *vals, bsp_mat **out)
{
    return from_coo_impl(ctx, rows, cols, k, 8, nnzb, brow, bcol, vals, out);
}

int bsp_mat_from_coo_f32(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int64_t nnzb, const int32_t *brow,
                         const int32_t *bcol, const float *vals, bsp_mat **out)
{
    return from_coo_impl(ctx, rows, cols, k, 4, nnzb, brow, bcol, vals, out);
}

int bsp_mat_is_f32(const bsp_mat *m) { return m && m->esize == 4; }

int bsp_ctx_set_f32_path(bsp_ctx *ctx, int path)
{
    if (!ctx || path < 0 || path > 2) return BSP_ERR_ARG;
    ctx->f32_path = path;
    return BSP_OK;
}

int bsp_mat_info(const bsp_mat *m, int32_t *rows, int32_t *cols, int32_t *k, int64_t *nnzb)
{
    if (!m) return BSP_ERR_ARG;
    if (rows) *rows = m->rows;
    if (cols) *cols = m->cols;
    if (k) *k = m->k;
    if (nnzb) *nnzb = m->nnzb;
    return BSP_OK;
}

static int to_coo_impl(bsp_ctx *ctx, const bsp_mat *m, int32_t esize, int32_t *brow, int32_t *bcol, void *vals)
{
    if (!ctx || !valid_mat(m)) return BSP_ERR_ARG;
    if (m->esize != esize) BSP_FAIL(ctx, BSP_ERR_ARG, "matrix holds %s entries", m->esize == 8 ? "uint64" : "float32");
    if (m->nnzb == 0) return BSP_OK;
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t tb = (size_t)m->k * m->k * m->esize;
    if (brow) {
        int32_t *d_brow = nullptr;
        BSP_TRY(bsp_alloc(ctx, &d_brow, (size_t)m->nnzb));
        BSP_TRY(bsp_expand_rows(ctx, m, d_brow));
        BSP_CUDA(ctx, cudaMemcpyAsync(brow, d_brow, (size_t)m->nnzb * 4, cudaMemcpyDeviceToHost, ctx->stream));
        bsp_release(ctx, d_brow);
    }
    if (bcol) BSP_CUDA(ctx, cudaMemcpyAsync(bcol, m->colidx, (size_t)m->nnzb * 4, cudaMemcpyDeviceToHost, ctx->stream));
    if (vals) BSP_CUDA(ctx, cudaMemcpyAsync(vals, m->vals, (size_t)m->nnzb * tb, cudaMemcpyDeviceToHost, ctx->stream));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BSP_OK;
}

int bsp_mat_to_coo(bsp_ctx *ctx, const bsp_mat *m, int32_t *brow, int32_t *bcol, uint64_t *vals)
{
    return to_coo_impl(ctx, m, 8, brow, bcol, vals);
}

int bsp_mat_to_coo_f32(bsp_ctx *ctx, const bsp_mat *m, int32_t *brow, int32_t *bcol, float *vals)
{
    return to_coo_impl(ctx, m, 4, brow, bcol, vals);
}

int bsp_mat_device_view(const bsp_mat *m, int32_t *nbrows, int32_t *nbcols, const int32_t **d_rowptr,
                        const int32_t **d_colidx, const uint64_t **d_vals)
{
    if (!valid_mat(m)) return BSP_ERR_ARG;
    if (nbrows) *nbrows = m->nbr;
    if (nbcols) *nbcols = m->nbc;
    if (d_rowptr) *d_rowptr = m->rowptr;
    if (d_colidx) *d_colidx = m->colidx;
    if (d_vals) *d_vals = m->vals;
    return BSP_OK;
}

int bsp_mat_export_device(bsp_ctx *ctx, const bsp_mat *m, int32_t *d_rowptr, int32_t *d_colidx, uint64_t *d_vals)
{
    if (!ctx || !valid_mat(m)) return BSP_ERR_ARG;
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    const size_t tb = (size_t)m->k * m->k * m->esize;
    if (d_rowptr)
        BSP_CUDA(ctx, cudaMemcpyAsync(d_rowptr, m->rowptr, ((size_t)m->nbr + 1) * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    if (d_colidx && m->nnzb)
        BSP_CUDA(ctx, cudaMemcpyAsync(d_colidx, m->colidx, (size_t)m->nnzb * 4, cudaMemcpyDeviceToDevice, ctx->stream));
    if (d_vals && m->nnzb)
        BSP_CUDA(ctx, cudaMemcpyAsync(d_vals, m->vals, (size_t)m->nnzb * tb, cudaMemcpyDeviceToDevice, ctx->stream));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BSP_OK;
}

int bsp_mat_copy_to_ctx(bsp_ctx *src_ctx, const bsp_mat *m, bsp_ctx *dst_ctx, bsp_mat **out)
{
    if (!src_ctx || !dst_ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (!valid_mat(m)) BSP_FAIL(dst_ctx, BSP_ERR_ARG, "bsp_mat_copy_to_ctx: null matrix");
    BSP_CUDA(src_ctx, cudaSetDevice(src_ctx->device));
    BSP_CUDA(src_ctx, cudaStreamSynchronize(src_ctx->stream)); // the source must be complete
    BSP_CUDA(dst_ctx, cudaSetDevice(dst_ctx->device));
    bsp_mat *c = nullptr;
    BSP_TRY(new_empty_like(dst_ctx, m->rows, m->cols, m->k, m->nbr, m->nbc, &c, m->esize));
    *out = c;
    c->nnzb = m->nnzb;
    const size_t tb = (size_t)m->k * m->k * m->esize;
    BSP_TRY(bsp_alloc(dst_ctx, &c->rowptr, (size_t)m->nbr + 1));
    BSP_TRY(bsp_alloc(dst_ctx, &c->colidx, (size_t)m->nnzb));
    BSP_TRY(bsp_alloc(dst_ctx, (char **)&c->vals, (size_t)m->nnzb * tb));
    const int sd = m->device, dd = dst_ctx->device;
    BSP_CUDA(dst_ctx, cudaMemcpyPeerAsync(c->rowptr, dd, m->rowptr, sd, ((size_t)m->nbr + 1) * 4, dst_ctx->stream));
    if (m->nnzb) {
        BSP_CUDA(dst_ctx, cudaMemcpyPeerAsync(c->colidx, dd, m->colidx, sd, (size_t)m->nnzb * 4, dst_ctx->stream));
        BSP_CUDA(dst_ctx, cudaMemcpyPeerAsync(c->vals, dd, m->vals, sd, (size_t)m->nnzb * tb, dst_ctx->stream));
    }
    BSP_CUDA(dst_ctx, cudaStreamSynchronize(dst_ctx->stream));
    return BSP_OK;
}

static int from_device_bcsr_impl(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t esize, int32_t nbrows,
                                 int32_t nbcols, int64_t nnzb, const int32_t *d_rowptr, const int32_t *d_colidx,
                                 const void *d_vals, bsp_mat **out)
{
    if (!ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (k < 1 || nbrows < 0 || nbcols < 0 || nnzb < 0 || !d_rowptr || (nnzb > 0 && (!d_colidx || !d_vals)))
        BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_from_device_bcsr: bad arguments");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    bsp_mat tmp;
    tmp.rows = rows; tmp.cols = cols; tmp.k = k; tmp.nbr = nbrows; tmp.nbc = nbcols; tmp.nnzb = nnzb;
    tmp.rowptr = const_cast<int32_t *>(d_rowptr);
    tmp.colidx = const_cast<int32_t *>(d_colidx);
    tmp.vals = reinterpret_cast<uint64_t *>(const_cast<void *>(d_vals));
    tmp.device = ctx->device;
    tmp.esize = esize;
    int rc = clone_mat(ctx, &tmp, out);
    // "copied, not adopted": the copies are complete when the call returns, so the caller may free or
    // overwrite its arrays on any stream
    if (rc == BSP_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "bsp_mat_from_device_bcsr: copy failed: %s", cudaGetErrorString(cudaGetLastError()));
        rc = BSP_ERR_CUDA;
    }
    if (rc != BSP_OK) { bsp_mat_free(ctx, *out); *out = nullptr; }
    return rc;
}

int bsp_mat_from_device_bcsr(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t nbrows, int32_t nbcols,
                             int64_t nnzb, const int32_t *d_rowptr, const int32_t *d_colidx, const uint64_t *d_vals,
                             bsp_mat **out)
{
    return from_device_bcsr_impl(ctx, rows, cols, k, 8, nbrows, nbcols, nnzb, d_rowptr, d_colidx, d_vals, out);
}

int bsp_mat_from_device_bcsr_f32(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t nbrows, int32_t nbcols,
                                 int64_t nnzb, const int32_t *d_rowptr, const int32_t *d_colidx, const float *d_vals,
                                 bsp_mat **out)
{
    return from_device_bcsr_impl(ctx, rows, cols, k, 4, nbrows, nbcols, nnzb, d_rowptr, d_colidx, d_vals, out);
}

void bsp_mat_free(bsp_ctx *ctx, bsp_mat *m)
{
    if (!m) return;
    if (ctx) {
        cudaSetDevice(ctx->device);
        bsp_release(ctx, m->rowptr); bsp_release(ctx, m->colidx); bsp_release(ctx, m->vals);
    } else {
        cudaFree(m->rowptr); cudaFree(m->colidx); cudaFree(m->vals);
    }
    delete m;
}

int bsp_spgemm(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, bsp_mat **C, bsp_stats *stats)
{
    return spgemm_impl(ctx, A, B, 0, INT32_MAX, C, stats);
}

int bsp_spgemm_panel(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t row_lo, int32_t row_hi, bsp_mat **C,
                     bsp_stats *stats)
{
    if (ctx && row_lo > row_hi) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_spgemm_panel: row_lo > row_hi");
    return spgemm_impl(ctx, A, B, row_lo, row_hi, C, stats);
}

int bsp_mat_concat_panels(bsp_ctx *ctx, bsp_mat *const *panels, int32_t n, bsp_mat **out)
{
    if (!ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (n < 1 || !panels) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_concat_panels: no panels");
    int64_t total = 0;
    for (int i = 0; i < n; i++) {
        if (!valid_mat(panels[i])) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_concat_panels: panel %d is null", i);
        if (panels[i]->k != panels[0]->k || panels[i]->nbr != panels[0]->nbr || panels[i]->esize != panels[0]->esize)
            BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_concat_panels: panel %d has another shape", i);
        total += panels[i]->nnzb;
    }
    if (total > 0x7fffffffLL) BSP_FAIL(ctx, BSP_ERR_LIMIT, "more than 2^31-1 tiles");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    bsp_mat *m = nullptr;
    int32_t nbc = 0;
    for (int i = 0; i < n; i++) nbc = panels[i]->nbc > nbc ? panels[i]->nbc : nbc;
    BSP_TRY(new_empty_like(ctx, panels[0]->rows, panels[0]->cols, panels[0]->k, panels[0]->nbr, nbc, &m, panels[0]->esize));
    *out = m;
    m->nnzb = total;
    BSP_TRY(bsp_alloc(ctx, &m->rowptr, (size_t)m->nbr + 1));
    BSP_TRY(bsp_alloc(ctx, &m->colidx, (size_t)total));
    BSP_TRY(bsp_alloc(ctx, (char **)&m->vals, (size_t)total * m->k * m->k * m->esize));
    return bsp_concat_panels_device(ctx, panels, n, m);
}

int bsp_chain(bsp_ctx *ctx, bsp_mat *const *mats, int32_t n, bsp_mat **out, bsp_stats *stats, char *log,
              size_t log_cap, int32_t log_base)
{
    if (!ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (n < 1 || !mats) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_chain: empty chain");
    for (int i = 0; i < n; i++)
        if (!valid_mat(mats[i])) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_chain: matrix %d is null", i);
    bsp_stats tot;
    memset(&tot, 0, sizeof(tot));
    size_t lp = 0;
    if (log && log_cap) log[0] = 0;
    // level 0 borrows the inputs; later levels own their matrices
    std::vector<bsp_mat *> cur(mats, mats + n);
    std::vector<char> owned((size_t)n, 0);
    int rc = BSP_OK;
    while (cur.size() > 1 && rc == BSP_OK) {
        const size_t len = cur.size();
        std::vector<bsp_mat *> nxt;
        std::vector<char> nxt_owned;
        for (size_t ind = 0; ind + 1 < len; ind += 2) // cu:297: adjacent pairs; cu:301: the reference's log line
            if (log && lp + 48 < log_cap)
                lp += (size_t)snprintf(log + lp, log_cap - lp, "multiplying %d %d\n", log_base + (int)ind,
                                       log_base + (int)ind + 1);
        // several small products on one level: run them as one stacked product
        const size_t npairs = len / 2;
        int64_t level_tiles = 0;
        bool same = true;
        for (size_t ind = 0; ind + 1 < len; ind += 2) {
            level_tiles += cur[ind]->nnzb + cur[ind + 1]->nnzb;
            same = same && cur[ind]->k == cur[0]->k && cur[ind + 1]->k == cur[0]->k &&
                   cur[ind]->esize == cur[0]->esize && cur[ind + 1]->esize == cur[0]->esize;
        }
        if (!ctx->no_batch && same && npairs >= 2 && level_tiles <= (int64_t)4 << 20) {
            std::vector<bsp_mat *> as, bs, cs;
            for (size_t ind = 0; ind + 1 < len; ind += 2) { as.push_back(cur[ind]); bs.push_back(cur[ind + 1]); }
            bsp_stats st;
            rc = batched_level(ctx, as, bs, &cs, &st);
            if (rc == BSP_OK) {
                accumulate(&tot, st);
                for (auto *c : cs) { nxt.push_back(c); nxt_owned.push_back(1); }
                for (size_t ind = 0; ind + 1 < len; ind += 2) {
                    if (owned[ind]) { bsp_mat_free(ctx, cur[ind]); cur[ind] = nullptr; }
                    if (owned[ind + 1]) { bsp_mat_free(ctx, cur[ind + 1]); cur[ind + 1] = nullptr; }
                }
            }
        } else
        for (size_t ind = 0; ind + 1 < len; ind += 2) {
            bsp_mat *c = nullptr;
            bsp_stats st;
            rc = bsp_spgemm(ctx, cur[ind], cur[ind + 1], &c, &st);
            if (rc != BSP_OK) { bsp_mat_free(ctx, c); break; }
            accumulate(&tot, st);
            nxt.push_back(c);
            nxt_owned.push_back(1);
            if (owned[ind]) { bsp_mat_free(ctx, cur[ind]); cur[ind] = nullptr; }
            if (owned[ind + 1]) { bsp_mat_free(ctx, cur[ind + 1]); cur[ind + 1] = nullptr; }
        }
        if (rc == BSP_OK && (len & 1)) { // cu:318-320: odd tail carried unchanged
            nxt.push_back(cur[len - 1]);
            nxt_owned.push_back(owned[len - 1]);
            cur[len - 1] = nullptr;
        }
        if (rc != BSP_OK) {
            for (size_t i = 0; i < nxt.size(); i++) if (nxt_owned[i]) bsp_mat_free(ctx, nxt[i]);
            for (size_t i = 0; i < cur.size(); i++) if (cur[i] && owned[i]) bsp_mat_free(ctx, cur[i]);
            return rc;
        }
        cur.swap(nxt);
        owned.swap(nxt_owned);
    }
    if (owned[0]) *out = cur[0];
    else rc = clone_mat(ctx, cur[0], out); // n == 1: the reference leaves arr[start] as it is
    if (stats) *stats = tot;
    return rc;
}

int bsp_prune_zero(bsp_ctx *ctx, bsp_mat *m)
{
    if (!ctx || !valid_mat(m)) return BSP_ERR_ARG;
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    return bsp_prune_device(ctx, m);
}

int bsp_multiply_host(bsp_ctx *ctx, int32_t k, int32_t rows_a, int32_t cols_a, int64_t nnzb_a, const int32_t *brow_a,
                      const int32_t *bcol_a, const uint64_t *vals_a, int32_t rows_b, int32_t cols_b, int64_t nnzb_b,
                      const int32_t *brow_b, const int32_t *bcol_b, const uint64_t *vals_b, int64_t cap_c,
                      int64_t *nnzb_c, int32_t *brow_c, int32_t *bcol_c, uint64_t *vals_c, bsp_stats *stats)
{
    if (!ctx || !nnzb_c) return BSP_ERR_ARG;
    bsp_mat *A = nullptr, *B = nullptr, *C = nullptr;
    auto now_ms = []() { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const char *th = getenv("BSP_TRACE_HOST"); // host-side timeline of this call on stderr (adds no device waits)
    const bool trace_host = ctx->trace || (th && th[0] == '1');
    const double t_start = now_ms();
    int rc = bsp_mat_from_coo(ctx, rows_a, cols_a, k, nnzb_a, brow_a, bcol_a, vals_a, &A);
    const double t_a = now_ms();
    if (rc == BSP_OK) rc = bsp_mat_from_coo(ctx, rows_b, cols_b, k, nnzb_b, brow_b, bcol_b, vals_b, &B);
    const double t_b = now_ms();
    const bool want_out = brow_c || bcol_c || vals_c;
    const char *nopipe = getenv("BSP_NO_PIPELINE");
    if (rc == BSP_OK && (!want_out || A->nnzb < 65536 || !ctx->copy_stream || (nopipe && nopipe[0] == '1'))) {
        // small problem or size query: one product, one download
        rc = bsp_spgemm(ctx, A, B, &C, stats);
        if (rc == BSP_OK) {
            *nnzb_c = C->nnzb;
            if (want_out) {
                if (C->nnzb > cap_c) {
                    snprintf(ctx->err, sizeof(ctx->err), "bsp_multiply_host: result has %lld tiles, capacity %lld",
                             (long long)C->nnzb, (long long)cap_c);
                    rc = BSP_ERR_LIMIT;
                } else {
                    rc = bsp_mat_to_coo(ctx, C, brow_c, bcol_c, vals_c);
                }
            }
        }
        bsp_mat_free(ctx, A); bsp_mat_free(ctx, B); bsp_mat_free(ctx, C);
        return rc;
    }
    // Large result: the download (PCIe, ~50 GB/s) dwarfs the product.  Compute block-row panels
    // one after the other and download panel p on a second stream while panel p+1 is computed.
    constexpr int NP = 8;
    bsp_mat *panels[NP] = {};
    int32_t *d_brow[NP] = {};
    bsp_stats tot;
    memset(&tot, 0, sizeof(tot));
    int64_t done = 0;
    if (rc == BSP_OK) {
        // panel bounds: equal shares of A's tiles (host histogram of the caller's coordinates)
        std::vector<int64_t> cnt((size_t)A->nbr + 1, 0);
        for (int64_t i = 0; i < nnzb_a; i++) cnt[(size_t)brow_a[i] + 1]++;
        for (int32_t r = 0; r < A->nbr; r++) cnt[(size_t)r + 1] += cnt[(size_t)r];
        int32_t lo = 0;
        const size_t tb = (size_t)k * k * sizeof(uint64_t);
        for (int p = 0; p < NP && rc == BSP_OK; p++) {
            int32_t hi = A->nbr;
            if (p + 1 < NP) {
                const int64_t target = nnzb_a * (p + 1) / NP;
                hi = (int32_t)(std::lower_bound(cnt.begin(), cnt.end(), target) - cnt.begin());
                if (hi < lo) hi = lo;
                if (hi > A->nbr) hi = A->nbr;
            }
            bsp_stats st;
            rc = spgemm_impl(ctx, A, B, lo, hi, &panels[p], &st);
            lo = hi;
            if (rc != BSP_OK) break;
            accumulate(&tot, st);
            tot.nnzb_b = st.nnzb_b; // B is the same in every panel: count it once
            bsp_mat *P = panels[p];
            if (done + P->nnzb > cap_c) {
                snprintf(ctx->err, sizeof(ctx->err), "bsp_multiply_host: result has more than %lld tiles (capacity)", (long long)cap_c);
                rc = BSP_ERR_LIMIT;
                break;
            }
            if (P->nnzb) {
                rc = bsp_alloc(ctx, &d_brow[p], (size_t)P->nnzb);
                if (rc == BSP_OK) rc = bsp_expand_rows(ctx, P, d_brow[p]);
                if (rc != BSP_OK) break;
                // copy stream waits for this panel, then downloads it while the next panel is computed
                cudaError_t ce = cudaEventRecord(ctx->ev[3], ctx->stream);
                if (ce == cudaSuccess) ce = cudaStreamWaitEvent(ctx->copy_stream, ctx->ev[3], 0);
                if (ce == cudaSuccess && brow_c)
                    ce = cudaMemcpyAsync(brow_c + done, d_brow[p], (size_t)P->nnzb * 4, cudaMemcpyDeviceToHost, ctx->copy_stream);
                if (ce == cudaSuccess && bcol_c)
                    ce = cudaMemcpyAsync(bcol_c + done, P->colidx, (size_t)P->nnzb * 4, cudaMemcpyDeviceToHost, ctx->copy_stream);
                if (ce == cudaSuccess && vals_c)
                    ce = cudaMemcpyAsync((char *)vals_c + (size_t)done * tb, P->vals, (size_t)P->nnzb * tb, cudaMemcpyDeviceToHost, ctx->copy_stream);
                if (ce != cudaSuccess) {
                    snprintf(ctx->err, sizeof(ctx->err), "bsp_multiply_host: queuing the download of panel %d failed: %s", p, cudaGetErrorString(ce));
                    cudaGetLastError();
                    rc = BSP_ERR_CUDA;
                    break;
                }
                done += P->nnzb;
            }
        }
    }
    const double t_panels = now_ms();
    if (ctx->copy_stream && cudaStreamSynchronize(ctx->copy_stream) != cudaSuccess && rc == BSP_OK) {
        snprintf(ctx->err, sizeof(ctx->err), "bsp_multiply_host: download failed: %s", cudaGetErrorString(cudaGetLastError()));
        rc = BSP_ERR_CUDA;
    }
    const double t_copied = now_ms();
    for (int p = 0; p < NP; p++) { bsp_release(ctx, d_brow[p]); bsp_mat_free(ctx, panels[p]); }
    bsp_mat_free(ctx, A); bsp_mat_free(ctx, B);
    if (trace_host)
        fprintf(stderr, "[bsp trace] multiply_host: A on the device after %.1f ms, B after %.1f, panels issued after %.1f, "
                        "downloads done after %.1f, buffers released after %.1f\n",
                t_a - t_start, t_b - t_start, t_panels - t_start, t_copied - t_start, now_ms() - t_start);
    if (rc == BSP_OK) {
        *nnzb_c = done;
        tot.nnzb_a = nnzb_a;
        if (stats) *stats = tot;
    }
    return rc;
}

} // extern "C"
