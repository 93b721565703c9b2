/*
 * bsp_oracle.c -- CPU restatement of the reference's block-sparse chain product.
 *
 * TEST INFRASTRUCTURE ONLY (see bsp_oracle.h).  Plain C, no dependence on the product.
 * "cu:N" = /root/reference/sparse_matrix_mult.cu line N.
 *
 * The reference stores a matrix as an ordered map (brow,bcol) -> k x k tile (cu:26-32)
 * and computes C = A*B as: for every A tile (i,j) in map order and every B tile (j,c),
 * append j to the list of output key (i,c) (cu:147-156); then one GPU block per output
 * key folds its list in order (cu:44-66).  Here the same lists are formed per block row
 * with a stable sort, and the same fold is run on the CPU.
 */
#define _GNU_SOURCE
#include "bsp_oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

static const uint64_t ORC_MAX = 18446744073709551615ULL; /* cu:48 */

/* cu:59: the product wraps in 64 bits, then "% MAX" (maps 2^64-1 to 0, nothing else). */
uint64_t orc_mulmod(uint64_t a, uint64_t b) { return (uint64_t)(a * b) % ORC_MAX; }
/* cu:61: the sum wraps in 64 bits, then "% MAX". */
uint64_t orc_addmod(uint64_t s, uint64_t p) { return (uint64_t)(s + p) % ORC_MAX; }

static inline uint64_t key_of(int32_t r, int32_t c)
{
    /* order-preserving for signed ints (std::pair<int,int> ordering, cu:28) */
    return ((uint64_t)((uint32_t)r ^ 0x80000000u) << 32) | (uint64_t)((uint32_t)c ^ 0x80000000u);
}

static orc_mat *orc_alloc(int32_t rows, int32_t cols, int32_t k, int64_t nnzb)
{
    orc_mat *m = (orc_mat *)calloc(1, sizeof(orc_mat));
    if (!m) return NULL;
    m->rows = rows; m->cols = cols; m->k = k; m->nnzb = nnzb;
    size_t n = (size_t)(nnzb > 0 ? nnzb : 1);
    m->brow = (int32_t *)malloc(n * sizeof(int32_t));
    m->bcol = (int32_t *)malloc(n * sizeof(int32_t));
    m->vals = (uint64_t *)malloc(n * (size_t)k * k * sizeof(uint64_t));
    if (!m->brow || !m->bcol || !m->vals) { orc_free(m); return NULL; }
    return m;
}

void orc_free(orc_mat *m)
{
    if (!m) return;
    free(m->brow); free(m->bcol); free(m->vals); free(m);
}

orc_mat *orc_clone(const orc_mat *s)
{
    orc_mat *m = orc_alloc(s->rows, s->cols, s->k, s->nnzb);
    if (!m) return NULL;
    memcpy(m->brow, s->brow, (size_t)s->nnzb * sizeof(int32_t));
    memcpy(m->bcol, s->bcol, (size_t)s->nnzb * sizeof(int32_t));
    memcpy(m->vals, s->vals, (size_t)s->nnzb * s->k * s->k * sizeof(uint64_t));
    return m;
}

/* ---- stable merge sort of (key, payload) records ------------------------------------ */
typedef struct { uint64_t key; int64_t a; int64_t b; } rec_t;

static void rec_sort(rec_t *v, rec_t *tmp, int64_t n)
{
    for (int64_t w = 1; w < n; w *= 2) {
        for (int64_t lo = 0; lo < n; lo += 2 * w) {
            int64_t mid = lo + w < n ? lo + w : n, hi = lo + 2 * w < n ? lo + 2 * w : n;
            int64_t i = lo, j = mid, o = lo;
            while (i < mid && j < hi) tmp[o++] = (v[j].key < v[i].key) ? v[j++] : v[i++];
            while (i < mid) tmp[o++] = v[i++];
            while (j < hi) tmp[o++] = v[j++];
        }
        memcpy(v, tmp, (size_t)n * sizeof(rec_t));
    }
}

orc_mat *orc_from_coo(int32_t rows, int32_t cols, int32_t k, int64_t nnzb,
                      const int32_t *brow, const int32_t *bcol, const uint64_t *vals)
{
    size_t kk = (size_t)k * k;
    rec_t *r = (rec_t *)malloc((size_t)(nnzb + 1) * sizeof(rec_t));
    rec_t *t = (rec_t *)malloc((size_t)(nnzb + 1) * sizeof(rec_t));
    if (!r || !t) { free(r); free(t); return NULL; }
    for (int64_t i = 0; i < nnzb; i++) { r[i].key = key_of(brow[i], bcol[i]); r[i].a = i; r[i].b = 0; }
    rec_sort(r, t, nnzb);
    /* m[{row,col}] = v overwrites an existing entry (cu:383): keep the last of each run */
    int64_t uniq = 0;
    for (int64_t i = 0; i < nnzb; i++)
        if (i + 1 == nnzb || r[i + 1].key != r[i].key) t[uniq++] = r[i];
    orc_mat *m = orc_alloc(rows, cols, k, uniq);
    if (m) {
        for (int64_t i = 0; i < uniq; i++) {
            int64_t s = t[i].a;
            m->brow[i] = brow[s]; m->bcol[i] = bcol[s];
            memcpy(m->vals + (size_t)i * kk, vals + (size_t)s * kk, kk * sizeof(uint64_t));
        }
    }
    free(r); free(t);
    return m;
}

/* first index p in m with brow[p] >= row (m sorted by (brow,bcol)) */
static int64_t row_lower(const orc_mat *m, int32_t row)
{
    int64_t lo = 0, hi = m->nnzb;
    while (lo < hi) {
        int64_t mid = (lo + hi) / 2;
        if (m->brow[mid] < row) lo = mid + 1; else hi = mid;
    }
    return lo;
}

/* Fold one output tile exactly as the kernel does (cu:53-65): thread (ty,tx) walks the
 * pair list in order, and inside a pair the inner index j ascending. */
static void fold_tile(uint64_t *out, const rec_t *pairs, int64_t np,
                      const uint64_t *av, const uint64_t *bv, int k)
{
    size_t kk = (size_t)k * k;
    for (int ty = 0; ty < k; ty++)
        for (int tx = 0; tx < k; tx++) {
            uint64_t total = 0;
            for (int64_t p = 0; p < np; p++) {
                const uint64_t *A = av + (size_t)pairs[p].a * kk;
                const uint64_t *B = bv + (size_t)pairs[p].b * kk;
                for (int j = 0; j < k; j++) {
                    uint64_t t1 = orc_mulmod(A[ty * k + j], B[j * k + tx]);
                    total = orc_addmod(total, t1);
                }
            }
            out[ty * k + tx] = total;
        }
}

typedef struct { int64_t a_lo, a_hi; int64_t nprod; int64_t nout; int64_t out_off; } rowjob_t;

/* products of one block row of A, sorted by output column (stable => j ascending) */
static int64_t row_products(const orc_mat *a, const orc_mat *b, const rowjob_t *job,
                            rec_t *rec, rec_t *tmp)
{
    int64_t n = 0;
    for (int64_t p = job->a_lo; p < job->a_hi; p++) {
        int32_t j = a->bcol[p];
        for (int64_t q = row_lower(b, j); q < b->nnzb && b->brow[q] == j; q++) {
            rec[n].key = key_of(0, b->bcol[q]); rec[n].a = p; rec[n].b = q; n++;
        }
    }
    rec_sort(rec, tmp, n);
    return n;
}

orc_mat *orc_multiply_mt(const orc_mat *a, const orc_mat *b, int64_t *nprod_out, int threads)
{
    if (a->k != b->k) return NULL;
    int k = a->k;
    size_t kk = (size_t)k * k;
    /* block rows of A */
    int64_t nrows = 0;
    for (int64_t p = 0; p < a->nnzb; p++) if (p == 0 || a->brow[p] != a->brow[p - 1]) nrows++;
    rowjob_t *jobs = (rowjob_t *)calloc((size_t)nrows + 1, sizeof(rowjob_t));
    int64_t r = -1;
    for (int64_t p = 0; p < a->nnzb; p++) {
        if (p == 0 || a->brow[p] != a->brow[p - 1]) { r++; jobs[r].a_lo = p; }
        jobs[r].a_hi = p + 1;
    }
    int64_t maxprod = 0, totprod = 0;
    for (int64_t i = 0; i < nrows; i++) {
        int64_t n = 0;
        for (int64_t p = jobs[i].a_lo; p < jobs[i].a_hi; p++) {
            int32_t j = a->bcol[p];
            int64_t q = row_lower(b, j), q2 = (j == INT32_MAX) ? b->nnzb : row_lower(b, j + 1);
            n += q2 - q;
        }
        jobs[i].nprod = n; totprod += n;
        if (n > maxprod) maxprod = n;
    }
    if (threads < 1) threads = 1;
    /* pass 1: count output tiles per row */
#ifdef _OPENMP
#pragma omp parallel num_threads(threads)
#endif
    {
        rec_t *rec = (rec_t *)malloc((size_t)(maxprod + 1) * sizeof(rec_t));
        rec_t *tmp = (rec_t *)malloc((size_t)(maxprod + 1) * sizeof(rec_t));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 4)
#endif
        for (int64_t i = 0; i < nrows; i++) {
            int64_t n = row_products(a, b, &jobs[i], rec, tmp), no = 0;
            for (int64_t t = 0; t < n; t++) if (t == 0 || rec[t].key != rec[t - 1].key) no++;
            jobs[i].nout = no;
        }
        free(rec); free(tmp);
    }
    int64_t nout = 0;
    for (int64_t i = 0; i < nrows; i++) { jobs[i].out_off = nout; nout += jobs[i].nout; }
    orc_mat *c = orc_alloc(a->rows, b->cols, k, nout); /* cu:280-283 */
    if (!c) { free(jobs); return NULL; }
    /* pass 2: fold */
#ifdef _OPENMP
#pragma omp parallel num_threads(threads)
#endif
    {
        rec_t *rec = (rec_t *)malloc((size_t)(maxprod + 1) * sizeof(rec_t));
        rec_t *tmp = (rec_t *)malloc((size_t)(maxprod + 1) * sizeof(rec_t));
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 4)
#endif
        for (int64_t i = 0; i < nrows; i++) {
            int64_t n = row_products(a, b, &jobs[i], rec, tmp);
            int64_t o = jobs[i].out_off, t = 0;
            while (t < n) {
                int64_t e = t + 1;
                while (e < n && rec[e].key == rec[t].key) e++;
                c->brow[o] = a->brow[jobs[i].a_lo];
                c->bcol[o] = b->bcol[rec[t].b];
                fold_tile(c->vals + (size_t)o * kk, rec + t, e - t, a->vals, b->vals, k);
                o++; t = e;
            }
        }
        free(rec); free(tmp);
    }
    free(jobs);
    if (nprod_out) *nprod_out = totprod;
    return c;
}

orc_mat *orc_multiply(const orc_mat *a, const orc_mat *b, int64_t *nprod_out)
{
    /* ORC_THREADS=n spreads the emulated kernel loop of every product (also those inside orc_chain /
     * orc_chain_ranks) over n threads; each output tile is still folded by one thread in reference
     * order, so the result does not depend on n (used by tests/golden/make_full_size.py). */
    const char *e = getenv("ORC_THREADS");
    int nt = e ? atoi(e) : 1;
    return orc_multiply_mt(a, b, nprod_out, nt > 0 ? nt : 1);
}

/* helper2 (cu:287-327): adjacent pairs at each level, odd tail carried unchanged. */
static orc_mat *chain_impl(orc_mat *const *arr, int n, int base, int64_t *nprod_total,
                           char *log, int64_t log_cap)
{
    if (n <= 0) return NULL;
    orc_mat **cur = (orc_mat **)malloc((size_t)n * sizeof(orc_mat *));
    for (int i = 0; i < n; i++) cur[i] = orc_clone(arr[i]);
    int len = n;
    int64_t lp = log ? (int64_t)strlen(log) : 0;
    while (len > 1) {
        int nn = len / 2 + (len & 1);
        orc_mat **nxt = (orc_mat **)malloc((size_t)nn * sizeof(orc_mat *));
        for (int ind = 0; ind + 1 < len; ind += 2) {
            if (log && lp + 48 < log_cap)
                lp += snprintf(log + lp, (size_t)(log_cap - lp), "multiplying %d %d\n", base + ind, base + ind + 1);
            int64_t np = 0;
            nxt[ind / 2] = orc_multiply(cur[ind], cur[ind + 1], &np);
            if (nprod_total) *nprod_total += np;
            orc_free(cur[ind]); orc_free(cur[ind + 1]);
        }
        if (len & 1) nxt[nn - 1] = cur[len - 1];
        free(cur); cur = nxt; len = nn;
    }
    orc_mat *res = cur[0];
    free(cur);
    return res;
}

orc_mat *orc_chain(orc_mat *const *arr, int n, int64_t *nprod_total, char *log, int64_t log_cap)
{
    return chain_impl(arr, n, 0, nprod_total, log, log_cap);
}

orc_mat *orc_chain_ranks(orc_mat *const *arr, int n, int P, int64_t *nprod_total)
{
    int per = n / P; /* cu:439 */
    if (per == 0) return chain_impl(arr, n, 0, nprod_total, NULL, 0); /* cu:612-616 */
    orc_mat **part = (orc_mat **)calloc((size_t)P, sizeof(orc_mat *));
    for (int r = 0; r < P; r++) {
        int s = r * per, e = (r == P - 1) ? n - 1 : (r + 1) * per - 1; /* cu:441-452 */
        part[r] = chain_impl(arr + s, e - s + 1, s, nprod_total, NULL, 0);
    }
    orc_mat *res = chain_impl(part, P, 0, nprod_total, NULL, 0); /* cu:569-571 */
    for (int r = 0; r < P; r++) orc_free(part[r]);
    free(part);
    return res;
}

void orc_prune_zero(orc_mat *m)
{
    size_t kk = (size_t)m->k * m->k;
    int64_t o = 0;
    for (int64_t p = 0; p < m->nnzb; p++) {
        int nz = 0;
        for (size_t e = 0; e < kk; e++) if (m->vals[(size_t)p * kk + e] > 0) { nz = 1; break; } /* cu:582 */
        if (!nz) continue;
        if (o != p) {
            m->brow[o] = m->brow[p]; m->bcol[o] = m->bcol[p];
            memmove(m->vals + (size_t)o * kk, m->vals + (size_t)p * kk, kk * sizeof(uint64_t));
        }
        o++;
    }
    m->nnzb = o;
}

/* ---- text format --------------------------------------------------------------------- */
static const char *skip_ws(const char *p, const char *e)
{
    while (p < e && (*p == ' ' || *p == '\n' || *p == '\r' || *p == '\t' || *p == '\v' || *p == '\f')) p++;
    return p;
}
static const char *rd_i64(const char *p, const char *e, int64_t *v, int *ok)
{
    p = skip_ws(p, e);
    int neg = 0; uint64_t x = 0; int nd = 0;
    if (p < e && (*p == '-' || *p == '+')) { neg = (*p == '-'); p++; }
    while (p < e && *p >= '0' && *p <= '9') { x = x * 10 + (uint64_t)(*p - '0'); p++; nd++; }
    if (!nd) *ok = 0;
    *v = neg ? -(int64_t)x : (int64_t)x;
    return p;
}
static const char *rd_u64(const char *p, const char *e, uint64_t *v, int *ok)
{
    p = skip_ws(p, e);
    uint64_t x = 0; int nd = 0;
    while (p < e && *p >= '0' && *p <= '9') { x = x * 10 + (uint64_t)(*p - '0'); p++; nd++; }
    if (!nd) *ok = 0;
    *v = x;
    return p;
}

orc_mat *orc_read(const char *path, int32_t k)
{
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    fseek(f, 0, SEEK_END);
    long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    char *buf = (char *)malloc((size_t)sz + 1);
    if (!buf || fread(buf, 1, (size_t)sz, f) != (size_t)sz) { fclose(f); free(buf); return NULL; }
    fclose(f);
    const char *p = buf, *e = buf + sz;
    int ok = 1;
    int64_t rows, cols, blocks;
    p = rd_i64(p, e, &rows, &ok); p = rd_i64(p, e, &cols, &ok); p = rd_i64(p, e, &blocks, &ok); /* cu:353,363 */
    if (!ok || blocks < 0) { free(buf); return NULL; }
    size_t kk = (size_t)k * k;
    int32_t *br = (int32_t *)malloc((size_t)(blocks + 1) * sizeof(int32_t));
    int32_t *bc = (int32_t *)malloc((size_t)(blocks + 1) * sizeof(int32_t));
    uint64_t *v = (uint64_t *)malloc((size_t)(blocks + 1) * kk * sizeof(uint64_t));
    for (int64_t b = 0; b < blocks && ok; b++) {
        int64_t r, c;
        p = rd_i64(p, e, &r, &ok); p = rd_i64(p, e, &c, &ok); /* cu:366 */
        br[b] = (int32_t)r; bc[b] = (int32_t)c;
        for (size_t t = 0; t < kk; t++) p = rd_u64(p, e, &v[(size_t)b * kk + t], &ok); /* cu:372-380 */
    }
    orc_mat *m = ok ? orc_from_coo((int32_t)rows, (int32_t)cols, k, blocks, br, bc, v) : NULL;
    free(br); free(bc); free(v); free(buf);
    return m;
}

int orc_write(const orc_mat *m, const char *path)
{
    FILE *f = fopen(path, "wb");
    if (!f) return -1;
    int k = m->k;
    fprintf(f, "%d %d\n%lld\n", m->rows, m->cols, (long long)m->nnzb); /* cu:596-597 */
    for (int64_t p = 0; p < m->nnzb; p++) {
        fprintf(f, "%d %d\n", m->brow[p], m->bcol[p]); /* cu:599 */
        const uint64_t *t = m->vals + (size_t)p * k * k;
        for (int r = 0; r < k; r++) {
            for (int c = 0; c < k; c++) /* cu:601-605: single spaces, none trailing */
                fprintf(f, c + 1 < k ? "%llu " : "%llu\n", (unsigned long long)t[r * k + c]);
        }
    }
    return fclose(f);
}
