// bsp_dist.cu -- multi-GPU layer of the C ABI: NCCL transport of matrices and the distributed
// chain product.
//
// Replaces the MPI half of the reference's main (cu:438-571): every rank reduces its contiguous
// sub-chain (cu:438-456), the partial products travel as header / keys / values messages to rank 0
// (cu:471-552), and rank 0 multiplies the P partials with helper2 (cu:569-571).  Here one process
// (or host thread) per GPU holds an NCCL communicator; matrices travel as CSR-of-blocks arrays with
// ncclSend / ncclRecv over NVLink, and the merge keeps the reference's association tree (the tree is
// part of the answer) but shards every merge product by block rows over the
// ranks of its subtree, so that no GPU idles where the chain is most expensive.
//
// Per tree level the ranks exchange ONE all-gather of fixed-size headers (tile counts and the split
// point of every held panel, computed on the device), wait once on the host to size their buffers,
// and then move every right-operand panel and every left half in ONE NCCL group -- received straight
// into the arrays of the operand they become part of, no staging copy, no sub-communicators.
//
// NCCL is bound at run time (dlopen of libnccl.so.2): the library loads without it and these entry
// points then fail with BSP_ERR_CUDA and a message.  Host threads of one process that drive different
// GPUs may each hold a communicator (a4 --devices D).
#include <dlfcn.h>
#include <nccl.h>

#include <mutex>
#include <new>
#include <vector>

#include "bsp_internal.h"

struct bsp_comm {
    ncclComm_t comm;
    int nranks, rank;
    bool owned;
    int64_t *d_hdr;  // device: [0, HW) this rank's header, [HW, HW*(1+nranks)) everybody's
    int64_t *h_hdr;  // pinned: HW * nranks
    int64_t bytes_sent;
};

namespace {

// ---- NCCL entry points, bound once ------------------------------------------------------------
struct nccl_api {
    ncclResult_t (*GetUniqueId)(ncclUniqueId *);
    ncclResult_t (*CommInitRank)(ncclComm_t *, int, ncclUniqueId, int);
    ncclResult_t (*CommDestroy)(ncclComm_t);
    ncclResult_t (*Send)(const void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*Recv)(void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*AllGather)(const void *, void *, size_t, ncclDataType_t, ncclComm_t, cudaStream_t);
    ncclResult_t (*Broadcast)(const void *, void *, size_t, ncclDataType_t, int, ncclComm_t, cudaStream_t);
    ncclResult_t (*GroupStart)();
    ncclResult_t (*GroupEnd)();
    const char *(*GetErrorString)(ncclResult_t);
    bool ok;
    char why[200];
};

nccl_api *nccl()
{
    static nccl_api api;
    static std::once_flag once;
    std::call_once(once, []() {
        api.ok = false;
        void *h = nullptr;
        for (const char *name : {"libnccl.so.2", "libnccl.so"}) {
            h = dlopen(name, RTLD_NOW | RTLD_GLOBAL);
            if (h) break;
        }
        if (!h) { snprintf(api.why, sizeof(api.why), "NCCL not found (dlopen libnccl.so.2: %s)", dlerror()); return; }
        bool all = true;
        auto bind = [&](auto &fn, const char *sym) {
            fn = reinterpret_cast<std::remove_reference_t<decltype(fn)>>(dlsym(h, sym));
            if (!fn) { all = false; snprintf(api.why, sizeof(api.why), "NCCL symbol %s missing", sym); }
        };
        bind(api.GetUniqueId, "ncclGetUniqueId");
        bind(api.CommInitRank, "ncclCommInitRank");
        bind(api.CommDestroy, "ncclCommDestroy");
        bind(api.Send, "ncclSend");
        bind(api.Recv, "ncclRecv");
        bind(api.AllGather, "ncclAllGather");
        bind(api.Broadcast, "ncclBroadcast");
        bind(api.GroupStart, "ncclGroupStart");
        bind(api.GroupEnd, "ncclGroupEnd");
        bind(api.GetErrorString, "ncclGetErrorString");
        api.ok = all;
    });
    return &api;
}

#define BSP_NCCL(ctx, call)                                                                                \
    do {                                                                                                   \
        ncclResult_t r_ = (call);                                                                          \
        if (r_ != ncclSuccess) {                                                                           \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d: %s failed: %s", __FILE__, __LINE__, #call,    \
                     nccl()->GetErrorString(r_));                                                          \
            return BSP_ERR_CUDA;                                                                           \
        }                                                                                                  \
    } while (0)

// ---- headers ----------------------------------------------------------------------------------
// What every rank tells the others before a level: the shape of the panel it holds and where it
// would cut it in two halves of equal tile count (rows [0, mid) stay, rows [mid, nbr) travel).
enum { H_VALID = 0, H_ROWS, H_COLS, H_K, H_ESIZE, H_NBR, H_NBC, H_NNZB, H_MID, H_S0, H_NLOCAL, HW = 12 };

struct hdr_fields { int64_t v[HW]; };

__global__ void header_kernel(const __grid_constant__ hdr_fields f, const int32_t *__restrict__ rowptr, int64_t *__restrict__ dst)
{
    if (threadIdx.x != 0) return;
    for (int i = 0; i < HW; i++) dst[i] = f.v[i];
    if (rowptr) {
        // smallest m >= 1 with rowptr[m] >= ceil(nnzb / 2): the first m rows hold half of the tiles
        const int nbr = (int)f.v[H_NBR];
        const int64_t half = (f.v[H_NNZB] + 1) / 2;
        int lo = 1, hi = nbr;
        if (f.v[H_NNZB] == 0 || nbr == 0) { lo = 0; }
        else while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (rowptr[mid] >= half) hi = mid; else lo = mid + 1;
        }
        dst[H_MID] = lo;
        dst[H_S0] = rowptr[lo];
    }
}

constexpr int SUM_MAX = 16;
struct rowptr_table { const int32_t *p[SUM_MAX]; int n; };

// dst[i] (+)= sum of the panels' row pointers: panels of one product over disjoint row ranges, each with a
// full-length row pointer, add up to the row pointer of the whole (same rule as bsp_mat_concat_panels)
__global__ void sum_rowptr_kernel(const __grid_constant__ rowptr_table tb, int n, int accumulate, int32_t *__restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    int32_t s = accumulate ? dst[i] : 0;
    for (int q = 0; q < tb.n; q++) s += tb.p[q][i];
    dst[i] = s;
}

// row pointer of "rows [mid, nbr) only" from the row pointer of the whole panel: tiles below s0 are gone
__global__ void clamp_rowptr_kernel(int32_t *__restrict__ rp, int n, int32_t s0)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) rp[i] = (rp[i] > s0 ? rp[i] : s0) - s0;
}

// tile products of every block row of A*B (warp per row): the balance measure of bsp_panel_bounds
__global__ void row_work_kernel(int a_nbr, const int32_t *__restrict__ a_rowptr, const int32_t *__restrict__ a_colidx,
                                int b_nbr, const int32_t *__restrict__ b_rowptr, int long_row, int long_pct,
                                int32_t *__restrict__ work)
{
    const int lane = threadIdx.x & 31;
    const int row = (int)(((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    if (row > a_nbr) return;
    long long w = 0;
    if (row < a_nbr) {
        for (int p = a_rowptr[row] + lane; p < a_rowptr[row + 1]; p += 32) {
            const int j = a_colidx[p];
            if (j < b_nbr) w += b_rowptr[j + 1] - b_rowptr[j];
        }
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) w += __shfl_xor_sync(0xffffffffu, w, d);
    if (w > long_row) w += w * long_pct / 100; // long rows cost more per tile product in the symbolic pass
    if (lane == 0) work[row] = row < a_nbr ? (int32_t)(w > 0x7fffffffLL ? 0x7fffffffLL : w) : 0;
}

inline size_t tile_bytes(int64_t k, int64_t esize) { return (size_t)(k * k * esize); }

int fill_header(bsp_ctx *ctx, bsp_comm *c, const bsp_mat *held, int64_t n_local)
{
    hdr_fields f;
    memset(&f, 0, sizeof(f));
    f.v[H_NLOCAL] = n_local;
    if (held) {
        f.v[H_VALID] = 1; f.v[H_ROWS] = held->rows; f.v[H_COLS] = held->cols; f.v[H_K] = held->k; f.v[H_ESIZE] = held->esize;
        f.v[H_NBR] = held->nbr; f.v[H_NBC] = held->nbc; f.v[H_NNZB] = held->nnzb;
    }
    header_kernel<<<1, 32, 0, ctx->stream>>>(f, held ? held->rowptr : nullptr, c->d_hdr);
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}

// one all-gather of everybody's header and the one host wait of a level
int exchange_headers(bsp_ctx *ctx, bsp_comm *c, const bsp_mat *held, int64_t n_local)
{
    BSP_TRY(fill_header(ctx, c, held, n_local));
    BSP_NCCL(ctx, nccl()->AllGather(c->d_hdr, c->d_hdr + HW, HW, ncclInt64, c->comm, ctx->stream));
    BSP_CUDA(ctx, cudaMemcpyAsync(c->h_hdr, c->d_hdr + HW, (size_t)HW * c->nranks * sizeof(int64_t), cudaMemcpyDeviceToHost,
                                  ctx->stream));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BSP_OK;
}

inline const int64_t *hdr_of(const bsp_comm *c, int r) { return c->h_hdr + (size_t)r * HW; }

int new_mat(bsp_ctx *ctx, const int64_t *h, int64_t nnzb, bsp_mat **out)
{
    *out = nullptr;
    bsp_mat *m = new (std::nothrow) bsp_mat();
    if (!m) BSP_FAIL(ctx, BSP_ERR_NOMEM, "host allocation failed");
    m->rows = (int32_t)h[H_ROWS]; m->cols = (int32_t)h[H_COLS]; m->k = (int32_t)h[H_K]; m->esize = (int32_t)h[H_ESIZE];
    m->nbr = (int32_t)h[H_NBR]; m->nbc = (int32_t)h[H_NBC]; m->nnzb = nnzb; m->device = ctx->device;
    m->rowptr = nullptr; m->colidx = nullptr; m->vals = nullptr;
    *out = m;
    BSP_TRY(bsp_alloc(ctx, &m->rowptr, (size_t)m->nbr + 1));
    BSP_TRY(bsp_alloc(ctx, &m->colidx, (size_t)nnzb));
    BSP_TRY(bsp_alloc(ctx, (char **)&m->vals, (size_t)nnzb * tile_bytes(m->k, m->esize)));
    return BSP_OK;
}

// send / receive the three arrays of (a slice of) a matrix; to be called inside an NCCL group
int post_send(bsp_ctx *ctx, bsp_comm *c, const bsp_mat *m, int64_t t0, int64_t n, int peer)
{
    const size_t tb = tile_bytes(m->k, m->esize);
    BSP_NCCL(ctx, nccl()->Send(m->rowptr, (size_t)m->nbr + 1, ncclInt32, peer, c->comm, ctx->stream));
    c->bytes_sent += 4 * ((int64_t)m->nbr + 1);
    if (n > 0) {
        BSP_NCCL(ctx, nccl()->Send(m->colidx + t0, (size_t)n, ncclInt32, peer, c->comm, ctx->stream));
        BSP_NCCL(ctx, nccl()->Send((const char *)m->vals + (size_t)t0 * tb, (size_t)n * tb, ncclChar, peer, c->comm, ctx->stream));
        c->bytes_sent += n * (int64_t)(4 + tb);
    }
    return BSP_OK;
}
int post_recv(bsp_ctx *ctx, bsp_comm *c, int32_t *rowptr, int nbr, int32_t *colidx, char *vals, size_t tb, int64_t n, int peer)
{
    BSP_NCCL(ctx, nccl()->Recv(rowptr, (size_t)nbr + 1, ncclInt32, peer, c->comm, ctx->stream));
    if (n > 0) {
        BSP_NCCL(ctx, nccl()->Recv(colidx, (size_t)n, ncclInt32, peer, c->comm, ctx->stream));
        BSP_NCCL(ctx, nccl()->Recv(vals, (size_t)n * tb, ncclChar, peer, c->comm, ctx->stream));
    }
    return BSP_OK;
}

int sum_rowptrs(bsp_ctx *ctx, const std::vector<const int32_t *> &src, int n, int32_t *dst)
{
    for (size_t i0 = 0; i0 < src.size(); i0 += SUM_MAX) {
        rowptr_table tb;
        tb.n = (int)(src.size() - i0 < SUM_MAX ? src.size() - i0 : SUM_MAX);
        for (int i = 0; i < tb.n; i++) tb.p[i] = src[i0 + (size_t)i];
        sum_rowptr_kernel<<<(n + 255) / 256, 256, 0, ctx->stream>>>(tb, n, i0 > 0 ? 1 : 0, dst);
        BSP_LAUNCH_CHECK(ctx);
    }
    return BSP_OK;
}

void add_stats(bsp_stats *dst, const bsp_stats &s)
{
    dst->nnzb_a += s.nnzb_a; dst->nnzb_b += s.nnzb_b; dst->nnzb_c += s.nnzb_c; dst->nprod += s.nprod;
    dst->ms_symbolic += s.ms_symbolic; dst->ms_numeric += s.ms_numeric; dst->ms_total += s.ms_total;
    dst->launches += s.launches; dst->exact_refolds += s.exact_refolds;
    if (s.numeric_path) dst->numeric_path = s.numeric_path;
}

// Panels of one matrix held by the ranks in `order` (ascending rows) -> one matrix on this rank, in three
// steps so that several assemblies and other messages can share ONE NCCL group: prepare (sizes from the
// headers, allocation), post (sends of this rank's panel to `send_to`, receives of the others' panels
// straight into the result's arrays; inside the caller's group), finish (own panel copied in, row
// pointers summed; after the group has ended, i.e. after the transfers are on the stream).
struct assembly {
    std::vector<int> order;
    const bsp_mat *mine = nullptr; // this rank's panel, or null when it holds none
    bsp_mat *R = nullptr;          // the result (null on ranks that only send)
    std::vector<int32_t *> tmp;    // received row pointers, one per foreign panel
    int nbr = 0;
    size_t tb = 0;
};

void asm_release(bsp_ctx *ctx, assembly *a)
{
    for (auto *p : a->tmp) bsp_release(ctx, p);
    a->tmp.clear();
    bsp_mat_free(ctx, a->R);
    a->R = nullptr;
}

int asm_prepare(bsp_ctx *ctx, bsp_comm *c, const std::vector<int> &order, const bsp_mat *mine, bool want, assembly *a)
{
    a->order = order; a->mine = mine; a->R = nullptr; a->tmp.assign(order.size(), nullptr);
    const int64_t *h0 = hdr_of(c, order[0]);
    a->tb = tile_bytes(h0[H_K], h0[H_ESIZE]);
    a->nbr = (int)h0[H_NBR];
    if (!want) return BSP_OK;
    int64_t total = 0;
    for (int q : order) total += hdr_of(c, q)[H_NNZB];
    if (total > 0x7fffffffLL) BSP_FAIL(ctx, BSP_ERR_LIMIT, "more than 2^31-1 tiles");
    int64_t hh[HW];
    memcpy(hh, h0, sizeof(hh));
    for (int q : order) if (hdr_of(c, q)[H_NBC] > hh[H_NBC]) hh[H_NBC] = hdr_of(c, q)[H_NBC];
    int rc = new_mat(ctx, hh, total, &a->R);
    for (size_t i = 0; i < order.size() && rc == BSP_OK; i++)
        if (order[i] != c->rank) rc = bsp_alloc(ctx, &a->tmp[i], (size_t)a->nbr + 1);
    if (rc != BSP_OK) asm_release(ctx, a);
    return rc;
}

int asm_post(bsp_ctx *ctx, bsp_comm *c, assembly *a, const std::vector<int> &send_to)
{
    if (a->mine)
        for (int dst : send_to)
            if (dst != c->rank) BSP_TRY(post_send(ctx, c, a->mine, 0, a->mine->nnzb, dst));
    if (a->R) {
        int64_t off = 0;
        for (size_t i = 0; i < a->order.size(); i++) {
            const int q = a->order[i];
            const int64_t n = hdr_of(c, q)[H_NNZB];
            if (q != c->rank)
                BSP_TRY(post_recv(ctx, c, a->tmp[i], a->nbr, a->R->colidx + off, (char *)a->R->vals + (size_t)off * a->tb, a->tb, n, q));
            off += n;
        }
    }
    return BSP_OK;
}

// hands the result over (*out) and releases the temporaries
int asm_finish(bsp_ctx *ctx, bsp_comm *c, assembly *a, bsp_mat **out)
{
    int rc = BSP_OK;
    if (a->R) {
        std::vector<const int32_t *> rps;
        int64_t off = 0;
        for (size_t i = 0; i < a->order.size() && rc == BSP_OK; i++) {
            const int q = a->order[i];
            const int64_t n = hdr_of(c, q)[H_NNZB];
            if (q == c->rank) {
                rps.push_back(a->mine->rowptr);
                if (n > 0 &&
                    (cudaMemcpyAsync(a->R->colidx + off, a->mine->colidx, (size_t)n * 4, cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess ||
                     cudaMemcpyAsync((char *)a->R->vals + (size_t)off * a->tb, a->mine->vals, (size_t)n * a->tb, cudaMemcpyDeviceToDevice, ctx->stream) != cudaSuccess)) {
                    snprintf(ctx->err, sizeof(ctx->err), "panel copy failed: %s", cudaGetErrorString(cudaGetLastError()));
                    rc = BSP_ERR_CUDA;
                }
            } else rps.push_back(a->tmp[i]);
            off += n;
        }
        if (rc == BSP_OK) rc = sum_rowptrs(ctx, rps, a->nbr + 1, a->R->rowptr);
    }
    if (rc == BSP_OK && out) { *out = a->R; a->R = nullptr; }
    asm_release(ctx, a);
    return rc;
}

// prepare + one group + finish
int assemble_from(bsp_ctx *ctx, bsp_comm *c, const std::vector<int> &order, const bsp_mat *mine,
                  const std::vector<int> &send_to, bsp_mat **out)
{
    assembly a;
    BSP_TRY(asm_prepare(ctx, c, order, mine, out != nullptr, &a));
    ncclResult_t g = nccl()->GroupStart();
    int rc = BSP_OK;
    if (g != ncclSuccess) { snprintf(ctx->err, sizeof(ctx->err), "ncclGroupStart failed: %s", nccl()->GetErrorString(g)); rc = BSP_ERR_CUDA; }
    if (rc == BSP_OK) {
        rc = asm_post(ctx, c, &a, send_to);
        g = nccl()->GroupEnd();
        if (rc == BSP_OK && g != ncclSuccess) { snprintf(ctx->err, sizeof(ctx->err), "ncclGroupEnd failed: %s", nccl()->GetErrorString(g)); rc = BSP_ERR_CUDA; }
    }
    if (rc != BSP_OK) { asm_release(ctx, &a); return rc; }
    return asm_finish(ctx, c, &a, out);
}

} // namespace

extern "C" {

int bsp_comm_unique_id(void *id, size_t cap)
{
    if (!id || cap < BSP_COMM_ID_BYTES || !nccl()->ok) return BSP_ERR_CUDA;
    static_assert(sizeof(ncclUniqueId) <= BSP_COMM_ID_BYTES, "unique id size");
    ncclUniqueId u;
    if (nccl()->GetUniqueId(&u) != ncclSuccess) return BSP_ERR_CUDA;
    memset(id, 0, cap);
    memcpy(id, &u, sizeof(u));
    return BSP_OK;
}

static int comm_finish(bsp_ctx *ctx, bsp_comm *c)
{
    BSP_CUDA(ctx, cudaMalloc(&c->d_hdr, (size_t)HW * (1 + c->nranks) * sizeof(int64_t)));
    BSP_CUDA(ctx, cudaMallocHost(&c->h_hdr, (size_t)HW * c->nranks * sizeof(int64_t)));
    return BSP_OK;
}

int bsp_comm_create(bsp_ctx *ctx, const void *id, int nranks, int rank, bsp_comm **out)
{
    if (!ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (!id || nranks < 1 || rank < 0 || rank >= nranks) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_comm_create: bad arguments");
    if (!nccl()->ok) BSP_FAIL(ctx, BSP_ERR_CUDA, "%s", nccl()->why);
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    bsp_comm *c = new (std::nothrow) bsp_comm();
    if (!c) BSP_FAIL(ctx, BSP_ERR_NOMEM, "host allocation failed");
    memset(c, 0, sizeof(*c));
    c->nranks = nranks; c->rank = rank; c->owned = true;
    ncclUniqueId u;
    memcpy(&u, id, sizeof(u));
    ncclResult_t r = nccl()->CommInitRank(&c->comm, nranks, u, rank);
    if (r != ncclSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "ncclCommInitRank failed: %s", nccl()->GetErrorString(r));
        delete c;
        return BSP_ERR_CUDA;
    }
    *out = c;
    return comm_finish(ctx, c);
}

int bsp_comm_adopt(bsp_ctx *ctx, void *nccl_comm, int nranks, int rank, bsp_comm **out)
{
    if (!ctx || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (!nccl_comm || nranks < 1 || rank < 0 || rank >= nranks) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_comm_adopt: bad arguments");
    if (!nccl()->ok) BSP_FAIL(ctx, BSP_ERR_CUDA, "%s", nccl()->why);
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    bsp_comm *c = new (std::nothrow) bsp_comm();
    if (!c) BSP_FAIL(ctx, BSP_ERR_NOMEM, "host allocation failed");
    memset(c, 0, sizeof(*c));
    c->comm = (ncclComm_t)nccl_comm; c->nranks = nranks; c->rank = rank; c->owned = false;
    *out = c;
    return comm_finish(ctx, c);
}

void bsp_comm_destroy(bsp_ctx *ctx, bsp_comm *c)
{
    if (!c) return;
    if (ctx) { cudaSetDevice(ctx->device); cudaStreamSynchronize(ctx->stream); }
    if (c->d_hdr) cudaFree(c->d_hdr);
    if (c->h_hdr) cudaFreeHost(c->h_hdr);
    if (c->owned && c->comm && nccl()->ok) nccl()->CommDestroy(c->comm);
    delete c;
}

int bsp_comm_info(const bsp_comm *c, int *nranks, int *rank, int64_t *bytes_sent)
{
    if (!c) return BSP_ERR_ARG;
    if (nranks) *nranks = c->nranks;
    if (rank) *rank = c->rank;
    if (bytes_sent) *bytes_sent = c->bytes_sent;
    return BSP_OK;
}

// ---- point to point and broadcast of one matrix ---------------------------------------------------

int bsp_mat_send(bsp_ctx *ctx, bsp_comm *c, const bsp_mat *m, int peer)
{
    if (!ctx || !c) return BSP_ERR_ARG;
    if (!m || !m->rowptr || peer < 0 || peer >= c->nranks || peer == c->rank) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_send: bad arguments");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    BSP_TRY(fill_header(ctx, c, m, 0));
    BSP_NCCL(ctx, nccl()->Send(c->d_hdr, HW, ncclInt64, peer, c->comm, ctx->stream));
    BSP_NCCL(ctx, nccl()->GroupStart());
    int rc = post_send(ctx, c, m, 0, m->nnzb, peer);
    ncclResult_t g = nccl()->GroupEnd();
    if (rc == BSP_OK && g != ncclSuccess) BSP_FAIL(ctx, BSP_ERR_CUDA, "ncclGroupEnd failed: %s", nccl()->GetErrorString(g));
    if (rc == BSP_OK) BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // m may be freed or changed on return
    return rc;
}

int bsp_mat_recv(bsp_ctx *ctx, bsp_comm *c, int peer, bsp_mat **out)
{
    if (!ctx || !c || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (peer < 0 || peer >= c->nranks || peer == c->rank) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_recv: bad peer");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    BSP_NCCL(ctx, nccl()->Recv(c->d_hdr, HW, ncclInt64, peer, c->comm, ctx->stream));
    BSP_CUDA(ctx, cudaMemcpyAsync(c->h_hdr, c->d_hdr, HW * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // the body is sized from the header (cu:513-520 does the same with MPI_Recv)
    int64_t h[HW];
    memcpy(h, c->h_hdr, sizeof(h));
    bsp_mat *m = nullptr;
    int rc = new_mat(ctx, h, h[H_NNZB], &m);
    if (rc == BSP_OK) {
        ncclResult_t g = nccl()->GroupStart();
        if (g == ncclSuccess) {
            rc = post_recv(ctx, c, m->rowptr, m->nbr, m->colidx, (char *)m->vals, tile_bytes(m->k, m->esize), m->nnzb, peer);
            g = nccl()->GroupEnd();
        }
        if (rc == BSP_OK && g != ncclSuccess) {
            snprintf(ctx->err, sizeof(ctx->err), "NCCL group failed: %s", nccl()->GetErrorString(g));
            rc = BSP_ERR_CUDA;
        }
    }
    if (rc != BSP_OK) { bsp_mat_free(ctx, m); return rc; }
    *out = m;
    return BSP_OK;
}

int bsp_mat_bcast(bsp_ctx *ctx, bsp_comm *c, int root, const bsp_mat *in, bsp_mat **out)
{
    if (!ctx || !c) return BSP_ERR_ARG;
    if (out) *out = nullptr;
    if (root < 0 || root >= c->nranks) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_bcast: bad root");
    const bool is_root = c->rank == root;
    if (is_root && (!in || !in->rowptr)) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_bcast: the root passes the matrix");
    if (!is_root && !out) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_bcast: nowhere to put the matrix");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    if (is_root) BSP_TRY(fill_header(ctx, c, in, 0));
    BSP_NCCL(ctx, nccl()->Broadcast(c->d_hdr, c->d_hdr, HW, ncclInt64, root, c->comm, ctx->stream));
    bsp_mat *m = nullptr;
    if (!is_root) {
        BSP_CUDA(ctx, cudaMemcpyAsync(c->h_hdr, c->d_hdr, HW * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
        BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        int64_t h[HW];
        memcpy(h, c->h_hdr, sizeof(h));
        int rc = new_mat(ctx, h, h[H_NNZB], &m);
        if (rc != BSP_OK) { bsp_mat_free(ctx, m); return rc; }
    }
    const bsp_mat *w = is_root ? in : m;
    const size_t tb = tile_bytes(w->k, w->esize);
    int rc = BSP_OK;
    ncclResult_t g = nccl()->GroupStart();
    if (g == ncclSuccess) g = nccl()->Broadcast(w->rowptr, w->rowptr, (size_t)w->nbr + 1, ncclInt32, root, c->comm, ctx->stream);
    if (g == ncclSuccess && w->nnzb > 0) {
        g = nccl()->Broadcast(w->colidx, w->colidx, (size_t)w->nnzb, ncclInt32, root, c->comm, ctx->stream);
        if (g == ncclSuccess) g = nccl()->Broadcast(w->vals, w->vals, (size_t)w->nnzb * tb, ncclChar, root, c->comm, ctx->stream);
    }
    ncclResult_t g2 = nccl()->GroupEnd();
    if (g != ncclSuccess || g2 != ncclSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "ncclBroadcast failed: %s", nccl()->GetErrorString(g != ncclSuccess ? g : g2));
        rc = BSP_ERR_CUDA;
    }
    if (is_root) c->bytes_sent += 4 * ((int64_t)w->nbr + 1) + w->nnzb * (int64_t)(4 + tb);
    if (rc == BSP_OK) BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (rc != BSP_OK) { bsp_mat_free(ctx, m); return rc; }
    if (!is_root) *out = m;
    return BSP_OK;
}

// ---- gather of row panels ----------------------------------------------------------------------

int bsp_gather_panels(bsp_ctx *ctx, bsp_comm *c, const bsp_mat *panel, int root, bsp_mat **out)
{
    if (!ctx || !c) return BSP_ERR_ARG;
    if (out) *out = nullptr;
    if (root < 0 || root >= c->nranks || !panel || !panel->rowptr) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_gather_panels: bad arguments");
    if (c->rank == root && !out) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_gather_panels: the root needs an output");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    BSP_TRY(exchange_headers(ctx, c, panel, 1));
    std::vector<int> order;
    for (int r = 0; r < c->nranks; r++) order.push_back(r);
    int rc = assemble_from(ctx, c, order, panel, std::vector<int>{root}, c->rank == root ? out : nullptr);
    if (rc == BSP_OK) BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return rc;
}

// ---- balanced row panels of one product ---------------------------------------------------------------

int bsp_panel_bounds(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t nparts, int32_t *bounds)
{
    if (!ctx || !bounds) return BSP_ERR_ARG;
    if (!A || !A->rowptr || !B || !B->rowptr || nparts < 1) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_panel_bounds: bad arguments");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    const int nbr = A->nbr;
    bsp_scratch sc(ctx);
    int32_t *work = nullptr;
    int64_t *cum = nullptr;
    BSP_TRY(sc.alloc(&work, (size_t)nbr + 1));
    BSP_TRY(sc.alloc(&cum, (size_t)nbr + 1));
    // Rows above the shared-memory classes of the symbolic pass (one CTA walks the whole row) are charged 40 % more per
    // tile product (BSP_PANEL_LONG_W overrides).  Measured with tools/rmat_panels_1gpu.py on the R-MAT pattern, max/mean of
    // the panel times at weights 0 / 20 / 40 / 60 %: 2 panels 1.014 / 1.012 / 1.023 / -, 4 panels 1.067 / 1.030 / 1.008 /
    // 1.017, 8 panels 1.158 / 1.132 / 1.120 / 1.111.
    int long_pct = 40;
    if (const char *e = getenv("BSP_PANEL_LONG_W")) { const int v = atoi(e); if (v >= 0 && v <= 400) long_pct = v; }
    row_work_kernel<<<(unsigned)(((int64_t)(nbr + 1) * 32 + 255) / 256), 256, 0, ctx->stream>>>(nbr, A->rowptr, A->colidx, B->nbr,
                                                                                                B->rowptr, 12288, long_pct, work);
    BSP_LAUNCH_CHECK(ctx);
    BSP_TRY(bsp_scan_i32_to_i64(ctx, work, cum, (int64_t)nbr + 1)); // cum[r] = work of the rows below r
    std::vector<int64_t> h((size_t)nbr + 1);
    BSP_CUDA(ctx, cudaMemcpyAsync(h.data(), cum, ((size_t)nbr + 1) * sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t total = h[(size_t)nbr];
    bounds[0] = 0;
    for (int p = 1; p < nparts; p++) {
        // first row r with cum[r] >= total*p/nparts
        const int64_t want = (int64_t)((__int128)total * p / nparts);
        int lo = 0, hi = nbr;
        while (lo < hi) {
            const int mid = (lo + hi) >> 1;
            if (h[(size_t)mid] >= want) hi = mid; else lo = mid + 1;
        }
        bounds[p] = lo < bounds[p - 1] ? bounds[p - 1] : lo;
    }
    bounds[nparts] = nbr;
    return BSP_OK;
}

// ---- the chain over several GPUs ------------------------------------------------------------------

int bsp_chain_distributed(bsp_ctx *ctx, bsp_comm *c, bsp_mat *const *mats, int32_t n_local, int32_t log_base,
                          bsp_mat **out, bsp_stats *stats, char *log, size_t log_cap)
{
    if (!ctx || !c || !out) return BSP_ERR_ARG;
    *out = nullptr;
    if (n_local < 0 || (n_local > 0 && !mats)) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_chain_distributed: bad arguments");
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    bsp_stats tot;
    memset(&tot, 0, sizeof(tot));
    if (log && log_cap) log[0] = 0;

    // 1. this rank's sub-chain by the pairwise tree (cu:457-467)
    bsp_mat *held = nullptr;
    if (n_local > 0) {
        bsp_stats st;
        BSP_TRY(bsp_chain(ctx, mats, n_local, &held, &st, log, log_cap, log_base));
        add_stats(&tot, st);
    }
    auto fail = [&](int rc) { bsp_mat_free(ctx, held); return rc; };
    if (c->nranks == 1) {
        *out = held;
        if (stats) *stats = tot;
        return BSP_OK;
    }

    // 2. who takes part: the reference gives every rank N/P matrices, or all of them to rank 0 when N < P
    int rc = exchange_headers(ctx, c, held, n_local);
    if (rc != BSP_OK) return fail(rc);
    int nparts = 0;
    while (nparts < c->nranks && hdr_of(c, nparts)[H_NLOCAL] > 0) nparts++;
    for (int r = nparts; r < c->nranks; r++)
        if (hdr_of(c, r)[H_NLOCAL] > 0) { snprintf(ctx->err, sizeof(ctx->err), "bsp_chain_distributed: working ranks must be 0..n-1"); return fail(BSP_ERR_ARG); }
    if (nparts == 0) { snprintf(ctx->err, sizeof(ctx->err), "bsp_chain_distributed: empty chain"); return fail(BSP_ERR_ARG); }
    const int me = c->rank;

    // 3. step-doubling tree with the association of helper2 over the partials (cu:569-571):
    //    at step s the groups led by g and g+s (g a multiple of 2s) merge, left = g's product.
    //    order[g] = ranks that hold the row panels of g's product, ascending rows.
    std::vector<std::vector<int>> order((size_t)nparts);
    for (int r = 0; r < nparts; r++) order[(size_t)r] = {r};
    bool first = true;
    for (int s = 1; s < nparts; s *= 2) {
        if (!first) { // the headers of step 1 came with the census above
            rc = exchange_headers(ctx, c, held, n_local);
            if (rc != BSP_OK) return fail(rc);
        }
        first = false;
        // my group at this step
        int g = -1;
        for (int gg = 0; gg + s < nparts; gg += 2 * s)
            if (me >= gg && me < gg + 2 * s && me < nparts) g = gg;
        std::vector<std::vector<int>> next = order;
        for (int gg = 0; gg + s < nparts; gg += 2 * s) {
            const int last = (gg + 2 * s < nparts ? gg + 2 * s : nparts) - 1;
            std::vector<int> merged;
            for (int p : order[(size_t)gg]) { merged.push_back(p); if (p + s <= last) merged.push_back(p + s); }
            next[(size_t)gg] = merged;
            next[(size_t)(gg + s)].clear();
        }
        if (g >= 0) {
            const int last = (g + 2 * s < nparts ? g + 2 * s : nparts) - 1;
            const std::vector<int> &left = order[(size_t)g], &right = order[(size_t)(g + s)];
            bool in_left = false, in_right = false;
            for (int p : left) in_left = in_left || p == me;
            for (int p : right) in_right = in_right || p == me;
            const bool has_partner = in_left && me + s <= last;
            const bool gets_half = !in_left; // every rank of [g+s, last] has a partner in the left group
            std::vector<int> members;
            for (int r = g; r <= last; r++) members.push_back(r);

            // the left half this rank receives (sized from its partner's header)
            bsp_mat *half = nullptr;
            int64_t half_s0 = 0;
            if (gets_half) {
                const int64_t *hp = hdr_of(c, me - s);
                half_s0 = hp[H_S0];
                rc = new_mat(ctx, hp, hp[H_NNZB] - half_s0, &half);
                if (rc != BSP_OK) { bsp_mat_free(ctx, half); return fail(rc); }
            }
            // everything of this level in flight at once: right panels to every member, left halves to the partners
            bsp_mat *R = nullptr;
            {
                assembly ra;
                rc = asm_prepare(ctx, c, right, in_right ? held : nullptr, true, &ra);
                if (rc != BSP_OK) { bsp_mat_free(ctx, half); return fail(rc); }
                ncclResult_t gs = nccl()->GroupStart();
                if (gs != ncclSuccess) { snprintf(ctx->err, sizeof(ctx->err), "ncclGroupStart failed: %s", nccl()->GetErrorString(gs)); rc = BSP_ERR_CUDA; }
                if (rc == BSP_OK) {
                    rc = asm_post(ctx, c, &ra, members);
                    if (rc == BSP_OK && has_partner) {
                        const int64_t s0 = hdr_of(c, me)[H_S0];
                        rc = post_send(ctx, c, held, s0, held->nnzb - s0, me + s);
                    }
                    if (rc == BSP_OK && gets_half)
                        rc = post_recv(ctx, c, half->rowptr, half->nbr, half->colidx, (char *)half->vals,
                                       tile_bytes(half->k, half->esize), half->nnzb, me - s);
                    gs = nccl()->GroupEnd();
                    if (rc == BSP_OK && gs != ncclSuccess) { snprintf(ctx->err, sizeof(ctx->err), "ncclGroupEnd failed: %s", nccl()->GetErrorString(gs)); rc = BSP_ERR_CUDA; }
                }
                if (rc == BSP_OK) rc = asm_finish(ctx, c, &ra, &R);
                else asm_release(ctx, &ra);
            }
            if (rc == BSP_OK && gets_half) {
                clamp_rowptr_kernel<<<(half->nbr + 1 + 255) / 256, 256, 0, ctx->stream>>>(half->rowptr, half->nbr + 1, (int32_t)half_s0);
                ctx->launches++;
                if (cudaGetLastError() != cudaSuccess) { snprintf(ctx->err, sizeof(ctx->err), "clamp_rowptr_kernel launch failed"); rc = BSP_ERR_CUDA; }
            }
            // this rank's panel of the product
            bsp_mat *prod = nullptr;
            if (rc == BSP_OK) {
                bsp_stats st;
                if (gets_half) rc = bsp_spgemm(ctx, half, R, &prod, &st);
                else rc = bsp_spgemm_panel(ctx, held, R, 0, has_partner ? (int32_t)hdr_of(c, me)[H_MID] : INT32_MAX, &prod, &st);
                if (rc == BSP_OK) add_stats(&tot, st);
            }
            bsp_mat_free(ctx, R);
            bsp_mat_free(ctx, half);
            if (rc != BSP_OK) { bsp_mat_free(ctx, prod); return fail(rc); }
            bsp_mat_free(ctx, held);
            held = prod;
        }
        order.swap(next);
    }

    // 4. the panels of the final product -> rank 0
    const std::vector<int> &fin = order[0];
    if (fin.size() > 1) {
        rc = exchange_headers(ctx, c, held, n_local);
        if (rc != BSP_OK) return fail(rc);
        bool mine_in = false;
        for (int p : fin) mine_in = mine_in || p == me;
        bsp_mat *full = nullptr;
        rc = assemble_from(ctx, c, fin, mine_in ? held : nullptr, std::vector<int>{0}, me == 0 ? &full : nullptr);
        if (rc == BSP_OK && cudaStreamSynchronize(ctx->stream) != cudaSuccess) { snprintf(ctx->err, sizeof(ctx->err), "final gather failed"); rc = BSP_ERR_CUDA; }
        if (rc != BSP_OK) { bsp_mat_free(ctx, full); return fail(rc); }
        bsp_mat_free(ctx, held);
        held = full;
    } else if (me != 0) {
        bsp_mat_free(ctx, held);
        held = nullptr;
    }
    *out = me == 0 ? held : nullptr;
    if (stats) *stats = tot;
    return BSP_OK;
}

} // extern "C"
