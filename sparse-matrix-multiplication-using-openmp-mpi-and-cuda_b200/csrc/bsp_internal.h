// bsp_internal.h -- shared declarations of libbspgemm (not part of the public ABI).
//
// Data layout in HBM: a matrix is CSR-of-blocks,
//   rowptr [nbr+1] int32   block-row r owns tiles rowptr[r] .. rowptr[r+1]-1
//   colidx [nnzb]  int32   block-column of each tile, ascending inside a row
//   vals   [nnzb*k*k] u64  tiles back to back, each row-major (512 B for k=8, 128 B for k=4)
// which is the reference's ordered std::map<(brow,bcol), tile> (cu:26-32) flattened in
// iteration order.  nbr = 1 + largest block-row coordinate seen (block coordinates are
// opaque ints in the reference, cu:150-151; they are never scaled by k).
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include "../../include/bspgemm.h"

struct bsp_ctx {
    int device;
    cudaStream_t stream;
    cudaStream_t copy_stream; // second stream: device->host copies of bsp_multiply_host overlap the next panel
    bool owns_stream;
    int sm_count;
    int64_t launches;
    cudaEvent_t ev[4];
    void *pinned;        // small pinned staging area for scalar read-backs
    void *pinned_dev;    // the same area as the device sees it (mapped): bsp_publish stores into it from a kernel
    size_t pinned_bytes;
    char *io_buf[2];     // pinned staging pieces of the binary reader/writer, allocated on first use
    cudaEvent_t io_ev[2];
    void *d_counters;    // 256 B of device counters (exact-refold count of the numeric pass)
    void *f32_pack[2];   // float32 tensor-core path: operands repacked into the instruction's layout (grow-only, bsp_numeric_f32_tc.cu)
    size_t f32_pack_bytes[2];
    bool force_generic;  // BSP_FORCE_GENERIC=1: always use the any-k exact numeric kernel (tests)
    int f32_path;        // float32 products: 0 = auto (tensor cores for k = 32 / 64), 1 = SIMT, 2 = tensor cores
    int64_t pair_budget; // tile products whose pair lists one symbolic pass may hold (BSP_PAIR_BUDGET); more -> row panels
    bool no_batch;       // BSP_NO_BATCH=1: a chain level is never run as one stacked product (tests, comparisons)
    bool trace;          // BSP_TRACE=1: per-sub-phase device timings of every product on stderr
    int ntrace;
    cudaEvent_t trace_ev[32];
    const char *trace_label[32];
    char err[512];
};

struct bsp_mat {
    int32_t rows, cols, k;
    int32_t nbr, nbc;    // block-row / block-column index space (max coordinate + 1)
    int64_t nnzb;
    int32_t *rowptr;     // device
    int32_t *colidx;     // device
    uint64_t *vals;      // device; raw tile storage (uint64 entries, or float32 entries when esize == 4)
    int device;
    int32_t esize;       // bytes per entry: 8 = uint64 (the reference's type), 4 = float32 (extension)
};

#define BSP_FAIL(ctx, code, ...)                                  \
    do {                                                          \
        snprintf((ctx)->err, sizeof((ctx)->err), __VA_ARGS__);    \
        return (code);                                            \
    } while (0)

#define BSP_CUDA(ctx, call)                                                                  \
    do {                                                                                     \
        cudaError_t e_ = (call);                                                             \
        if (e_ != cudaSuccess) {                                                             \
            snprintf((ctx)->err, sizeof((ctx)->err), "%s:%d: %s failed: %s", __FILE__,       \
                     __LINE__, #call, cudaGetErrorString(e_));                               \
            return (e_ == cudaErrorMemoryAllocation) ? BSP_ERR_NOMEM : BSP_ERR_CUDA;         \
        }                                                                                    \
    } while (0)

#define BSP_TRY(expr)                  \
    do {                               \
        int rc_ = (expr);              \
        if (rc_ != BSP_OK) return rc_; \
    } while (0)

// stream-ordered allocation on the context's stream
template <typename T>
static inline int bsp_alloc(bsp_ctx *ctx, T **p, size_t count)
{
    *p = nullptr;
    size_t bytes = (count ? count : 1) * sizeof(T);
    cudaError_t e = cudaMallocAsync((void **)p, bytes, ctx->stream);
    if (e != cudaSuccess) {
        snprintf(ctx->err, sizeof(ctx->err), "device allocation of %zu bytes failed: %s", bytes,
                 cudaGetErrorString(e));
        cudaGetLastError();
        return BSP_ERR_NOMEM;
    }
    return BSP_OK;
}
static inline void bsp_release(bsp_ctx *ctx, void *p)
{
    if (p) cudaFreeAsync(p, ctx->stream);
}

// Scalar read-backs (tile totals, class counts, refold counter): up to four runs of 32-bit words are stored into
// ctx->pinned at the given byte offsets by a one-warp kernel on ctx->stream; the caller reads them after it has
// waited on the stream.  Not a device->host copy on purpose: those queue in the copy engine behind a bulk
// download running on ctx->copy_stream (bsp_multiply_host), and the next panel's symbolic pass would wait for it.
struct bsp_publish_list {
    const void *src[4];
    uint32_t offset[4]; // bytes into ctx->pinned, multiple of 4
    uint32_t words[4];
    int n;
};
int bsp_publish(bsp_ctx *ctx, const bsp_publish_list &list);

// BSP_TRACE=1: mark a point on the stream; bsp_trace_dump prints the gaps between marks
static inline void bsp_trace_mark(bsp_ctx *ctx, const char *label)
{
    if (!ctx->trace || ctx->ntrace >= 32) return;
    if (!ctx->trace_ev[ctx->ntrace]) cudaEventCreate(&ctx->trace_ev[ctx->ntrace]);
    cudaEventRecord(ctx->trace_ev[ctx->ntrace], ctx->stream);
    ctx->trace_label[ctx->ntrace++] = label;
}
static inline void bsp_trace_dump(bsp_ctx *ctx)
{
    if (!ctx->trace || ctx->ntrace < 2) { ctx->ntrace = 0; return; }
    cudaEventSynchronize(ctx->trace_ev[ctx->ntrace - 1]);
    fprintf(stderr, "[bsp trace]");
    for (int i = 1; i < ctx->ntrace; i++) {
        float ms = 0;
        cudaEventElapsedTime(&ms, ctx->trace_ev[i - 1], ctx->trace_ev[i]);
        fprintf(stderr, " %s=%.3f", ctx->trace_label[i], ms);
    }
    fprintf(stderr, "\n");
    ctx->ntrace = 0;
}

// Scratch allocations of one call: everything still registered is released when the guard
// goes out of scope, so early returns on errors do not strand pool memory.  keep(p) hands a
// buffer over to the caller (result arrays).
struct bsp_scratch {
    bsp_ctx *ctx;
    void *ptrs[32];
    int n;
    explicit bsp_scratch(bsp_ctx *c) : ctx(c), n(0) {}
    ~bsp_scratch() { for (int i = 0; i < n; i++) bsp_release(ctx, ptrs[i]); }
    template <typename T> int alloc(T **p, size_t count)
    {
        const int rc = bsp_alloc(ctx, p, count);
        if (rc == BSP_OK && n < 32) ptrs[n++] = *p;
        return rc;
    }
    void keep(void *p)
    {
        for (int i = 0; i < n; i++)
            if (ptrs[i] == p) { ptrs[i] = ptrs[--n]; return; }
    }
    bsp_scratch(const bsp_scratch &) = delete;
    bsp_scratch &operator=(const bsp_scratch &) = delete;
};

#define BSP_LAUNCH_CHECK(ctx)                  \
    do {                                       \
        (ctx)->launches++;                     \
        BSP_CUDA(ctx, cudaGetLastError());     \
    } while (0)

// ---- phases (each defined in its own .cu) ---------------------------------------------------

// exclusive scans on the context's stream; out may alias in; out[n] is NOT written (n values in, n out)
int bsp_scan_i32(bsp_ctx *ctx, const int32_t *in, int32_t *out, int64_t n, int32_t *total_dev);
int bsp_scan_i32_to_i64(bsp_ctx *ctx, const int32_t *in, int64_t *out, int64_t n);

struct bsp_symbolic_out {
    int32_t *c_rowptr;   // [nbr+1]
    int32_t *c_colidx;   // [nnzb_c]
    int64_t *pair_ptr;   // [nnzb_c+1]
    int2 *pairs;         // [nprod]  (index of A tile, index of B tile | sign bit on the last pair of a tile),
                         //          ascending inner block index inside a tile
    int64_t nnzb_c;
    int64_t nprod;
    // When the pair lists of the requested rows would exceed pair_budget entries and the range holds
    // more than one non-empty row, nothing is filled: n_split > 0 and split[0..n_split] are row
    // boundaries of panels whose pair lists fit (a single row above the budget gets a panel of its own).
    int64_t pair_budget; // in: 0 = no limit
    int32_t n_split;     // out
    int32_t *split;      // out: malloc'ed [n_split+1], the caller frees
};

// GPU symbolic pass (replaces the host hash maps of cu:141-176 and the pack loop cu:189-218)
int bsp_symbolic(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t row_lo, int32_t row_hi,
                 bsp_symbolic_out *out);

// numeric pass (replaces matrix_multiplyKernel cu:44-66 and the 500-key rounds cu:181-270)
int bsp_numeric(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym,
                uint64_t *c_vals, int64_t *exact_refolds, int32_t *path);

int bsp_prune_device(bsp_ctx *ctx, bsp_mat *m);
int bsp_concat_panels_device(bsp_ctx *ctx, bsp_mat *const *panels, int32_t n, bsp_mat *out);
int bsp_stack_rows_device(bsp_ctx *ctx, const bsp_mat *const *mats, int n, const int32_t *shifts, const int32_t *limits,
                          int32_t sentinel, bsp_mat *out);
int bsp_split_rows_device(bsp_ctx *ctx, const bsp_mat *big, int n, const int32_t *row_off, const int64_t *tile_off,
                          bsp_mat *const *parts);
int bsp_gather_i32_to_host(bsp_ctx *ctx, const int32_t *src, const int32_t *idx_host, int n, int32_t *dst_host);
int bsp_expand_rows(bsp_ctx *ctx, const bsp_mat *m, int32_t *d_brow);
int bsp_ingest_coo(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t esize, int64_t nnzb,
                   const int32_t *d_brow, const int32_t *d_bcol, const void *d_vals,
                   int32_t max_row, int32_t max_col, bsp_mat **out);
// float32 numeric pass (extension; no reference counterpart)
int bsp_numeric_f32(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, float *c_vals,
                    int32_t *path);
