// bsp_matrix.cu -- building, exporting and compacting CSR-of-blocks matrices on device.
//
//   ingest : coordinate list (file order, maybe unsorted, maybe duplicate keys) -> CSR of
//            blocks.  Stands in for the reader's map insert m[{row,col}] = v (cu:383) and
//            the flatten step of helper (cu:102-138).  The key sort uses CUB's radix sort
//            (library plumbing on the ingest side, not on the multiply path).
//   expand : rowptr -> one block-row coordinate per tile, for the sorted coordinate export
//            that the writer needs (cu:598-599).
//   prune  : drop all-zero tiles of the final product (intended semantics of cu:577-592):
//            per-tile predicate, exclusive scan, compaction of colidx / vals / rowptr.
#include <cub/device/device_radix_sort.cuh>

#include <new>

#include "bsp_internal.h"

namespace {

__global__ void make_keys_kernel(const int32_t *__restrict__ brow, const int32_t *__restrict__ bcol, int64_t n,
                                 uint64_t *__restrict__ keys, int32_t *__restrict__ idx)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        keys[i] = ((uint64_t)(uint32_t)brow[i] << 32) | (uint32_t)bcol[i];
        idx[i] = (int32_t)i;
    }
}

// keep the LAST entry of each run of equal keys (cu:383: later insert overwrites)
__global__ void flag_last_kernel(const uint64_t *__restrict__ keys, int64_t n, int32_t *__restrict__ keep)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) keep[i] = (i + 1 == n || keys[i + 1] != keys[i]) ? 1 : 0;
    if (i == n) keep[i] = 0;
}

// one thread per sorted entry: write colidx and the rowptr boundaries it is responsible for
__global__ void scatter_index_kernel(const uint64_t *__restrict__ keys, const int32_t *__restrict__ keep,
                                     const int32_t *__restrict__ pos, int64_t n, int32_t nbr,
                                     int32_t *__restrict__ rowptr, int32_t *__restrict__ colidx)
{
    int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n || !keep[i]) return;
    const int32_t o = pos[i];
    const int32_t row = (int32_t)(keys[i] >> 32);
    colidx[o] = (int32_t)(keys[i] & 0xffffffffu);
    // previous kept entry's row (kept entries have distinct keys; find the previous distinct key)
    int32_t prev_row = -1;
    int64_t j = i - 1;
    while (j >= 0 && keys[j] == keys[i]) j--; // cannot happen for a kept (last) entry's successors, only predecessors
    if (j >= 0) prev_row = (int32_t)(keys[j] >> 32);
    for (int32_t r = prev_row + 1; r <= row; r++) rowptr[r] = o;
    if (o == pos[n] - 1)
        for (int32_t r = row + 1; r <= nbr; r++) rowptr[r] = o + 1;
}

// gather tiles: one CTA-strided loop over (tile, 16-byte chunk)
// (tiles are moved as 32-bit words so that the same kernels serve uint64 and float32 entries)
__global__ void gather_tiles_kernel(const uint32_t *__restrict__ src, const int32_t *__restrict__ idx,
                                    const int32_t *__restrict__ keep, const int32_t *__restrict__ pos, int64_t n,
                                    int kk, uint32_t *__restrict__ dst)
{
    const int64_t total = n * kk;
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (int64_t)gridDim.x * blockDim.x) {
        const int64_t i = g / kk;
        const int e = (int)(g - i * kk);
        if (keep[i]) dst[(size_t)pos[i] * kk + e] = src[(size_t)idx[i] * kk + e];
    }
}

__global__ void expand_rows_kernel(const int32_t *__restrict__ rowptr, int32_t nbr, int64_t nnzb,
                                   int32_t *__restrict__ brow)
{
    int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= nnzb) return;
    int lo = 0, hi = nbr; // largest r with rowptr[r] <= t
    while (hi - lo > 1) {
        int mid = (lo + hi) >> 1;
        if (rowptr[mid] <= t) lo = mid; else hi = mid;
    }
    brow[t] = lo;
}

// one warp per tile: keep = any entry non-zero (cu:580-586)
__global__ void tile_nonzero_kernel(const uint32_t *__restrict__ vals, int64_t nnzb, int kk, int32_t *__restrict__ keep)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t t = warp; t < nnzb; t += nwarps) {
        const uint32_t *v = vals + (size_t)t * kk;
        uint32_t acc = 0;
        for (int e = lane; e < kk; e += 32) acc |= v[e];
        const unsigned any = __ballot_sync(0xffffffffu, acc != 0);
        if (lane == 0) keep[t] = any ? 1 : 0;
    }
    if (warp == 0 && lane == 0) keep[nnzb] = 0;
}

// kept tiles move to their new place, `gs` threads (a power of two <= 32) per tile, `units` pieces of
// VEC 32-bit words per tile; dropped tiles are not read at all
template <int VEC>
__global__ void __launch_bounds__(256)
compact_tiles_kernel(const uint32_t *__restrict__ vals, const int32_t *__restrict__ colidx,
                     const int32_t *__restrict__ keep, const int32_t *__restrict__ pos, int64_t nnzb, int units,
                     int gs, uint32_t *__restrict__ vals_out, int32_t *__restrict__ colidx_out)
{
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t ngroups = ((int64_t)gridDim.x * blockDim.x) / gs;
    const int lig = (int)(tid & (gs - 1));
    // four tiles per round: the flag, position and data loads of the four are independent, so their
    // latencies overlap (one tile per round ran at 1.8 TB/s: three dependent memory round trips each)
    constexpr int UN = 4;
    for (int64_t t0 = tid / gs; t0 < nnzb; t0 += UN * ngroups) {
        int kp[UN];
        int64_t ps[UN];
#pragma unroll
        for (int i = 0; i < UN; i++) {
            const int64_t t = t0 + i * ngroups;
            kp[i] = t < nnzb ? keep[t] : 0;
            ps[i] = t < nnzb ? pos[t] : 0;
        }
        if (VEC == 4) {
            for (int u = lig; u < units; u += gs) {
                uint4 v[UN];
#pragma unroll
                for (int i = 0; i < UN; i++)
                    if (kp[i]) v[i] = (reinterpret_cast<const uint4 *>(vals) + (t0 + i * ngroups) * units)[u];
#pragma unroll
                for (int i = 0; i < UN; i++)
                    if (kp[i]) (reinterpret_cast<uint4 *>(vals_out) + ps[i] * units)[u] = v[i];
            }
        } else {
            for (int u = lig; u < units; u += gs) {
                uint32_t v[UN];
#pragma unroll
                for (int i = 0; i < UN; i++)
                    if (kp[i]) v[i] = (vals + (t0 + i * ngroups) * units)[u];
#pragma unroll
                for (int i = 0; i < UN; i++)
                    if (kp[i]) (vals_out + ps[i] * units)[u] = v[i];
            }
        }
        if (lig == 0) {
#pragma unroll
            for (int i = 0; i < UN; i++)
                if (kp[i]) colidx_out[ps[i]] = colidx[t0 + i * ngroups];
        }
    }
}

__global__ void remap_rowptr_kernel(const int32_t *__restrict__ rowptr, int32_t nbr, const int32_t *__restrict__ pos,
                                    int32_t *__restrict__ rowptr_out)
{
    int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r <= nbr) rowptr_out[r] = pos[rowptr[r]];
}

__global__ void add_rowptr_kernel(const int32_t *__restrict__ src, int32_t n, int32_t *__restrict__ dst)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r < n) dst[r] += src[r];
}

inline unsigned blocks_for(int64_t n, int threads) { return (unsigned)((n + threads - 1) / threads > 0 ? (n + threads - 1) / threads : 1); }

} // namespace

int bsp_ingest_coo(bsp_ctx *ctx, int32_t rows, int32_t cols, int32_t k, int32_t esize, int64_t nnzb,
                   const int32_t *d_brow, const int32_t *d_bcol, const void *d_vals, int32_t max_row, int32_t max_col,
                   bsp_mat **out)
{
    *out = nullptr;
    bsp_mat *m = new (std::nothrow) bsp_mat();
    if (!m) BSP_FAIL(ctx, BSP_ERR_NOMEM, "host allocation failed");
    m->rows = rows; m->cols = cols; m->k = k; m->device = ctx->device; m->esize = esize;
    m->nbr = nnzb > 0 ? max_row + 1 : 0;
    m->nbc = nnzb > 0 ? max_col + 1 : 0;
    m->nnzb = 0; m->rowptr = nullptr; m->colidx = nullptr; m->vals = nullptr;
    *out = m;
    const int kk = k * k * (esize / 4); // 32-bit words per tile
    BSP_TRY(bsp_alloc(ctx, &m->rowptr, (size_t)m->nbr + 1));
    if (nnzb == 0) {
        BSP_CUDA(ctx, cudaMemsetAsync(m->rowptr, 0, sizeof(int32_t), ctx->stream));
        BSP_TRY(bsp_alloc(ctx, &m->colidx, 1));
        BSP_TRY(bsp_alloc(ctx, &m->vals, 1));
        return BSP_OK;
    }
    uint64_t *keys = nullptr, *keys2 = nullptr;
    int32_t *idx = nullptr, *idx2 = nullptr, *keep = nullptr, *pos = nullptr;
    BSP_TRY(bsp_alloc(ctx, &keys, (size_t)nnzb));
    BSP_TRY(bsp_alloc(ctx, &keys2, (size_t)nnzb));
    BSP_TRY(bsp_alloc(ctx, &idx, (size_t)nnzb));
    BSP_TRY(bsp_alloc(ctx, &idx2, (size_t)nnzb));
    BSP_TRY(bsp_alloc(ctx, &keep, (size_t)nnzb + 1));
    BSP_TRY(bsp_alloc(ctx, &pos, (size_t)nnzb + 1));
    make_keys_kernel<<<blocks_for(nnzb, 256), 256, 0, ctx->stream>>>(d_brow, d_bcol, nnzb, keys, idx);
    BSP_LAUNCH_CHECK(ctx);
    size_t tmp_bytes = 0;
    BSP_CUDA(ctx, cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, keys, keys2, idx, idx2, (int)nnzb, 0, 64, ctx->stream));
    char *tmp = nullptr;
    BSP_TRY(bsp_alloc(ctx, &tmp, tmp_bytes));
    BSP_CUDA(ctx, cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, keys, keys2, idx, idx2, (int)nnzb, 0, 64, ctx->stream));
    ctx->launches += 1;
    flag_last_kernel<<<blocks_for(nnzb + 1, 256), 256, 0, ctx->stream>>>(keys2, nnzb, keep);
    BSP_LAUNCH_CHECK(ctx);
    BSP_TRY(bsp_scan_i32(ctx, keep, pos, nnzb + 1, nullptr));
    int32_t *h_uniq = (int32_t *)ctx->pinned;
    {
        bsp_publish_list pl = {};
        pl.n = 1;
        pl.src[0] = pos + nnzb; pl.offset[0] = 0; pl.words[0] = 1;
        BSP_TRY(bsp_publish(ctx, pl));
    }
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    m->nnzb = *h_uniq;
    BSP_TRY(bsp_alloc(ctx, &m->colidx, (size_t)m->nnzb));
    BSP_TRY(bsp_alloc(ctx, (uint32_t **)&m->vals, (size_t)m->nnzb * kk));
    scatter_index_kernel<<<blocks_for(nnzb, 256), 256, 0, ctx->stream>>>(keys2, keep, pos, nnzb, m->nbr, m->rowptr,
                                                                         m->colidx);
    BSP_LAUNCH_CHECK(ctx);
    {
        int64_t total = nnzb * kk;
        int64_t g = (total + 255) / 256;
        if (g > (int64_t)ctx->sm_count * 32) g = (int64_t)ctx->sm_count * 32;
        gather_tiles_kernel<<<(unsigned)g, 256, 0, ctx->stream>>>((const uint32_t *)d_vals, idx2, keep, pos, nnzb, kk,
                                                                  (uint32_t *)m->vals);
        BSP_LAUNCH_CHECK(ctx);
    }
    bsp_release(ctx, keys); bsp_release(ctx, keys2); bsp_release(ctx, idx); bsp_release(ctx, idx2);
    bsp_release(ctx, keep); bsp_release(ctx, pos); bsp_release(ctx, tmp);
    return BSP_OK;
}

int bsp_expand_rows(bsp_ctx *ctx, const bsp_mat *m, int32_t *d_brow)
{
    if (m->nnzb == 0) return BSP_OK;
    expand_rows_kernel<<<blocks_for(m->nnzb, 256), 256, 0, ctx->stream>>>(m->rowptr, m->nbr, m->nnzb, d_brow);
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}

int bsp_prune_device(bsp_ctx *ctx, bsp_mat *m)
{
    if (m->nnzb == 0) return BSP_OK;
    const int kk = m->k * m->k * (m->esize / 4); // 32-bit words per tile
    int32_t *keep = nullptr, *pos = nullptr;
    BSP_TRY(bsp_alloc(ctx, &keep, (size_t)m->nnzb + 1));
    BSP_TRY(bsp_alloc(ctx, &pos, (size_t)m->nnzb + 1));
    {
        int64_t g = (m->nnzb * 32 + 255) / 256;
        if (g > (int64_t)ctx->sm_count * 16) g = (int64_t)ctx->sm_count * 16;
        tile_nonzero_kernel<<<(unsigned)g, 256, 0, ctx->stream>>>((const uint32_t *)m->vals, m->nnzb, kk, keep);
        BSP_LAUNCH_CHECK(ctx);
    }
    BSP_TRY(bsp_scan_i32(ctx, keep, pos, m->nnzb + 1, nullptr));
    int32_t *h_kept = (int32_t *)ctx->pinned;
    {
        bsp_publish_list pl = {};
        pl.n = 1;
        pl.src[0] = pos + m->nnzb; pl.offset[0] = 0; pl.words[0] = 1;
        BSP_TRY(bsp_publish(ctx, pl));
    }
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    const int64_t kept = *h_kept;
    if (kept != m->nnzb) {
        int32_t *rowptr2 = nullptr, *colidx2 = nullptr;
        uint32_t *vals2 = nullptr;
        BSP_TRY(bsp_alloc(ctx, &rowptr2, (size_t)m->nbr + 1));
        BSP_TRY(bsp_alloc(ctx, &colidx2, (size_t)kept));
        BSP_TRY(bsp_alloc(ctx, &vals2, (size_t)kept * kk));
        const int vec = (kk % 4 == 0) ? 4 : 1;
        const int units = kk / vec;
        int gs = 1;
        while (gs < 32 && gs < units) gs <<= 1;
        int64_t g = (m->nnzb * gs + 255) / 256;
        if (g > (int64_t)ctx->sm_count * 32) g = (int64_t)ctx->sm_count * 32;
        if (vec == 4)
            compact_tiles_kernel<4><<<(unsigned)g, 256, 0, ctx->stream>>>((const uint32_t *)m->vals, m->colidx, keep, pos,
                                                                         m->nnzb, units, gs, vals2, colidx2);
        else
            compact_tiles_kernel<1><<<(unsigned)g, 256, 0, ctx->stream>>>((const uint32_t *)m->vals, m->colidx, keep, pos,
                                                                         m->nnzb, units, gs, vals2, colidx2);
        BSP_LAUNCH_CHECK(ctx);
        remap_rowptr_kernel<<<blocks_for((int64_t)m->nbr + 1, 256), 256, 0, ctx->stream>>>(m->rowptr, m->nbr, pos,
                                                                                            rowptr2);
        BSP_LAUNCH_CHECK(ctx);
        bsp_release(ctx, m->rowptr); bsp_release(ctx, m->colidx); bsp_release(ctx, m->vals);
        m->rowptr = rowptr2; m->colidx = colidx2; m->vals = (uint64_t *)vals2; m->nnzb = kept;
    }
    bsp_release(ctx, keep);
    bsp_release(ctx, pos);
    return BSP_OK;
}

int bsp_concat_panels_device(bsp_ctx *ctx, bsp_mat *const *panels, int32_t n, bsp_mat *out)
{
    // Panel i has tiles only in its own row range, so its rowptr is 0 before the range and
    // nnzb_i after it: the row pointers of the concatenation are the element-wise sum, and
    // colidx / vals are the panels' arrays back to back.
    const int nbr = out->nbr;
    const size_t tb = (size_t)out->k * out->k * out->esize; // bytes per tile
    BSP_CUDA(ctx, cudaMemsetAsync(out->rowptr, 0, ((size_t)nbr + 1) * sizeof(int32_t), ctx->stream));
    int64_t off = 0;
    for (int i = 0; i < n; i++) {
        const bsp_mat *p = panels[i];
        add_rowptr_kernel<<<blocks_for((int64_t)nbr + 1, 256), 256, 0, ctx->stream>>>(p->rowptr, nbr + 1, out->rowptr);
        BSP_LAUNCH_CHECK(ctx);
        if (p->nnzb) {
            BSP_CUDA(ctx, cudaMemcpyAsync(out->colidx + off, p->colidx, (size_t)p->nnzb * 4, cudaMemcpyDefault, ctx->stream));
            BSP_CUDA(ctx, cudaMemcpyAsync((char *)out->vals + (size_t)off * tb, p->vals, (size_t)p->nnzb * tb, cudaMemcpyDefault, ctx->stream));
        }
        off += p->nnzb;
    }
    return BSP_OK;
}

// ---- batching of independent products (one tree level of a chain) ------------------------------
// C_i = A_i * B_i for i = 0..m-1 equals the vertical stack of the C_i when the A_i are stacked
// with their column indices shifted by the row offset of B_i in the stack of the B_i:
//     [A_0 (cols + 0); A_1 (cols + nbr(B_0)); ...] * [B_0; B_1; ...] = [C_0; C_1; ...]
// Column indices of B (and C) stay local, so the symbolic bitmaps keep their width; every pair
// list is the same as in the separate product.  A column index that has no partner row in its
// own B_i is replaced by a sentinel beyond the stacked B so that it cannot hit B_{i+1}.
namespace {

// Building the stacks and cutting the result used to cost 9 launches or copies per product (about
// 1.5 ms of API calls for a level of 32 products).  Here the parts are described by a table passed BY
// VALUE as a kernel argument, every thread finds its part by a binary search in that table, and a
// level costs two launches per stack or split whatever the number of products.
constexpr int STACK_MAX = 32; // parts per launch (longer levels take several launches)

struct stack_table {
    int n;
    const int32_t *rowptr[STACK_MAX];
    const int32_t *colidx[STACK_MAX];
    const uint32_t *vals[STACK_MAX];
    int32_t shift[STACK_MAX], limit[STACK_MAX];
    int32_t row_off[STACK_MAX + 1];   // first row of part i, relative to the first part of this launch
    int64_t tile_off[STACK_MAX + 1];  // first tile of part i in the stack (absolute)
};

struct split_table {
    int n;
    int32_t *rowptr[STACK_MAX];
    int32_t *colidx[STACK_MAX];
    uint32_t *vals[STACK_MAX];
    int32_t row_off[STACK_MAX + 1];   // first row of part i in the big matrix
    int64_t tile_off[STACK_MAX + 1];  // first tile of part i in the big matrix
};

template <typename T, typename V>
__device__ __forceinline__ int find_part(const T *off, int n, V x) // last i < n with off[i] <= x
{
    int lo = 0, hi = n;
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if ((V)off[mid] <= x) lo = mid; else hi = mid;
    }
    return lo;
}

// row pointers and column indices of the stack (dst_rowptr / dst_colidx point at the first row / tile of this launch)
__global__ void stack_meta_kernel(const __grid_constant__ stack_table tb, int32_t sentinel, int closing,
                                  int32_t *__restrict__ dst_rowptr, int32_t *__restrict__ dst_colidx)
{
    const int rows = tb.row_off[tb.n];
    const int64_t tiles = tb.tile_off[tb.n] - tb.tile_off[0];
    const int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x < rows) {
        const int i = find_part(tb.row_off, tb.n, (int)x);
        dst_rowptr[x] = tb.rowptr[i][x - tb.row_off[i]] + (int32_t)tb.tile_off[i];
    } else if (x == rows) {
        if (closing) dst_rowptr[rows] = (int32_t)tb.tile_off[tb.n];
    } else if (x - rows - 1 < tiles) {
        const int64_t y = x - rows - 1 + tb.tile_off[0];
        const int i = find_part(tb.tile_off, tb.n, y);
        const int32_t j = tb.colidx[i][y - tb.tile_off[i]];
        dst_colidx[y - tb.tile_off[0]] = (tb.limit[i] >= 0 && j >= tb.limit[i]) ? sentinel : j + tb.shift[i];
    }
}

// tile values of the stack, `units` pieces of VEC words per tile (dst points at the first tile of this launch)
template <int VEC>
__global__ void stack_vals_kernel(const __grid_constant__ stack_table tb, int units, uint32_t *__restrict__ dst)
{
    const int64_t total = (tb.tile_off[tb.n] - tb.tile_off[0]) * units;
    for (int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; x < total; x += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = x / units + tb.tile_off[0];
        const int i = find_part(tb.tile_off, tb.n, t);
        const int64_t local = x - (tb.tile_off[i] - tb.tile_off[0]) * units;
        if (VEC == 4) reinterpret_cast<uint4 *>(dst)[x] = reinterpret_cast<const uint4 *>(tb.vals[i])[local];
        else dst[x] = tb.vals[i][local];
    }
}

// the reverse: block rows [row_off[i], row_off[i+1]) of `big` become part i
__global__ void split_meta_kernel(const __grid_constant__ split_table tb, const int32_t *__restrict__ big_rowptr,
                                  const int32_t *__restrict__ big_colidx)
{
    const int rows = tb.row_off[tb.n] - tb.row_off[0];
    const int64_t tiles = tb.tile_off[tb.n] - tb.tile_off[0];
    const int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (x < rows + tb.n) { // part i has nbr_i + 1 row pointers: entry x of the concatenation of those lists
        int lo = 0, hi = tb.n; // last i with row_off[i] - row_off[0] + i <= x
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (tb.row_off[mid] - tb.row_off[0] + mid <= x) lo = mid; else hi = mid;
        }
        const int r = (int)x - (tb.row_off[lo] - tb.row_off[0] + lo);
        tb.rowptr[lo][r] = big_rowptr[tb.row_off[lo] + r] - (int32_t)tb.tile_off[lo];
    } else if (x - rows - tb.n < tiles) {
        const int64_t y = x - rows - tb.n + tb.tile_off[0];
        const int i = find_part(tb.tile_off, tb.n, y);
        tb.colidx[i][y - tb.tile_off[i]] = big_colidx[y];
    }
}

template <int VEC>
__global__ void split_vals_kernel(const __grid_constant__ split_table tb, int units, const uint32_t *__restrict__ big_vals)
{
    const int64_t total = (tb.tile_off[tb.n] - tb.tile_off[0]) * units;
    const int64_t base = tb.tile_off[0] * units;
    for (int64_t x = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; x < total; x += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = x / units + tb.tile_off[0];
        const int i = find_part(tb.tile_off, tb.n, t);
        const int64_t local = x - (tb.tile_off[i] - tb.tile_off[0]) * units;
        if (VEC == 4) reinterpret_cast<uint4 *>(tb.vals[i])[local] = reinterpret_cast<const uint4 *>(big_vals)[base + x];
        else tb.vals[i][local] = big_vals[base + x];
    }
}

__global__ void gather_i32_kernel(const int32_t *__restrict__ src, const int32_t *__restrict__ idx, int n,
                                  int32_t *__restrict__ dst)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) dst[i] = src[idx[i]];
}

} // namespace

// shifts / limits may be null (no column shift: used for the B side)
int bsp_stack_rows_device(bsp_ctx *ctx, const bsp_mat *const *mats, int n, const int32_t *shifts, const int32_t *limits,
                          int32_t sentinel, bsp_mat *out /* arrays allocated, nbr / nnzb set */)
{
    const int kk = out->k * out->k * (out->esize / 4); // 32-bit words per tile
    const int vec = (kk % 4 == 0) ? 4 : 1;
    const int units = kk / vec;
    if (n == 0) BSP_CUDA(ctx, cudaMemsetAsync(out->rowptr, 0, sizeof(int32_t), ctx->stream));
    int64_t tile_off = 0;
    int32_t row_off = 0;
    for (int i0 = 0; i0 < n; i0 += STACK_MAX) {
        stack_table tb;
        tb.n = n - i0 < STACK_MAX ? n - i0 : STACK_MAX;
        int32_t r = 0;
        int64_t t = tile_off;
        for (int i = 0; i < tb.n; i++) {
            const bsp_mat *m = mats[i0 + i];
            tb.rowptr[i] = m->rowptr; tb.colidx[i] = m->colidx; tb.vals[i] = (const uint32_t *)m->vals;
            tb.shift[i] = shifts ? shifts[i0 + i] : 0;
            tb.limit[i] = limits ? limits[i0 + i] : -1;
            tb.row_off[i] = r; tb.tile_off[i] = t;
            r += m->nbr; t += m->nnzb;
        }
        tb.row_off[tb.n] = r; tb.tile_off[tb.n] = t;
        const int64_t items = (int64_t)r + 1 + (t - tile_off);
        stack_meta_kernel<<<blocks_for(items, 256), 256, 0, ctx->stream>>>(tb, sentinel, i0 + tb.n == n ? 1 : 0,
                                                                          out->rowptr + row_off, out->colidx + tile_off);
        BSP_LAUNCH_CHECK(ctx);
        if (t > tile_off) {
            int64_t g = ((t - tile_off) * units + 255) / 256;
            if (g > (int64_t)ctx->sm_count * 16) g = (int64_t)ctx->sm_count * 16;
            uint32_t *dst = (uint32_t *)out->vals + (size_t)tile_off * kk;
            if (vec == 4) stack_vals_kernel<4><<<(unsigned)g, 256, 0, ctx->stream>>>(tb, units, dst);
            else stack_vals_kernel<1><<<(unsigned)g, 256, 0, ctx->stream>>>(tb, units, dst);
            BSP_LAUNCH_CHECK(ctx);
        }
        tile_off = t;
        row_off += r;
    }
    return BSP_OK;
}

// parts[i] (arrays allocated by the caller, nbr / nnzb set) = block rows [row_off[i], row_off[i] + parts[i]->nbr) of
// `big`, whose tiles start at tile_off[i]; the parts are consecutive in `big`
int bsp_split_rows_device(bsp_ctx *ctx, const bsp_mat *big, int n, const int32_t *row_off, const int64_t *tile_off,
                          bsp_mat *const *parts)
{
    const int kk = big->k * big->k * (big->esize / 4);
    const int vec = (kk % 4 == 0) ? 4 : 1;
    const int units = kk / vec;
    for (int i0 = 0; i0 < n; i0 += STACK_MAX) {
        split_table tb;
        tb.n = n - i0 < STACK_MAX ? n - i0 : STACK_MAX;
        for (int i = 0; i < tb.n; i++) {
            bsp_mat *p = parts[i0 + i];
            tb.rowptr[i] = p->rowptr; tb.colidx[i] = p->colidx; tb.vals[i] = (uint32_t *)p->vals;
            tb.row_off[i] = row_off[i0 + i]; tb.tile_off[i] = tile_off[i0 + i];
        }
        const bsp_mat *last = parts[i0 + tb.n - 1];
        tb.row_off[tb.n] = tb.row_off[tb.n - 1] + last->nbr;
        tb.tile_off[tb.n] = tb.tile_off[tb.n - 1] + last->nnzb;
        const int64_t tiles = tb.tile_off[tb.n] - tb.tile_off[0];
        const int64_t items = (int64_t)(tb.row_off[tb.n] - tb.row_off[0]) + tb.n + tiles;
        split_meta_kernel<<<blocks_for(items, 256), 256, 0, ctx->stream>>>(tb, big->rowptr, big->colidx);
        BSP_LAUNCH_CHECK(ctx);
        if (tiles > 0) {
            int64_t g = (tiles * units + 255) / 256;
            if (g > (int64_t)ctx->sm_count * 16) g = (int64_t)ctx->sm_count * 16;
            if (vec == 4) split_vals_kernel<4><<<(unsigned)g, 256, 0, ctx->stream>>>(tb, units, (const uint32_t *)big->vals);
            else split_vals_kernel<1><<<(unsigned)g, 256, 0, ctx->stream>>>(tb, units, (const uint32_t *)big->vals);
            BSP_LAUNCH_CHECK(ctx);
        }
    }
    return BSP_OK;
}

// dst_host[i] = src[idx_host[i]]: one small kernel and one copy instead of n copies of 4 bytes
int bsp_gather_i32_to_host(bsp_ctx *ctx, const int32_t *src, const int32_t *idx_host, int n, int32_t *dst_host)
{
    bsp_scratch sc(ctx);
    int32_t *d_idx = nullptr, *d_out = nullptr;
    BSP_TRY(sc.alloc(&d_idx, (size_t)n));
    BSP_TRY(sc.alloc(&d_out, (size_t)n));
    BSP_CUDA(ctx, cudaMemcpyAsync(d_idx, idx_host, (size_t)n * 4, cudaMemcpyHostToDevice, ctx->stream));
    gather_i32_kernel<<<blocks_for(n, 256), 256, 0, ctx->stream>>>(src, d_idx, n, d_out);
    BSP_LAUNCH_CHECK(ctx);
    BSP_CUDA(ctx, cudaMemcpyAsync(dst_host, d_out, (size_t)n * 4, cudaMemcpyDeviceToHost, ctx->stream));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return BSP_OK;
}
