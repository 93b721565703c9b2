// bsp_symbolic.cu -- GPU symbolic pass of C = A*B on CSR-of-blocks.
//
// Replaces the reference's single-threaded host symbolic phase: the B row index
// (cu:141-145), the map d: (i,c) -> [j...] (cu:147-156), the key array (cu:167-176) and the
// per-round pack loop that copies every A/B tile pair into a staging buffer (cu:189-218).
// Here nothing is copied: the pass produces, per output tile, a list of (A tile index,
// B tile index) pairs in ascending inner block index j -- the order in which the reference's
// kernel folds them (cu:54-62) -- and the numeric pass gathers tiles by index.
//
// One warp per block row i of A:
//   count: set bit c in a per-warp shared-memory bitmap for every (j in A row i, c in B row j)
//          (warp lanes run over B row j, atomicOr into shared memory), popcount -> |C row i|;
//          the bitmap makes the merged column set sorted and duplicate-free for free.
//   fill : rebuild the bitmap, prefix-popcount its words with warp shuffles (rank structure),
//          emit C's sorted colidx, count pairs per output tile, scan them in place into
//          pair_ptr, then append the pairs walking A row i in ascending j (lanes over B row j;
//          inside one j all target tiles differ, so the cursor updates never collide and the
//          append order per tile is ascending j).
// HBM-bound integer work; algorithmic bytes:
//   4*nnzb(A) + 4*nprod + 4*nnzb(C) + 12*(nbr+1), plus the pair lists this design adds:
//   8*nprod + 8*(nnzb(C)+1).
#include "bsp_internal.h"

namespace {

constexpr int SYM_WARPS = 8;
constexpr int SYM_THREADS = SYM_WARPS * 32;

__device__ __forceinline__ int warp_sum(int v)
{
#pragma unroll
    for (int d = 16; d; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

__device__ __forceinline__ int warp_excl_scan(int v, int *total)
{
    const int lane = threadIdx.x & 31;
    int inc = v;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        int o = __shfl_up_sync(0xffffffffu, inc, d);
        if (lane >= d) inc += o;
    }
    *total = __shfl_sync(0xffffffffu, inc, 31);
    return inc - v;
}

// Set, in the warp's bitmap, the block columns reachable from block row `row` of A.
// Returns (warp-uniform) the number of tile products of the row.
__device__ __forceinline__ int build_row_bitmap(uint32_t *bm, int words, const int32_t *__restrict__ a_rowptr,
                                                const int32_t *__restrict__ a_colidx, int b_nbr,
                                                const int32_t *__restrict__ b_rowptr,
                                                const int32_t *__restrict__ b_colidx, int row)
{
    const int lane = threadIdx.x & 31;
    for (int w = lane; w < words; w += 32) bm[w] = 0u;
    __syncwarp();
    const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
    int nprod = 0;
    for (int p = a0; p < a1; p++) {
        const int j = a_colidx[p];
        if (j >= b_nbr) continue;
        const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
        nprod += b1 - b0;
        for (int q = b0 + lane; q < b1; q += 32) {
            const int c = b_colidx[q];
            atomicOr(&bm[c >> 5], 1u << (c & 31));
        }
    }
    __syncwarp();
    return nprod;
}

__global__ void __launch_bounds__(SYM_THREADS)
sym_count_kernel(int a_nbr, const int32_t *__restrict__ a_rowptr, const int32_t *__restrict__ a_colidx, int b_nbr,
                 const int32_t *__restrict__ b_rowptr, const int32_t *__restrict__ b_colidx, int words, int row_lo,
                 int row_hi, int32_t *__restrict__ row_cnt, int32_t *__restrict__ row_prod)
{
    extern __shared__ uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *bm = smem + (size_t)warp * words;
    for (int row = blockIdx.x * SYM_WARPS + warp; row < a_nbr; row += gridDim.x * SYM_WARPS) {
        int cnt = 0, nprod = 0;
        if (row >= row_lo && row < row_hi && a_rowptr[row + 1] > a_rowptr[row]) {
            nprod = build_row_bitmap(bm, words, a_rowptr, a_colidx, b_nbr, b_rowptr, b_colidx, row);
            for (int w = lane; w < words; w += 32) cnt += __popc(bm[w]);
            cnt = warp_sum(cnt);
            __syncwarp();
        }
        if (lane == 0) {
            row_cnt[row] = cnt;
            row_prod[row] = nprod;
        }
    }
}

__global__ void __launch_bounds__(SYM_THREADS)
sym_fill_kernel(int a_nbr, const int32_t *__restrict__ a_rowptr, const int32_t *__restrict__ a_colidx, int b_nbr,
                const int32_t *__restrict__ b_rowptr, const int32_t *__restrict__ b_colidx, int words,
                const int32_t *__restrict__ c_rowptr, const int64_t *__restrict__ row_prod_ptr,
                int32_t *__restrict__ c_colidx, unsigned long long *pair_ptr /* zeroed, [nnzb_c+1] */,
                int2 *__restrict__ pairs)
{
    extern __shared__ uint32_t smem[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    uint32_t *bm = smem + (size_t)warp * 2 * words;
    uint32_t *wrank = bm + words;
    for (int row = blockIdx.x * SYM_WARPS + warp; row < a_nbr; row += gridDim.x * SYM_WARPS) {
        const int cbase = c_rowptr[row], cend = c_rowptr[row + 1];
        if (cend == cbase) continue; // warp-uniform
        build_row_bitmap(bm, words, a_rowptr, a_colidx, b_nbr, b_rowptr, b_colidx, row);
        // word-level exclusive prefix of popcounts: rank(c) = wrank[c>>5] + popc(bits below c)
        int carry = 0;
        for (int w0 = 0; w0 < words; w0 += 32) {
            const int w = w0 + lane;
            const uint32_t bits = w < words ? bm[w] : 0u;
            int tot;
            const int ex = warp_excl_scan(__popc(bits), &tot);
            if (w < words) {
                wrank[w] = (uint32_t)(carry + ex);
                // sorted, duplicate-free block columns of C row
                uint32_t b = bits;
                int o = cbase + carry + ex;
                while (b) {
                    const int bit = __ffs(b) - 1;
                    b &= b - 1;
                    c_colidx[o++] = w * 32 + bit;
                }
            }
            carry += tot;
        }
        __syncwarp();
        // pairs per output tile, counted at pair_ptr[tile+1]
        const int a0 = a_rowptr[row], a1 = a_rowptr[row + 1];
        for (int p = a0; p < a1; p++) {
            const int j = a_colidx[p];
            if (j >= b_nbr) continue;
            const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
            for (int q = b0 + lane; q < b1; q += 32) {
                const int c = b_colidx[q];
                const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
                atomicAdd(&pair_ptr[(size_t)cbase + t + 1], 1ull);
            }
        }
        __syncwarp();
        // in-place exclusive scan of the row's counts: pair_ptr[tile+1] = first slot of tile
        // (it is advanced while filling and ends as the first slot of tile+1, i.e. the CSR pointer)
        unsigned long long run = (unsigned long long)row_prod_ptr[row];
        for (int t0 = cbase; t0 < cend; t0 += 32) {
            const int t = t0 + lane;
            const unsigned long long v = t < cend ? pair_ptr[(size_t)t + 1] : 0ull;
            unsigned long long inc = v;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                unsigned long long o = __shfl_up_sync(0xffffffffu, inc, d);
                if (lane >= d) inc += o;
            }
            if (t < cend) pair_ptr[(size_t)t + 1] = run + inc - v;
            run += __shfl_sync(0xffffffffu, inc, 31);
        }
        __syncwarp();
        // ordered append: A row walked in ascending j; lanes over B row j (distinct tiles)
        for (int p = a0; p < a1; p++) {
            const int j = a_colidx[p];
            if (j >= b_nbr) continue;
            const int b0 = b_rowptr[j], b1 = b_rowptr[j + 1];
            for (int q = b0 + lane; q < b1; q += 32) {
                const int c = b_colidx[q];
                const int t = (int)wrank[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
                volatile unsigned long long *cur = &pair_ptr[(size_t)cbase + t + 1];
                const unsigned long long pos = *cur;
                *cur = pos + 1ull;
                pairs[pos] = make_int2(p, q);
            }
            __syncwarp();
        }
        __syncwarp();
    }
}

} // namespace

int bsp_symbolic(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, int32_t row_lo, int32_t row_hi,
                 bsp_symbolic_out *out)
{
    memset(out, 0, sizeof(*out));
    const int nbr = A->nbr;
    const int words = (B->nbc + 31) / 32 > 0 ? (B->nbc + 31) / 32 : 1;
    const size_t smem_count = (size_t)SYM_WARPS * words * sizeof(uint32_t);
    const size_t smem_fill = 2 * smem_count;
    if (smem_fill > 200 * 1024)
        BSP_FAIL(ctx, BSP_ERR_LIMIT,
                 "symbolic pass: block-column space %d needs %zu B of shared memory per CTA (limit 204800 B, "
                 "i.e. at most 102400 block columns)", B->nbc, smem_fill);

    int32_t *row_cnt = nullptr, *row_prod = nullptr;
    int64_t *row_prod_ptr = nullptr;
    BSP_TRY(bsp_alloc(ctx, &row_cnt, (size_t)nbr + 1));
    BSP_TRY(bsp_alloc(ctx, &row_prod, (size_t)nbr + 1));
    BSP_TRY(bsp_alloc(ctx, &row_prod_ptr, (size_t)nbr + 1));
    BSP_TRY(bsp_alloc(ctx, &out->c_rowptr, (size_t)nbr + 1));
    BSP_CUDA(ctx, cudaMemsetAsync(row_cnt + nbr, 0, sizeof(int32_t), ctx->stream));
    BSP_CUDA(ctx, cudaMemsetAsync(row_prod + nbr, 0, sizeof(int32_t), ctx->stream));

    int grid = (nbr + SYM_WARPS - 1) / SYM_WARPS;
    const int max_grid = ctx->sm_count * 8;
    if (grid > max_grid) grid = max_grid;
    if (grid < 1) grid = 1;
    if (nbr > 0) {
        BSP_CUDA(ctx, cudaFuncSetAttribute(sym_count_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem_count));
        sym_count_kernel<<<grid, SYM_THREADS, smem_count, ctx->stream>>>(
            nbr, A->rowptr, A->colidx, B->nbr, B->rowptr, B->colidx, words, row_lo, row_hi, row_cnt, row_prod);
        BSP_LAUNCH_CHECK(ctx);
    }
    BSP_TRY(bsp_scan_i32(ctx, row_cnt, out->c_rowptr, (int64_t)nbr + 1, nullptr));
    BSP_TRY(bsp_scan_i32_to_i64(ctx, row_prod, row_prod_ptr, (int64_t)nbr + 1));

    // totals -> host (the one unavoidable sync of the pass: C and the pair lists are sized from them)
    int32_t *h_nnzb = (int32_t *)ctx->pinned;
    int64_t *h_nprod = (int64_t *)((char *)ctx->pinned + 8);
    BSP_CUDA(ctx, cudaMemcpyAsync(h_nnzb, out->c_rowptr + nbr, sizeof(int32_t), cudaMemcpyDeviceToHost, ctx->stream));
    BSP_CUDA(ctx, cudaMemcpyAsync(h_nprod, row_prod_ptr + nbr, sizeof(int64_t), cudaMemcpyDeviceToHost, ctx->stream));
    BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    out->nnzb_c = *h_nnzb;
    out->nprod = *h_nprod;
    if (out->nnzb_c < 0) BSP_FAIL(ctx, BSP_ERR_LIMIT, "product has more than 2^31-1 tiles");

    BSP_TRY(bsp_alloc(ctx, &out->c_colidx, (size_t)out->nnzb_c));
    BSP_TRY(bsp_alloc(ctx, &out->pair_ptr, (size_t)out->nnzb_c + 1));
    BSP_TRY(bsp_alloc(ctx, &out->pairs, (size_t)out->nprod));
    BSP_CUDA(ctx, cudaMemsetAsync(out->pair_ptr, 0, ((size_t)out->nnzb_c + 1) * sizeof(int64_t), ctx->stream));
    if (out->nnzb_c > 0) {
        BSP_CUDA(ctx, cudaFuncSetAttribute(sym_fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)smem_fill));
        sym_fill_kernel<<<grid, SYM_THREADS, smem_fill, ctx->stream>>>(
            nbr, A->rowptr, A->colidx, B->nbr, B->rowptr, B->colidx, words, out->c_rowptr, row_prod_ptr,
            out->c_colidx, (unsigned long long *)out->pair_ptr, out->pairs);
        BSP_LAUNCH_CHECK(ctx);
    }
    bsp_release(ctx, row_cnt);
    bsp_release(ctx, row_prod);
    bsp_release(ctx, row_prod_ptr);
    return BSP_OK;
}
