// bsp_numeric.cu -- numeric pass of C = A*B: multiply-accumulate of k x k uint64 tiles.
//
// Replaces matrix_multiplyKernel (cu:44-66) and the host loop around it that stages tile
// pairs in rounds of 500 output keys (cu:181-270).  Output-stationary: the threads that own
// an output tile walk its pair list (built by the symbolic pass in ascending inner block
// index) and gather the A and B tiles by index straight from the CSR-of-blocks arrays, so no
// tile is ever copied into a staging buffer and no atomics are needed.
//
// Arithmetic (cu:59, cu:61), per output element, pairs in list order, inner index ascending:
//     p   = phi(a * b mod 2^64)        phi(x) = x % (2^64-1) = (x == 2^64-1 ? 0 : x)
//     acc = phi(acc + p mod 2^64)
// phi is not a ring homomorphism, so the fold order is part of the result (tests/golden t7/t8).
//
// Fast path (k = 4, 8, 16).  phi changes a value only when an intermediate equals 2^64-1.
// As long as that has not happened, the running value equals the plain wrapping sum
// s' = sum a*b mod 2^64 (3 integer multiply-adds per step instead of ~12 instructions).  The
// kernel therefore accumulates s' and, for every step, folds a conservative detector into one
// register:   flag = max(flag, lo32(s'_new), lo32(s'_new) - lo32(s'_old))
// lo32(s'_new) - lo32(s'_old) is the low word of the product, so flag reaches 0xffffffff
// whenever the product or the new sum has an all-ones low word -- a superset of "equals
// 2^64-1".  By induction s' is exact up to the first such step; if the flag fired anywhere in
// a thread's row of an output tile (probability ~2^-31 per step on random data, certain on the
// edge-case tests), that row is recomputed with the per-operation arithmetic in reference
// order.  The result is bit-exact in every case.
//
// Work distribution.  A group of K threads owns one output tile at a time, thread r holding
// row r of the tile (K accumulators).  Work is cut in the PAIR index space, not the tile
// space: chunk g owns the tiles whose first pair falls in [g*Q, (g+1)*Q), so every group
// performs the same number of tile products whatever the distribution of pairs per tile
// (1.5 on average for 1 % density, thousands in a dense chain level, power-law for R-MAT).
// All groups of a warp step through their pair lists in lockstep: one tile product per step.
// Per step a thread loads its A row (K u64) into registers and the group stages the B tile in
// shared memory with coalesced 128-bit loads (double buffered, one __syncwarp per step); the
// next step's A row and B chunk are prefetched into registers before the multiply-adds of the
// current one.
#include <cstdlib>

#include "bsp_internal.h"

namespace {

__device__ __forceinline__ uint64_t phi(uint64_t x) { return x == ~0ull ? 0ull : x; }

// exact per-operation fold of one output element
__device__ __forceinline__ uint64_t fold_exact(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                                               const int2 *__restrict__ pairs, int64_t p0, int64_t p1, int k, int r,
                                               int c)
{
    const size_t kk = (size_t)k * k;
    uint64_t acc = 0;
    for (int64_t p = p0; p < p1; p++) {
        const int2 pr = pairs[p];
        const uint64_t *A = a_vals + (size_t)pr.x * kk + (size_t)r * k;
        const uint64_t *B = b_vals + (size_t)pr.y * kk + c;
        for (int j = 0; j < k; j++) {
            const uint64_t t = phi(A[j] * B[(size_t)j * k]);
            acc = phi(acc + t);
        }
    }
    return acc;
}

// ---- any tile edge k: exact arithmetic throughout -------------------------------------------
// 256-thread CTAs; when k*k <= 256 a CTA works on 256/(k*k) output tiles at a time with one
// thread per element, otherwise on one tile with threads striding over its elements.
constexpr int GEN_THREADS = 256;

__global__ void __launch_bounds__(GEN_THREADS)
numeric_generic_kernel(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                       const int64_t *__restrict__ pair_ptr, const int2 *__restrict__ pairs, int64_t nnzb_c, int k,
                       uint64_t *__restrict__ c_vals)
{
    const int E = k * k;
    const int slots = E <= GEN_THREADS ? GEN_THREADS / E : 1;
    const int slot = E <= GEN_THREADS ? threadIdx.x / E : 0;
    const int e0 = E <= GEN_THREADS ? threadIdx.x % E : threadIdx.x;
    if (slot >= slots) return;
    for (int64_t t = (int64_t)blockIdx.x * slots + slot; t < nnzb_c; t += (int64_t)gridDim.x * slots) {
        const int64_t p0 = pair_ptr[t], p1 = pair_ptr[t + 1];
        for (int e = e0; e < E; e += GEN_THREADS) {
            c_vals[(size_t)t * E + e] = fold_exact(a_vals, b_vals, pairs, p0, p1, k, e / k, e % k);
            if (E <= GEN_THREADS) break;
        }
    }
}

// ---- fast path ------------------------------------------------------------------------------

// chunk g = tiles whose first pair index is in [g*Q, (g+1)*Q)
__global__ void chunk_bounds_kernel(const int64_t *__restrict__ pair_ptr, int64_t nnzb_c, int64_t nprod, int64_t Q,
                                    int nchunks, int32_t *__restrict__ chunk_tile)
{
    const int g = blockIdx.x * blockDim.x + threadIdx.x;
    if (g > nchunks) return;
    if (g == nchunks) { chunk_tile[g] = (int32_t)nnzb_c; return; }
    int64_t target = (int64_t)g * Q;
    if (target > nprod) target = nprod;
    int64_t lo = 0, hi = nnzb_c; // first t with pair_ptr[t] >= target
    while (lo < hi) {
        const int64_t mid = (lo + hi) >> 1;
        if (pair_ptr[mid] < target) lo = mid + 1; else hi = mid;
    }
    chunk_tile[g] = (int32_t)lo;
}

// s += a*b (mod 2^64) and the conservative 2^64-1 detector on the low words
__device__ __forceinline__ void mac_flag(uint64_t &s, uint32_t &flag, uint64_t a, uint64_t b)
{
    const uint32_t alo = (uint32_t)a, ahi = (uint32_t)(a >> 32), blo = (uint32_t)b, bhi = (uint32_t)(b >> 32);
    const uint32_t lo_old = (uint32_t)s;
    const uint64_t t = (uint64_t)alo * blo + s; // IMAD.WIDE.U32
    uint32_t thi = (uint32_t)(t >> 32);
    thi = alo * bhi + thi; // IMAD
    thi = ahi * blo + thi; // IMAD
    const uint32_t lo = (uint32_t)t;
    flag = max(flag, max(lo, lo - lo_old)); // VIADDMNMX / VIMNMX3
    s = ((uint64_t)thi << 32) | lo;
}

template <int K>
__device__ __noinline__ void refold_row_exact(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                                              const int2 *__restrict__ pairs, int64_t p0, int64_t p1, int r,
                                              uint64_t *out_row)
{
    for (int c = 0; c < K; c++) out_row[c] = fold_exact(a_vals, b_vals, pairs, p0, p1, K, r, c);
}

constexpr int ROWS_THREADS = 256;

template <int K>
struct RowsCfg {
    static constexpr int GROUPS = 32 / K;                 // groups per warp
    static constexpr int TILE_BYTES = K * K * 8;
    static constexpr int CHUNKS = TILE_BYTES / 16;        // 16-byte chunks per tile
    static constexpr int CPT = CHUNKS / K;                // chunks each thread moves per tile (= K/2)
    static constexpr int GROUP_STRIDE = TILE_BYTES + 16;  // +16 B: groups land in different banks
    static constexpr int BUF_BYTES = GROUPS * GROUP_STRIDE;
    static constexpr int WARP_BYTES = 2 * BUF_BYTES;      // double buffered
};

template <int K>
__global__ void __launch_bounds__(ROWS_THREADS)
numeric_rows_kernel(const uint64_t *__restrict__ a_vals, const uint64_t *__restrict__ b_vals,
                    const int64_t *__restrict__ pair_ptr, const int2 *__restrict__ pairs,
                    const int32_t *__restrict__ chunk_tile, int nchunks, uint64_t *__restrict__ c_vals,
                    unsigned long long *__restrict__ refold_count)
{
    using Cfg = RowsCfg<K>;
    __shared__ __align__(16) unsigned char smem[(ROWS_THREADS / 32) * Cfg::WARP_BYTES];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int grp = lane / K, r = lane % K;
    unsigned char *wbase = smem + warp * Cfg::WARP_BYTES + grp * Cfg::GROUP_STRIDE;

    const int g = (blockIdx.x * (ROWS_THREADS / 32) + warp) * Cfg::GROUPS + grp;
    int t = 0, t_end = 0;
    if (g < nchunks) { t = chunk_tile[g]; t_end = chunk_tile[g + 1]; }
    int64_t p = 0, p_end = 0;
    if (t < t_end) { p = pair_ptr[t]; p_end = pair_ptr[t_end]; }
    int64_t pe = (t < t_end) ? pair_ptr[t + 1] : 0; // end of the current tile's pairs
    int steps = (int)(p_end - p);
#pragma unroll
    for (int d = 16; d; d >>= 1) steps = max(steps, __shfl_xor_sync(0xffffffffu, steps, d));

    uint64_t acc[K], a_cur[K], a_nxt[K];
    ulonglong2 b_nxt[Cfg::CPT];
    uint32_t flag = 0;
#pragma unroll
    for (int c = 0; c < K; c++) { acc[c] = 0; a_nxt[c] = 0; }
#pragma unroll
    for (int i = 0; i < Cfg::CPT; i++) b_nxt[i] = make_ulonglong2(0, 0);

    auto prefetch = [&](int64_t pp) {
        const int2 pr = pairs[pp];
        const ulonglong2 *ap = reinterpret_cast<const ulonglong2 *>(a_vals + (size_t)pr.x * (K * K) + (size_t)r * K);
#pragma unroll
        for (int i = 0; i < K / 2; i++) {
            const ulonglong2 v = ap[i];
            a_nxt[2 * i] = v.x;
            a_nxt[2 * i + 1] = v.y;
        }
        const ulonglong2 *bp = reinterpret_cast<const ulonglong2 *>(b_vals + (size_t)pr.y * (K * K));
#pragma unroll
        for (int i = 0; i < Cfg::CPT; i++) b_nxt[i] = bp[r + K * i]; // chunk r + K*i: coalesced over the group
    };
    if (p < p_end) prefetch(p);

    for (int s = 0; s < steps; s++) {
        const bool active = p < p_end;
        unsigned char *buf = wbase + (s & 1) * Cfg::BUF_BYTES;
        if (active) {
#pragma unroll
            for (int c = 0; c < K; c++) a_cur[c] = a_nxt[c];
#pragma unroll
            for (int i = 0; i < Cfg::CPT; i++)
                *reinterpret_cast<ulonglong2 *>(buf + (r + K * i) * 16) = b_nxt[i]; // natural tile layout
        }
        __syncwarp();
        if (active) {
            if (p + 1 < p_end) prefetch(p + 1);
            const ulonglong2 *bs = reinterpret_cast<const ulonglong2 *>(buf);
#pragma unroll
            for (int j = 0; j < K; j++) {
#pragma unroll
                for (int c2 = 0; c2 < K / 2; c2++) {
                    const ulonglong2 bv = bs[j * (K / 2) + c2]; // broadcast inside the group
                    mac_flag(acc[2 * c2], flag, a_cur[j], bv.x);
                    mac_flag(acc[2 * c2 + 1], flag, a_cur[j], bv.y);
                }
            }
            p++;
            if (p == pe) { // last pair of the tile: write row r of the tile, move on
                if (flag == 0xffffffffu) {
                    refold_row_exact<K>(a_vals, b_vals, pairs, pair_ptr[t], pe, r, acc);
                    atomicAdd(refold_count, 1ull);
                }
                ulonglong2 *out = reinterpret_cast<ulonglong2 *>(c_vals + (size_t)t * (K * K) + (size_t)r * K);
#pragma unroll
                for (int i = 0; i < K / 2; i++) out[i] = make_ulonglong2(acc[2 * i], acc[2 * i + 1]);
#pragma unroll
                for (int c = 0; c < K; c++) acc[c] = 0;
                flag = 0;
                t++;
                if (t < t_end) pe = pair_ptr[t + 1];
            }
        }
    }
}

template <int K>
int launch_rows(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, uint64_t *c_vals)
{
    using Cfg = RowsCfg<K>;
    const int groups_per_cta = (ROWS_THREADS / 32) * Cfg::GROUPS;
    // about four waves of chunks, but never chunks smaller than 8 tile products
    int64_t want_groups = (int64_t)ctx->sm_count * 2 * groups_per_cta * 4;
    int64_t Q = (sym->nprod + want_groups - 1) / want_groups;
    if (Q < 8) Q = 8;
    int64_t nchunks64 = (sym->nprod + Q - 1) / Q;
    if (nchunks64 < 1) nchunks64 = 1;
    const int nchunks = (int)nchunks64;
    int32_t *chunk_tile = nullptr;
    BSP_TRY(bsp_alloc(ctx, &chunk_tile, (size_t)nchunks + 1));
    chunk_bounds_kernel<<<(nchunks + 1 + 255) / 256, 256, 0, ctx->stream>>>(sym->pair_ptr, sym->nnzb_c, sym->nprod, Q,
                                                                            nchunks, chunk_tile);
    BSP_LAUNCH_CHECK(ctx);
    const int grid = (nchunks + groups_per_cta - 1) / groups_per_cta;
    numeric_rows_kernel<K><<<grid, ROWS_THREADS, 0, ctx->stream>>>(A->vals, B->vals, sym->pair_ptr, sym->pairs,
                                                                   chunk_tile, nchunks, c_vals,
                                                                   (unsigned long long *)ctx->d_counters);
    BSP_LAUNCH_CHECK(ctx);
    bsp_release(ctx, chunk_tile);
    return BSP_OK;
}

} // namespace

int bsp_numeric(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym,
                uint64_t *c_vals, int64_t *exact_refolds, int32_t *path)
{
    if (exact_refolds) *exact_refolds = 0;
    if (path) *path = 0;
    if (sym->nnzb_c == 0) return BSP_OK;
    const int k = A->k;
    if (!ctx->force_generic && (k == 4 || k == 8)) {
        BSP_CUDA(ctx, cudaMemsetAsync(ctx->d_counters, 0, sizeof(unsigned long long), ctx->stream));
        if (k == 8) BSP_TRY(launch_rows<8>(ctx, A, B, sym, c_vals));
        else BSP_TRY(launch_rows<4>(ctx, A, B, sym, c_vals));
        if (path) *path = k;
        if (exact_refolds) {
            unsigned long long *h = (unsigned long long *)((char *)ctx->pinned + 64);
            // the caller reads *h after its event sync (bsp_capi.cu)
            BSP_CUDA(ctx, cudaMemcpyAsync(h, ctx->d_counters, sizeof(*h), cudaMemcpyDeviceToHost, ctx->stream));
        }
        return BSP_OK;
    }
    const int E = k * k;
    const int slots = E <= GEN_THREADS ? GEN_THREADS / E : 1;
    int64_t grid = (sym->nnzb_c + slots - 1) / slots;
    const int64_t max_grid = (int64_t)ctx->sm_count * 64;
    if (grid > max_grid) grid = max_grid;
    numeric_generic_kernel<<<(unsigned)grid, GEN_THREADS, 0, ctx->stream>>>(A->vals, B->vals, sym->pair_ptr,
                                                                             sym->pairs, sym->nnzb_c, k, c_vals);
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}
