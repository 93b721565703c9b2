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

// Any tile edge k.  256-thread CTAs; when k*k <= 256 a CTA works on 256/(k*k) output tiles
// at a time with one thread per element, otherwise on one tile with threads striding over
// its elements.
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

} // namespace

int bsp_numeric(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym,
                uint64_t *c_vals, int64_t *exact_refolds, int32_t *path)
{
    if (exact_refolds) *exact_refolds = 0;
    if (path) *path = 0;
    if (sym->nnzb_c == 0) return BSP_OK;
    const int k = A->k;
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
