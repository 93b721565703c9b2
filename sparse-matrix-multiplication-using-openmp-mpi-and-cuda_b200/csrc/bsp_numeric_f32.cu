// bsp_numeric_f32.cu -- numeric pass for float32 tiles (BASELINE.json configs[4]).
//
// An extension: the reference reads and multiplies uint64 only (cu:375-376, cu:44-66), so
// there is no reference path to be identical to.  Same pair lists as the integer pass (the
// symbolic pass is type-blind); per output tile the products A_tile * B_tile are summed in pair
// order with fp32 accumulation.  Two kernels:
//   f32_simt_kernel<K>  K = 32 / 64: one CTA of (K/4)^2 threads per output tile, 4x4 register
//                       block per thread, A and B tiles staged in shared memory;
//   f32_generic_kernel  any K: one thread per output element.
// Stated tolerance against a float64 evaluation (tests/test_gpu_f32.py): |err| <= 1e-5 *
// (|A||B|) row-column bound for the SIMT kernels (fp32 FMA rounding only).
#include "bsp_internal.h"

namespace {

__global__ void __launch_bounds__(256)
f32_generic_kernel(const float *__restrict__ a_vals, const float *__restrict__ b_vals,
                   const int64_t *__restrict__ pair_ptr, const int2 *__restrict__ pairs, int64_t nnzb_c, int k,
                   float *__restrict__ c_vals)
{
    const int E = k * k;
    for (int64_t t = blockIdx.x; t < nnzb_c; t += gridDim.x) {
        const int64_t p0 = pair_ptr[t], p1 = pair_ptr[t + 1];
        for (int e = threadIdx.x; e < E; e += blockDim.x) {
            const int r = e / k, c = e % k;
            float acc = 0.f;
            for (int64_t p = p0; p < p1; p++) {
                const int2 pr = pairs[p];
                const float *A = a_vals + (size_t)pr.x * E + (size_t)r * k;
                const float *B = b_vals + (size_t)(pr.y & 0x7fffffff) * E + c;
                for (int j = 0; j < k; j++) acc = fmaf(A[j], B[(size_t)j * k], acc);
            }
            c_vals[(size_t)t * E + e] = acc;
        }
    }
}

template <int K>
__global__ void __launch_bounds__((K / 4) * (K / 4))
f32_simt_kernel(const float *__restrict__ a_vals, const float *__restrict__ b_vals,
                const int64_t *__restrict__ pair_ptr, const int2 *__restrict__ pairs, int64_t nnzb_c,
                float *__restrict__ c_vals)
{
    constexpr int T = (K / 4) * (K / 4);           // threads per tile
    constexpr int V = K * K / 4 / T;               // float4 per thread per tile (= 4)
    __shared__ float As[K][K + 1];                 // +1: the column reads below hit distinct banks
    __shared__ __align__(16) float Bs[K][K];
    const int tid = threadIdx.x, tx = tid % (K / 4), ty = tid / (K / 4);
    for (int64_t t = blockIdx.x; t < nnzb_c; t += gridDim.x) {
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; i++)
#pragma unroll
            for (int j = 0; j < 4; j++) acc[i][j] = 0.f;
        const int64_t p0 = pair_ptr[t], p1 = pair_ptr[t + 1];
        for (int64_t p = p0; p < p1; p++) {
            const int2 pr = pairs[p];
            const float4 *A4 = reinterpret_cast<const float4 *>(a_vals + (size_t)pr.x * (K * K));
            const float4 *B4 = reinterpret_cast<const float4 *>(b_vals + (size_t)(pr.y & 0x7fffffff) * (K * K));
            __syncthreads(); // the previous product is done with the tiles
#pragma unroll
            for (int i = 0; i < V; i++) {
                const int x = tid + i * T; // float4 number inside the tile
                const float4 a = A4[x], b = B4[x];
                const int r = x / (K / 4), c4 = (x % (K / 4)) * 4;
                As[r][c4] = a.x; As[r][c4 + 1] = a.y; As[r][c4 + 2] = a.z; As[r][c4 + 3] = a.w;
                *reinterpret_cast<float4 *>(&Bs[r][c4]) = b;
            }
            __syncthreads();
#pragma unroll 8
            for (int j = 0; j < K; j++) {
                const float4 b = *reinterpret_cast<const float4 *>(&Bs[j][4 * tx]);
#pragma unroll
                for (int rr = 0; rr < 4; rr++) {
                    const float a = As[4 * ty + rr][j];
                    acc[rr][0] = fmaf(a, b.x, acc[rr][0]);
                    acc[rr][1] = fmaf(a, b.y, acc[rr][1]);
                    acc[rr][2] = fmaf(a, b.z, acc[rr][2]);
                    acc[rr][3] = fmaf(a, b.w, acc[rr][3]);
                }
            }
        }
        float *out = c_vals + (size_t)t * (K * K);
#pragma unroll
        for (int rr = 0; rr < 4; rr++)
            *reinterpret_cast<float4 *>(out + (size_t)(4 * ty + rr) * K + 4 * tx) =
                make_float4(acc[rr][0], acc[rr][1], acc[rr][2], acc[rr][3]);
    }
}

} // namespace

int bsp_numeric_f32_tc(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, float *c_vals,
                       int32_t *path); // bsp_numeric_f32_tc.cu; returns BSP_ERR_LIMIT when it has no kernel for this k

int bsp_numeric_f32(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, float *c_vals,
                    int32_t *path)
{
    if (path) *path = 100;
    if (sym->nnzb_c == 0) return BSP_OK;
    const int k = A->k;
    const float *a = (const float *)A->vals, *b = (const float *)B->vals;
    // auto: tensor cores for k = 32 and 64 (bench workloads: 4.7 against 7.9 ms SIMT, 2.4 against 6.4 ms); TF32 inputs --
    // path = 1 keeps fp32 products
    if ((ctx->f32_path == 2 || ctx->f32_path == 0) && (k == 32 || k == 64)) {
        const int rc = bsp_numeric_f32_tc(ctx, A, B, sym, c_vals, path);
        if (rc != BSP_ERR_LIMIT) return rc;
        if (ctx->f32_path == 2) return rc; // tensor cores were demanded
    } else if (ctx->f32_path == 2) {
        BSP_FAIL(ctx, BSP_ERR_LIMIT, "float32 tensor-core kernel exists for k = 32 and 64 only (k = %d)", k);
    }
    int64_t grid = sym->nnzb_c;
    const int64_t max_grid = (int64_t)ctx->sm_count * 32;
    if (grid > max_grid) grid = max_grid;
    if (k == 32)
        f32_simt_kernel<32><<<(unsigned)grid, 64, 0, ctx->stream>>>(a, b, sym->pair_ptr, sym->pairs, sym->nnzb_c, c_vals);
    else if (k == 64)
        f32_simt_kernel<64><<<(unsigned)grid, 256, 0, ctx->stream>>>(a, b, sym->pair_ptr, sym->pairs, sym->nnzb_c, c_vals);
    else
        f32_generic_kernel<<<(unsigned)grid, 256, 0, ctx->stream>>>(a, b, sym->pair_ptr, sym->pairs, sym->nnzb_c, k, c_vals);
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}
