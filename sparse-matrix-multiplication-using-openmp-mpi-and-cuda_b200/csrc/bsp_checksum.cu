// bsp_checksum.cu -- content checksum of a block-sparse matrix, computed where the matrix lives.
//
// SURVEY.md 8(f) row 3 ("binary on-disk format + checksum mode"); no counterpart in the
// reference, whose only way to compare two results is to write and diff the text files
// (35-50 GB for BASELINE.json configs[1]).  The checksum is a function of the matrix CONTENT --
// header, tile coordinates, tile values -- and not of how it was produced, so a product computed
// by row panels on eight GPUs and one computed by one GPU agree, and tests can compare a full-size
// result with a value obtained elsewhere without moving 19 GB to the host.
//
// Definition (tests/util.py restates it in numpy):
//   words  w_0..w_{W-1} of a tile = its bytes as little-endian 32-bit words, W = k*k*esize/4
//   s1 = fmix32(brow*0x9E3779B1 + bcol*0x85EBCA77 + 0x27D4EB2F)      (32-bit arithmetic)
//   s2 = fmix32(s1 ^ 0x165667B1)                                    fmix32 = MurmurHash3 finaliser
//   m_i = (s1 + i*0x9E3779B1) | 1                                    (32-bit, odd)
//   tile  = (s2 << 32 | s1) + sum_i  w_i * m_i                       (64-bit, the products are 32x32 -> 64)
//   body  = sum over tiles                                           (mod 2^64: order-independent)
//   checksum = splitmix64(body ^ header),  header = splitmix64 chain over rows, cols, k, esize, nnzb
// It is a position-keyed linear sum: a cheap integrity check, not a cryptographic hash.
//
// One streaming pass over vals (HBM-bound: 16 B loads, one IMAD.WIDE per word), plus the row
// expansion (4 B per tile).
#include "bsp_internal.h"

namespace {

__host__ __device__ inline uint32_t fmix32(uint32_t h)
{
    h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
    return h;
}

inline uint64_t splitmix64(uint64_t x)
{
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}

constexpr int CS_THREADS = 256;

// VEC = 4: W % 4 == 0, one uint4 per thread and iteration; VEC = 1: any W, one word.
// w4 = W / VEC; shift >= 0 when w4 is a power of two.
template <int VEC>
__global__ void __launch_bounds__(CS_THREADS)
checksum_kernel(const uint32_t *__restrict__ vals, const int32_t *__restrict__ brow, const int32_t *__restrict__ bcol,
                int64_t nnzb, uint32_t w4, int shift, unsigned long long *__restrict__ out)
{
    const uint64_t total = (uint64_t)nnzb * w4;
    uint64_t h = 0;
    for (uint64_t q = (uint64_t)blockIdx.x * CS_THREADS + threadIdx.x; q < total; q += (uint64_t)gridDim.x * CS_THREADS) {
        const uint64_t t = shift >= 0 ? q >> shift : q / w4;
        const uint32_t i0 = (uint32_t)(q - t * w4) * VEC;
        const uint32_t s1 = fmix32((uint32_t)brow[t] * 0x9E3779B1u + (uint32_t)bcol[t] * 0x85EBCA77u + 0x27D4EB2Fu);
        uint32_t m = s1 + i0 * 0x9E3779B1u;
        if (i0 == 0) h += ((uint64_t)fmix32(s1 ^ 0x165667B1u) << 32) | s1;
        if (VEC == 4) {
            const uint4 v = reinterpret_cast<const uint4 *>(vals)[q];
            h += (uint64_t)v.x * (m | 1u); m += 0x9E3779B1u;
            h += (uint64_t)v.y * (m | 1u); m += 0x9E3779B1u;
            h += (uint64_t)v.z * (m | 1u); m += 0x9E3779B1u;
            h += (uint64_t)v.w * (m | 1u);
        } else {
            h += (uint64_t)vals[q] * (m | 1u);
        }
    }
#pragma unroll
    for (int d = 16; d; d >>= 1) h += __shfl_xor_sync(0xffffffffu, h, d);
    __shared__ unsigned long long part[CS_THREADS / 32];
    if ((threadIdx.x & 31) == 0) part[threadIdx.x >> 5] = h;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long s = 0;
        for (int i = 0; i < CS_THREADS / 32; i++) s += part[i];
        atomicAdd(out, s);
    }
}

} // namespace

extern "C" uint64_t bsp_checksum_finish(int32_t rows, int32_t cols, int32_t k, int32_t esize, int64_t nnzb, uint64_t body)
{
    uint64_t header = splitmix64((uint64_t)(uint32_t)rows);
    header = splitmix64(header ^ (uint64_t)(uint32_t)cols);
    header = splitmix64(header ^ (uint64_t)(uint32_t)k);
    header = splitmix64(header ^ (uint64_t)(uint32_t)esize);
    header = splitmix64(header ^ (uint64_t)nnzb);
    return splitmix64(body ^ header);
}

extern "C" int bsp_mat_checksum_body(bsp_ctx *ctx, const bsp_mat *m, uint64_t *out)
{
    if (!ctx || !m || !out) return BSP_ERR_ARG;
    if (m->device != ctx->device) BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_mat_checksum: matrix lives on device %d, context on %d", m->device, ctx->device);
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));
    uint64_t body = 0;
    if (m->nnzb > 0) {
        const uint32_t W = (uint32_t)(m->k * m->k * (m->esize / 4));
        const int vec = (W % 4 == 0) ? 4 : 1;
        const uint32_t w4 = W / vec;
        int shift = -1;
        if ((w4 & (w4 - 1)) == 0) { shift = 0; while ((1u << shift) < w4) shift++; }
        bsp_scratch sc(ctx);
        int32_t *d_brow = nullptr;
        unsigned long long *d_sum = nullptr;
        BSP_TRY(sc.alloc(&d_brow, (size_t)m->nnzb));
        BSP_TRY(sc.alloc(&d_sum, 1));
        BSP_TRY(bsp_expand_rows(ctx, m, d_brow));
        BSP_CUDA(ctx, cudaMemsetAsync(d_sum, 0, sizeof(unsigned long long), ctx->stream));
        const uint64_t total = (uint64_t)m->nnzb * w4;
        uint64_t grid = (total + CS_THREADS - 1) / CS_THREADS;
        const uint64_t max_grid = (uint64_t)ctx->sm_count * 16;
        if (grid > max_grid) grid = max_grid;
        if (vec == 4)
            checksum_kernel<4><<<(unsigned)grid, CS_THREADS, 0, ctx->stream>>>((const uint32_t *)m->vals, d_brow, m->colidx,
                                                                              m->nnzb, w4, shift, d_sum);
        else
            checksum_kernel<1><<<(unsigned)grid, CS_THREADS, 0, ctx->stream>>>((const uint32_t *)m->vals, d_brow, m->colidx,
                                                                              m->nnzb, w4, shift, d_sum);
        BSP_LAUNCH_CHECK(ctx);
        unsigned long long *h = (unsigned long long *)((char *)ctx->pinned + 128);
        BSP_CUDA(ctx, cudaMemcpyAsync(h, d_sum, sizeof(unsigned long long), cudaMemcpyDeviceToHost, ctx->stream));
        BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        body = *h;
    }
    *out = body;
    return BSP_OK;
}

extern "C" int bsp_mat_checksum(bsp_ctx *ctx, const bsp_mat *m, uint64_t *out)
{
    if (!ctx || !m || !out) return BSP_ERR_ARG;
    uint64_t body = 0;
    BSP_TRY(bsp_mat_checksum_body(ctx, m, &body));
    *out = bsp_checksum_finish(m->rows, m->cols, m->k, m->esize, m->nnzb, body);
    return BSP_OK;
}
