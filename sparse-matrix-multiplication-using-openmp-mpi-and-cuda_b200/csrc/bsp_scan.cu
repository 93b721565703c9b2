// bsp_scan.cu -- exclusive prefix sums used by the symbolic pass (row pointers of C, per-row
// product offsets) and by the zero-tile prune (CSR-of-blocks compaction on device).
// Three launches: per-CTA sums, scan of the sums by one CTA, per-CTA scan + offset.
// HBM-bound: 2 reads + 1 write of the array.
#include "bsp_internal.h"

namespace {

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_TILE = SCAN_THREADS * SCAN_ITEMS;

template <typename T>
__device__ __forceinline__ T warp_incl_scan(T v)
{
    const int lane = threadIdx.x & 31;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        T o = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += o;
    }
    return v;
}

// inclusive scan across the CTA of one value per thread; returns inclusive value, total in *total
template <typename T>
__device__ __forceinline__ T block_incl_scan(T v, T *smem /* >= 33 entries */, T *total)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarp = blockDim.x >> 5;
    T inc = warp_incl_scan(v);
    if (lane == 31) smem[warp] = inc;
    __syncthreads();
    if (warp == 0) {
        T w = lane < nwarp ? smem[lane] : T(0);
        T ws = warp_incl_scan(w);
        smem[lane] = ws - w; // exclusive offset of each warp
        if (lane == 31) smem[32] = ws;
    }
    __syncthreads();
    T res = inc + smem[warp];
    *total = smem[32];
    __syncthreads();
    return res;
}

template <typename Tout>
__global__ void __launch_bounds__(SCAN_THREADS) scan_sums_kernel(const int32_t *__restrict__ in, int64_t n,
                                                                 Tout *__restrict__ sums)
{
    __shared__ Tout sm[33];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    Tout s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++)
        if (base + i < n) s += (Tout)in[base + i];
    Tout tot;
    block_incl_scan(s, sm, &tot);
    if (threadIdx.x == 0) sums[blockIdx.x] = tot;
}

template <typename Tout>
__global__ void __launch_bounds__(SCAN_THREADS) scan_top_kernel(Tout *__restrict__ sums, int64_t nb,
                                                                Tout *__restrict__ total_out)
{
    __shared__ Tout sm[33];
    Tout carry = 0;
    for (int64_t b0 = 0; b0 < nb; b0 += SCAN_THREADS) {
        int64_t i = b0 + threadIdx.x;
        Tout v = i < nb ? sums[i] : Tout(0);
        Tout tot;
        Tout inc = block_incl_scan(v, sm, &tot);
        if (i < nb) sums[i] = carry + inc - v;
        carry += tot;
    }
    if (threadIdx.x == 0 && total_out) *total_out = carry;
}

template <typename Tout>
__global__ void __launch_bounds__(SCAN_THREADS) scan_final_kernel(const int32_t *in, int64_t n,
                                                                  const Tout *__restrict__ sums,
                                                                  Tout *out /* may alias in */)
{
    __shared__ Tout sm[33];
    int64_t base = (int64_t)blockIdx.x * SCAN_TILE + (int64_t)threadIdx.x * SCAN_ITEMS;
    Tout v[SCAN_ITEMS];
    Tout s = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        v[i] = (base + i < n) ? (Tout)in[base + i] : Tout(0);
        s += v[i];
    }
    Tout tot;
    Tout inc = block_incl_scan(s, sm, &tot);
    Tout run = sums[blockIdx.x] + inc - s;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; i++) {
        if (base + i < n) out[base + i] = run;
        run += v[i];
    }
}

template <typename Tout>
int scan_impl(bsp_ctx *ctx, const int32_t *in, Tout *out, int64_t n, Tout *total_dev)
{
    if (n <= 0) {
        if (total_dev) BSP_CUDA(ctx, cudaMemsetAsync(total_dev, 0, sizeof(Tout), ctx->stream));
        return BSP_OK;
    }
    int64_t nb = (n + SCAN_TILE - 1) / SCAN_TILE;
    Tout *sums = nullptr;
    BSP_TRY(bsp_alloc(ctx, &sums, (size_t)nb));
    scan_sums_kernel<Tout><<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums);
    BSP_LAUNCH_CHECK(ctx);
    scan_top_kernel<Tout><<<1, SCAN_THREADS, 0, ctx->stream>>>(sums, nb, total_dev);
    BSP_LAUNCH_CHECK(ctx);
    scan_final_kernel<Tout><<<(unsigned)nb, SCAN_THREADS, 0, ctx->stream>>>(in, n, sums, out);
    BSP_LAUNCH_CHECK(ctx);
    bsp_release(ctx, sums);
    return BSP_OK;
}

} // namespace

int bsp_scan_i32(bsp_ctx *ctx, const int32_t *in, int32_t *out, int64_t n, int32_t *total_dev)
{
    return scan_impl<int32_t>(ctx, in, out, n, total_dev);
}

int bsp_scan_i32_to_i64(bsp_ctx *ctx, const int32_t *in, int64_t *out, int64_t n)
{
    return scan_impl<int64_t>(ctx, in, out, n, nullptr);
}

// ---- scalar read-backs without the copy engine (bsp_internal.h) -------------------------------------
namespace {
__global__ void publish_kernel(const __grid_constant__ bsp_publish_list list, uint32_t *__restrict__ host_words)
{
    for (int s = 0; s < list.n; s++) {
        const uint32_t *src = (const uint32_t *)list.src[s];
        uint32_t *dst = host_words + list.offset[s] / 4;
        for (uint32_t i = threadIdx.x; i < list.words[s]; i += blockDim.x) dst[i] = src[i];
    }
    __threadfence_system();
}
} // namespace

int bsp_publish(bsp_ctx *ctx, const bsp_publish_list &list)
{
    for (int s = 0; s < list.n; s++)
        if ((list.offset[s] & 3) || (size_t)list.offset[s] + 4 * (size_t)list.words[s] > ctx->pinned_bytes)
            BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_publish: run %d does not fit the pinned area", s);
    publish_kernel<<<1, 32, 0, ctx->stream>>>(list, (uint32_t *)ctx->pinned_dev);
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}
