// bsp_noalloc.cu -- bsp_prune_zero_noalloc: the zero-tile compaction for callers that must
// not allocate device memory.
//
// Who this is for.  Some hosts of this library hand the whole GPU to their own arena when
// they start (a serving process with a fixed device-memory budget, an engine that sub-allocates
// one big slab itself).  Inside such a process a cudaMalloc from a library either fails outright
// or eats into memory the host believes it owns, and cudaMalloc/cudaFree also stall every stream
// of the device while they run -- not acceptable in the middle of a latency-bound loop.  Those
// callers already own their input and output buffers; the only thing bsp_prune_zero needs
// beyond them is scratch (one keep flag and one output position per tile).  Here that scratch
// is a fixed array that is part of the library image, so the call needs nothing from any
// allocator: no cudaMalloc, no cudaFree, no stream-ordered pool, no library that allocates.
//
// Cost and limits, so you can decide whether to link it in:
//   * the array is BSP_NOALLOC_MAX_TILES tiles deep (about 8 MiB); every GPU on which this
//     library is loaded carries it from load until the process exits;
//   * larger inputs are refused with BSP_ERR_LIMIT -- use bsp_prune_zero for those;
//   * there is ONE array per GPU, shared by every call and every context on that GPU.  Two
//     calls whose GPU work can overlap (different streams, different host threads) race on
//     it and corrupt each other's result; run them one after the other on one stream, or wait
//     for the first to finish.  The array keeps whatever the previous call left in it; the
//     call clears the small header it accumulates into, and rewrites every flag/position slot
//     it later reads, so stale contents never leak into a result;
//   * the first call on each GPU looks up the device address of the array
//     (cudaGetSymbolAddress, once per device, with that device current).  Make that first call
//     during start-up, before you begin any stream capture.
#include <mutex>

#include "bsp_internal.h"

namespace {

constexpr int64_t NA_MAX_TILES = BSP_NOALLOC_MAX_TILES;
constexpr size_t na_round256(size_t x) { return (x + 255) & ~(size_t)255; }
constexpr size_t NA_HDR_BYTES = 256;                                   // [0]: kept-tile total (int32)
constexpr size_t NA_KEEP_OFF = NA_HDR_BYTES;                           // int32 keep[NA_MAX_TILES + 1]
constexpr size_t NA_POS_OFF = NA_KEEP_OFF + na_round256((size_t)(NA_MAX_TILES + 1) * 4); // int32 pos[...]
constexpr size_t NA_SCRATCH_BYTES = NA_POS_OFF + na_round256((size_t)(NA_MAX_TILES + 1) * 4);

// The scratch itself.  In host code this name is only a handle for cudaGetSymbolAddress -- it
// is not a device pointer and must never be passed to a kernel as one.
static __device__ __align__(256) unsigned char g_na_scratch[NA_SCRATCH_BYTES];

constexpr int NA_MAX_DEVICES = 64;
std::mutex g_na_mutex;
unsigned char *g_na_addr[NA_MAX_DEVICES]; // device address of g_na_scratch, per device, 0 = not looked up yet

// keep[t] = 1 if tile t has any non-zero entry (the test of cu:580-586); one warp per tile
__global__ void __launch_bounds__(256) na_flag_kernel(const uint64_t *__restrict__ vals, int64_t nnzb, int kk,
                                                      int32_t *__restrict__ keep)
{
    const int lane = threadIdx.x & 31;
    const int64_t warp = ((int64_t)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int64_t nwarps = ((int64_t)gridDim.x * blockDim.x) >> 5;
    for (int64_t t = warp; t < nnzb; t += nwarps) {
        const uint64_t *v = vals + (size_t)t * kk;
        uint64_t acc = 0;
        for (int e = lane; e < kk; e += 32) acc |= v[e];
        const unsigned any = __ballot_sync(0xffffffffu, acc != 0);
        if (lane == 0) keep[t] = any ? 1 : 0;
    }
    if (warp == 0 && lane == 0) keep[nnzb] = 0; // sentinel so that pos[nnzb] = number kept
}

// exclusive scan of keep[0..n) into pos[0..n) by ONE CTA (n <= NA_MAX_TILES + 1); total -> *total
constexpr int NA_SCAN_THREADS = 1024;
constexpr int NA_SCAN_ITEMS = 8;
__global__ void __launch_bounds__(NA_SCAN_THREADS) na_scan_kernel(const int32_t *__restrict__ keep, int64_t n,
                                                                  int32_t *__restrict__ pos, int32_t *total)
{
    __shared__ int32_t warp_sums[32];
    __shared__ int32_t chunk_total;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int32_t carry = 0;
    for (int64_t base = 0; base < n; base += (int64_t)NA_SCAN_THREADS * NA_SCAN_ITEMS) {
        const int64_t i0 = base + (int64_t)threadIdx.x * NA_SCAN_ITEMS;
        int32_t v[NA_SCAN_ITEMS], s = 0;
#pragma unroll
        for (int i = 0; i < NA_SCAN_ITEMS; i++) {
            v[i] = (i0 + i < n) ? keep[i0 + i] : 0;
            s += v[i];
        }
        int32_t inc = s;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int32_t o = __shfl_up_sync(0xffffffffu, inc, d);
            if (lane >= d) inc += o;
        }
        if (lane == 31) warp_sums[warp] = inc;
        __syncthreads();
        if (warp == 0) {
            const int32_t w = warp_sums[lane];
            int32_t ws = w;
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const int32_t o = __shfl_up_sync(0xffffffffu, ws, d);
                if (lane >= d) ws += o;
            }
            warp_sums[lane] = ws - w;
            if (lane == 31) chunk_total = ws;
        }
        __syncthreads();
        int32_t run = carry + warp_sums[warp] + inc - s;
#pragma unroll
        for (int i = 0; i < NA_SCAN_ITEMS; i++) {
            if (i0 + i < n) pos[i0 + i] = run;
            run += v[i];
        }
        carry += chunk_total;
        __syncthreads();
    }
    if (threadIdx.x == 0) *total = carry;
}

__global__ void __launch_bounds__(256)
na_compact_kernel(const uint64_t *__restrict__ vals, const int32_t *__restrict__ colidx,
                  const int32_t *__restrict__ keep, const int32_t *__restrict__ pos, int64_t nnzb, int kk,
                  uint64_t *__restrict__ vals_out, int32_t *__restrict__ colidx_out)
{
    const int64_t total = nnzb * kk;
    for (int64_t g = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; g < total; g += (int64_t)gridDim.x * blockDim.x) {
        const int64_t t = g / kk;
        const int e = (int)(g - t * kk);
        if (keep[t]) {
            vals_out[(size_t)pos[t] * kk + e] = vals[g];
            if (e == 0) colidx_out[pos[t]] = colidx[t];
        }
    }
}

__global__ void __launch_bounds__(256) na_rowptr_kernel(const int32_t *__restrict__ rowptr, int32_t nbr,
                                                        const int32_t *__restrict__ pos,
                                                        int32_t *__restrict__ rowptr_out)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r <= nbr) rowptr_out[r] = pos[rowptr[r]];
}

} // namespace

extern "C" int bsp_prune_zero_noalloc(bsp_ctx *ctx, int32_t k, int32_t nbrows, int64_t nnzb,
                                      const int32_t *d_rowptr, const int32_t *d_colidx, const uint64_t *d_vals,
                                      int32_t *d_rowptr_out, int32_t *d_colidx_out, uint64_t *d_vals_out,
                                      int64_t *nnzb_out)
{
    if (!ctx || !nnzb_out) return BSP_ERR_ARG;
    if (k < 1 || nbrows < 0 || nnzb < 0 || !d_rowptr || !d_rowptr_out ||
        (nnzb > 0 && (!d_colidx || !d_vals || !d_colidx_out || !d_vals_out)))
        BSP_FAIL(ctx, BSP_ERR_ARG, "bsp_prune_zero_noalloc: bad arguments (k=%d nbrows=%d nnzb=%lld)", k, nbrows,
                 (long long)nnzb);
    if (nnzb > NA_MAX_TILES)
        BSP_FAIL(ctx, BSP_ERR_LIMIT,
                 "bsp_prune_zero_noalloc: %lld tiles, but the built-in scratch holds at most %lld "
                 "(BSP_NOALLOC_MAX_TILES); use bsp_prune_zero, which allocates, for larger matrices",
                 (long long)nnzb, (long long)NA_MAX_TILES);
    if (ctx->device < 0 || ctx->device >= NA_MAX_DEVICES)
        BSP_FAIL(ctx, BSP_ERR_LIMIT, "bsp_prune_zero_noalloc: device index %d not supported", ctx->device);
    BSP_CUDA(ctx, cudaSetDevice(ctx->device));

    // once per device, with that device current: device address of the scratch array
    unsigned char *scratch = nullptr;
    {
        std::lock_guard<std::mutex> lock(g_na_mutex);
        if (!g_na_addr[ctx->device]) {
            void *p = nullptr;
            BSP_CUDA(ctx, cudaGetSymbolAddress(&p, g_na_scratch));
            g_na_addr[ctx->device] = (unsigned char *)p;
        }
        scratch = g_na_addr[ctx->device];
    }
    int32_t *total = (int32_t *)scratch;
    int32_t *keep = (int32_t *)(scratch + NA_KEEP_OFF);
    int32_t *pos = (int32_t *)(scratch + NA_POS_OFF);
    const int kk = k * k;
    cudaStream_t s = ctx->stream;

    // the header is the only part accumulated into; keep/pos slots are all rewritten below
    BSP_CUDA(ctx, cudaMemsetAsync(scratch, 0, NA_HDR_BYTES, s));
    {
        int64_t g = (nnzb * 32 + 255) / 256;
        if (g > (int64_t)ctx->sm_count * 16) g = (int64_t)ctx->sm_count * 16;
        if (g < 1) g = 1;
        na_flag_kernel<<<(unsigned)g, 256, 0, s>>>(d_vals, nnzb, kk, keep);
        BSP_LAUNCH_CHECK(ctx);
    }
    na_scan_kernel<<<1, NA_SCAN_THREADS, 0, s>>>(keep, nnzb + 1, pos, total);
    BSP_LAUNCH_CHECK(ctx);
    if (nnzb > 0) {
        int64_t g = (nnzb * kk + 255) / 256;
        if (g > (int64_t)ctx->sm_count * 32) g = (int64_t)ctx->sm_count * 32;
        na_compact_kernel<<<(unsigned)g, 256, 0, s>>>(d_vals, d_colidx, keep, pos, nnzb, kk, d_vals_out, d_colidx_out);
        BSP_LAUNCH_CHECK(ctx);
    }
    na_rowptr_kernel<<<(unsigned)(((int64_t)nbrows + 1 + 255) / 256), 256, 0, s>>>(d_rowptr, nbrows, pos, d_rowptr_out);
    BSP_LAUNCH_CHECK(ctx);
    int32_t *h_total = (int32_t *)ctx->pinned; // pinned word owned by the context since bsp_ctx_create
    BSP_CUDA(ctx, cudaMemcpyAsync(h_total, total, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
    BSP_CUDA(ctx, cudaStreamSynchronize(s));
    *nnzb_out = *h_total;
    return BSP_OK;
}
