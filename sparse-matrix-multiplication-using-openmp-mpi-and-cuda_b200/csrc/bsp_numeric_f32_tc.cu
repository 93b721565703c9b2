// bsp_numeric_f32_tc.cu -- float32 tile products on the 5th-generation tensor cores.
//
// BASELINE.json configs[4] (k = 32 / 64 float tiles; an extension, the reference has no float
// path).  Per output tile the products A_tile * B_tile of its pair list are accumulated by
// tcgen05.mma (kind::tf32: fp32 storage read as TF32, fp32 accumulate) into one TMEM
// accumulator and read back once with tcgen05.ld -- the tile never leaves the tensor memory
// between pairs.
//
// One CTA (4 warps) walks a stream of tile products, two shared-memory stages and two TMEM accumulators:
//   repacking   : both operands are first rewritten, tile by tile, into the canonical no-swizzle K-major UMMA layout
//                 (8-row x 16-byte core matrices), A as it is and B transposed:
//                   A: element (m, k) at (k/4)*(K*16) + m*16 + (k%4)*4      LBO = K*16, SBO = 128
//                   B: element (n, k) at (k/4)*(K*16) + n*16 + (k%4)*4      (the instruction wants B as N x K, K contiguous)
//                 (An MN-major B descriptor would avoid the transpose; with no-swizzle layouts it
//                 produced zeros in a probe on this toolchain -- tools/scratch/umma_probe.cu.)
//   one thread  : starts the two bulk copies of product i+1 (a packed tile is one contiguous run), waits for those of
//                 product i on the stage's mbarrier, issues K/8 tcgen05.mma (M=64, N=K, K=8 each) and commits them
//   all threads : read an accumulator back one product after its tile closed -- while the next tile is multiplied
//                 into the other accumulator -- with the 16-lane load shape and store it in whole 32-byte sectors.
// The pair of product i+2 and the pair_ptr bounds of the next chunk of tiles are fetched ahead, so the loop waits for
// no index load.  For k = 32 the M=64 instruction runs with its upper 32 rows unused.
// BSP_F32_PROF=1 prints where thread 0 of every CTA spends its cycles (the numbers DESIGN.md section 8 quotes).
#include <algorithm>
#include <vector>

#include "bsp_internal.h"

namespace {

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity)
{
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(bar), "r"(parity) : "memory");
}
// shared-memory matrix descriptor, no swizzle (cute::UMMA::SmemDescriptor): start >> 4,
// leading byte offset >> 4 at bit 16, stride byte offset >> 4 at bit 32, version 1 at bit 46
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo)
{
    return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) |
           ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | (1ull << 46);
}

// instruction descriptor (cute::UMMA::InstrDescriptor): D = F32, A = B = TF32, both K-major,
// N >> 3 at bit 17, M >> 4 at bit 24
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N)
{
    return (1u << 4) | (2u << 7) | (2u << 10) | (0u << 15) | (0u << 16) | ((uint32_t)(N >> 3) << 17) |
           ((uint32_t)(M >> 4) << 24);
}

// 16 TMEM lanes x (8 * n) columns per warp, load and wait in one statement (no use of the registers can be scheduled
// between them).  Lane i gets, for every group j of 8 columns: v[4j], v[4j+1] = row i/4, columns 8j + 2(i%4) + {0, 1};
// v[4j+2], v[4j+3] = row i/4 + 8, same columns -- four lanes hold 32 contiguous bytes of a row.
__device__ __forceinline__ void tmem_ld_16x256b_x4(uint32_t taddr, uint32_t *v)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x4.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tmem_ld_16x256b_x8(uint32_t taddr, uint32_t *v)
{
    asm volatile(
        "tcgen05.ld.sync.aligned.16x256b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n\t"
        "tcgen05.wait::ld.sync.aligned;"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}

// One entry of the CTA's product stream.  tb: -2 = the stream has ended, -1 = next product of the current chunk,
// >= 0 = first product of a new chunk whose first output tile is tb.
struct tc_entry {
    int2 pr;
    int64_t tb;
};

// ---- operand tiles packed in the UMMA layout ----------------------------------------------------------------------
// Two small kernels first rewrite the operands tile by tile into the layout the instruction reads (A: [k/4][m] 16-byte
// chunks; B: transposed, [k/4][n] chunks), so that a tile is one contiguous run of 4*K*K bytes and ONE thread fetches it
// with a single cp.async.bulk that signals the stage's mbarrier.  (The version before this one staged the tiles with all
// threads -- 16 LDGSTS/LDG and 8 STS per thread and product, B transposed in registers: 3.66 ms on the k = 64 bench
// workload against 2.15 ms + 0.23 ms of repacking here; the bytes an SM keeps in flight through its load/store unit
// were the limit, more resident CTAs only made each one slower.)

template <int K>
__global__ void __launch_bounds__(256)
f32_pack_a_kernel(const float4 *__restrict__ src, float4 *__restrict__ dst, int64_t nchunks_total)
{
    constexpr int CH = K * K / 4; // 16-byte chunks per tile
    for (int64_t g = (int64_t)blockIdx.x * 256 + threadIdx.x; g < nchunks_total; g += (int64_t)gridDim.x * 256) {
        const int64_t tile = g / CH;
        const int o = (int)(g % CH);   // output chunk: kc * K + m
        const int kc = o / K, m = o % K;
        dst[g] = src[tile * CH + m * (K / 4) + kc]; // rows of one tile stay in L1 between the K/4 passes over them
    }
}

template <int K>
__global__ void __launch_bounds__(256)
f32_pack_b_kernel(const float *__restrict__ src, float4 *__restrict__ dst, int64_t nchunks_total)
{
    constexpr int CH = K * K / 4;
    for (int64_t g = (int64_t)blockIdx.x * 256 + threadIdx.x; g < nchunks_total; g += (int64_t)gridDim.x * 256) {
        const int64_t tile = g / CH;
        const int o = (int)(g % CH);   // output chunk: kc * K + n  =  B[4kc .. 4kc+3][n]
        const int kc = o / K, n = o % K;
        const float *s = src + tile * (K * K) + (size_t)(4 * kc) * K + n;
        dst[g] = make_float4(s[0], s[K], s[2 * K], s[3 * K]);
    }
}

__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void *src, uint32_t bytes, uint32_t bar)
{
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
                 "l"(src), "r"(bytes), "r"(bar) : "memory");
}

template <int K, bool PROF>
__global__ void __launch_bounds__(128)
f32_tcb_kernel(const float *__restrict__ a_pack, const float *__restrict__ b_pack, const int64_t *__restrict__ pair_ptr,
               const int2 *__restrict__ pairs, int64_t nnzb_c, int tch, float *__restrict__ c_vals, long long *__restrict__ prof)
{
    constexpr int M = 64, N = K;
    constexpr int TILE = K * K * 4;   // bytes of one packed tile
    constexpr int STAGE = 2 * TILE;   // A then B.  k = 32: the M=64 instruction reads 32 rows past the A tile's, i.e. the
                                      // next chunk column or the head of B -- finite values in accumulator rows nobody reads
    constexpr int TMEM_COLS = 2 * N;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t mbar[4]; // [0..1] stage filled (bulk copies landed), [2..3] stage's MMAs retired
    __shared__ uint32_t tmem_base_sh;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t barf = smem_u32(&mbar[0]), bard = smem_u32(&mbar[2]);

    if (tid == 0) {
        for (int i = 0; i < 4; i++) mbar_init(barf + 8 * i, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_sh)),
                     "n"(TMEM_COLS) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem = tmem_base_sh;
    constexpr uint32_t idesc = umma_idesc_tf32(M, N);

    // The CTA's work is one stream of tile products.  Output tiles are dealt to the CTAs in chunks of `tch`
    // consecutive tiles (chunk c = blockIdx.x, += gridDim.x): the products of a chunk are one contiguous run of the
    // pair stream, the sign bit of pair.y ends a tile, so inside a chunk the stream is walked without pair_ptr.
    // The generator below runs two products ahead of the multiplier, and the pair_ptr bounds of the following chunk
    // are fetched a whole chunk ahead: no index load is ever waited for.
    const int64_t nchunks = (nnzb_c + tch - 1) / tch;
    int64_t g_chunk = blockIdx.x, g_p = 0, g_pend = 0;
    int64_t n_chunk = g_chunk + gridDim.x, n_b = 0, n_e = 0;
    bool g_new = true;
    auto chunk_bounds = [&](int64_t c, int64_t &b, int64_t &e) {
        const int64_t t0 = c * tch, t1 = (c + 1) * tch < nnzb_c ? (c + 1) * tch : nnzb_c;
        b = pair_ptr[t0];
        e = pair_ptr[t1];
    };
    if (g_chunk < nchunks) chunk_bounds(g_chunk, g_p, g_pend); else g_chunk = -1;
    if (g_chunk >= 0 && n_chunk < nchunks) chunk_bounds(n_chunk, n_b, n_e);
    auto gen = [&]() -> tc_entry {
        tc_entry e;
        e.pr = make_int2(0, 0);
        e.tb = -2;
        if (g_chunk < 0) return e;
        if (g_p >= g_pend) {
            if (n_chunk >= nchunks) { g_chunk = -1; return e; }
            g_chunk = n_chunk; g_p = n_b; g_pend = n_e;
            n_chunk += gridDim.x;
            if (n_chunk < nchunks) chunk_bounds(n_chunk, n_b, n_e);
            g_new = true;
        }
        e.tb = g_new ? g_chunk * tch : -1;
        g_new = false;
        e.pr = pairs[g_p++];
        return e;
    };

    auto fill = [&](int st, const int2 pr) { // one thread: both tiles of a product into stage st
        const uint32_t fb = barf + 8 * st, dst = sbase + st * STAGE;
        mbar_expect_tx(fb, 2 * TILE);
        bulk_g2s(dst, a_pack + (size_t)pr.x * (K * K), TILE, fb);
        bulk_g2s(dst + TILE, b_pack + (size_t)(pr.y & 0x7fffffff) * (K * K), TILE, fb);
    };
    // read accumulator `ab` back and store it as output tile `tile`: the M=64 accumulator layout keeps tile rows
    // 16w .. 16w+15 in TMEM lanes 32w .. 32w+15, which is what warp w reads with the 16-lane load shape; four lanes then
    // hold 32 contiguous bytes of a row, so every store instruction writes whole 32-byte sectors with all lanes
    auto epilogue = [&](int64_t tile, int ab) {
        float *out = c_vals + (size_t)tile * (K * K) + (size_t)(16 * warp + (lane >> 2)) * K + 2 * (lane & 3);
        uint32_t v[N / 2];
        const uint32_t taddr = tmem + ((uint32_t)(32 * warp) << 16) + (uint32_t)(ab * N);
        if constexpr (N == 64) tmem_ld_16x256b_x8(taddr, v); else tmem_ld_16x256b_x4(taddr, v);
        if (16 * warp < K) {
#pragma unroll
            for (int j = 0; j < N / 8; j++) {
                __stcs(reinterpret_cast<float2 *>(out + 8 * j), make_float2(__uint_as_float(v[4 * j]), __uint_as_float(v[4 * j + 1])));
                __stcs(reinterpret_cast<float2 *>(out + 8 * K + 8 * j),
                       make_float2(__uint_as_float(v[4 * j + 2]), __uint_as_float(v[4 * j + 3])));
            }
        }
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    };

    // Product i (stage s = i & 1, accumulator ab):
    //   1  CTA barrier (orders the last read-back before MMAs into that accumulator)
    //   2  one thread waits for the stage's bulk copies, issues the MMAs of product i and commits them
    //   3  all wait for the MMAs of product i-1: stage s^1 is free, the accumulator of a tile closed by i-1 is complete
    //   4  one thread starts the bulk copies of product i+1 into stage s^1
    //   5  if product i-1 closed a tile: read it back and store it, while product i is multiplied
    uint32_t uses0 = 0, uses1 = 0, fills0 = 0, fills1 = 0;
    tc_entry cur = gen();
    tc_entry nx1 = gen();
    int64_t t = cur.tb;
    int s = 0, ab = 0;
    bool first_of_tile = true, pend = false;
    int64_t pend_t = 0;
    int pend_ab = 0;
    if (cur.tb != -2) {
        if (tid == 0) fill(0, cur.pr);
        fills0++;
    }
    long long pc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    long long tk = 0;
    auto tick = [&](int slot) {
        if (PROF && tid == 0) { const long long now = clock64(); pc[slot] += now - tk; tk = now; }
    };
    if (PROF && tid == 0) tk = clock64();
    while (cur.tb != -2) {
        const bool have_next = nx1.tb != -2;
        const bool last_of_tile = cur.pr.y < 0;
        const tc_entry nx2 = gen();
        tick(0);
        __syncthreads();
        tick(1);
        if (tid == 0) {
            mbar_wait(barf + 8 * s, ((s ? fills1 : fills0) - 1) & 1);
            if (PROF) tick(2);
            const uint32_t a_s = sbase + s * STAGE, b_s = a_s + TILE;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int ks = 0; ks < K / 8; ks++) {
                const uint64_t da = umma_desc(a_s + ks * 2 * (K * 16), K * 16, 128);
                const uint64_t db = umma_desc(b_s + ks * 2 * (K * 16), K * 16, 128);
                const uint32_t accumulate = (!first_of_tile || ks > 0) ? 1u : 0u;
                asm volatile(
                    "{\n\t"
                    ".reg .pred p;\n\t"
                    "setp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
                    "}" ::"r"(tmem + (uint32_t)(ab * N)), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bard + 8 * s)
                         : "memory");
        }
        if (s) uses1++; else uses0++;
        tick(3);
        {
            const uint32_t uprev = s ? uses0 : uses1;
            if (uprev > 0) mbar_wait(bard + 8 * (s ^ 1), (uprev - 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        }
        tick(4);
        if (have_next) {
            if (tid == 0) fill(s ^ 1, nx1.pr);
            if (s) fills0++; else fills1++;
        }
        tick(5);
        if (pend) {
            epilogue(pend_t, pend_ab);
            pend = false;
        }
        tick(6);
        if (last_of_tile) {
            pend = true; pend_t = t; pend_ab = ab;
            ab ^= 1;
        }
        first_of_tile = last_of_tile;
        if (have_next) t = nx1.tb >= 0 ? nx1.tb : (last_of_tile ? t + 1 : t);
        cur = nx1;
        nx1 = nx2;
        s ^= 1;
    }
    if (pend) {
        const uint32_t ulast = s ? uses0 : uses1;
        mbar_wait(bard + 8 * (s ^ 1), (ulast - 1) & 1);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        epilogue(pend_t, pend_ab);
    }
    if (PROF && tid == 0)
        for (int i = 0; i < 8; i++) prof[8 * (size_t)blockIdx.x + i] = pc[i];
    __syncthreads();
    if (warp == 0)
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(TMEM_COLS) : "memory");
}

} // namespace

int bsp_numeric_f32_tc(bsp_ctx *ctx, const bsp_mat *A, const bsp_mat *B, const bsp_symbolic_out *sym, float *c_vals,
                       int32_t *path)
{
    const int k = A->k;
    if (k != 32 && k != 64) {
        snprintf(ctx->err, sizeof(ctx->err), "no float32 tensor-core kernel for k = %d", k);
        return BSP_ERR_LIMIT;
    }
    const float *a = (const float *)A->vals, *b = (const float *)B->vals;
    if (sym->nnzb_c == 0) { if (path) *path = 100 + k; return BSP_OK; }
    const int smem = 4 * k * k * 4; // two stages of (A tile, B tile)
    const char *pe = getenv("BSP_F32_PROF"); // per-step cycle counts of thread 0 of every CTA on stderr (diagnosis)
    const bool profiled = pe && pe[0] == '1';
    const void *fn = profiled ? (k == 32 ? (const void *)f32_tcb_kernel<32, true> : (const void *)f32_tcb_kernel<64, true>)
                              : (k == 32 ? (const void *)f32_tcb_kernel<32, false> : (const void *)f32_tcb_kernel<64, false>);
    BSP_CUDA(ctx, cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    { // operands rewritten tile by tile into the layout the instruction reads
        // the two buffers live on the context and only grow: taking 2 x 344 MB from the stream-ordered pool cost 3.4 ms of
        // idle GPU per product on the bench workload, thirty times the repacking itself
        const size_t need[2] = {(size_t)A->nnzb * k * k * sizeof(float), (size_t)B->nnzb * k * k * sizeof(float)};
        for (int i = 0; i < 2; i++) {
            if (ctx->f32_pack_bytes[i] >= need[i]) continue;
            BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream)); // earlier products may still read the old buffer
            if (ctx->f32_pack[i]) cudaFree(ctx->f32_pack[i]);
            ctx->f32_pack[i] = nullptr;
            ctx->f32_pack_bytes[i] = 0;
            const size_t bytes = need[i] + need[i] / 8 + 256;
            BSP_CUDA(ctx, cudaMalloc(&ctx->f32_pack[i], bytes));
            ctx->f32_pack_bytes[i] = bytes;
        }
        float *ap = (float *)ctx->f32_pack[0], *bp = (float *)ctx->f32_pack[1];
        const int64_t ca = A->nnzb * (int64_t)(k * k / 4), cb = B->nnzb * (int64_t)(k * k / 4);
        const int64_t cap = (int64_t)ctx->sm_count * 32;
        const unsigned ga = (unsigned)std::min<int64_t>((ca + 255) / 256, cap), gb = (unsigned)std::min<int64_t>((cb + 255) / 256, cap);
        if (k == 32) {
            if (ca) f32_pack_a_kernel<32><<<ga, 256, 0, ctx->stream>>>((const float4 *)a, (float4 *)ap, ca);
            if (cb) f32_pack_b_kernel<32><<<gb, 256, 0, ctx->stream>>>(b, (float4 *)bp, cb);
        } else {
            if (ca) f32_pack_a_kernel<64><<<ga, 256, 0, ctx->stream>>>((const float4 *)a, (float4 *)ap, ca);
            if (cb) f32_pack_b_kernel<64><<<gb, 256, 0, ctx->stream>>>(b, (float4 *)bp, cb);
        }
        BSP_LAUNCH_CHECK(ctx);
        ctx->launches++;
        bsp_trace_mark(ctx, "f32_pack");
        a = ap; b = bp;
    }
    // one wave: as many CTAs per SM as registers, shared memory and the 512 TMEM columns (2k per CTA) allow.  Worked out
    // by hand: cudaOccupancyMaxActiveBlocksPerMultiprocessor answers 1 for these kernels on this toolchain, while three
    // (k = 64) and five (k = 32) CTAs are resident in fact (ncu: sm__warps_active).
    cudaFuncAttributes fa;
    BSP_CUDA(ctx, cudaFuncGetAttributes(&fa, fn));
    const int regs = ((fa.numRegs + 7) / 8) * 8;
    int per_sm = 65536 / (128 * (regs > 0 ? regs : 8));
    const int by_smem = (int)((228 * 1024) / ((size_t)smem + fa.sharedSizeBytes + 1024));
    if (per_sm > by_smem) per_sm = by_smem;
    if (per_sm > 512 / (2 * k)) per_sm = 512 / (2 * k);
    if (per_sm < 1) per_sm = 1;
    const int64_t max_grid = (int64_t)ctx->sm_count * per_sm;
    // output tiles are dealt in chunks of up to 8 consecutive tiles, fewer when that would leave CTAs without work
    int tch = (int)(sym->nnzb_c / max_grid);
    tch = tch < 1 ? 1 : (tch > 8 ? 8 : tch);
    if (const char *e = getenv("BSP_F32_TCH")) { const int v = atoi(e); if (v >= 1 && v <= 64) tch = v; } // comparison hook
    int64_t grid = (sym->nnzb_c + tch - 1) / tch;
    if (grid > max_grid) grid = max_grid;
    long long *prof = nullptr;
    if (profiled) {
        BSP_TRY(bsp_alloc(ctx, &prof, (size_t)grid * 8));
        BSP_CUDA(ctx, cudaMemsetAsync(prof, 0, (size_t)grid * 8 * sizeof(long long), ctx->stream));
    }
    {
        void *args[] = {(void *)&a, (void *)&b, (void *)&sym->pair_ptr, (void *)&sym->pairs, (void *)&sym->nnzb_c, (void *)&tch,
                        (void *)&c_vals, (void *)&prof};
        BSP_CUDA(ctx, cudaLaunchKernel(fn, dim3((unsigned)grid), dim3(128), args, (size_t)smem, ctx->stream));
    }
    if (prof) {
        std::vector<long long> h((size_t)grid * 8);
        BSP_CUDA(ctx, cudaMemcpyAsync(h.data(), prof, h.size() * sizeof(long long), cudaMemcpyDeviceToHost, ctx->stream));
        BSP_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        bsp_release(ctx, prof);
        double sum[8] = {0, 0, 0, 0, 0, 0, 0, 0};
        for (int64_t c = 0; c < grid; c++) for (int i = 0; i < 8; i++) sum[i] += (double)h[(size_t)c * 8 + i];
        static const char *names[8] = {"generator", "cta_barrier", "wait_copies", "mma_issue", "wait_mma_prev", "start_copies", "read_back", "-"};
        fprintf(stderr, "[bsp f32 prof] k=%d grid=%lld tch=%d per_sm=%d  cycles per product (nprod=%lld):", k, (long long)grid, tch, per_sm, (long long)sym->nprod);
        for (int i = 0; i < 8; i++) fprintf(stderr, " %s=%.0f", names[i], sum[i] * (double)1 / ((double)sym->nprod));
        fprintf(stderr, "\n");
    }
    if (path) *path = 100 + k;
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}
