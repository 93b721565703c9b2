// bsp_numeric_f32_tc.cu -- float32 tile products on the 5th-generation tensor cores.
//
// BASELINE.json configs[4] (k = 32 / 64 float tiles; an extension, the reference has no float
// path).  Per output tile the products A_tile * B_tile of its pair list are accumulated by
// tcgen05.mma (kind::tf32: fp32 storage read as TF32, fp32 accumulate) into one TMEM
// accumulator and read back once with tcgen05.ld -- the tile never leaves the tensor memory
// between pairs.
//
// One CTA (4 warps) per output tile at a time, two shared-memory stages:
//   all threads : stage the tiles of pair p into stage p&1 in the canonical no-swizzle K-major
//                 UMMA layout (8-row x 16-byte core matrices):
//                   A: element (m, k)  at (k/4)*(64*16) + m*16 + (k%4)*4     LBO = 1024, SBO = 128
//                      a 16-byte chunk of the row-major tile is already one core-matrix row, so
//                      A is a pure chunk permutation done by cp.async;
//                   B: the instruction wants B as N x K with K contiguous, the tile is K x N with
//                      N contiguous, so each thread transposes 4x4 blocks in registers
//                      (4 LDG.128 -> 4 STS.128): element (n, k) at (k/4)*(K*16) + n*16 + (k%4)*4
//                      LBO = K*16, SBO = 128.
//                 (An MN-major B descriptor would avoid the transpose; with no-swizzle layouts it
//                 produced zeros in a probe on this toolchain -- tools/scratch/umma_probe.cu.)
//   one thread  : K/8 tcgen05.mma (M=64, N=K, K=8 each), then tcgen05.commit -> the stage's mbarrier
//   epilogue    : warp w reads TMEM lanes 32w..32w+15 = tile rows 16w..16w+15 (the M=64
//                 accumulator layout uses 16 lanes of every sub-partition) and stores them.
// For k = 32 the M=64 instruction runs with its upper 32 rows unused (the tile product is far
// below the tensor pipe's rate either way: 8 flop per byte moved).
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
__device__ __forceinline__ void cp16(uint32_t dst, const void *src)
{
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst), "l"(src) : "memory");
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

__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float *v)
{
    uint32_t r[16];
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
          "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}

template <int K>
__global__ void __launch_bounds__(128)
f32_tc_kernel(const float *__restrict__ a_vals, const float *__restrict__ b_vals, const int64_t *__restrict__ pair_ptr,
              const int2 *__restrict__ pairs, int64_t nnzb_c, float *__restrict__ c_vals)
{
    constexpr int M = 64, N = K;
    constexpr int A_BYTES = (K / 4) * M * 16;  // chunk-major A: [kc][m] 16-byte chunks (64 rows, K=32 uses 32)
    constexpr int B_BYTES = (K / 4) * N * 16;  // chunk-major B^T: [kc][n] 16-byte chunks
    constexpr int STAGE = A_BYTES + B_BYTES;
    constexpr int CHUNKS = K * K / 4;          // 16-byte chunks per tile
    constexpr int TMEM_COLS = N < 32 ? 32 : N;
    extern __shared__ __align__(128) unsigned char smem[];
    __shared__ __align__(8) uint64_t mbar[2];
    __shared__ uint32_t tmem_base_sh;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const uint32_t sbase = smem_u32(smem);
    const uint32_t bar0 = smem_u32(&mbar[0]);

    if (tid == 0) {
        mbar_init(bar0, 1);
        mbar_init(bar0 + 8, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (K == 32) // rows 32..63 of the A stages are never written: clear them once so they hold finite values
        for (int i = tid; i < 2 * STAGE / 16; i += 128) reinterpret_cast<uint4 *>(smem)[i] = make_uint4(0, 0, 0, 0);
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

    // The CTA's work is one stream of tile products: for every output tile it owns (t = blockIdx.x, += gridDim.x),
    // the pairs of its list.  Product i+1 is fetched while product i is multiplied and, when i closes its tile,
    // while the accumulator is read back and stored: A by cp.async (a chunk permutation), B into registers with
    // plain loads that are transposed into the stage only after the MMAs that last read that stage have retired.
    constexpr int BBLK = (CHUNKS / 4 + 127) / 128; // 4x4 blocks of B per thread (2 for k = 64, 1 for k = 32 on half the threads)
    uint32_t uses0 = 0, uses1 = 0;                 // commits issued per stage (all threads track them identically)
    int64_t t = blockIdx.x;
    int64_t p = 0, p1 = 0;
    if (t < nnzb_c) { p = pair_ptr[t]; p1 = pair_ptr[t + 1]; }
    int2 pr_next = make_int2(0, 0);
    float4 breg[BBLK][4];

    auto load_a = [&](int s, int ai) {
        const unsigned char *ag = reinterpret_cast<const unsigned char *>(a_vals + (size_t)ai * (K * K));
        const uint32_t a_s = sbase + s * STAGE;
#pragma unroll
        for (int x = tid; x < CHUNKS; x += 128) {
            const int row = x / (K / 4), col = x % (K / 4); // chunk `col` of row `row` of the row-major tile
            cp16(a_s + col * (M * 16) + row * 16, ag + (size_t)x * 16);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    auto load_b = [&](int bi) {
        const float *bg = b_vals + (size_t)bi * (K * K);
#pragma unroll
        for (int u = 0; u < BBLK; u++) {
            const int x = tid + 128 * u;
            if (x < CHUNKS / 4) {
                const int kq = x / (K / 4), nq = x % (K / 4);
#pragma unroll
                for (int r = 0; r < 4; r++) breg[u][r] = *reinterpret_cast<const float4 *>(bg + (size_t)(4 * kq + r) * K + 4 * nq);
            }
        }
    };
    auto store_b = [&](int s) { // 4x4 blocks transposed in registers: element (n, k) at (k/4)*(K*16) + n*16 + (k%4)*4
#pragma unroll
        for (int u = 0; u < BBLK; u++) {
            const int x = tid + 128 * u;
            if (x < CHUNKS / 4) {
                const int kq = x / (K / 4), nq = x % (K / 4);
                float4 *dst = reinterpret_cast<float4 *>(smem + s * STAGE + A_BYTES + kq * (K * 16) + (4 * nq) * 16);
                dst[0] = make_float4(breg[u][0].x, breg[u][1].x, breg[u][2].x, breg[u][3].x);
                dst[1] = make_float4(breg[u][0].y, breg[u][1].y, breg[u][2].y, breg[u][3].y);
                dst[2] = make_float4(breg[u][0].z, breg[u][1].z, breg[u][2].z, breg[u][3].z);
                dst[3] = make_float4(breg[u][0].w, breg[u][1].w, breg[u][2].w, breg[u][3].w);
            }
        }
    };

    if (t < nnzb_c) { // first product: nothing to hide behind
        const int2 pr = pairs[p];
        load_a(0, pr.x);
        load_b(pr.y & 0x7fffffff);
        store_b(0);
    }
    int s = 0;
    bool first_of_tile = true;
    while (t < nnzb_c) {
        // the product after this one (same tile, or the first of the CTA's next tile)
        int64_t tn = t, pn = p + 1, pn1 = p1;
        if (pn >= p1) {
            tn = t + gridDim.x;
            if (tn < nnzb_c) { pn = pair_ptr[tn]; pn1 = pair_ptr[tn + 1]; }
        }
        const bool have_next = tn < nnzb_c;
        const bool last_of_tile = p + 1 >= p1;
        if (have_next) {
            const int s1 = s ^ 1;
            const uint32_t uses = s1 ? uses1 : uses0;
            if (uses > 0) mbar_wait(bar0 + 8 * s1, (uses - 1) & 1); // the MMAs that read stage s1 are done
            pr_next = pairs[pn];
            load_a(s1, pr_next.x);
            load_b(pr_next.y & 0x7fffffff);
            asm volatile("cp.async.wait_group 1;" ::: "memory"); // this product's A chunks have landed
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); // generic-proxy writes -> visible to the tensor core
        __syncthreads();
        if (tid == 0) {
            const uint32_t a_s = sbase + s * STAGE, b_s = a_s + A_BYTES;
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
            for (int ks = 0; ks < K / 8; ks++) {
                const uint64_t da = umma_desc(a_s + ks * 2 * (M * 16), M * 16, 128);
                const uint64_t db = umma_desc(b_s + ks * 2 * (K * 16), K * 16, 128);
                const uint32_t accumulate = (!first_of_tile || ks > 0) ? 1u : 0u;
                asm volatile(
                    "{\n\t"
                    ".reg .pred p;\n\t"
                    "setp.ne.b32 p, %4, 0;\n\t"
                    "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
                    "}" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(accumulate) : "memory");
            }
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar0 + 8 * s)
                         : "memory");
        }
        if (s) uses1++; else uses0++;
        if (last_of_tile) {
            // accumulator complete when this product's commit has arrived (tcgen05 ops retire in order)
            mbar_wait(bar0 + 8 * s, ((s ? uses1 : uses0) - 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            // epilogue: TMEM lanes 32w + i (i < 16) hold tile row 16w + i
            const int row = 16 * warp + lane;
            const bool valid = lane < 16 && row < K;
            float *out = c_vals + (size_t)t * (K * K) + (size_t)row * K;
#pragma unroll
            for (int c0 = 0; c0 < N; c0 += 16) {
                float v[16];
                tmem_ld16(tmem + ((uint32_t)(32 * warp) << 16) + c0, v);
                if (valid) {
#pragma unroll
                    for (int i = 0; i < 16; i += 4)
                        __stcs(reinterpret_cast<float4 *>(out + c0 + i), make_float4(v[i], v[i + 1], v[i + 2], v[i + 3]));
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            // the next product's __syncthreads orders these TMEM reads before its first MMA
        }
        if (have_next) store_b(s ^ 1); // the B loads have had the MMA issue and the read-back of the accumulator to hide behind
        first_of_tile = last_of_tile;
        t = tn; p = pn; p1 = pn1;
        s ^= 1;
    }
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
    // latency-bound per CTA (load -> MMA -> load ...): fill the SM with as many CTAs as shared memory allows
    int64_t grid = sym->nnzb_c;
    const int64_t max_grid = (int64_t)ctx->sm_count * (k == 32 ? 8 : 3);
    if (grid > max_grid) grid = max_grid;
    if (k == 32) {
        constexpr int smem = 2 * ((32 / 4) * 64 * 16 + (32 / 4) * 32 * 16);
        BSP_CUDA(ctx, cudaFuncSetAttribute(f32_tc_kernel<32>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        f32_tc_kernel<32><<<(unsigned)grid, 128, smem, ctx->stream>>>(a, b, sym->pair_ptr, sym->pairs, sym->nnzb_c, c_vals);
        if (path) *path = 132;
    } else {
        constexpr int smem = 2 * ((64 / 4) * 64 * 16 + (64 / 4) * 64 * 16);
        BSP_CUDA(ctx, cudaFuncSetAttribute(f32_tc_kernel<64>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        f32_tc_kernel<64><<<(unsigned)grid, 128, smem, ctx->stream>>>(a, b, sym->pair_ptr, sym->pairs, sym->nnzb_c, c_vals);
        if (path) *path = 164;
    }
    BSP_LAUNCH_CHECK(ctx);
    return BSP_OK;
}
