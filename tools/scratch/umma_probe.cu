// Probe for the no-swizzle UMMA smem layouts (tf32, M=64): one tile product, dump all 128 TMEM lanes.
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint64_t umma_desc(uint32_t addr, uint32_t lbo, uint32_t sbo)
{
    return (uint64_t)((addr >> 4) & 0x3fff) | ((uint64_t)((lbo >> 4) & 0x3fff) << 16) | ((uint64_t)((sbo >> 4) & 0x3fff) << 32) | (1ull << 46);
}
template <int K>
__global__ void probe(const float *A, const float *B, float *out, int swapA, int swapB, int bmajor_mn, int bsw128)
{
    constexpr int M = 64, N = K;
    extern __shared__ __align__(1024) unsigned char smem[];
    __shared__ __align__(8) uint64_t mbar;
    __shared__ uint32_t tmem_sh;
    const int tid = threadIdx.x, warp = tid >> 5;
    const uint32_t sbase = smem_u32(smem), bar = smem_u32(&mbar);
    constexpr int A_BYTES = (K / 4) * M * 16;
    for (int i = tid; i < (A_BYTES + K * K * 4) / 4; i += 128) ((float *)smem)[i] = 0.f;
    __syncthreads();
    if (tid == 0) { asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar)); asm volatile("fence.mbarrier_init.release.cluster;"); }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_sh)), "n"(64));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;");
    const uint32_t tmem = tmem_sh;
    // stage with plain stores: A chunk (m,kc) -> kc*(M*16)+m*16 ; B chunk (kk,nc) -> nc*(K*16)+kk*16
    for (int x = tid; x < K * K / 4; x += 128) {
        const int row = x / (K / 4), col = x % (K / 4);
        *(float4 *)(smem + col * (M * 16) + row * 16) = ((const float4 *)A)[x];
        if (bsw128) // row-major [k][128 B] with the 128-byte swizzle: chunk n4 of row k at (n4 ^ (k % 8)); atoms of 32 columns LBO apart
            *(float4 *)(smem + A_BYTES + (col / 8) * (K * 128) + (row / 8) * 1024 + (row % 8) * 128 + (((col % 8) ^ (row % 8)) * 16)) = ((const float4 *)B)[x];
        else
            *(float4 *)(smem + A_BYTES + col * (K * 16) + row * 16) = ((const float4 *)B)[x];
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;");
        const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | (0u << 15) | ((uint32_t)bmajor_mn << 16) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
        for (int ks = 0; ks < K / 8; ks++) {
            const uint32_t a_addr = sbase + ks * 2 * (M * 16), b_addr = sbase + A_BYTES + ks * (bsw128 ? 1024 : 128);
            const uint64_t da = swapA ? umma_desc(a_addr, 128, M * 16) : umma_desc(a_addr, M * 16, 128);
            uint64_t db = swapB ? umma_desc(b_addr, K * 16, 128) : umma_desc(b_addr, 128, K * 16);
            if (bsw128) db = umma_desc(b_addr, K * 128, 1024) | (2ull << 61); // SWIZZLE_128B, MN atoms K*128 apart, k-groups 1024 apart
            const uint32_t acc = ks > 0;
            asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem), "l"(da), "l"(db), "r"(idesc), "r"(acc) : "memory");
        }
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(bar) : "memory");
    }
    asm volatile("{\n\t.reg .pred p;\n\tW: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], 0;\n\t@p bra D;\n\tbra W;\n\tD:\n\t}" ::"r"(bar) : "memory");
    asm volatile("tcgen05.fence::after_thread_sync;");
    for (int c0 = 0; c0 < N; c0 += 16) {
        uint32_t r[16];
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
                     : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
                     : "r"(tmem + ((uint32_t)(32 * warp) << 16) + c0));
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        for (int i = 0; i < 16; i++) out[tid * N + c0 + i] = __uint_as_float(r[i]);
    }
    asm volatile("tcgen05.fence::before_thread_sync;");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "n"(64));
}
int main(int argc, char **argv)
{
    constexpr int K = 32;
    float hA[K * K], hB[K * K], hO[128 * K];
    for (int cfg = atoi(argv[1]); cfg <= atoi(argv[1]); cfg++) {
        int swapA = cfg & 1, swapB = (cfg >> 1) & 1, bmn = (cfg >> 2) & 1;
        for (int m = 0; m < K; m++) for (int k = 0; k < K; k++) { hA[m * K + k] = (m == k); hB[m * K + k] = 100.f * m + k; }
        float *dA, *dB, *dO;
        cudaMalloc(&dA, sizeof hA); cudaMalloc(&dB, sizeof hB); cudaMalloc(&dO, sizeof hO);
        cudaMemcpy(dA, hA, sizeof hA, cudaMemcpyHostToDevice); cudaMemcpy(dB, hB, sizeof hB, cudaMemcpyHostToDevice);
        cudaMemset(dO, 0, sizeof hO);
        const int smem = (K / 4) * 64 * 16 + K * K * 4 + 1024;
        cudaFuncSetAttribute(probe<K>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem);
        probe<K><<<1, 128, smem>>>(dA, dB, dO, swapA, swapB, bmn, (cfg >> 3) & 1);
        cudaError_t e = cudaDeviceSynchronize();
        cudaMemcpy(hO, dO, sizeof hO, cudaMemcpyDeviceToHost);
        int nz = 0; double err = 0;
        for (int i = 0; i < 128 * K; i++) nz += hO[i] != 0;
        for (int w = 0; w < 2; w++) for (int i = 0; i < 16; i++) for (int n = 0; n < K; n++) err += fabs(hO[(32 * w + i) * K + n] - (100.f * (16 * w + i) + n));
        printf("cfg%d swapA=%d swapB=%d bMN=%d : %s nonzeros=%d err(I*R, lanes 32w+i)=%.0f | lane0: %.0f %.0f %.0f %.0f | lane1: %.0f %.0f %.0f | lane16: %.0f %.0f | lane32: %.0f %.0f\n", cfg, swapA, swapB, bmn,
               cudaGetErrorString(e), nz, err, hO[0], hO[1], hO[2], hO[3], hO[K], hO[K + 1], hO[K + 2], hO[16 * K], hO[16 * K + 1], hO[32 * K], hO[32 * K + 1]);
        cudaFree(dA); cudaFree(dB); cudaFree(dO);
    }
    return 0;
}
