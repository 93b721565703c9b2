// u64mac_peak.cu -- micro-benchmark: how many uint64 multiply-accumulates per second can a
// B200 issue, for the formulations the numeric pass can choose from?  MEASURED_PEAKS.json has
// no integer figure, so this is the denominator for "integer-pipe utilisation" in DESIGN.md.
//   plain   : s += a*b                      (wrapping; what a flag-free fast path would cost)
//   flagged : plain + conservative 2^64-1 detector on the low words (the shipped fast path)
//   exact   : phi after the multiply and after the add (the reference's per-op arithmetic)
//   split   : flagged, with the low x low products in the 64-bit accumulator (IMAD.WIDE) and the cross terms in
//             a separate 32-bit accumulator (two IMAD), i.e. without the add into the high word
//   splitp  : split without the detector
// Each thread keeps 8 accumulators (one output row of a k=8 tile), A row in registers, B tile
// read from shared memory with 128-bit broadcast loads -- the inner loop of the real kernel
// without its global loads.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o u64mac_peak u64mac_peak.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>

__device__ __forceinline__ uint64_t phi(uint64_t x) { return x == ~0ull ? 0ull : x; }

template <int MODE>
__global__ void __launch_bounds__(256) mac_kernel(const uint64_t *__restrict__ src, uint64_t *__restrict__ out, int iters)
{
    __shared__ __align__(16) uint64_t tile[8][64]; // 8 different B tiles
    for (int i = threadIdx.x; i < 512; i += 256) tile[i >> 6][i & 63] = src[i] | 1;
    __syncthreads();
    uint64_t a[8], acc[8];
    uint32_t xs[8];
#pragma unroll
    for (int j = 0; j < 8; j++) { a[j] = src[(threadIdx.x * 8 + j) & 511] + blockIdx.x; acc[j] = 0; xs[j] = 0; }
    uint32_t flag = MODE == 8 ? 0xffffffffu : 0, flag2 = MODE >= 7 ? 0xffffffffu : 0;
    bool hit = false;
    for (int it = 0; it < iters; it++) {
        const uint64_t *B = tile[it & 7];
#pragma unroll
        for (int j = 0; j < 8; j++) {
#pragma unroll
            for (int c = 0; c < 8; c += 2) {
                const ulonglong2 b2 = *reinterpret_cast<const ulonglong2 *>(&B[j * 8 + c]);
                const uint64_t bb[2] = {b2.x, b2.y};
#pragma unroll
                for (int h = 0; h < 2; h++) {
                    if (MODE == 0) {
                        acc[c + h] += a[j] * bb[h];
                    } else if (MODE == 1) {
                        const uint32_t lo_old = (uint32_t)acc[c + h];
                        acc[c + h] += a[j] * bb[h];
                        const uint32_t lo = (uint32_t)acc[c + h];
                        flag = max(flag, max(lo, lo - lo_old));
                    } else if (MODE == 2) {
                        acc[c + h] = phi(acc[c + h] + phi(a[j] * bb[h]));
                    } else {
                        const uint32_t al = (uint32_t)a[j], ah = (uint32_t)(a[j] >> 32), bl = (uint32_t)bb[h], bh = (uint32_t)(bb[h] >> 32);
                        const uint32_t lo_old = (uint32_t)acc[c + h];
                        uint64_t t;
                        asm("mul.wide.u32 %0, %1, %2;" : "=l"(t) : "r"(al), "r"(bl));
                        acc[c + h] += t;
                        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(xs[c + h]) : "r"(al), "r"(bh));
                        asm("mad.lo.u32 %0, %1, %2, %0;" : "+r"(xs[c + h]) : "r"(ah), "r"(bl));
                        const uint32_t lo = (uint32_t)acc[c + h];
                        if (MODE == 3) flag = max(flag, max(lo, lo - lo_old));
                        if (MODE == 5) hit = hit | (lo == 0xffffffffu) | (lo_old - lo == 1u);          // compare + predicate OR
                        if (MODE == 6) { flag = max(flag, lo); flag2 = max(flag2, lo - lo_old); }     // two-input max only
                        if (MODE == 7) { flag |= ~lo - 1u < 0x7fffffffu ? 0u : 0u; flag2 = min(flag2, min(~lo, lo_old + ~lo)); } // min3 form
                        if (MODE == 8) { flag = min(flag, ~lo); flag2 = min(flag2, lo_old + ~lo); }   // two-input min of "distance to the event"
                    }
                }
            }
        }
        // fresh A values every iteration, as in the real kernel (and nothing the compiler can strength-reduce)
        const ulonglong2 *an = reinterpret_cast<const ulonglong2 *>(&tile[(it + 3) & 7][(threadIdx.x & 7) * 8]);
#pragma unroll
        for (int j = 0; j < 8; j += 2) { const ulonglong2 v = an[j / 2]; a[j] = v.x; a[j + 1] = v.y; }
    }
    uint64_t r = flag + flag2 + (hit ? 12345u : 0u);
#pragma unroll
    for (int j = 0; j < 8; j++) r ^= acc[j] + ((uint64_t)xs[j] << 32);
    out[blockIdx.x * 256 + threadIdx.x] = r;
}

template <int MODE>
double run(const char *name, int grid, int iters, uint64_t *src, uint64_t *out)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    mac_kernel<MODE><<<grid, 256>>>(src, out, iters);
    cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 5; r++) {
        cudaEventRecord(e0);
        mac_kernel<MODE><<<grid, 256>>>(src, out, iters);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    double macs = (double)grid * 256 * iters * 64;
    printf("{\"variant\": \"%s\", \"grid\": %d, \"ms\": %.3f, \"u64_mac_per_s\": %.4g}\n", name, grid, best, macs / (best * 1e-3));
    return macs / (best * 1e-3);
}

int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    int sms = p.multiProcessorCount;
    uint64_t *src, *out;
    cudaMalloc(&src, 512 * 8); cudaMalloc(&out, (size_t)sms * 8 * 256 * 8);
    uint64_t h[512];
    uint64_t z = 0x9E3779B97F4A7C15ull;
    for (int i = 0; i < 512; i++) { z ^= z << 13; z ^= z >> 7; z ^= z << 17; h[i] = z; }
    cudaMemcpy(src, h, sizeof h, cudaMemcpyHostToDevice);
    printf("{\"device\": \"%s\", \"sms\": %d, \"clock_khz\": %d}\n", p.name, sms, p.clockRate);
    for (int occ = 4; occ <= 4; occ *= 2) {
        run<0>("plain", sms * occ, 2000, src, out);
        run<1>("flagged", sms * occ, 2000, src, out);
        run<2>("exact", sms * occ, 2000, src, out);
        run<3>("split", sms * occ, 2000, src, out);
        run<4>("splitp", sms * occ, 2000, src, out);
        run<5>("split_setp", sms * occ, 2000, src, out);
        run<6>("split_max2", sms * occ, 2000, src, out);
        run<8>("split_min2", sms * occ, 2000, src, out);
    }
    return 0;
}
