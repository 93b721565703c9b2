// int_pipe_rates.cu -- issue rates of the integer instructions the numeric pass is made of, on one B200:
// warp-instructions per cycle per SM sub-partition for IMAD, IMAD.WIDE.U32, IADD3, VIADDMNMX and mixes.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o int_pipe_rates int_pipe_rates.cu
#include <cstdint>
#include <cstdio>
#include <cuda_runtime.h>
template <int MODE>
__global__ void __launch_bounds__(256) k(uint32_t *out, int iters, uint32_t seed)
{
    uint32_t a[8], b[8];
    uint64_t w[8];
    for (int i = 0; i < 8; i++) { a[i] = seed + threadIdx.x * 8 + i; b[i] = seed * 3 + i; w[i] = a[i]; }
    const uint32_t m = seed | 1, n = seed * 7 + 3;
    for (int it = 0; it < iters; it++) {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int i = 0; i < 8; i++) {
                if (MODE == 0) a[i] = a[i] * m + n;                                   // IMAD
                if (MODE == 1) w[i] = (uint64_t)(uint32_t)w[i] * m + w[i];            // IMAD.WIDE.U32 accumulate
                if (MODE == 2) a[i] = a[i] + b[i] + n;                                // IADD3
                if (MODE == 3) a[i] = max(a[i] - b[i], a[i]) + 1;                     // VIADDMNMX (+1 add)
                if (MODE == 4) { a[i] = a[i] * m + n; b[i] = b[i] + a[i] + n; }       // IMAD + IADD3
                if (MODE == 5) { w[i] = (uint64_t)(uint32_t)w[i] * m + w[i]; b[i] = b[i] + (uint32_t)w[i] + n; } // WIDE + IADD3
                if (MODE == 7) { const uint64_t t = (uint64_t)a[i] * m; a[i] = (uint32_t)t ^ (uint32_t)(t >> 32); }  // IMAD.WIDE (no addend) + LOP3
                if (MODE == 8) { uint32_t lo = (uint32_t)w[i], hi = (uint32_t)(w[i] >> 32), tl, th;                // WIDE (no addend) + add.cc/addc + 2 IMAD
                                 asm("mul.wide.u32 %0, %1, %2;" : "=l"(w[i]) : "r"(a[i]), "r"(m));
                                 tl = (uint32_t)w[i]; th = (uint32_t)(w[i] >> 32);
                                 asm("add.cc.u32 %0, %0, %2; addc.u32 %1, %1, %3;" : "+r"(lo), "+r"(hi) : "r"(tl), "r"(th));
                                 hi = a[i] * n + hi; hi = b[i] * m + hi; a[i] = lo; b[i] = hi; }
                if (MODE == 9) { w[i] = (uint64_t)a[i] * m + w[i]; uint32_t hi = (uint32_t)(w[i] >> 32);                // WIDE accumulate + 2 IMAD on hi
                                 hi = a[i] * n + hi; hi = b[i] * m + hi; w[i] = ((uint64_t)hi << 32) | (uint32_t)w[i]; a[i] += 1; }
                if (MODE == 6) { w[i] = (uint64_t)(uint32_t)w[i] * m + w[i]; a[i] = a[i] * m + n; b[i] = b[i] * n + m; } // WIDE + 2 IMAD
            }
    }
    uint32_t r = 0;
    for (int i = 0; i < 8; i++) r ^= a[i] ^ b[i] ^ (uint32_t)w[i] ^ (uint32_t)(w[i] >> 32);
    out[blockIdx.x * 256 + threadIdx.x] = r;
}
template <int MODE>
void run(const char *name, int ops_per_inner, uint32_t *out, int sms, double clk_hz)
{
    const int iters = 4000, grid = sms * 4;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<MODE><<<grid, 256>>>(out, iters, 12345u); cudaDeviceSynchronize();
    float best = 1e30f;
    for (int r = 0; r < 3; r++) { cudaEventRecord(e0); k<MODE><<<grid, 256>>>(out, iters, 12345u); cudaEventRecord(e1); cudaEventSynchronize(e1); float ms; cudaEventElapsedTime(&ms, e0, e1); if (ms < best) best = ms; }
    const double warp_instr = (double)grid * 8 * iters * 32.0 * ops_per_inner; // 8 warps per CTA, 4*8 inner per iter
    const double per_cycle_per_smsp = warp_instr / (best * 1e-3) / clk_hz / (sms * 4.0);
    printf("{\"mix\": \"%s\", \"ms\": %.3f, \"warp_instr_per_cycle_per_smsp\": %.3f}\n", name, best, per_cycle_per_smsp);
}
int main()
{
    cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
    uint32_t *out; cudaMalloc(&out, (size_t)p.multiProcessorCount * 4 * 256 * 4);
    const double clk = p.clockRate * 1e3;
    run<0>("IMAD", 1, out, p.multiProcessorCount, clk);
    run<1>("IMAD.WIDE.U32 (accumulate)", 1, out, p.multiProcessorCount, clk);
    run<2>("IADD3", 1, out, p.multiProcessorCount, clk);
    run<3>("VIADDMNMX + IADD", 2, out, p.multiProcessorCount, clk);
    run<4>("IMAD + IADD3", 2, out, p.multiProcessorCount, clk);
    run<5>("IMAD.WIDE + IADD3", 2, out, p.multiProcessorCount, clk);
    run<6>("IMAD.WIDE + 2 IMAD", 3, out, p.multiProcessorCount, clk);
    run<7>("IMAD.WIDE (no addend) + LOP3", 2, out, p.multiProcessorCount, clk);
    run<8>("MAC: WIDE(no addend) + add.cc/addc + 2 IMAD  [count as 1]", 1, out, p.multiProcessorCount, clk);
    run<9>("MAC: WIDE accumulate + 2 IMAD  [count as 1]", 1, out, p.multiProcessorCount, clk);
    return 0;
}
