/*
 * cuda_runtime.h stand-in for building the reference translation unit host-only.
 * TEST INFRASTRUCTURE ONLY: lets oracle/build_ref.sh compile /root/reference's single
 * source file with g++ so that its own code becomes the parity oracle and the timed CPU
 * baseline.  Device memory becomes host memory, a launch becomes a loop nest over
 * (block, thread.y, thread.x).  With -DSHIM_OMP_LAUNCH the block loop is spread over
 * OpenMP threads (the kernel body is unchanged; every output element is still folded by
 * one thread in the same order).
 */
#ifndef BSP_SHIM_CUDA_RUNTIME_H
#define BSP_SHIM_CUDA_RUNTIME_H
#include <cstdlib>
#include <cstring>
#include <cstddef>

#define __global__
#define __device__
#define __host__

struct dim3 {
    unsigned x, y, z;
    dim3(unsigned a = 1, unsigned b = 1, unsigned c = 1) : x(a), y(b), z(c) {}
};

typedef int cudaError_t;
enum { cudaSuccess = 0, cudaErrorMemoryAllocation = 2 };
enum cudaMemcpyKind { cudaMemcpyHostToHost, cudaMemcpyHostToDevice, cudaMemcpyDeviceToHost, cudaMemcpyDeviceToDevice };

static thread_local dim3 blockIdx, threadIdx;
static dim3 gridDim, blockDim;

static inline const char *cudaGetErrorString(cudaError_t e) { return e ? "shim allocation failure" : "no error"; }
template <typename T> static inline cudaError_t cudaMalloc(T **p, size_t n)
{
    *p = (T *)std::malloc(n ? n : 1);
    return *p ? cudaSuccess : cudaErrorMemoryAllocation;
}
static inline cudaError_t cudaMemcpy(void *d, const void *s, size_t n, cudaMemcpyKind) { std::memcpy(d, s, n); return cudaSuccess; }
static inline cudaError_t cudaFree(void *p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }
static inline void __syncthreads() {}

#ifdef SHIM_OMP_LAUNCH
#define SHIM_PRAGMA _Pragma("omp parallel for schedule(static)")
#else
#define SHIM_PRAGMA
#endif

#define SHIM_LAUNCH(kern, g, b, ...)                                               \
    do {                                                                           \
        dim3 g_ = (g), b_ = (b);                                                   \
        gridDim = g_; blockDim = b_;                                               \
        SHIM_PRAGMA                                                                \
        for (long bx_ = 0; bx_ < (long)g_.x; ++bx_)                                \
            for (unsigned ty_ = 0; ty_ < b_.y; ++ty_)                              \
                for (unsigned tx_ = 0; tx_ < b_.x; ++tx_) {                        \
                    blockIdx.x = (unsigned)bx_; threadIdx.y = ty_; threadIdx.x = tx_; \
                    kern(__VA_ARGS__);                                             \
                }                                                                  \
    } while (0)
#endif
