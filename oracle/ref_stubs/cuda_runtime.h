/*
 * ref_stubs/cuda_runtime.h — TEST INFRASTRUCTURE.  Just enough of the CUDA runtime
 * for g++ to build the UNMODIFIED reference source (/root/reference/main.cu) as a
 * host program: device memory is host memory, a kernel launch is a serial double
 * loop over blocks x threads.  The emulation is exact because the reference
 * kernel's result does not depend on thread order (every racing writer stores the
 * same level+1, and a vertex set in this launch is not expanded in this launch:
 * main.cu:21, 30-33).  The one `<<<...>>>` launch (main.cu:66) is rewritten to the
 * LAUNCH1D macro by the sed line in oracle/Makefile; nothing else is touched.
 */
#ifndef REF_STUB_CUDA_RUNTIME_H
#define REF_STUB_CUDA_RUNTIME_H

#include <cstdlib>
#include <cstring>

#define __global__
#define __device__
#define __host__

struct ref_stub_dim3 { unsigned x, y, z; };
static ref_stub_dim3 blockDim = {1, 1, 1}, blockIdx = {0, 0, 0}, threadIdx = {0, 0, 0}, gridDim = {1, 1, 1};

typedef int cudaError_t;
enum { cudaSuccess = 0 };
enum cudaMemcpyKind { cudaMemcpyHostToDevice = 1, cudaMemcpyDeviceToHost = 2, cudaMemcpyDeviceToDevice = 3 };

template <class T> static inline cudaError_t cudaMalloc(T** p, size_t bytes) { *p = (T*)std::malloc(bytes ? bytes : 1); return cudaSuccess; }
static inline cudaError_t cudaFree(void* p) { std::free(p); return cudaSuccess; }
static inline cudaError_t cudaMemset(void* p, int v, size_t bytes) { std::memset(p, v, bytes); return cudaSuccess; }
static inline cudaError_t cudaMemcpy(void* d, const void* s, size_t bytes, cudaMemcpyKind) { std::memcpy(d, s, bytes); return cudaSuccess; }
static inline cudaError_t cudaSetDevice(int) { return cudaSuccess; }
static inline cudaError_t cudaDeviceSynchronize() { return cudaSuccess; }

#define LAUNCH1D(kernel, nblocks, nthreads, ...)                                   \
    do {                                                                           \
        gridDim.x = (unsigned)(nblocks); blockDim.x = (unsigned)(nthreads);        \
        for (unsigned _b = 0; _b < (unsigned)(nblocks); ++_b) {                    \
            blockIdx.x = _b;                                                       \
            for (unsigned _t = 0; _t < (unsigned)(nthreads); ++_t) {               \
                threadIdx.x = _t;                                                  \
                kernel(__VA_ARGS__);                                               \
            }                                                                      \
        }                                                                          \
    } while (0)

#endif
