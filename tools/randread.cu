// Experiment helper: achieved bandwidth of random aligned reads of 32 / 64 / 128 bytes from a
// 4 GiB array (far beyond L2).  Answers "does a wider lane word (more query groups per vertex)
// buy DRAM efficiency on the dense bottom-up levels?".  Build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int BYTES> struct alignas(BYTES >= 32 ? 32 : BYTES) Chunk { unsigned long long w[BYTES / 8]; };

__device__ __forceinline__ uint64_t mix(uint64_t x) { x ^= x >> 30; x *= 0xBF58476D1CE4E5B9ULL; x ^= x >> 27; x *= 0x94D049BB133111EBULL; x ^= x >> 31; return x; }

template <int BYTES, int U>
__global__ void __launch_bounds__(256, 4) rr(const unsigned long long* base, size_t nelem, size_t per_thread, unsigned long long* sink) {
    size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    unsigned long long acc = 0;
    for (size_t i = 0; i < per_thread; i += U) {
        size_t idx[U];
#pragma unroll
        for (int k = 0; k < U; ++k) idx[k] = mix(tid * 1315423911ull + i + k) & (nelem - 1);
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const Chunk<BYTES>* p = reinterpret_cast<const Chunk<BYTES>*>(base) + idx[k];
            if (BYTES >= 32) {
#pragma unroll
                for (int j = 0; j < BYTES / 32; ++j) {
                    Chunk<32> c = reinterpret_cast<const Chunk<32>*>(p)[j];
                    acc ^= c.w[0] ^ c.w[1] ^ c.w[2] ^ c.w[3];
                }
            } else {
                Chunk<BYTES> c = *p;
#pragma unroll
                for (int j = 0; j < BYTES / 8; ++j) acc ^= c.w[j];
            }
        }
    }
    if (acc == 0x1234567) sink[0] = acc;
}

template <int BYTES, int U> void run(const unsigned long long* d, size_t bytes, unsigned long long* sink) { /* bytes: power of two */
    size_t nelem = bytes / BYTES, per_thread = 256;
    int blocks = 148 * 4 * 8;
    cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
    rr<BYTES, U><<<blocks, 256>>>(d, nelem, per_thread, sink);
    cudaEventRecord(a);
    rr<BYTES, U><<<blocks, 256>>>(d, nelem, per_thread, sink);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b);
    double gb = (double)blocks * 256 * per_thread * BYTES / 1e9;
    printf("array %6.0f MiB  random %3d-byte reads, %d in flight per thread: %8.1f GB/s  (%.2f G accesses/s)\n", bytes / 1048576.0, BYTES, U, gb / (ms / 1e3), gb / BYTES / (ms / 1e3));
}

int main() {
    size_t maxbytes = 16ull << 30;
    unsigned long long *d, *sink;
    cudaMalloc(&d, maxbytes); cudaMalloc(&sink, 8); cudaMemset(d, 1, maxbytes);
    for (size_t bytes : {256ull << 20, 1ull << 30, 4ull << 30, 16ull << 30}) {
        run<32, 4>(d, bytes, sink); run<32, 8>(d, bytes, sink);
        run<64, 4>(d, bytes, sink); run<128, 4>(d, bytes, sink);
    }
    cudaError_t e = cudaDeviceSynchronize();
    printf("%s\n", cudaGetErrorString(e));
    return 0;
}
