/*
 * noalloc.cu — msbfs_small_run_noalloc: distance sums of up to 64 query groups on a small
 * graph without allocating any device memory.
 *
 * Who needs this: code that sits where cudaMalloc / cudaFree are off limits — a library
 * called from inside a latency-critical loop (an allocation may synchronise the whole
 * device), a host whose memory pool owns every byte of the GPU and forbids side
 * allocations, or a process running under a fixed memory budget that was carved up at
 * start-up.  The regular entry points size their traversal state with cudaMalloc on
 * first use; this one keeps its state (visited / frontier / next lane words) in a
 * file-scope __device__ array of fixed size instead, so it can serve graphs of at most
 * MSBFS_NOALLOC_MAX_VERTICES vertices and batches of at most MSBFS_NOALLOC_MAX_GROUPS
 * groups.  Every GPU that loads libmsbfs.so carries that array (1.5 MiB) until exit.
 *
 * The kernel is the top-down level step of traverse.cu (frontier vertices push their
 * lane word to neighbours, then new = next & ~visited) folded into a single thread
 * block that loops over the BFS levels itself, so one launch does a whole batch.
 * It has no counterpart in the reference.
 */
#include "internal.h"

#include <mutex>

#define MSBFS_NOALLOC_MAX_VERTICES 65536
#define MSBFS_NOALLOC_MAX_GROUPS   64

namespace {

typedef unsigned long long ull;

/* visited | frontier | next, MSBFS_NOALLOC_MAX_VERTICES lane words each */
static __device__ __align__(256) ull g_noalloc_scratch[3 * MSBFS_NOALLOC_MAX_VERTICES];

constexpr int SMALL_CTA = 1024;

__global__ void __launch_bounds__(SMALL_CTA) small_batch_kernel(ull* __restrict__ scratch, int n,
                                                                const long long* __restrict__ row,
                                                                const int* __restrict__ col, int num_groups,
                                                                const long long* __restrict__ q_off,
                                                                const int* __restrict__ q_src,
                                                                long long* __restrict__ out_sums) {
    ull* visited = scratch;
    ull* frontier = scratch + MSBFS_NOALLOC_MAX_VERTICES;
    ull* next = scratch + 2 * MSBFS_NOALLOC_MAX_VERTICES;
    __shared__ ull s_sum[MSBFS_NOALLOC_MAX_GROUPS];
    __shared__ int s_any;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    /* the array still holds the previous call's words: everything this call reads starts at zero */
    for (int v = tid; v < n; v += SMALL_CTA) { visited[v] = 0ull; frontier[v] = 0ull; next[v] = 0ull; }
    if (tid < MSBFS_NOALLOC_MAX_GROUPS) s_sum[tid] = 0ull;
    __syncthreads();

    /* sources: ids outside [0, n) are ignored (main.cu:49) */
    for (int g = warp; g < num_groups; g += SMALL_CTA / 32) {
        for (long long i = q_off[g] + lane; i < q_off[g + 1]; i += 32) {
            const int s = q_src[i];
            if ((unsigned)s < (unsigned)n) {
                atomicOr(&frontier[s], 1ull << g);
                atomicOr(&visited[s], 1ull << g);
            }
        }
    }
    __syncthreads();

    for (int level = 0;; ++level) {
        if (tid == 0) s_any = 0;
        /* expand: one warp per frontier vertex, lanes stride its row */
        for (int v = warp; v < n; v += SMALL_CTA / 32) {
            const ull f = frontier[v];
            if (f == 0ull) continue;
            for (long long e = row[v] + lane; e < row[v + 1]; e += 32) {
                const int u = col[e];
                if ((unsigned)u >= (unsigned)n) continue; /* a bad entry must not index past the shared array */
                const ull d = f & ~visited[u];
                if (d) atomicOr(&next[u], d);
            }
        }
        __syncthreads();
        /* finalize: a vertex first reached while `level` is expanded lies at distance level + 1 */
        int any = 0;
        for (int v = tid; v < n; v += SMALL_CTA) {
            ull x = next[v] & ~visited[v];
            next[v] = 0ull;
            frontier[v] = x;
            if (x) {
                visited[v] |= x;
                any = 1;
                while (x) {
                    const int j = __ffsll((long long)x) - 1;
                    x &= x - 1ull;
                    atomicAdd(&s_sum[j], (ull)(level + 1));
                }
            }
        }
        if (any) s_any = 1;
        __syncthreads();
        if (!s_any) break;
        __syncthreads(); /* everyone has read s_any before thread 0 clears it again */
    }
    if (tid < num_groups) out_sums[tid] = (long long)s_sum[tid];
}

constexpr int MAX_DEVICES = 64;
std::mutex g_setup_mutex;
ull* g_scratch_addr[MAX_DEVICES]; /* device address of g_noalloc_scratch, per device */

} // namespace

extern "C" int msbfs_small_run_noalloc(int32_t n, const int64_t* d_row_offsets, const int32_t* d_col_indices,
                                       int32_t num_groups, const int64_t* d_group_offsets, const int32_t* d_sources,
                                       int64_t* d_out_sums, void* cuda_stream) {
    if (n < 0 || num_groups < 0)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_small_run_noalloc: negative size");
    if (n > MSBFS_NOALLOC_MAX_VERTICES)
        return msbfs_fail(MSBFS_ERR_RANGE,
                          "msbfs_small_run_noalloc: %d vertices exceed the fixed scratch of %d; use msbfs_run", n,
                          MSBFS_NOALLOC_MAX_VERTICES);
    if (num_groups > MSBFS_NOALLOC_MAX_GROUPS)
        return msbfs_fail(MSBFS_ERR_RANGE,
                          "msbfs_small_run_noalloc: %d groups exceed the limit of %d per call; use msbfs_run",
                          num_groups, MSBFS_NOALLOC_MAX_GROUPS);
    if (num_groups == 0) return MSBFS_OK;
    /* d_col_indices / d_sources may only be NULL when the graph has no entries / the groups have
     * no sources; their lengths live on the device, so the caller vouches for that — a NULL
     * here with entries to read would fault inside the kernel, hence the flat refusal */
    if (!d_row_offsets || !d_group_offsets || !d_out_sums || !d_col_indices || !d_sources)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_small_run_noalloc: null pointer");

    int dev = -1;
    MSBFS_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= MAX_DEVICES)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_small_run_noalloc: device %d not supported", dev);
    ull* scratch = nullptr;
    {
        /* first call on this device: resolve the array's device address.  In host code the
         * symbol's name is only a stand-in; the kernel must be given this address. */
        std::lock_guard<std::mutex> lock(g_setup_mutex);
        if (!g_scratch_addr[dev]) {
            void* p = nullptr;
            MSBFS_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
            g_scratch_addr[dev] = static_cast<ull*>(p);
        }
        scratch = g_scratch_addr[dev];
    }
    /* The address is looked up once per device and kept.  cudaDeviceReset() in this process
     * unloads the module the array belongs to; the kept address is stale after that
     * (msbfs.h says so) and there is no call that would tell us. */
    /* An error some earlier call of the caller's left pending is not this call's failure:
     * report only what the launch itself raises. */
    const cudaError_t before = cudaPeekAtLastError();
    small_batch_kernel<<<1, SMALL_CTA, 0, static_cast<cudaStream_t>(cuda_stream)>>>(
        scratch, n, reinterpret_cast<const long long*>(d_row_offsets), d_col_indices, num_groups,
        reinterpret_cast<const long long*>(d_group_offsets), d_sources, reinterpret_cast<long long*>(d_out_sums));
    const cudaError_t after = cudaPeekAtLastError();
    if (after != cudaSuccess && after != before) {
        cudaGetLastError();
        return msbfs_fail(MSBFS_ERR_CUDA, "msbfs_small_run_noalloc: launch failed: %s", cudaGetErrorString(after));
    }
    return MSBFS_OK;
}
