/*
 * gen.cu — device side of the synthetic-input recipes (gen_common.h): writes the
 * (u, v) records of an R-MAT / Kronecker / lattice / Erdos-Renyi graph straight
 * into HBM, so that the 8 GiB edge list of the Kronecker scale-26 config never has
 * to cross PCIe or a file system.  The reference has no generator;
 * the record layout is its graph.bin payload (/root/reference/main.cu:108-116).
 */
#include "internal.h"
#include "gen_common.h"

namespace {

__global__ void gen_edges_kernel(int kind, int p0, int p1, unsigned n, long long m, unsigned long long seed,
                                 int2* __restrict__ uv) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < m; i += stride) {
        int32_t u, v;
        switch (kind) {
            case MSBFS_GEN_RMAT: msbfs_gen_rmat_edge(seed, p0, (uint64_t)i, 0, &u, &v); break;
            case MSBFS_GEN_KRON: msbfs_gen_rmat_edge(seed, p0, (uint64_t)i, 1, &u, &v); break;
            case MSBFS_GEN_LATTICE: msbfs_gen_lattice_edge((uint32_t)p0, (uint32_t)p1, (uint64_t)i, &u, &v); break;
            default: msbfs_gen_er_edge(seed, n, (uint64_t)i, &u, &v); break;
        }
        uv[i] = make_int2(u, v);
    }
}

} // namespace

extern "C" int msbfs_gen_edges_device(int device, int kind, int32_t p0, int32_t p1, int64_t m, uint64_t seed,
                                      int32_t* d_uv_out) {
    int32_t n;
    int rc = msbfs_gen_num_vertices(kind, p0, p1, &n);
    if (rc) return rc;
    if (m < 0 || (m > 0 && !d_uv_out)) return msbfs_fail(MSBFS_ERR_INVALID, "bad edge buffer");
    if (kind == MSBFS_GEN_LATTICE && (uint64_t)m != msbfs_lattice_num_edges((uint32_t)p0, (uint32_t)p1))
        return msbfs_fail(MSBFS_ERR_INVALID, "lattice %dx%d has %llu records, not %lld", p0, p1,
                          (unsigned long long)msbfs_lattice_num_edges((uint32_t)p0, (uint32_t)p1), (long long)m);
    MSBFS_CUDA(cudaSetDevice(device));
    if (m == 0) return MSBFS_OK;
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device);
    const int threads = 256;
    long long blocks = (m + threads - 1) / threads;
    if (blocks > (long long)sms * 32) blocks = (long long)sms * 32;
    gen_edges_kernel<<<(int)blocks, threads>>>(kind, p0, p1, (unsigned)n, m, seed, reinterpret_cast<int2*>(d_uv_out));
    MSBFS_CUDA(cudaGetLastError());
    MSBFS_CUDA(cudaDeviceSynchronize());
    return MSBFS_OK;
}
