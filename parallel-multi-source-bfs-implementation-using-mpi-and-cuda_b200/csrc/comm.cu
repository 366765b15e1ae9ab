/*
 * comm.cu — the NCCL side of the result reduction: the (F, query id) pairs that the reference
 * ships to rank 0 with MPI_Gather + MPI_Gatherv of a struct datatype and then scans
 * (/root/reference/main.cu:166-193, 324-397) become ONE 8-byte unsigned-min all-reduce over
 * NCCL (NVLink / NVSwitch) of the packed word (F << 24 | id).
 *
 * NCCL is resolved at run time (dlopen) so that libmsbfs.so loads on a box without it; the
 * types and enum values come from <nccl.h>.  MPI (or any other launcher) only carries the
 * 128-byte ncclUniqueId from rank 0 to the others (host/main.cpp).  One rank per GPU is a
 * NCCL requirement: with several ranks on one device the caller keeps the MPI reduction.
 */
#include "engine.h"

#include <dlfcn.h>
#include <nccl.h>

#include <unistd.h>

#include <algorithm>
#include <cstring>
#include <mutex>

namespace {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
    ncclResult_t (*AllGather)(const void*, void*, size_t, ncclDataType_t, ncclComm_t, cudaStream_t) = nullptr;
    const char* (*GetErrorString)(ncclResult_t) = nullptr;
    bool ok = false;
};

std::mutex g_mutex;
NcclApi g_api;

const NcclApi* api() {
    std::lock_guard<std::mutex> lock(g_mutex);
    if (g_api.ok) return &g_api;
    void* lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_NOLOAD); /* reuse a copy the process already has */
    if (!lib) lib = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) lib = dlopen("libnccl.so", RTLD_NOW | RTLD_GLOBAL);
    if (!lib) { msbfs_fail(MSBFS_ERR_INVALID, "NCCL not found: %s", dlerror()); return nullptr; }
    g_api.GetUniqueId = (decltype(g_api.GetUniqueId))dlsym(lib, "ncclGetUniqueId");
    g_api.CommInitRank = (decltype(g_api.CommInitRank))dlsym(lib, "ncclCommInitRank");
    g_api.CommDestroy = (decltype(g_api.CommDestroy))dlsym(lib, "ncclCommDestroy");
    g_api.AllReduce = (decltype(g_api.AllReduce))dlsym(lib, "ncclAllReduce");
    g_api.AllGather = (decltype(g_api.AllGather))dlsym(lib, "ncclAllGather");
    g_api.GetErrorString = (decltype(g_api.GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!g_api.GetUniqueId || !g_api.CommInitRank || !g_api.CommDestroy || !g_api.AllReduce || !g_api.AllGather) {
        msbfs_fail(MSBFS_ERR_INVALID, "libnccl lacks a required symbol");
        return nullptr;
    }
    g_api.ok = true;
    return &g_api;
}

int nccl_fail(const NcclApi* a, const char* what, ncclResult_t r) {
    return msbfs_fail(MSBFS_ERR_CUDA, "%s failed: %s", what, a->GetErrorString ? a->GetErrorString(r) : "NCCL error");
}

} // namespace

static_assert(sizeof(ncclUniqueId) == MSBFS_NCCL_ID_BYTES, "msbfs.h states the size of an ncclUniqueId");

extern "C" int msbfs_nccl_unique_id(void* id_out) {
    if (!id_out) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_nccl_unique_id: null argument");
    const NcclApi* a = api();
    if (!a) return MSBFS_ERR_INVALID;
    ncclUniqueId id;
    const ncclResult_t r = a->GetUniqueId(&id);
    if (r != ncclSuccess) return nccl_fail(a, "ncclGetUniqueId", r);
    std::memcpy(id_out, &id, sizeof id);
    return MSBFS_OK;
}

extern "C" int msbfs_nccl_comm_create(msbfs_t* h, const void* id, int32_t world, int32_t rank, void** comm_out) {
    if (!h || !id || !comm_out || world < 1 || rank < 0 || rank >= world)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_nccl_comm_create: bad argument");
    const NcclApi* a = api();
    if (!a) return MSBFS_ERR_INVALID;
    MSBFS_CUDA(cudaSetDevice(h->device));
    ncclUniqueId uid;
    std::memcpy(&uid, id, sizeof uid);
    ncclComm_t comm = nullptr;
    const ncclResult_t r = a->CommInitRank(&comm, world, uid, rank);
    if (r != ncclSuccess) return nccl_fail(a, "ncclCommInitRank", r);
    *comm_out = comm;
    return MSBFS_OK;
}

extern "C" int msbfs_nccl_comm_destroy(void* comm) {
    if (!comm) return MSBFS_OK;
    const NcclApi* a = api();
    if (!a) return MSBFS_ERR_INVALID;
    const ncclResult_t r = a->CommDestroy(static_cast<ncclComm_t>(comm));
    if (r != ncclSuccess) return nccl_fail(a, "ncclCommDestroy", r);
    return MSBFS_OK;
}

extern "C" int msbfs_allreduce_min_u64(msbfs_t* h, void* nccl_comm, uint64_t* inout) {
    if (!h || !nccl_comm || !inout) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_allreduce_min_u64: null argument");
    const NcclApi* a = api();
    if (!a) return MSBFS_ERR_INVALID;
    MSBFS_CUDA(cudaSetDevice(h->device));
    if (!h->d_scalar) MSBFS_CUDA(cudaMalloc(&h->d_scalar, 8));
    uint64_t* d = reinterpret_cast<uint64_t*>(h->d_scalar);
    MSBFS_CUDA(cudaMemcpyAsync(d, inout, 8, cudaMemcpyHostToDevice, h->stream));
    const ncclResult_t r = a->AllReduce(d, d, 1, ncclUint64, ncclMin, static_cast<ncclComm_t>(nccl_comm), h->stream);
    if (r != ncclSuccess) return nccl_fail(a, "ncclAllReduce", r);
    MSBFS_CUDA(cudaMemcpyAsync(inout, d, 8, cudaMemcpyDeviceToHost, h->stream));
    MSBFS_CUDA(cudaStreamSynchronize(h->stream));
    return MSBFS_OK;
}

/* The graph file read ONCE per box instead of once per rank: rank r reads records
 * [r * chunk, (r + 1) * chunk) through the pinned staging pipeline into its slice of a
 * device buffer, one ncclAllGather over NVLink completes every rank's copy of the edge list,
 * and each rank builds its CSR on its own GPU.  Replaces rank 0's LoadGraphBin + host-memory
 * MPI_Bcast of the CSR + per-rank upload (main.cu:237-255, 283-291); with 8 ranks the host
 * side moves 8 GiB instead of 64 GiB for Kronecker-26. */
extern "C" int msbfs_load_graph_file_sharded(msbfs_t* h, const char* path, void* nccl_comm, int32_t rank, int32_t world,
                                             int32_t* n_out, int64_t* m_out) {
    if (!h || !path || !nccl_comm || world < 1 || rank < 0 || rank >= world)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_load_graph_file_sharded: bad argument");
    const NcclApi* a = api();
    if (!a) return MSBFS_ERR_INVALID;
    MsbfsNvtxRange nvtx("msbfs_load_graph_file_sharded");
    MSBFS_CUDA(cudaSetDevice(h->device));
    int fd = -1;
    int32_t n; int64_t m;
    int rc = msbfs_open_graph_file(path, &fd, &n, &m);
    if (rc) return rc; /* every rank sees the same file: all fail here together, before the collective */
    struct Closer { int fd; ~Closer() { close(fd); } } closer{fd};
    const int64_t chunk = (m + world - 1) / world;             /* records per rank, the last slice padded */
    const int64_t first = std::min<int64_t>(m, (int64_t)rank * chunk);
    const int64_t count = std::min<int64_t>(m - first, chunk);
    char* d_uv = nullptr;
    const size_t slice = 8 * (size_t)chunk;
    MSBFS_CUDA(cudaMalloc(&d_uv, std::max<size_t>(slice * (size_t)world, 8)));
    char* mine = d_uv + slice * (size_t)rank;
    cudaError_t ce = cudaSuccess;
    if (count < chunk) ce = cudaMemsetAsync(mine + 8 * (size_t)count, 0, 8 * (size_t)(chunk - count), h->stream);
    rc = (ce == cudaSuccess) ? msbfs_stage_records(h, fd, first, count, mine, path) : msbfs_fail(MSBFS_ERR_CUDA, "cudaMemsetAsync failed");
    /* a rank whose read failed still takes part in the collective (its peers are waiting in it) */
    ncclResult_t r = ncclSuccess;
    if (slice > 0) r = a->AllGather(mine, d_uv, slice, ncclChar, static_cast<ncclComm_t>(nccl_comm), h->stream);
    ce = cudaStreamSynchronize(h->stream);
    if (rc == MSBFS_OK && r != ncclSuccess) rc = nccl_fail(a, "ncclAllGather", r);
    if (rc == MSBFS_OK && ce != cudaSuccess) rc = msbfs_fail(MSBFS_ERR_CUDA, "edge all-gather failed: %s", cudaGetErrorString(ce));
    if (rc == MSBFS_OK) rc = msbfs_build_from_device_edges(h, n, m, reinterpret_cast<const int32_t*>(d_uv));
    cudaFree(d_uv);
    if (rc == MSBFS_OK) {
        if (n_out) *n_out = n;
        if (m_out) *m_out = m;
    }
    return rc;
}
