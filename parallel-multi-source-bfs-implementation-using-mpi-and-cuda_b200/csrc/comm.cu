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

#include <cstring>
#include <mutex>

namespace {

struct NcclApi {
    ncclResult_t (*GetUniqueId)(ncclUniqueId*) = nullptr;
    ncclResult_t (*CommInitRank)(ncclComm_t*, int, ncclUniqueId, int) = nullptr;
    ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
    ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
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
    g_api.GetErrorString = (decltype(g_api.GetErrorString))dlsym(lib, "ncclGetErrorString");
    if (!g_api.GetUniqueId || !g_api.CommInitRank || !g_api.CommDestroy || !g_api.AllReduce) {
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
