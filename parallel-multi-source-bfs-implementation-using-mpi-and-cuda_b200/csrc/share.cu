/*
 * share.cu — one CSR per GPU when several ranks share it.
 *
 * The reference gives every rank its own copy of the CSR on its device (cudaMalloc + H2D per
 * rank, /root/reference/main.cu:283-291, device = rank % numGPU, main.cu:227): with
 * `mpirun -np 2 ... -gn 1` the graph sits twice in the memory of GPU 0 and is uploaded twice.
 * Here the lowest rank on a device loads the file and builds the CSR once; the others map
 * the same arrays read-only through CUDA IPC handles:
 *
 *   owner:  msbfs_export_graph  -> an opaque blob (sizes, bin limits, the two small level
 *                                  tables and three cudaIpcMemHandle_t: perm, row_ptr, col)
 *   host:   ships the blob to the peers (MPI_Bcast of bytes in host/main.cpp)
 *   peer:   msbfs_import_graph  -> a handle whose graph arrays are the owner's
 *
 * The traversal state (lane words, bitmaps) stays private to each handle.  The owner must
 * keep its handle alive until every peer has destroyed (or re-loaded) its own.
 */
#include "engine.h"

#include <cstring>
#include <vector>

namespace {

constexpr uint32_t BLOB_MAGIC = 0x4853424du; /* "MBSH" */

struct BlobHeader {
    uint32_t magic, version;
    int32_t device;
    int32_t n, n_active, n_pad;
    int64_t m, nnz;
    int32_t blk_heavy, blk_warp, blk_g8, blk_end;
    int32_t n_chunk_levels, n_seg_levels, blk_multi, seg;
    int64_t n_chunks, n_seg_items;
    uint32_t max_degree, pad;
    cudaIpcMemHandle_t perm, row_ptr, col;
};

} // namespace

extern "C" int msbfs_export_graph(msbfs_t* h, void* blob, int64_t capacity, int64_t* size_out) {
    if (!h || !size_out || capacity < 0 || (capacity > 0 && !blob))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_export_graph: bad argument");
    if (!h->has_graph) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_export_graph: no graph loaded");
    if (h->graph_imported) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_export_graph: this handle's graph is itself imported");
    const size_t n_chunk_tab = (size_t)h->n_chunk_levels + 1, n_seg_tab = (size_t)h->n_seg_levels + 2;
    const int64_t need = (int64_t)(sizeof(BlobHeader) + sizeof(long long) * (n_chunk_tab + n_seg_tab));
    *size_out = need;
    if (capacity < need) return capacity == 0 ? MSBFS_OK : msbfs_fail(MSBFS_ERR_RANGE, "msbfs_export_graph: blob needs %lld bytes", (long long)need);
    MSBFS_CUDA(cudaSetDevice(h->device));
    BlobHeader hd;
    std::memset(&hd, 0, sizeof hd);
    hd.magic = BLOB_MAGIC; hd.version = 1; hd.device = h->device;
    hd.n = h->n; hd.n_active = h->n_active; hd.n_pad = h->n_pad; hd.m = h->m; hd.nnz = h->nnz;
    hd.blk_heavy = h->blk_heavy; hd.blk_warp = h->blk_warp; hd.blk_g8 = h->blk_g8; hd.blk_end = h->blk_end;
    hd.n_chunk_levels = h->n_chunk_levels; hd.n_seg_levels = h->n_seg_levels; hd.blk_multi = h->blk_multi; hd.seg = h->seg;
    hd.n_chunks = h->n_chunks; hd.n_seg_items = h->n_seg_items; hd.max_degree = h->max_degree;
    MSBFS_CUDA(cudaIpcGetMemHandle(&hd.perm, h->d_perm));
    MSBFS_CUDA(cudaIpcGetMemHandle(&hd.row_ptr, h->d_row_ptr));
    MSBFS_CUDA(cudaIpcGetMemHandle(&hd.col, h->d_col));
    char* out = static_cast<char*>(blob);
    std::memcpy(out, &hd, sizeof hd);
    MSBFS_CUDA(cudaMemcpy(out + sizeof hd, h->d_chunk_level_base, sizeof(long long) * n_chunk_tab, cudaMemcpyDeviceToHost));
    MSBFS_CUDA(cudaMemcpy(out + sizeof hd + sizeof(long long) * n_chunk_tab, h->d_seg_level_base, sizeof(long long) * n_seg_tab,
                          cudaMemcpyDeviceToHost));
    return MSBFS_OK;
}

extern "C" int msbfs_import_graph(msbfs_t* h, const void* blob, int64_t size) {
    if (!h || !blob || size < (int64_t)sizeof(BlobHeader)) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_import_graph: bad argument");
    BlobHeader hd;
    std::memcpy(&hd, blob, sizeof hd);
    if (hd.magic != BLOB_MAGIC || hd.version != 1) return msbfs_fail(MSBFS_ERR_FORMAT, "msbfs_import_graph: not a graph blob");
    if (hd.device != h->device)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_import_graph: blob is for device %d, this handle is on device %d", hd.device, h->device);
    const size_t n_chunk_tab = (size_t)hd.n_chunk_levels + 1, n_seg_tab = (size_t)hd.n_seg_levels + 2;
    if (size < (int64_t)(sizeof hd + sizeof(long long) * (n_chunk_tab + n_seg_tab)))
        return msbfs_fail(MSBFS_ERR_FORMAT, "msbfs_import_graph: blob is truncated");
    MSBFS_CUDA(cudaSetDevice(h->device));
    msbfs_release_graph(h);
    msbfs_release_traversal(h);
    void *perm = nullptr, *row = nullptr, *col = nullptr;
    cudaError_t e = cudaIpcOpenMemHandle(&perm, hd.perm, cudaIpcMemLazyEnablePeerAccess);
    if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&row, hd.row_ptr, cudaIpcMemLazyEnablePeerAccess);
    if (e == cudaSuccess) e = cudaIpcOpenMemHandle(&col, hd.col, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        if (perm) cudaIpcCloseMemHandle(perm);
        if (row) cudaIpcCloseMemHandle(row);
        if (col) cudaIpcCloseMemHandle(col);
        cudaGetLastError();
        return msbfs_fail(MSBFS_ERR_CUDA, "msbfs_import_graph: cudaIpcOpenMemHandle failed: %s", cudaGetErrorString(e));
    }
    h->graph_imported = true;
    h->d_perm = static_cast<int32_t*>(perm);
    h->d_row_ptr = static_cast<long long*>(row);
    h->d_col = static_cast<int*>(col);
    const char* in = static_cast<const char*>(blob) + sizeof hd;
    MSBFS_CUDA(cudaMalloc(&h->d_chunk_level_base, sizeof(long long) * n_chunk_tab));
    MSBFS_CUDA(cudaMalloc(&h->d_seg_level_base, sizeof(long long) * n_seg_tab));
    MSBFS_CUDA(cudaMemcpy(h->d_chunk_level_base, in, sizeof(long long) * n_chunk_tab, cudaMemcpyHostToDevice));
    MSBFS_CUDA(cudaMemcpy(h->d_seg_level_base, in + sizeof(long long) * n_chunk_tab, sizeof(long long) * n_seg_tab, cudaMemcpyHostToDevice));
    h->n = hd.n; h->n_active = hd.n_active; h->n_pad = hd.n_pad; h->m = hd.m; h->nnz = hd.nnz;
    h->blk_heavy = hd.blk_heavy; h->blk_warp = hd.blk_warp; h->blk_g8 = hd.blk_g8; h->blk_end = hd.blk_end;
    h->n_chunk_levels = hd.n_chunk_levels; h->n_seg_levels = hd.n_seg_levels; h->blk_multi = hd.blk_multi; h->seg = hd.seg;
    h->n_chunks = hd.n_chunks; h->n_seg_items = hd.n_seg_items; h->max_degree = hd.max_degree;
    h->build_ms = 0.0;
    h->has_graph = true;
    msbfs_prealloc_traversal(h);
    return MSBFS_OK;
}
