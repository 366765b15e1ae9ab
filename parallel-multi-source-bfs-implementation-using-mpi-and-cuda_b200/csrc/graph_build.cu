/*
 * graph_build.cu — device-side CSR construction.
 *
 * Replaces the host adjacency build of the reference (LoadGraphBin,
 * /root/reference/main.cu:106-129: vector<vector<int>> push_back per record, then a
 * flatten with int32 offsets) and the blocking CSR upload (main.cu:283-291).  The
 * raw (u, v) records arrive in HBM (streamed by api.cu or written by gen.cu) and
 * everything else happens on the GPU:
 *
 *   1. validate ids and histogram degrees                      (degree_kernel)
 *   2. stable radix sort of vertices by degree, descending     (cub)
 *      -> internal ids: hubs first, isolated vertices last
 *   3. 64-bit exclusive scan of the sorted degrees -> row_ptr  (cub)
 *   4. emit both directions of every record as a 64-bit key
 *      (row << 32 | neighbour) in internal ids                 (emit_keys_kernel)
 *   5. radix sort the keys; the low halves are col             (cub + split kernel)
 *
 * The result has the same multiset of neighbours per vertex as the reference CSR
 * (both directions, duplicates and self-loops kept, main.cu:114-115); only the
 * labelling and the order inside a row differ, and neither can change a hop
 * distance or a distance sum.  Sorting rows by internal id puts the highest-degree
 * neighbour first, which is what lets the bottom-up kernels stop early.
 */
#include "engine.h"

#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

#include <algorithm>

namespace {

__global__ void degree_kernel(const int2* __restrict__ uv, long long m, int n,
                              unsigned int* __restrict__ deg, int* __restrict__ bad) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < m; i += stride) {
        const int2 e = uv[i];
        if ((unsigned)e.x >= (unsigned)n || (unsigned)e.y >= (unsigned)n) { *bad = 1; continue; }
        atomicAdd(&deg[e.x], 1u);
        atomicAdd(&deg[e.y], 1u);
    }
}

__global__ void iota_kernel(int* __restrict__ ids, int n) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) ids[i] = i;
}

/* order[new] = old  ->  perm[old] = new; also 64-bit degrees for the scan and the
 * count of non-isolated vertices (degrees are sorted descending). */
__global__ void perm_kernel(const int* __restrict__ order, const unsigned int* __restrict__ deg_sorted,
                            int n, int n_scan, int* __restrict__ perm, long long* __restrict__ deg64,
                            int* __restrict__ n_active) {
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) {
        perm[order[i]] = i;
        const unsigned d = deg_sorted[i];
        if (d > 0 && (i == n - 1 || deg_sorted[i + 1] == 0)) *n_active = i + 1;
    }
    if (i < n_scan) deg64[i] = (i < n) ? (long long)deg_sorted[i] : 0;
}

__global__ void emit_keys_kernel(const int2* __restrict__ uv, long long m, const int* __restrict__ perm,
                                 unsigned long long* __restrict__ keys) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < m; i += stride) {
        const int2 e = uv[i];
        const unsigned long long pu = (unsigned)perm[e.x], pv = (unsigned)perm[e.y];
        ulonglong2 k;
        k.x = (pu << 32) | pv;   /* row u gets v   (main.cu:114) */
        k.y = (pv << 32) | pu;   /* row v gets u   (main.cu:115) */
        reinterpret_cast<ulonglong2*>(keys)[i] = k;
    }
}

__global__ void split_keys_kernel(const unsigned long long* __restrict__ keys, long long nnz, int* __restrict__ col) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const long long stride = (long long)gridDim.x * blockDim.x;
    for (; i < nnz; i += stride) col[i] = (int)(unsigned)(keys[i] & 0xffffffffull);
}

/* first index whose (descending) degree is <= limit[t]; one thread per threshold */
__global__ void bin_bounds_kernel(const unsigned int* __restrict__ deg_sorted, int n,
                                  const unsigned int* __restrict__ limit, int nlimits, int* __restrict__ out) {
    int t = threadIdx.x;
    if (t >= nlimits) return;
    int lo = 0, hi = n;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if (deg_sorted[mid] > limit[t]) lo = mid + 1; else hi = mid;
    }
    out[t] = lo;
}

/* cnt[k] = number of vertices (of the first n_multi, degrees descending) with more than
 * k * seg entries, for k in [0, levels]; cnt[levels] = 0 */
__global__ void seg_level_count_kernel(const unsigned int* __restrict__ deg_sorted, int n_multi, int seg, int levels,
                                       long long* __restrict__ cnt) {
    int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k > levels) return;
    const unsigned long long limit = (unsigned long long)k * (unsigned long long)seg;
    int lo = 0, hi = n_multi;
    while (lo < hi) {
        int mid = (lo + hi) >> 1;
        if ((unsigned long long)deg_sorted[mid] > limit) lo = mid + 1; else hi = mid;
    }
    cnt[k] = (k == levels) ? 0 : lo;
}

struct DevBuf {
    void* p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    cudaError_t alloc(size_t bytes) { return cudaMalloc(&p, bytes ? bytes : 1); }
    template <class T> T* as() { return static_cast<T*>(p); }
};

inline int grid_for(long long items, int threads, int sm_count) {
    long long blocks = (items + threads - 1) / threads;
    long long cap = (long long)sm_count * 16;
    return (int)std::max<long long>(1, std::min(blocks, cap));
}

} // namespace

void msbfs_release_graph(msbfs_handle* h) {
    if (h->graph_imported) { /* mapped from the owning process: unmap */
        if (h->d_perm) cudaIpcCloseMemHandle(h->d_perm);
        if (h->d_row_ptr) cudaIpcCloseMemHandle(h->d_row_ptr);
        if (h->d_col) cudaIpcCloseMemHandle(h->d_col);
        cudaGetLastError();
    } else {
        cudaFree(h->d_perm);
        cudaFree(h->d_row_ptr);
        cudaFree(h->d_col);
    }
    h->graph_imported = false;
    h->d_perm = nullptr; h->d_row_ptr = nullptr; h->d_col = nullptr;
    cudaFree(h->d_chunk_level_base); h->d_chunk_level_base = nullptr;
    cudaFree(h->d_seg_level_base); h->d_seg_level_base = nullptr;
    h->has_graph = false;
    h->n = h->n_active = h->n_pad = 0;
    h->m = h->nnz = 0;
}

int msbfs_build_from_device_edges(msbfs_handle* h, int32_t n, int64_t m, const int32_t* d_uv) {
    if (n < 0 || m < 0) return msbfs_fail(MSBFS_ERR_INVALID, "negative graph size");
    if (m > MSBFS_MAX_RECORDS) return msbfs_fail(MSBFS_ERR_RANGE, "graph too large (%lld records)", (long long)m);
    MsbfsNvtxRange nvtx("msbfs_build_csr");
    cudaStream_t s = h->stream;
    const int T = 256;

    struct EventPair { /* destroyed on every return path */
        cudaEvent_t a = nullptr, b = nullptr;
        ~EventPair() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    } evp;
    MSBFS_CUDA(cudaEventCreate(&evp.a));
    MSBFS_CUDA(cudaEventCreate(&evp.b));
    cudaEvent_t ev0 = evp.a, ev1 = evp.b;
    MSBFS_CUDA(cudaEventRecord(ev0, s));

    const int64_t nnz = 2 * m;
    const int n_alloc = std::max(n, 1);

    DevBuf deg, deg_sorted, ids, order, flags, tmp;
    MSBFS_CUDA(deg.alloc(sizeof(unsigned) * (size_t)n_alloc));
    MSBFS_CUDA(deg_sorted.alloc(sizeof(unsigned) * (size_t)n_alloc));
    MSBFS_CUDA(ids.alloc(sizeof(int) * (size_t)n_alloc));
    MSBFS_CUDA(order.alloc(sizeof(int) * (size_t)n_alloc));
    MSBFS_CUDA(flags.alloc(64));
    MSBFS_CUDA(cudaMemsetAsync(deg.p, 0, sizeof(unsigned) * (size_t)n_alloc, s));
    MSBFS_CUDA(cudaMemsetAsync(flags.p, 0, 64, s));
    int* d_bad = flags.as<int>();
    int* d_nact = flags.as<int>() + 1;

    /* 1. degrees + validation */
    if (m > 0) {
        degree_kernel<<<grid_for(m, T, h->sm_count), T, 0, s>>>(reinterpret_cast<const int2*>(d_uv), m, n,
                                                               deg.as<unsigned>(), d_bad);
        MSBFS_CUDA(cudaGetLastError());
    }
    int h_flags[2] = {0, 0};
    MSBFS_CUDA(cudaMemcpyAsync(h_flags, flags.p, 8, cudaMemcpyDeviceToHost, s));
    MSBFS_CUDA(cudaStreamSynchronize(s));
    if (h_flags[0]) return msbfs_fail(MSBFS_ERR_FORMAT, "edge record with a vertex id outside [0, %d)", n);
    /* the new edge list is valid: only now does the old graph go (a bad file leaves it in place) */
    msbfs_release_graph(h);
    msbfs_release_traversal(h);

    /* 2. vertices by degree, descending (stable: ties keep ascending file id) */
    if (n > 0) {
        iota_kernel<<<(n + T - 1) / T, T, 0, s>>>(ids.as<int>(), n);
        size_t tb = 0;
        MSBFS_CUDA(cub::DeviceRadixSort::SortPairsDescending(nullptr, tb, deg.as<unsigned>(), deg_sorted.as<unsigned>(),
                                                            ids.as<int>(), order.as<int>(), n, 0, 32, s));
        MSBFS_CUDA(tmp.alloc(tb));
        MSBFS_CUDA(cub::DeviceRadixSort::SortPairsDescending(tmp.p, tb, deg.as<unsigned>(), deg_sorted.as<unsigned>(),
                                                            ids.as<int>(), order.as<int>(), n, 0, 32, s));
    }

    /* 3. perm, n_active, row_ptr */
    MSBFS_CUDA(cudaMalloc(&h->d_perm, sizeof(int) * (size_t)n_alloc));
    const int n_pad_max = ((n + 31) / 32) * 32;
    const int n_scan = n_pad_max + 1;
    DevBuf deg64;
    MSBFS_CUDA(deg64.alloc(sizeof(long long) * (size_t)n_scan));
    perm_kernel<<<(n_scan + T - 1) / T, T, 0, s>>>(order.as<int>(), deg_sorted.as<unsigned>(), n, n_scan, h->d_perm,
                                                   deg64.as<long long>(), d_nact);
    MSBFS_CUDA(cudaGetLastError());
    MSBFS_CUDA(cudaMemcpyAsync(h_flags, flags.p, 8, cudaMemcpyDeviceToHost, s));
    MSBFS_CUDA(cudaStreamSynchronize(s));
    const int n_active = h_flags[1];
    const int n_pad = ((n_active + 31) / 32) * 32;

    MSBFS_CUDA(cudaMalloc(&h->d_row_ptr, sizeof(long long) * (size_t)(n_pad + 1)));
    {
        DevBuf tmp2;
        size_t tb = 0;
        MSBFS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, deg64.as<long long>(), h->d_row_ptr, n_pad + 1, s));
        MSBFS_CUDA(tmp2.alloc(tb));
        MSBFS_CUDA(cub::DeviceScan::ExclusiveSum(tmp2.p, tb, deg64.as<long long>(), h->d_row_ptr, n_pad + 1, s));
        MSBFS_CUDA(cudaStreamSynchronize(s));
    }

    /* degree bins over the sorted order (block = 32 consecutive internal ids) */
    {
        DevBuf lim, out;
        const unsigned limits[3] = {(unsigned)h->opt_deg_heavy - 1u, (unsigned)h->opt_deg_warp, (unsigned)h->opt_deg_thread};
        int bounds[3] = {0, 0, 0};
        MSBFS_CUDA(lim.alloc(sizeof limits));
        MSBFS_CUDA(out.alloc(sizeof bounds));
        MSBFS_CUDA(cudaMemcpyAsync(lim.p, limits, sizeof limits, cudaMemcpyHostToDevice, s));
        bin_bounds_kernel<<<1, 32, 0, s>>>(deg_sorted.as<unsigned>(), n_active, lim.as<unsigned>(), 3, out.as<int>());
        MSBFS_CUDA(cudaGetLastError());
        MSBFS_CUDA(cudaMemcpyAsync(bounds, out.p, sizeof bounds, cudaMemcpyDeviceToHost, s));
        MSBFS_CUDA(cudaStreamSynchronize(s));
        const int nblk = n_pad / 32;
        auto up = [&](int v) { return std::min(nblk, (v + 31) / 32); };
        h->blk_heavy = up(bounds[0]);
        h->blk_warp = std::max(h->blk_heavy, up(bounds[1]));
        h->blk_g8 = std::max(h->blk_warp, up(bounds[2]));
        h->blk_end = nblk;
    }

    /* heavy rows: chunk c of level k belongs to vertex (c - base[k]); level k holds
     * every vertex with more than k * CHUNK entries — a prefix, degrees being sorted.
     * Ordering chunks level-major makes the first chunks of all hubs run first, so
     * later chunks of a row usually find its lanes already complete and exit. */
    {
        const int nh = std::min(h->blk_heavy * 32, n_active);
        std::vector<unsigned> hdeg((size_t)nh);
        if (nh > 0) {
            MSBFS_CUDA(cudaMemcpyAsync(hdeg.data(), deg_sorted.p, sizeof(unsigned) * (size_t)nh, cudaMemcpyDeviceToHost, s));
            MSBFS_CUDA(cudaStreamSynchronize(s));
        }
        std::vector<long long> base(1, 0);
        if (nh > 0) {
            const long long levels = ((long long)hdeg[0] + MSBFS_CHUNK - 1) / MSBFS_CHUNK;
            int cnt = nh;
            for (long long k = 0; k < levels; ++k) {
                while (cnt > 0 && (long long)hdeg[(size_t)cnt - 1] <= k * MSBFS_CHUNK) --cnt;
                base.push_back(base.back() + cnt);
            }
        }
        h->n_chunk_levels = (int)base.size() - 1;
        h->n_chunks = base.back();
        MSBFS_CUDA(cudaMalloc(&h->d_chunk_level_base, sizeof(long long) * base.size()));
        MSBFS_CUDA(cudaMemcpyAsync(h->d_chunk_level_base, base.data(), sizeof(long long) * base.size(),
                                   cudaMemcpyHostToDevice, s));
        MSBFS_CUDA(cudaStreamSynchronize(s));
    }

    /* bottom-up segments of the long rows (deg > seg), level-major */
    {
        /* segment length: 256 entries on small graphs (more segments = more threads), up to 1024 on
         * multi-billion-entry ones (fewer atomics and look-ups); measured in profiles/r1/sweep_r2a/b.txt */
        const int seg = h->opt_seg > 0 ? h->opt_seg : (int)std::min<long long>(1024, std::max<long long>(256, nnz >> 21));
        DevBuf lim, out;
        const unsigned limit = (unsigned)seg;
        int first_short = 0;
        unsigned maxdeg = 0;
        MSBFS_CUDA(lim.alloc(sizeof limit));
        MSBFS_CUDA(out.alloc(sizeof first_short));
        MSBFS_CUDA(cudaMemcpyAsync(lim.p, &limit, sizeof limit, cudaMemcpyHostToDevice, s));
        bin_bounds_kernel<<<1, 32, 0, s>>>(deg_sorted.as<unsigned>(), n_active, lim.as<unsigned>(), 1, out.as<int>());
        MSBFS_CUDA(cudaGetLastError());
        MSBFS_CUDA(cudaMemcpyAsync(&first_short, out.p, sizeof first_short, cudaMemcpyDeviceToHost, s));
        if (n_active > 0) MSBFS_CUDA(cudaMemcpyAsync(&maxdeg, deg_sorted.p, sizeof maxdeg, cudaMemcpyDeviceToHost, s));
        MSBFS_CUDA(cudaStreamSynchronize(s));
        const int nblk = n_pad / 32;
        h->seg = seg;
        h->max_degree = maxdeg;
        h->blk_multi = std::min(nblk, (first_short + 31) / 32);
        const int n_multi = std::min(h->blk_multi * 32, n_active);
        const int levels = n_multi > 0 ? (int)(((long long)maxdeg + seg - 1) / seg) : 0;
        h->n_seg_levels = levels;
        MSBFS_CUDA(cudaMalloc(&h->d_seg_level_base, sizeof(long long) * (size_t)(levels + 2)));
        DevBuf cnt, tmp4;
        MSBFS_CUDA(cnt.alloc(sizeof(long long) * (size_t)(levels + 2)));
        seg_level_count_kernel<<<(levels + 1 + T - 1) / T, T, 0, s>>>(deg_sorted.as<unsigned>(), n_multi, seg, levels, cnt.as<long long>());
        MSBFS_CUDA(cudaGetLastError());
        size_t tb = 0;
        MSBFS_CUDA(cub::DeviceScan::ExclusiveSum(nullptr, tb, cnt.as<long long>(), h->d_seg_level_base, levels + 1, s));
        MSBFS_CUDA(tmp4.alloc(tb));
        MSBFS_CUDA(cub::DeviceScan::ExclusiveSum(tmp4.p, tb, cnt.as<long long>(), h->d_seg_level_base, levels + 1, s));
        long long total = 0;
        MSBFS_CUDA(cudaMemcpyAsync(&total, h->d_seg_level_base + levels, sizeof total, cudaMemcpyDeviceToHost, s));
        MSBFS_CUDA(cudaStreamSynchronize(s));
        h->n_seg_items = total;
    }

    /* the degree scratch is no longer needed: release before the big key buffers */
    cudaFree(deg.p); deg.p = nullptr;
    cudaFree(ids.p); ids.p = nullptr;
    cudaFree(order.p); order.p = nullptr;
    cudaFree(deg64.p); deg64.p = nullptr;
    cudaFree(tmp.p); tmp.p = nullptr;
    cudaFree(deg_sorted.p); deg_sorted.p = nullptr;

    /* 4 + 5. keys -> sort -> col */
    MSBFS_CUDA(cudaMalloc(&h->d_col, sizeof(int) * (size_t)(nnz + 8))); /* +8: 128-bit loads of the last entries */
    if (nnz > 0) {
        DevBuf k0, k1, tmp3;
        MSBFS_CUDA(k0.alloc(sizeof(unsigned long long) * (size_t)nnz));
        MSBFS_CUDA(k1.alloc(sizeof(unsigned long long) * (size_t)nnz));
        emit_keys_kernel<<<grid_for(m, T, h->sm_count), T, 0, s>>>(reinterpret_cast<const int2*>(d_uv), m, h->d_perm,
                                                                  k0.as<unsigned long long>());
        MSBFS_CUDA(cudaGetLastError());
        int id_bits = 1;
        while (id_bits < 32 && ((long long)1 << id_bits) < (long long)std::max(n_active, 1)) ++id_bits;
        cub::DoubleBuffer<unsigned long long> db(k0.as<unsigned long long>(), k1.as<unsigned long long>());
        size_t tb = 0;
        MSBFS_CUDA(cub::DeviceRadixSort::SortKeys(nullptr, tb, db, (long long)nnz, 0, 32 + id_bits, s));
        MSBFS_CUDA(tmp3.alloc(tb));
        MSBFS_CUDA(cub::DeviceRadixSort::SortKeys(tmp3.p, tb, db, (long long)nnz, 0, 32 + id_bits, s));
        split_keys_kernel<<<grid_for(nnz, T, h->sm_count), T, 0, s>>>(db.Current(), nnz, h->d_col);
        MSBFS_CUDA(cudaGetLastError());
        MSBFS_CUDA(cudaStreamSynchronize(s));
    }

    MSBFS_CUDA(cudaEventRecord(ev1, s));
    MSBFS_CUDA(cudaEventSynchronize(ev1));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ev0, ev1);

    h->n = n; h->m = m; h->nnz = nnz;
    h->n_active = n_active; h->n_pad = n_pad;
    h->has_graph = true;
    h->build_ms = ms;
    msbfs_prealloc_traversal(h);
    return MSBFS_OK;
}
