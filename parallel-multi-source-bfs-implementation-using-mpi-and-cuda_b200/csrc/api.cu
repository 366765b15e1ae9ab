/*
 * api.cu — the extern "C" entry points of include/msbfs.h that touch the GPU:
 * handle life cycle, the streaming loader, query upload / run / download (the NCCL
 * min-reduction lives in comm.cu).  Reference lines replaced are cited in include/msbfs.h.
 */
#include "engine.h"

#include <fcntl.h>
#include <sys/stat.h>
#include <unistd.h>

#include <algorithm>
#include <atomic>
#include <cerrno>
#include <cstring>
#include <thread>
#include <vector>

namespace {

constexpr size_t STAGE_BYTES = 16u << 20; /* one pinned staging buffer */
constexpr int STAGE_THREADS = 4;          /* filler threads (two buffers each) for loads below 1 GiB */
constexpr int STAGE_THREADS_MAX = 16;

/* Pinned staging pipeline: STAGE_THREADS host threads fill pinned buffers (pread from the
 * file, or memcpy from the caller's pageable array) and queue asynchronous H2D copies on
 * the handle's stream; each thread double-buffers, so filling, PCIe transfer and the other
 * threads' fills overlap.  One thread reading the page cache tops out near 4 GB/s; four
 * keep PCIe busy.  The buffers live in the handle and are allocated once.
 * Replaces the per-field fread loop and the pageable blocking cudaMemcpy of the reference
 * (main.cu:108-116, 290-291).  fill(dst, offset, len) returns 0 or an errno-style code. */
int ensure_stage(msbfs_handle* h, int slot, size_t bytes) {
    if (h->stage_buf[slot] && h->stage_cap[slot] < bytes) {
        cudaFreeHost(h->stage_buf[slot]);
        h->stage_buf[slot] = nullptr;
        h->stage_cap[slot] = 0;
    }
    if (!h->stage_buf[slot]) {
        MSBFS_CUDA(cudaHostAlloc((void**)&h->stage_buf[slot], bytes, cudaHostAllocDefault));
        h->stage_cap[slot] = bytes;
    }
    if (!h->stage_ev[slot]) MSBFS_CUDA(cudaEventCreateWithFlags(&h->stage_ev[slot], cudaEventDisableTiming));
    return MSBFS_OK;
}

template <class Fill>
int staged_upload(msbfs_handle* h, char* d_dst, size_t total, Fill fill, const char* what) {
    if (total == 0) return MSBFS_OK;
    const size_t nchunks = (total + STAGE_BYTES - 1) / STAGE_BYTES;
    /* one thread copies ~4 GB/s out of the page cache: four keep up with a small file, a multi-GiB
     * one wants more (option "stage_threads"; by default 8 from 1 GiB on, capped by the cores) */
    int want = h->opt_stage_threads > 0 ? h->opt_stage_threads : (total >= ((size_t)1 << 30) ? 8 : STAGE_THREADS);
    const unsigned cores = std::thread::hardware_concurrency();
    if (h->opt_stage_threads == 0 && cores > 0) want = std::min<int>(want, std::max(2u, cores / 2));
    const int nthreads = (int)std::min<size_t>((size_t)std::min(want, STAGE_THREADS_MAX), nchunks);
    /* only the buffers this upload will use, and no larger than it needs: pinning memory costs
     * milliseconds per buffer, which is most of the preprocessing window of a small graph */
    const size_t one = nchunks == 1 ? ((total + 4095) / 4096) * 4096 : STAGE_BYTES;
    for (int t = 0; t < nthreads; ++t) {
        int rc = ensure_stage(h, 2 * t, one);
        if (rc) return rc;
        if (nchunks > (size_t)(t + nthreads)) { /* this thread gets a second chunk: double-buffer */
            rc = ensure_stage(h, 2 * t + 1, one);
            if (rc) return rc;
        }
    }
    std::atomic<int> failed{0};
    std::atomic<int> cuda_err{0};
    auto worker = [&](int t) {
        cudaSetDevice(h->device);
        int use = 0;
        bool pending[2] = {false, false};
        for (size_t c = (size_t)t; c < nchunks && !failed.load(); c += (size_t)nthreads, use ^= 1) {
            const int slot = 2 * t + use;
            if (pending[use] && cudaEventSynchronize(h->stage_ev[slot]) != cudaSuccess) { cuda_err = 1; failed = 1; break; }
            const size_t off = c * STAGE_BYTES, len = std::min(STAGE_BYTES, total - off);
            if (fill(h->stage_buf[slot], off, len) != 0) { failed = 1; break; }
            if (cudaMemcpyAsync(d_dst + off, h->stage_buf[slot], len, cudaMemcpyHostToDevice, h->stream) != cudaSuccess ||
                cudaEventRecord(h->stage_ev[slot], h->stream) != cudaSuccess) { cuda_err = 1; failed = 1; break; }
            pending[use] = true;
        }
    };
    if (nthreads == 1) {
        worker(0);
    } else {
        std::vector<std::thread> pool;
        for (int t = 0; t < nthreads; ++t) pool.emplace_back(worker, t);
        for (auto& th : pool) th.join();
    }
    const cudaError_t e = cudaStreamSynchronize(h->stream);
    if (cuda_err.load() || e != cudaSuccess)
        return msbfs_fail(MSBFS_ERR_CUDA, "%s: upload failed: %s", what, cudaGetErrorString(e != cudaSuccess ? e : cudaGetLastError()));
    if (failed.load()) return msbfs_fail(MSBFS_ERR_IO, "short read on %s", what);
    return MSBFS_OK;
}

__global__ void degree_lookup_kernel(const int* __restrict__ ids, long long count, const int* __restrict__ perm, int n,
                                     int n_active, const long long* __restrict__ row_ptr, long long* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int s = ids[i];
    long long d = 0;
    if ((unsigned)s < (unsigned)n) {
        const int v = perm[s];
        if (v < n_active) d = row_ptr[v + 1] - row_ptr[v];
    }
    out[i] = d;
}

/* Weight of a source for the run order: its degree plus the degrees of its first 256
 * neighbours (rows are sorted hubs-first, so these are its largest ones) — roughly the size of
 * the frontier two levels after the source, i.e. how fast its BFS reaches the hubs. */
__global__ void degree2_lookup_kernel(const int* __restrict__ ids, long long count, const int* __restrict__ perm, int n,
                                      int n_active, const long long* __restrict__ row_ptr, const int* __restrict__ col,
                                      long long* __restrict__ out) {
    long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    const int s = ids[i];
    long long d = 0;
    if ((unsigned)s < (unsigned)n) {
        const int v = perm[s];
        if (v < n_active) {
            const long long b = row_ptr[v], e = row_ptr[v + 1];
            d = e - b;
            const long long lim = b + 256 < e ? b + 256 : e;
            for (long long k = b; k < lim; ++k) { const int u = col[k]; d += row_ptr[u + 1] - row_ptr[u]; }
        }
    }
    out[i] = d;
}

int set_device(msbfs_handle* h) {
    MSBFS_CUDA(cudaSetDevice(h->device));
    return MSBFS_OK;
}

int ensure_query_capacity(msbfs_handle* h, int64_t groups, int64_t sources) {
    if (groups + 1 > h->q_cap_groups) {
        cudaFree(h->d_q_offsets); cudaFree(h->d_sums);
        h->d_q_offsets = nullptr; h->d_sums = nullptr; h->q_cap_groups = 0;
        const int64_t cap = std::max<int64_t>(groups + 1, 1024);
        MSBFS_CUDA(cudaMalloc(&h->d_q_offsets, sizeof(long long) * (size_t)cap));
        MSBFS_CUDA(cudaMalloc(&h->d_sums, sizeof(long long) * (size_t)cap));
        h->q_cap_groups = cap;
    }
    if (sources > h->q_cap_sources) {
        cudaFree(h->d_q_sources);
        h->d_q_sources = nullptr; h->q_cap_sources = 0;
        const int64_t cap = std::max<int64_t>(sources, 4096);
        MSBFS_CUDA(cudaMalloc(&h->d_q_sources, sizeof(int) * (size_t)cap));
        h->q_cap_sources = cap;
    }
    const size_t need = sizeof(long long) * (size_t)(groups + 1) + sizeof(int) * (size_t)(std::max<int64_t>(sources, 1) + 2)
                        + sizeof(long long) * (size_t)std::max<int64_t>(sources, 1)   /* source degrees (run order) */
                        + sizeof(long long) * (size_t)std::max<int64_t>(groups, 1);
    if (need > h->h_pinned_io_bytes) {
        if (h->h_pinned_io) cudaFreeHost(h->h_pinned_io);
        h->h_pinned_io = nullptr; h->h_pinned_io_bytes = 0;
        const size_t cap = std::max<size_t>(need, 1u << 20);
        MSBFS_CUDA(cudaHostAlloc((void**)&h->h_pinned_io, cap, cudaHostAllocDefault));
        h->h_pinned_io_bytes = cap;
    }
    return MSBFS_OK;
}

} // namespace

extern "C" int msbfs_create(int device, msbfs_t** out) {
    if (!out) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_create: null output pointer");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return msbfs_fail(MSBFS_ERR_CUDA, "no CUDA device available (%s); this engine has no CPU fallback",
                          e != cudaSuccess ? cudaGetErrorString(e) : "device count is 0");
    if (device < 0 || device >= count)
        return msbfs_fail(MSBFS_ERR_INVALID, "device %d out of range [0, %d)", device, count);
    MSBFS_CUDA(cudaSetDevice(device));
    msbfs_handle* h = new (std::nothrow) msbfs_handle();
    if (!h) return msbfs_fail(MSBFS_ERR_NOMEM, "out of host memory");
    h->device = device;
    cudaDeviceGetAttribute(&h->sm_count, cudaDevAttrMultiProcessorCount, device);
    if (h->sm_count <= 0) h->sm_count = 148;
    e = cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking);
    if (e != cudaSuccess) {
        delete h;
        return msbfs_fail(MSBFS_ERR_CUDA, "cudaStreamCreate failed: %s", cudaGetErrorString(e));
    }
    std::memset(&h->stats, 0, sizeof h->stats);
    *out = h;
    return MSBFS_OK;
}

extern "C" int msbfs_destroy(msbfs_t* h) {
    if (!h) return MSBFS_OK;
    cudaSetDevice(h->device);
    if (h->stream) cudaStreamSynchronize(h->stream);
    msbfs_release_traversal(h);
    msbfs_release_graph(h);
    cudaFree(h->d_q_offsets); cudaFree(h->d_q_sources); cudaFree(h->d_sums); cudaFree(h->d_scalar); cudaFree(h->d_deg);
    if (h->h_pinned_io) cudaFreeHost(h->h_pinned_io);
    for (cudaEvent_t e : h->ev_pool) cudaEventDestroy(e);
    for (char* b : h->stage_buf) if (b) cudaFreeHost(b);
    for (cudaEvent_t e : h->stage_ev) if (e) cudaEventDestroy(e);
    if (h->stream) cudaStreamDestroy(h->stream);
    delete h;
    return MSBFS_OK;
}

extern "C" int msbfs_set_option(msbfs_t* h, const char* key, int64_t value) {
    if (!h || !key) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_set_option: null argument");
    if (!std::strcmp(key, "lanes")) {
        if (value != 0 && value != 64 && value != 128 && value != 256 && value != 512)
            return msbfs_fail(MSBFS_ERR_INVALID, "lanes must be 0, 64, 128, 256 or 512");
        h->opt_lanes = (int)value;
    } else if (!std::strcmp(key, "direction")) {
        if (value < 0 || value > 2) return msbfs_fail(MSBFS_ERR_INVALID, "direction must be 0, 1 or 2");
        h->opt_direction = (int)value;
    } else if (!std::strcmp(key, "alpha")) {
        if (value < 1) return msbfs_fail(MSBFS_ERR_INVALID, "alpha must be >= 1");
        h->opt_alpha = value;
    } else if (!std::strcmp(key, "l2_persist")) {
        h->opt_l2_persist = value != 0;
    } else if (!std::strcmp(key, "deg_thread")) {
        if (value < 0 || value > 1024) return msbfs_fail(MSBFS_ERR_INVALID, "deg_thread out of range");
        h->opt_deg_thread = (int)value;
    } else if (!std::strcmp(key, "deg_warp")) {
        if (value < 0 || value > 65536) return msbfs_fail(MSBFS_ERR_INVALID, "deg_warp out of range");
        h->opt_deg_warp = (int)value;
    } else if (!std::strcmp(key, "deg_heavy")) {
        if (value < MSBFS_CHUNK || value > (1 << 30)) return msbfs_fail(MSBFS_ERR_INVALID, "deg_heavy out of range");
        h->opt_deg_heavy = (int)value;
    } else if (!std::strcmp(key, "streak_max_deg")) {
        if (value < 1 || value > 4096) return msbfs_fail(MSBFS_ERR_INVALID, "streak_max_deg out of range");
        h->opt_streak_max_deg = (int)value;
    } else if (!std::strcmp(key, "streak_levels")) {
        if (value < 1) return msbfs_fail(MSBFS_ERR_INVALID, "streak_levels must be >= 1");
        h->opt_streak_levels = (int)std::min<int64_t>(value, 1 << 20);
    } else if (!std::strcmp(key, "reorder")) {
        h->opt_reorder = value != 0;
    } else if (!std::strcmp(key, "streak")) {
        h->opt_streak = value != 0;
    } else if (!std::strcmp(key, "seg")) {
        if (value != 0 && (value < 8 || value > (1 << 20))) return msbfs_fail(MSBFS_ERR_INVALID, "seg out of range");
        h->opt_seg = (int)value;
    } else if (!std::strcmp(key, "expect_groups")) {
        if (value < 0) return msbfs_fail(MSBFS_ERR_INVALID, "expect_groups must be >= 0");
        h->opt_expect_groups = (int)std::min<int64_t>(value, 1 << 30);
    } else if (!std::strcmp(key, "order_key")) {
        if (value != 0 && value != 1) return msbfs_fail(MSBFS_ERR_INVALID, "order_key must be 0 or 1");
        h->opt_order_key = (int)value;
    } else if (!std::strcmp(key, "lane_major")) {
        if (value < -1 || value > 1) return msbfs_fail(MSBFS_ERR_INVALID, "lane_major must be -1, 0 or 1");
        h->opt_lane_major = (int)value;
    } else if (!std::strcmp(key, "lm_ell")) {
        h->opt_lm_ell = value != 0;
        msbfs_lm_release(h); /* rebuilt with or without the records at the next run */
    } else if (!std::strcmp(key, "stage_threads")) {
        if (value < 0 || value > STAGE_THREADS_MAX) return msbfs_fail(MSBFS_ERR_INVALID, "stage_threads out of range [0, 16]");
        h->opt_stage_threads = (int)value;
    } else if (!std::strcmp(key, "tail")) {
        if (value < 0 || value > 2) return msbfs_fail(MSBFS_ERR_INVALID, "tail must be 0, 1 or 2");
        h->opt_tail = (int)value;
    } else if (!std::strcmp(key, "count_missing")) {
        h->opt_count_missing = value != 0;
    } else if (!std::strcmp(key, "col_width")) {
        if (value != 0 && value != 4 && value != 8) return msbfs_fail(MSBFS_ERR_INVALID, "col_width must be 0, 4 or 8");
        h->opt_col_width = (int)value;
    } else if (!std::strcmp(key, "profile")) {
        h->opt_profile = value != 0;
    } else {
        return msbfs_fail(MSBFS_ERR_INVALID, "unknown option '%s'", key);
    }
    return MSBFS_OK;
}

/* ----------------------------------------------------------------- graph --- */

extern "C" int msbfs_set_graph_edges_device(msbfs_t* h, int32_t n, int64_t m, const int32_t* d_uv_pairs) {
    if (!h || n < 0 || m < 0 || (m > 0 && !d_uv_pairs))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_set_graph_edges_device: bad argument");
    if (m > MSBFS_MAX_RECORDS) return msbfs_fail(MSBFS_ERR_RANGE, "graph too large (%lld records)", (long long)m);
    int rc = set_device(h);
    if (rc) return rc;
    /* the caller's buffer may have been written on another stream */
    MSBFS_CUDA(cudaDeviceSynchronize());
    return msbfs_build_from_device_edges(h, n, m, d_uv_pairs);
}

extern "C" int msbfs_set_graph_edges(msbfs_t* h, int32_t n, int64_t m, const int32_t* uv_pairs) {
    if (!h || n < 0 || m < 0 || (m > 0 && !uv_pairs))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_set_graph_edges: bad argument");
    if (m > MSBFS_MAX_RECORDS) return msbfs_fail(MSBFS_ERR_RANGE, "graph too large (%lld records)", (long long)m);
    int rc = set_device(h);
    if (rc) return rc;
    int32_t* d_uv = nullptr;
    const size_t total = sizeof(int32_t) * 2 * (size_t)m;
    MSBFS_CUDA(cudaMalloc(&d_uv, std::max<size_t>(total, 8)));
    const char* src = reinterpret_cast<const char*>(uv_pairs);
    rc = staged_upload(h, reinterpret_cast<char*>(d_uv), total,
                       [&](char* p, size_t off, size_t len) { std::memcpy(p, src + off, len); return 0; }, "edge array");
    if (rc == MSBFS_OK) rc = msbfs_build_from_device_edges(h, n, m, d_uv);
    cudaFree(d_uv);
    return rc;
}

/* Opens a graph file and validates its header (main.cu:94-104): fd, n and m on success. */
int msbfs_open_graph_file(const char* path, int* fd_out, int32_t* n_out, int64_t* m_out) {
    const int fd = open(path, O_RDONLY);
    if (fd < 0) return msbfs_fail(MSBFS_ERR_IO, "Could not open graph file %s", path); /* main.cu:95-99 */
    unsigned char hdr[12];
    if (read(fd, hdr, 12) != 12) { close(fd); return msbfs_fail(MSBFS_ERR_IO, "graph file %s is shorter than its 12-byte header", path); }
    int32_t n; int64_t m;
    std::memcpy(&n, hdr, 4);       /* main.cu:102 */
    std::memcpy(&m, hdr + 4, 8);   /* main.cu:104 */
    if (n < 0 || m < 0) { close(fd); return msbfs_fail(MSBFS_ERR_FORMAT, "graph file %s has a negative size field", path); }
    /* before m enters any product: 12 + 8 m and 2 m wrap for a crafted header */
    if (m > MSBFS_MAX_RECORDS) {
        close(fd);
        return msbfs_fail(MSBFS_ERR_RANGE, "graph file %s declares %lld records; the limit is %lld", path, (long long)m,
                          (long long)MSBFS_MAX_RECORDS);
    }
    struct stat stt;
    if (fstat(fd, &stt) == 0 && S_ISREG(stt.st_mode) && (int64_t)stt.st_size < 12 + 8 * m) {
        close(fd);
        return msbfs_fail(MSBFS_ERR_IO, "graph file %s is truncated: %lld records declared, %lld bytes present", path,
                          (long long)m, (long long)stt.st_size);
    }
#ifdef POSIX_FADV_SEQUENTIAL
    posix_fadvise(fd, 0, 0, POSIX_FADV_SEQUENTIAL);
#endif
    *fd_out = fd; *n_out = n; *m_out = m;
    return MSBFS_OK;
}

/* Records [first, first + count) of an open graph file -> d_dst, through the pinned staging pipeline. */
int msbfs_stage_records(msbfs_handle* h, int fd, int64_t first, int64_t count, char* d_dst, const char* what) {
    const off_t base = (off_t)(12 + 8 * first); /* records start after the 12-byte header */
    return staged_upload(h, d_dst, 8 * (size_t)count,
                         [&](char* p, size_t off, size_t len) {
                             size_t got = 0;
                             while (got < len) {
                                 const ssize_t r = pread(fd, p + got, len - got, base + (off_t)(off + got));
                                 if (r < 0 && errno == EINTR) continue;
                                 if (r <= 0) return 1;
                                 got += (size_t)r;
                             }
                             return 0;
                         }, what);
}

extern "C" int msbfs_load_graph_file(msbfs_t* h, const char* path, int32_t* n_out, int64_t* m_out) {
    if (!h || !path) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_load_graph_file: null argument");
    MsbfsNvtxRange nvtx("msbfs_load_graph_file");
    int rc = set_device(h);
    if (rc) return rc;
    int fd = -1;
    int32_t n; int64_t m;
    rc = msbfs_open_graph_file(path, &fd, &n, &m);
    if (rc) return rc;
    struct Closer { int fd; ~Closer() { close(fd); } } closer{fd};
    int32_t* d_uv = nullptr;
    const size_t total = 8 * (size_t)m;
    MSBFS_CUDA(cudaMalloc(&d_uv, std::max<size_t>(total, 8)));
    rc = msbfs_stage_records(h, fd, 0, m, reinterpret_cast<char*>(d_uv), path);
    if (rc == MSBFS_OK) rc = msbfs_build_from_device_edges(h, n, m, d_uv);
    cudaFree(d_uv);
    if (rc == MSBFS_OK) {
        if (n_out) *n_out = n;
        if (m_out) *m_out = m;
    }
    return rc;
}

/* --------------------------------------------------------------- queries --- */

extern "C" int msbfs_upload_queries(msbfs_t* h, int32_t num_groups, const int64_t* group_offsets,
                                    const int32_t* sources) {
    if (!h || num_groups < 0 || (num_groups > 0 && !group_offsets))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_upload_queries: bad argument");
    MsbfsNvtxRange nvtx("msbfs_upload_queries");
    int rc = set_device(h);
    if (rc) return rc;
    const int64_t base = num_groups > 0 ? group_offsets[0] : 0;
    const int64_t total = num_groups > 0 ? group_offsets[num_groups] - base : 0;
    if (total < 0 || (total > 0 && !sources)) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_upload_queries: bad offsets");
    for (int32_t i = 0; i < num_groups; ++i)
        if (group_offsets[i + 1] < group_offsets[i]) return msbfs_fail(MSBFS_ERR_INVALID, "group offsets must not decrease");
    /* from here on the previous upload is being overwritten: a failure below must not leave its
     * group count behind (download_sums / run_resident would hand out stale sums) */
    h->q_groups = 0;
    h->q_sources = 0;
    h->q_order.clear();
    rc = ensure_query_capacity(h, num_groups, total);
    if (rc) return rc;
    /* stage through pinned memory so both copies are truly asynchronous */
    long long* p_off = h->h_pinned_io;
    int* p_src = reinterpret_cast<int*>(p_off + num_groups + 1);
    for (int32_t i = 0; i <= num_groups; ++i) p_off[i] = num_groups > 0 ? group_offsets[i] - base : 0;
    if (total > 0) std::memcpy(p_src, sources + base, sizeof(int) * (size_t)total);
    if (total > 0)
        MSBFS_CUDA(cudaMemcpyAsync(h->d_q_sources, p_src, sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, h->stream));

    /* Run order: heaviest groups first, weight = sum over a group's sources of their two-hop
     * weight (degree2_lookup_kernel) — a cheap proxy for how fast a group's BFS reaches the hubs.  Batches cut from consecutive
     * runs of that order hold lanes that advance together: they reach the dense levels at the
     * same depth, so the bottom-up kernels' early exit ("every live lane this vertex misses has
     * been found") fires for the whole batch instead of waiting for a few slow lanes.
     * Kronecker-26, 4096 x 16: scanned entries per step 19.5 G -> 14.9 G, lane words fetched
     * 7.9 G -> 4.3 G, 198 -> 146 ms.  The sums go back to the caller's order in
     * msbfs_download_sums; no result depends on the order. */
    h->q_order.clear();
    if (h->opt_reorder && num_groups > 1 && total > 0 && h->has_graph && h->n_active > 0) {
        if (total > h->deg_cap) {
            cudaFree(h->d_deg); h->d_deg = nullptr; h->deg_cap = 0;
            MSBFS_CUDA(cudaMalloc(&h->d_deg, sizeof(long long) * (size_t)total));
            h->deg_cap = total;
        }
        degree2_lookup_kernel<<<(unsigned)((total + 255) / 256), 256, 0, h->stream>>>(h->d_q_sources, total, h->d_perm, h->n,
                                                                                       h->n_active, h->d_row_ptr, h->d_col, h->d_deg);
        long long* p_deg = reinterpret_cast<long long*>(p_src + total + (total & 1));
        MSBFS_CUDA(cudaMemcpyAsync(p_deg, h->d_deg, sizeof(long long) * (size_t)total, cudaMemcpyDeviceToHost, h->stream));
        MSBFS_CUDA(cudaStreamSynchronize(h->stream));
        std::vector<long long> cost((size_t)num_groups, 0);
        for (int32_t g = 0; g < num_groups; ++g)
            for (long long i = p_off[g]; i < p_off[g + 1]; ++i) cost[(size_t)g] += p_deg[i];
        if (h->opt_order_key == 1) { /* spatial proxy: groups ordered by their smallest source id */
            for (int32_t g = 0; g < num_groups; ++g) {
                long long lo = INT64_MAX;
                for (long long i = p_off[g]; i < p_off[g + 1]; ++i) lo = std::min<long long>(lo, p_src[i]);
                cost[(size_t)g] = -lo;
            }
        }
        std::vector<int32_t> order((size_t)num_groups);
        for (int32_t g = 0; g < num_groups; ++g) order[(size_t)g] = g;
        std::stable_sort(order.begin(), order.end(), [&](int32_t x, int32_t y) { return cost[(size_t)x] > cost[(size_t)y]; });
        bool identity = true;
        for (int32_t g = 0; g < num_groups && identity; ++g) identity = order[(size_t)g] == g;
        if (!identity) {
            std::vector<long long> off2((size_t)num_groups + 1, 0);
            std::vector<int> src2((size_t)total);
            for (int32_t p = 0; p < num_groups; ++p) {
                const int32_t g = order[(size_t)p];
                const long long len = p_off[g + 1] - p_off[g];
                std::memcpy(src2.data() + off2[(size_t)p], p_src + p_off[g], sizeof(int) * (size_t)len);
                off2[(size_t)p + 1] = off2[(size_t)p] + len;
            }
            std::memcpy(p_off, off2.data(), sizeof(long long) * ((size_t)num_groups + 1));
            std::memcpy(p_src, src2.data(), sizeof(int) * (size_t)total);
            MSBFS_CUDA(cudaMemcpyAsync(h->d_q_sources, p_src, sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, h->stream));
            h->q_order.swap(order);
        }
    }
    MSBFS_CUDA(cudaMemcpyAsync(h->d_q_offsets, p_off, sizeof(long long) * (size_t)(num_groups + 1), cudaMemcpyHostToDevice, h->stream));
    MSBFS_CUDA(cudaStreamSynchronize(h->stream));
    h->q_groups = num_groups;
    h->q_sources = total;
    return MSBFS_OK;
}

extern "C" int msbfs_run_resident(msbfs_t* h) {
    if (!h) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_run_resident: null handle");
    MsbfsNvtxRange nvtx("msbfs_run_resident");
    int rc = set_device(h);
    if (rc) return rc;
    return msbfs_traverse(h);
}

extern "C" int msbfs_download_sums(msbfs_t* h, int64_t* out_sums, int32_t count) {
    if (!h || count < 0 || (count > 0 && !out_sums)) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_download_sums: bad argument");
    if (count > h->q_groups) return msbfs_fail(MSBFS_ERR_INVALID, "only %d sums are resident", h->q_groups);
    if (count == 0) return MSBFS_OK;
    int rc = set_device(h);
    if (rc) return rc;
    long long* p = h->h_pinned_io;
    const int32_t all = h->q_order.empty() ? count : h->q_groups; /* run order -> caller's order needs every sum */
    MSBFS_CUDA(cudaMemcpyAsync(p, h->d_sums, sizeof(long long) * (size_t)all, cudaMemcpyDeviceToHost, h->stream));
    MSBFS_CUDA(cudaStreamSynchronize(h->stream));
    if (h->q_order.empty()) {
        std::memcpy(out_sums, p, sizeof(long long) * (size_t)count);
    } else {
        for (int32_t pos = 0; pos < all; ++pos) {
            const int32_t g = h->q_order[(size_t)pos];
            if (g < count) out_sums[g] = p[pos];
        }
    }
    return MSBFS_OK;
}

extern "C" int msbfs_run(msbfs_t* h, int32_t num_groups, const int64_t* group_offsets, const int32_t* sources,
                         int64_t* out_sums) {
    if (!h) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_run: null handle");
    if (!h->has_graph) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_run: no graph loaded");
    if (num_groups > 0 && !out_sums) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_run: null output");
    int rc = msbfs_upload_queries(h, num_groups, group_offsets, sources);
    if (rc) return rc;
    rc = msbfs_run_resident(h);
    if (rc) return rc;
    return msbfs_download_sums(h, out_sums, num_groups);
}

extern "C" int msbfs_vertex_degrees(msbfs_t* h, int64_t count, const int32_t* vertex_ids, int64_t* out_degrees) {
    if (!h || count < 0 || (count > 0 && (!vertex_ids || !out_degrees)))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_vertex_degrees: bad argument");
    if (!h->has_graph) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_vertex_degrees: no graph loaded");
    if (count == 0) return MSBFS_OK;
    int rc = set_device(h);
    if (rc) return rc;
    int* d_ids = nullptr;
    long long* d_out = nullptr;
    MSBFS_CUDA(cudaMalloc(&d_ids, sizeof(int) * (size_t)count));
    cudaError_t e = cudaMalloc(&d_out, sizeof(long long) * (size_t)count);
    if (e != cudaSuccess) { cudaFree(d_ids); return msbfs_fail(MSBFS_ERR_CUDA, "cudaMalloc failed: %s", cudaGetErrorString(e)); }
    e = cudaMemcpyAsync(d_ids, vertex_ids, sizeof(int) * (size_t)count, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) {
        degree_lookup_kernel<<<(unsigned)((count + 255) / 256), 256, 0, h->stream>>>(d_ids, count, h->d_perm, h->n, h->n_active,
                                                                                      h->d_row_ptr, d_out);
        e = cudaMemcpyAsync(out_degrees, d_out, sizeof(long long) * (size_t)count, cudaMemcpyDeviceToHost, h->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d_ids);
    cudaFree(d_out);
    if (e != cudaSuccess) return msbfs_fail(MSBFS_ERR_CUDA, "msbfs_vertex_degrees: %s", cudaGetErrorString(e));
    return MSBFS_OK;
}

extern "C" int msbfs_group_weights(msbfs_t* h, int32_t num_groups, const int64_t* group_offsets, const int32_t* sources,
                                   int64_t* out_weights) {
    if (!h || num_groups < 0 || (num_groups > 0 && (!group_offsets || !out_weights)))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_group_weights: bad argument");
    if (!h->has_graph) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_group_weights: no graph loaded");
    if (num_groups == 0) return MSBFS_OK;
    const int64_t base = group_offsets[0], total = group_offsets[num_groups] - base;
    if (total < 0 || (total > 0 && !sources)) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_group_weights: bad offsets");
    for (int32_t g = 0; g < num_groups; ++g) out_weights[g] = 0;
    if (total == 0 || h->n_active == 0) return MSBFS_OK;
    int rc = set_device(h);
    if (rc) return rc;
    int* d_ids = nullptr;
    long long* d_out = nullptr;
    MSBFS_CUDA(cudaMalloc(&d_ids, sizeof(int) * (size_t)total));
    cudaError_t e = cudaMalloc(&d_out, sizeof(long long) * (size_t)total);
    if (e != cudaSuccess) { cudaFree(d_ids); return msbfs_fail(MSBFS_ERR_CUDA, "cudaMalloc failed: %s", cudaGetErrorString(e)); }
    std::vector<long long> w((size_t)total);
    e = cudaMemcpyAsync(d_ids, sources + base, sizeof(int) * (size_t)total, cudaMemcpyHostToDevice, h->stream);
    if (e == cudaSuccess) {
        degree2_lookup_kernel<<<(unsigned)((total + 255) / 256), 256, 0, h->stream>>>(d_ids, total, h->d_perm, h->n, h->n_active,
                                                                                       h->d_row_ptr, h->d_col, d_out);
        e = cudaMemcpyAsync(w.data(), d_out, sizeof(long long) * (size_t)total, cudaMemcpyDeviceToHost, h->stream);
    }
    if (e == cudaSuccess) e = cudaStreamSynchronize(h->stream);
    cudaFree(d_ids);
    cudaFree(d_out);
    if (e != cudaSuccess) return msbfs_fail(MSBFS_ERR_CUDA, "msbfs_group_weights: %s", cudaGetErrorString(e));
    for (int32_t g = 0; g < num_groups; ++g)
        for (int64_t i = group_offsets[g] - base; i < group_offsets[g + 1] - base; ++i) out_weights[g] += w[(size_t)i];
    if (h->opt_order_key == 1) {
        for (int32_t g = 0; g < num_groups; ++g) {
            int64_t lo = INT64_MAX;
            for (int64_t i = group_offsets[g]; i < group_offsets[g + 1]; ++i) lo = std::min<int64_t>(lo, sources[i]);
            out_weights[g] = -lo;
        }
    }
    return MSBFS_OK;
}

extern "C" int msbfs_reached_entries(msbfs_t* h, int64_t* out_entries, int32_t count) {
    if (!h || count < 0 || (count > 0 && !out_entries)) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_reached_entries: bad argument");
    if (count > h->q_groups) return msbfs_fail(MSBFS_ERR_INVALID, "only %d groups are resident", h->q_groups);
    int rc = set_device(h);
    if (rc) return rc;
    if (h->q_order.empty()) return msbfs_reached_entries_impl(h, out_entries, count);
    std::vector<int64_t> run_order((size_t)h->q_groups, 0);
    rc = msbfs_reached_entries_impl(h, run_order.data(), h->q_groups);
    if (rc) return rc;
    for (int32_t pos = 0; pos < h->q_groups; ++pos) {
        const int32_t g = h->q_order[(size_t)pos];
        if (g < count) out_entries[g] = run_order[(size_t)pos];
    }
    return MSBFS_OK;
}

extern "C" int msbfs_get_stats(msbfs_t* h, msbfs_stats* out) {
    if (!h || !out) return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_get_stats: null argument");
    h->stats.n = h->n; h->stats.n_active = h->n_active; h->stats.m = h->m; h->stats.csr_entries = h->nnz;
    h->stats.build_ms = h->build_ms;
    *out = h->stats;
    return MSBFS_OK;
}

extern "C" int msbfs_get_level_log(msbfs_t* h, msbfs_level_record* out, int32_t capacity, int32_t* count_out) {
    if (!h || !count_out || capacity < 0 || (capacity > 0 && !out))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_get_level_log: bad argument");
    static_assert(sizeof(msbfs_level_record) == sizeof(LevelLog), "level record layouts must match");
    const int32_t total = (int32_t)h->level_log.size();
    const int32_t ncopy = total < capacity ? total : capacity;
    if (ncopy > 0) std::memcpy(out, h->level_log.data(), sizeof(LevelLog) * (size_t)ncopy);
    *count_out = total;
    return MSBFS_OK;
}

extern "C" int msbfs_get_launch_log(msbfs_t* h, msbfs_launch_record* out, int32_t capacity, int32_t* count_out) {
    if (!h || !count_out || capacity < 0 || (capacity > 0 && !out))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_get_launch_log: bad argument");
    const int32_t total = (int32_t)h->launch_log.size();
    const int32_t ncopy = total < capacity ? total : capacity;
    if (ncopy > 0) std::memcpy(out, h->launch_log.data(), sizeof(msbfs_launch_record) * (size_t)ncopy);
    *count_out = total;
    return MSBFS_OK;
}

extern "C" int msbfs_get_batch_ms(msbfs_t* h, double* out, int32_t capacity, int32_t* count_out) {
    if (!h || !count_out || capacity < 0 || (capacity > 0 && !out))
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_get_batch_ms: bad argument");
    const int32_t total = (int32_t)h->batch_ms.size();
    const int32_t ncopy = total < capacity ? total : capacity;
    if (ncopy > 0) std::memcpy(out, h->batch_ms.data(), sizeof(double) * (size_t)ncopy);
    *count_out = total;
    return MSBFS_OK;
}

extern "C" int msbfs_num_kernel_slots(void) { return MSBFS_NUM_SLOTS; }

extern "C" int msbfs_get_kernel_profile(msbfs_t* h, int slot, msbfs_kernel_profile* out) {
    if (!h || !out || slot < 0 || slot >= MSBFS_NUM_SLOTS)
        return msbfs_fail(MSBFS_ERR_INVALID, "msbfs_get_kernel_profile: bad argument");
    *out = h->slot_profile[slot];
    return MSBFS_OK;
}
