/*
 * lane_major.cu — the second traversal engine, for road-like graphs (every row short, thousands of
 * BFS levels): one sparse queue-driven BFS per query group, all groups of a batch advancing in
 * one persistent cooperative kernel.
 *
 * Why a second engine.  The bit-parallel engine of traverse.cu shares one pass over the CSR
 * among all lanes of a batch; that pays when the lanes' frontiers overlap (power-law and random
 * graphs: everything meets at the hubs within three levels).  On a lattice with a few random
 * sources per group the lanes' frontiers are thin rings that almost never coincide: every
 * (group, vertex) visit is separate work whatever the layout, and the vertex-major layout then
 * moves ~38 32-byte sectors of 64-byte lane words per visit to set ONE bit.  Here
 * the state is lane-major instead: a visited BITMAP per group (n / 8 bytes) and a queue of
 * (group, vertex) pairs per level; a visit reads one queue entry, two row offsets, the row, and
 * touches ~3 bitmap words: 5 - 6 sectors.
 *
 * Replaces the same reference code as traverse.cu — GPUMultiSourceBFS + BFSKernal + ComputeFofU,
 * /root/reference/main.cu:16-89 — with the same results: a vertex first reached while level L is
 * expanded adds L + 1 to its group's F; unreachable vertices add nothing; source ids outside
 * [0, n) are ignored.
 *
 * The queue holds at most `cap` entries per level (2 x the non-isolated vertices).  A batch whose
 * level outgrows it — frontiers that explode, i.e. a graph on which the bit-parallel engine is
 * the right one — sets `overflow`, and the caller runs that batch through traverse.cu instead.
 */
#include "engine.h"

#include <cooperative_groups.h>

#include <algorithm>

namespace cg = cooperative_groups;

namespace {

typedef unsigned long long ull;
#ifndef MSBFS_LM_CTA
#define MSBFS_LM_CTA 256
#endif
constexpr int LM_CTA = MSBFS_LM_CTA;
#ifndef MSBFS_LM_MINB
#define MSBFS_LM_MINB 4 /* resident blocks per SM asked of the BFS kernel (register cap) */
#endif

struct LmCtrl {
    ull tail[2];                       /* entries in queue 0 / 1 */
    ull expanded, scanned, fresh;      /* whole batch: (group, vertex) visits, CSR entries read, newly visited */
    ull lvl_scanned;                   /* CSR entries read by the level in flight */
    int levels;                        /* expand passes run */
    int overflow;                      /* a level did not fit the queue */
    int log_pos;
    ull sums[MSBFS_MAX_LANES];         /* F per group of the batch */
    ull reach[MSBFS_MAX_LANES];        /* CSR entries of the vertices each group reached */
};

__device__ __forceinline__ ull pack(int lane, int v) { return ((ull)(unsigned)lane << 32) | (unsigned)v; }

/* Append the calling threads' entries (flag = this thread has one) to queue `q`: one atomicAdd
 * on the tail per warp. */
__device__ __forceinline__ void warp_append(ull* q, ull* tail, ull cap, int* overflow, bool flag, ull entry) {
    const unsigned mask = __ballot_sync(0xffffffffu, flag);
    if (mask == 0u) return;
    const int lane = threadIdx.x & 31;
    ull base = 0;
    if (lane == (__ffs(mask) - 1)) base = atomicAdd(tail, (ull)__popc(mask));
    base = __shfl_sync(0xffffffffu, base, __ffs(mask) - 1);
    if (flag) {
        const ull pos = base + (ull)__popc(mask & ((1u << lane) - 1u));
        if (pos < cap) q[pos] = entry;
        else *overflow = 1;
    }
}

/* Sources of the batch's groups: mark and enqueue (level 0: they add 0 to F).  Ids outside
 * [0, n) are dropped (main.cu:49); isolated sources reach nothing. */
__global__ void __launch_bounds__(LM_CTA) lm_inject_kernel(LmCtrl* c, unsigned* __restrict__ bits, long long nwords, ull* __restrict__ q0,
                                                           ull cap, const int* __restrict__ perm, int n, int n_active,
                                                           const long long* __restrict__ q_off, const int* __restrict__ q_src,
                                                           int group_base, int lanes) {
    const int g = (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5);
    const int lane = threadIdx.x & 31;
    if (g >= lanes) return; /* whole warp */
    const long long beg = q_off[group_base + g], end = q_off[group_base + g + 1];
    for (long long i0 = beg; i0 < end; i0 += 32) {
        const long long i = i0 + lane;
        bool fresh = false;
        int v = 0;
        if (i < end) {
            const int s = q_src[i];
            if ((unsigned)s < (unsigned)n) {
                v = perm[s];
                if (v < n_active) {
                    const unsigned bit = 1u << (v & 31);
                    const unsigned old = atomicOr(&bits[(long long)g * nwords + (v >> 5)], bit);
                    fresh = !(old & bit); /* a source listed twice enters once */
                }
            }
        }
        warp_append(q0, &c->tail[0], cap, &c->overflow, fresh, pack(g, v));
    }
}

/* All levels of a batch in one cooperative launch.  Per level: every queue entry (group, v) reads
 * v's row and tries each neighbour u in the group's bitmap — a plain load first (most neighbours
 * of a ring vertex are already in), an atomicOr only for candidates — and enqueues the ones it
 * set.  Per-group counts of the level live in shared memory; F(group) += (level + 1) x count once
 * per level and block, one global atomicAdd per group and block at the very end. */
/* ELL = width of the fixed-width copy of the rows (0: read the CSR).  Road-like graphs have rows of
 * at most 4 or 8 entries: one 16- or 32-byte record per vertex (-1 padded) replaces two row offsets
 * and the row, three sectors, by one. */
template <bool REACH, int ELL>
__global__ void __launch_bounds__(LM_CTA, MSBFS_LM_MINB) lm_bfs_kernel(LmCtrl* c, unsigned* __restrict__ bits, long long nwords, ull* q0, ull* q1, ull cap,
                                                        const long long* __restrict__ row_ptr, const int* __restrict__ col,
                                                        const int* __restrict__ ell, int lanes, int max_degree, LevelLog* log, int log_cap) {
    cg::grid_group grid = cg::this_grid();
    __shared__ unsigned s_cnt[MSBFS_MAX_LANES];   /* newly visited this level, per group */
    __shared__ unsigned s_deg[MSBFS_MAX_LANES];   /* CSR entries of the vertices expanded this level, per group */
    __shared__ ull s_sum[MSBFS_MAX_LANES];
    __shared__ ull s_reach[MSBFS_MAX_LANES];
    __shared__ int s_warp[2][LM_CTA / 32]; /* double-buffered by trip: two barriers per trip instead of three */
    __shared__ ull s_base[2];
    __shared__ ull s_lvl;
    int trip = 0;
    if (threadIdx.x == 0) s_lvl = 0ull;
    for (int j = threadIdx.x; j < MSBFS_MAX_LANES; j += LM_CTA) { s_cnt[j] = 0u; s_deg[j] = 0u; s_sum[j] = 0ull; s_reach[j] = 0ull; }
    __syncthreads();
    const ull gtid = (ull)blockIdx.x * blockDim.x + threadIdx.x, gsize = (ull)gridDim.x * blockDim.x;
    ull expanded = 0, scanned = 0, fresh_total = 0;
    if (blockIdx.x == 0 && threadIdx.x == 0) { /* the injection record, as the other engine writes one per batch */
        if (c->log_pos < log_cap) {
            LevelLog& l = log[c->log_pos];
            l.level = -1; l.direction = 0; l.frontier_vertices = min(c->tail[0], cap); l.frontier_entries = 0; l.unexplored_entries = 0;
            l.scanned = 0; l.words_fetched = 0; l.vertices_examined = 0; l.words_written = 0;
        }
        c->log_pos += 1;
    }
    int level = 0;
    for (;; ++level) {
        const int cur = level & 1;
        const ull n = min(*((volatile ull*)&c->tail[cur]), cap);
        if (n == 0 || *((volatile int*)&c->overflow)) break;
        ull* qc = cur ? q1 : q0;
        ull* qn = cur ? q0 : q1;
        /* whole blocks iterate together (the queue reservation synchronises the block) */
        const ull rounds = (n + gsize - 1) / gsize;
        ull lvl_sc = 0;
        for (ull r = 0; r < rounds; ++r) {
            const ull i = r * gsize + gtid;
            const bool on = i < n;
            int lane_id = 0, v = 0;
            long long beg = 0, end = 0;
            int row[ELL > 0 ? ELL : 1];
#pragma unroll
            for (int k = 0; k < (ELL > 0 ? ELL : 1); ++k) row[k] = -1;
            if (on) {
                const ull e = qc[i];
                lane_id = (int)(e >> 32);
                v = (int)(unsigned)e;
                if (ELL > 0) {
#pragma unroll
                    for (int k4 = 0; k4 < ELL / 4; ++k4) {
                        const int4 r4 = __ldg(reinterpret_cast<const int4*>(ell) + (long long)v * (ELL / 4) + k4);
                        row[4 * k4] = r4.x; row[4 * k4 + 1] = r4.y; row[4 * k4 + 2] = r4.z; row[4 * k4 + 3] = r4.w;
                    }
                    int deg = 0;
#pragma unroll
                    for (int k = 0; k < ELL; ++k) deg += row[k] >= 0;
                    end = deg; /* beg = 0: only the difference is used below */
                } else {
                    beg = row_ptr[v];
                    end = row_ptr[v + 1];
                }
                ++expanded;
                scanned += (ull)(end - beg);
                lvl_sc += (ull)(end - beg);
                if (REACH) atomicAdd(&s_deg[lane_id], (unsigned)(end - beg)); /* only the traversed-edges pass needs it per group */
            }
            unsigned* mybits = bits + (long long)lane_id * nwords;
            /* four neighbours per trip; the trip count is the same for the whole block (block_append
             * synchronises), so it comes from the graph's longest row */
            const int trips_end = ELL > 0 ? ELL : max_degree;
#pragma unroll
            for (int k0 = 0; k0 < trips_end; k0 += 4) {
                int u[4];
                bool cand[4];
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (ELL > 0) {
                        u[k] = row[(k0 + k) < ELL ? (k0 + k) : 0];
                        cand[k] = on && u[k] >= 0;
                        if (!cand[k]) u[k] = 0;
                    } else {
                        const long long e = beg + k0 + k;
                        cand[k] = on && e < end;
                        u[k] = cand[k] ? __ldg(col + e) : 0;
                    }
                }
#pragma unroll
                for (int k = 0; k < 4; ++k)
                    if (cand[k]) cand[k] = !((mybits[u[k] >> 5] >> (u[k] & 31)) & 1u);
                int mine = 0;
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (cand[k]) {
                        const unsigned bit = 1u << (u[k] & 31);
                        const unsigned old = atomicOr(&mybits[u[k] >> 5], bit);
                        cand[k] = !(old & bit); /* now: this thread set it */
                    }
                    mine += cand[k];
                }
                if (mine) { atomicAdd(&s_cnt[lane_id], (unsigned)mine); fresh_total += (ull)mine; }
                /* one reservation on the queue tail per block and trip (a tail bumped once per warp
                 * and neighbour serialises a billion atomics on one address) */
                int incl = mine;
#pragma unroll
                for (int d = 1; d < 32; d <<= 1) {
                    const int t = __shfl_up_sync(0xffffffffu, incl, d);
                    if ((threadIdx.x & 31) >= d) incl += t;
                }
                const int pb = trip & 1;
                ++trip;
                if ((threadIdx.x & 31) == 31) s_warp[pb][threadIdx.x >> 5] = incl;
                __syncthreads();
                if (threadIdx.x == 0) {
                    int total = 0;
#pragma unroll
                    for (int w = 0; w < LM_CTA / 32; ++w) { const int t = s_warp[pb][w]; s_warp[pb][w] = total; total += t; }
                    s_base[pb] = total ? atomicAdd(&c->tail[cur ^ 1], (ull)total) : 0ull;
                }
                __syncthreads();
                ull pos = s_base[pb] + (ull)s_warp[pb][threadIdx.x >> 5] + (ull)(incl - mine);
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    if (cand[k]) {
                        if (pos < cap) qn[pos] = pack(lane_id, u[k]);
                        else c->overflow = 1;
                        ++pos;
                    }
                }
            }
        }
        __syncthreads();
        for (int j = threadIdx.x; j < lanes; j += LM_CTA) {
            s_sum[j] += (ull)(level + 1) * (ull)s_cnt[j];   /* main.cu:81-87, level by level */
            s_cnt[j] = 0u;
            if (REACH) { s_reach[j] += (ull)s_deg[j]; s_deg[j] = 0u; }
        }
        for (int d = 16; d > 0; d >>= 1) lvl_sc += __shfl_xor_sync(0xffffffffu, lvl_sc, d);
        if ((threadIdx.x & 31) == 0 && lvl_sc) atomicAdd(&s_lvl, lvl_sc);
        __syncthreads();
        if (threadIdx.x == 0) { if (s_lvl) atomicAdd(&c->lvl_scanned, s_lvl); s_lvl = 0ull; } /* one global atomic per block and level */
        grid.sync(); /* everyone is done reading queue `cur` and appending to the other one */
        if (blockIdx.x == 0 && threadIdx.x == 0) {
            const ull next_n = c->tail[cur ^ 1];
            if (c->log_pos < log_cap) {
                LevelLog& l = log[c->log_pos];
                l.level = level; l.direction = 3;
                l.frontier_vertices = min(next_n, cap); l.frontier_entries = 0; l.unexplored_entries = 0;
                l.scanned = c->lvl_scanned; l.words_fetched = c->lvl_scanned; l.vertices_examined = n; l.words_written = min(next_n, cap);
            }
            c->log_pos += 1;
            c->lvl_scanned = 0ull;
            c->tail[cur] = 0ull; /* this queue is the next level's output */
            __threadfence();
        }
        grid.sync();
    }
    for (int j = threadIdx.x; j < lanes; j += LM_CTA) {
        if (s_sum[j]) atomicAdd(&c->sums[j], s_sum[j]);
        if (s_reach[j]) atomicAdd(&c->reach[j], s_reach[j]);
    }
    /* batch totals: one atomic per warp */
    for (int d = 16; d > 0; d >>= 1) {
        expanded += __shfl_xor_sync(0xffffffffu, expanded, d);
        scanned += __shfl_xor_sync(0xffffffffu, scanned, d);
        fresh_total += __shfl_xor_sync(0xffffffffu, fresh_total, d);
    }
    if ((threadIdx.x & 31) == 0 && expanded) {
        atomicAdd(&c->expanded, expanded); atomicAdd(&c->scanned, scanned); atomicAdd(&c->fresh, fresh_total);
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) c->levels = level;
}

/* CSR rows -> fixed-width records of W ids, -1 padded (every row has at most W entries) */
__global__ void lm_build_ell_kernel(const long long* __restrict__ row_ptr, const int* __restrict__ col, int n_active, int W,
                                    int* __restrict__ ell) {
    const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (long long)n_active * W) return;
    const int v = (int)(i / W), k = (int)(i % W);
    const long long e = row_ptr[v] + k;
    ell[i] = e < row_ptr[v + 1] ? col[e] : -1;
}

__global__ void lm_store_kernel(const LmCtrl* c, long long* __restrict__ sums, int group_base, int lanes) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < lanes) sums[group_base + j] = (long long)c->sums[j];
}

} // namespace

void msbfs_lm_release(msbfs_handle* h) {
    cudaFree(h->d_lm_bits); h->d_lm_bits = nullptr;
    cudaFree(h->d_lm_q[0]); cudaFree(h->d_lm_q[1]); h->d_lm_q[0] = h->d_lm_q[1] = nullptr;
    cudaFree(h->d_lm_ctrl); h->d_lm_ctrl = nullptr;
    cudaFree(h->d_lm_ell); h->d_lm_ell = nullptr; h->lm_ell_width = 0;
    h->lm_lanes = 0; h->lm_cap = 0; h->lm_grid = 0;
}

/* Sizes the lane-major state for batches of up to `lanes` groups; MSBFS_OK with h->lm_grid == 0
 * when this device cannot co-schedule the grid (the caller then stays with traverse.cu). */
int msbfs_lm_ensure(msbfs_handle* h, int lanes) {
    if (h->d_lm_bits && h->lm_lanes >= lanes) return MSBFS_OK;
    msbfs_lm_release(h);
    int coop = 0, per_sm = 0;
    cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device);
    if (!coop || cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, lm_bfs_kernel<true, 8>, LM_CTA, 0) != cudaSuccess || per_sm <= 0) {
        cudaGetLastError();
        return MSBFS_OK;
    }
    const long long nwords = std::max(h->n_pad, 32) / 32;
    const size_t cap = (size_t)std::max<long long>(2ll * h->n_active, 1 << 16);
    MSBFS_CUDA(cudaMalloc(&h->d_lm_bits, sizeof(unsigned) * (size_t)nwords * (size_t)lanes));
    MSBFS_CUDA(cudaMalloc(&h->d_lm_q[0], sizeof(ull) * cap));
    MSBFS_CUDA(cudaMalloc(&h->d_lm_q[1], sizeof(ull) * cap));
    MSBFS_CUDA(cudaMalloc(&h->d_lm_ctrl, sizeof(LmCtrl)));
    if (h->opt_lm_ell && h->max_degree <= 8u && h->n_active > 0) {
        const int W = h->max_degree <= 4u ? 4 : 8;
        MSBFS_CUDA(cudaMalloc(&h->d_lm_ell, sizeof(int) * (size_t)h->n_active * W));
        const long long items = (long long)h->n_active * W;
        lm_build_ell_kernel<<<(unsigned)((items + 255) / 256), 256, 0, h->stream>>>(h->d_row_ptr, h->d_col, h->n_active, W, h->d_lm_ell);
        MSBFS_CUDA(cudaGetLastError());
        h->lm_ell_width = W;
    }
    h->lm_lanes = lanes;
    h->lm_cap = (long long)cap;
    h->lm_grid = per_sm * h->sm_count;
    return MSBFS_OK;
}

/* One batch (groups [group_base, group_base + lanes) of the resident queries).  On return
 * out->overflow says whether the batch has to be redone by the bit-parallel engine; otherwise its
 * sums are in h->d_sums and its counters in *out. */
int msbfs_lm_run_batch(msbfs_handle* h, int group_base, int lanes, msbfs_lm_result* out) {
    cudaStream_t s = h->stream;
    LmCtrl* c = static_cast<LmCtrl*>(h->d_lm_ctrl);
    long long nwords = std::max(h->n_pad, 32) / 32;
    ull cap = (ull)h->lm_cap;
    MSBFS_CUDA(cudaMemsetAsync(c, 0, sizeof(LmCtrl), s));
    MSBFS_CUDA(cudaMemsetAsync(h->d_lm_bits, 0, sizeof(unsigned) * (size_t)nwords * (size_t)lanes, s));
    lm_inject_kernel<<<(lanes * 32 + LM_CTA - 1) / LM_CTA, LM_CTA, 0, s>>>(c, h->d_lm_bits, nwords, h->d_lm_q[0], cap, h->d_perm, h->n,
                                                                          h->n_active, h->d_q_offsets, h->d_q_sources, group_base, lanes);
    MSBFS_CUDA(cudaGetLastError());
    unsigned* bits = h->d_lm_bits;
    ull *q0 = h->d_lm_q[0], *q1 = h->d_lm_q[1];
    const long long* row_ptr = h->d_row_ptr;
    const int* col = h->d_col;
    LevelLog* log = h->d_log;
    int log_cap = h->log_cap;
    /* the level records of this batch go after the ones the run has written so far */
    int log_pos = out->log_pos_in;
    MSBFS_CUDA(cudaMemcpyAsync(&c->log_pos, &log_pos, sizeof(int), cudaMemcpyHostToDevice, s));
    int max_degree = (int)h->max_degree;
    const int* ell = h->d_lm_ell;
    void* args[] = {&c, &bits, &nwords, &q0, &q1, &cap, &row_ptr, &col, &ell, &lanes, &max_degree, &log, &log_cap};
    const void* fn;
    const bool wr = out->want_reach != 0;
    if (h->lm_ell_width == 4) fn = wr ? (const void*)lm_bfs_kernel<true, 4> : (const void*)lm_bfs_kernel<false, 4>;
    else if (h->lm_ell_width == 8) fn = wr ? (const void*)lm_bfs_kernel<true, 8> : (const void*)lm_bfs_kernel<false, 8>;
    else fn = wr ? (const void*)lm_bfs_kernel<true, 0> : (const void*)lm_bfs_kernel<false, 0>;
    const cudaError_t ce = cudaLaunchCooperativeKernel(fn, dim3(h->lm_grid), dim3(LM_CTA), args, 0, s);
    if (ce != cudaSuccess) return msbfs_fail(MSBFS_ERR_CUDA, "cooperative launch failed: %s", cudaGetErrorString(ce));
    lm_store_kernel<<<(lanes + 255) / 256, 256, 0, s>>>(c, h->d_sums, group_base, lanes);
    LmCtrl hc;
    MSBFS_CUDA(cudaMemcpyAsync(&hc, c, sizeof hc, cudaMemcpyDeviceToHost, s));
    MSBFS_CUDA(cudaStreamSynchronize(s));
    MSBFS_CUDA(cudaGetLastError());
    out->overflow = hc.overflow;
    out->levels = hc.levels;
    out->expanded = (int64_t)hc.expanded;
    out->scanned = (int64_t)hc.scanned;
    out->fresh = (int64_t)hc.fresh;
    out->log_pos_out = hc.log_pos;
    for (int j = 0; j < lanes; ++j) out->reach[j] = hc.reach[j];
    return MSBFS_OK;
}
