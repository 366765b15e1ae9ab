/*
 * traverse.cu — the bit-parallel, direction-optimizing multi-source BFS kernels
 * (sm_100a) and the level loop that drives them.
 *
 * Replaces, for a whole batch of query groups at once, what the reference does one
 * group at a time: GPUMultiSourceBFS (/root/reference/main.cu:40-73, one launch of
 * BFSKernal main.cu:16-38 and three blocking runtime calls per BFS level) and
 * ComputeFofU (main.cu:75-89, a D2H copy of all distances and a serial host sum).
 *
 * Up to 512 query groups ("lanes") advance together.  Per vertex, bit j of a lane
 * word says something about group j: visited, in the current frontier, in the next
 * frontier.  A BFS level is then, for every vertex, an OR over neighbours' frontier
 * words; one pass over the CSR serves all lanes.  Distances are never stored: a
 * vertex first reached by lane j while level L is expanded adds (L + 1) to F(j), so
 * F(j) = sum over levels of (L + 1) x (vertices newly reached by j), which is
 * exactly the reference's sum over non-negative distances (main.cu:81-87);
 * unreachable vertices add nothing (main.cu:84-85) and ids outside [0, n) never
 * enter a frontier (main.cu:49).
 *
 * Kernels (NW = 64-bit words per lane word; G = threads cooperating on one vertex):
 *   inject      sources -> acc words + touched bitmap
 *   td_expand   top-down: frontier vertices push their word to neighbours (RED.OR)
 *   td_heavy    same for chunked hub rows
 *   td_finalize acc -> new frontier word, visited |= new
 *   bu_thread   bottom-up, rows of at most `seg` entries: one thread per not-yet-full
 *               vertex ORs its neighbours' frontier words and stops as soon as all the
 *               live lanes it misses are found; fused finalize
 *   bu_segment  bottom-up, longer rows: one thread per `seg`-entry segment, level-major,
 *               segments of a row co-operate through hub_acc
 *   bu_heavy_finalize  hub_acc -> new frontier words of the long rows
 *   td_streak   persistent cooperative top-down loop for short-row (road-like) graphs
 *   tail_list / tail_levels   once few vertices are left unfinished: list them, then run every
 *               remaining level bottom-up over the list in one cooperative launch
 *   count       positional popcount (per-lane counts) with bit-sliced carry-save counters and
 *               warp bit-matrix transposes: of the new frontier words, or, after a dense
 *               bottom-up level, of the lanes the unfinished vertices still miss
 *   epilogue    F += (L+1) x counts, live-lane mask, direction for the next level
 * Road-like graphs (every row short) go to the second engine in lane_major.cu instead.
 */
#include "engine.h"

#include <cooperative_groups.h>

#include <algorithm>
#include <cstring>

namespace cg = cooperative_groups;

namespace {

typedef unsigned long long ull;

constexpr int CTA = 256;
#ifndef MSBFS_NW8_BLOCKS
#define MSBFS_NW8_BLOCKS 2 /* resident 256-thread blocks per SM asked of the 512-lane bottom-up kernels */
#endif

/* ------------------------------------------------------------ lane words --- */

template <int NW> struct alignas(8 * NW) LW { ull w[NW]; };

template <int NW> __device__ __forceinline__ LW<NW> lw_zero() {
    LW<NW> r;
#pragma unroll
    for (int i = 0; i < NW; ++i) r.w[i] = 0ull;
    return r;
}
template <int NW> __device__ __forceinline__ bool lw_any(const LW<NW>& a) {
    ull r = 0ull;
#pragma unroll
    for (int i = 0; i < NW; ++i) r |= a.w[i];
    return r != 0ull;
}
template <int NW> __device__ __forceinline__ LW<NW> lw_or(const LW<NW>& a, const LW<NW>& b) {
    LW<NW> r;
#pragma unroll
    for (int i = 0; i < NW; ++i) r.w[i] = a.w[i] | b.w[i];
    return r;
}
template <int NW> __device__ __forceinline__ LW<NW> lw_and(const LW<NW>& a, const LW<NW>& b) {
    LW<NW> r;
#pragma unroll
    for (int i = 0; i < NW; ++i) r.w[i] = a.w[i] & b.w[i];
    return r;
}
template <int NW> __device__ __forceinline__ LW<NW> lw_andn(const LW<NW>& a, const LW<NW>& b) { /* a & ~b */
    LW<NW> r;
#pragma unroll
    for (int i = 0; i < NW; ++i) r.w[i] = a.w[i] & ~b.w[i];
    return r;
}
/* (a & b) == b : every bit of b is present in a */
template <int NW> __device__ __forceinline__ bool lw_covers(const LW<NW>& a, const LW<NW>& b) {
    ull r = 0ull;
#pragma unroll
    for (int i = 0; i < NW; ++i) r |= (b.w[i] & ~a.w[i]);
    return r == 0ull;
}
/* at NW = 4 these compile to one LDG.E.256 / STG.E.256 (32-byte sector) */
template <int NW> __device__ __forceinline__ LW<NW> lw_load(const ull* base, long long idx) {
    return *reinterpret_cast<const LW<NW>*>(base + idx * NW);
}
template <int NW> __device__ __forceinline__ LW<NW> lw_load_volatile(const ull* base, long long idx) {
    LW<NW> r;
    const volatile ull* p = base + idx * NW;
#pragma unroll
    for (int i = 0; i < NW; ++i) r.w[i] = p[i];
    return r;
}
template <int NW> __device__ __forceinline__ void lw_store(ull* base, long long idx, const LW<NW>& v) {
    *reinterpret_cast<LW<NW>*>(base + idx * NW) = v;
}
template <int NW> __device__ __forceinline__ LW<NW> lw_from(const ull* p) {
    LW<NW> r;
#pragma unroll
    for (int i = 0; i < NW; ++i) r.w[i] = p[i];
    return r;
}

/* OR across the G threads named in gmask (redux.sync) */
template <int NW, int G> __device__ __forceinline__ LW<NW> group_or(const LW<NW>& a, unsigned gmask) {
    if (G == 1) return a;
    LW<NW> r;
#pragma unroll
    for (int i = 0; i < NW; ++i) {
        unsigned lo = (unsigned)a.w[i], hi = (unsigned)(a.w[i] >> 32);
        lo = __reduce_or_sync(gmask, lo);
        hi = __reduce_or_sync(gmask, hi);
        r.w[i] = ((ull)hi << 32) | lo;
    }
    return r;
}

__device__ __forceinline__ ull warp_sum(ull v) {
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
    return v;
}

/* A vertex that has just been visited by every lane of the batch leaves the "unexplored"
 * totals the direction heuristic works with: one vertex and deg CSR entries.  Both ride in
 * one 64-bit accumulator (count << 40 | entries) until the warp flushes them. */
constexpr int FULL_SHIFT = 40;
__device__ __forceinline__ void note_full(ull& acc, long long deg) { acc += (1ull << FULL_SHIFT) + (ull)deg; }
__device__ __forceinline__ void commit_full(LevelCtrl* c, ull packed) {
    if (!packed) return;
    atomicAdd((ull*)&c->m_u, (ull)(-(long long)(packed & ((1ull << FULL_SHIFT) - 1ull))));
    atomicAdd((ull*)&c->n_u, (ull)(-(long long)(packed >> FULL_SHIFT)));
}
__device__ __forceinline__ ull warp_sum(ull v);
__device__ __forceinline__ void flush_full(LevelCtrl* c, ull acc, int lane) {
    acc = warp_sum(acc);
    if (lane == 0) commit_full(c, acc);
}

/* The bottom-up kernels know their new frontier as they write it: vertices, their CSR entries and
 * whether a chunked (hub) row is among them go to the control block here, so that the count pass
 * that follows is free to count something else (see count_kernel). */
__device__ __forceinline__ void flush_frontier_totals(LevelCtrl* c, ull nf, ull mf, unsigned heavy, int lane) {
    nf = warp_sum(nf);
    mf = warp_sum(mf);
    if (lane == 0 && nf) { atomicAdd(&c->n_f, nf); atomicAdd(&c->m_f, mf); }
    if (__any_sync(0xffffffffu, heavy != 0u) && lane == 0) c->heavy_frontier = 1u;
}

__device__ __forceinline__ bool test_bit(const unsigned* bitmap, int v) {
    return (bitmap[v >> 5] >> (v & 31)) & 1u;
}

/* warps of a launch walk 32-vertex blocks with this stride */
__device__ __forceinline__ int global_warp() { return (int)((blockIdx.x * blockDim.x + threadIdx.x) >> 5); }
__device__ __forceinline__ int total_warps() { return (int)((gridDim.x * blockDim.x) >> 5); }

/* ---------------------------------------------------------------- inject --- */

/* Sources of the batch's groups become bits in acc (all-zero on entry) and in the
 * touched bitmap; td_finalize then turns them into the level-0 frontier.  Ids
 * outside [0, n) are dropped (main.cu:49); isolated sources reach nothing and add
 * 0 to F, so they are dropped as well. */
template <int NW>
__global__ void __launch_bounds__(CTA) inject_kernel(TravArgs a, const int* __restrict__ perm, int n, int n_active,
                                                     const long long* __restrict__ q_off, const int* __restrict__ q_src,
                                                     int group_base, int lanes) {
    const int g = global_warp();
    const int lane = threadIdx.x & 31;
    if (g >= lanes) return;
    const long long beg = q_off[group_base + g], end = q_off[group_base + g + 1];
    for (long long i = beg + lane; i < end; i += 32) {
        const int s = q_src[i];
        if ((unsigned)s >= (unsigned)n) continue;
        const int v = perm[s];
        if (v >= n_active) continue;
        atomicOr(&a.acc[(long long)v * NW + (g >> 6)], 1ull << (g & 63));
        atomicOr(&a.nxt_nz[v >> 5], 1u << (v & 31));
    }
}

/* ------------------------------------------------------------- top-down --- */

/* One frontier vertex v pushes cur[v] to every neighbour u that still misses some
 * of those lanes.  U entries are staged per thread: U independent col loads, then U
 * bitmap tests, then U visited-word loads, so a thread keeps up to U random
 * requests in flight. */
template <int NW, int U>
__device__ __forceinline__ void td_push(const TravArgs& a, const LW<NW>& f, long long e, long long end, int stride,
                                        ull& scanned, ull& fetched, ull& written) {
    for (; e < end; e += (long long)stride * U) {
        int u[U];
        bool ok[U];
#pragma unroll
        for (int k = 0; k < U; ++k) {
            const long long ee = e + (long long)k * stride;
            ok[k] = ee < end;
            u[k] = ok[k] ? __ldg(a.col + ee) : 0;
        }
#pragma unroll
        for (int k = 0; k < U; ++k) {
            scanned += ok[k];
            ok[k] = ok[k] && !test_bit(a.vis_full, u[k]);
        }
        LW<NW> vis[U];
#pragma unroll
        for (int k = 0; k < U; ++k) vis[k] = ok[k] ? lw_load<NW>(a.visited, u[k]) : lw_zero<NW>();
#pragma unroll
        for (int k = 0; k < U; ++k) {
            if (!ok[k]) continue;
            ++fetched;
            const LW<NW> d = lw_andn(f, vis[k]);
            if (!lw_any(d)) continue;
#pragma unroll
            for (int i = 0; i < NW; ++i)
                if (d.w[i]) atomicOr(&a.acc[(long long)u[k] * NW + i], d.w[i]);
            ++written;
            if (!test_bit(a.nxt_nz, u[k])) atomicOr(&a.nxt_nz[u[k] >> 5], 1u << (u[k] & 31));
        }
    }
}

/* Variant of td_push for G co-operating threads with 128-bit col loads: thread `sub` takes
 * entries [4 sub, 4 sub + 4) of every 4 G-entry stretch of the row (experiment MSBFS_TD_INT4,
 * profiles/r2/sweep_f_*: no gain — top-down is bound by the random visited reads and the
 * atomics, not by the col stream, which the scalar form already reads coalesced). */
template <int NW, int G>
__device__ __forceinline__ void td_push_vec(const TravArgs& a, const LW<NW>& f, long long beg, long long end, int sub,
                                            ull& scanned, ull& fetched, ull& written) {
    for (long long e4 = (beg & ~3ll) + 4ll * sub; e4 < end; e4 += 4ll * G) {
        const int4 c = __ldg(reinterpret_cast<const int4*>(a.col + e4));
        const int u[4] = {c.x, c.y, c.z, c.w};
        bool ok[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            ok[k] = (e4 + k >= beg) && (e4 + k < end);
            scanned += ok[k];
            ok[k] = ok[k] && !test_bit(a.vis_full, u[k]);
        }
        LW<NW> vis[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) vis[k] = ok[k] ? lw_load<NW>(a.visited, u[k]) : lw_zero<NW>();
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (!ok[k]) continue;
            ++fetched;
            const LW<NW> d = lw_andn(f, vis[k]);
            if (!lw_any(d)) continue;
#pragma unroll
            for (int i = 0; i < NW; ++i)
                if (d.w[i]) atomicOr(&a.acc[(long long)u[k] * NW + i], d.w[i]);
            ++written;
            if (!test_bit(a.nxt_nz, u[k])) atomicOr(&a.nxt_nz[u[k] >> 5], 1u << (u[k] & 31));
        }
    }
}

__device__ __forceinline__ void flush_counters(LevelCtrl* c, int slot, ull scanned, ull fetched, ull examined, ull written) { /* per-thread counts may be 32-bit: a thread never scans 2^32 entries in one launch */
    scanned = warp_sum(scanned);
    fetched = warp_sum(fetched);
    examined = warp_sum(examined);
    written = warp_sum(written);
    if ((threadIdx.x & 31) == 0) {
        if (scanned) { atomicAdd(&c->scanned, scanned); atomicAdd(&c->k_scanned[slot], scanned); }
        if (fetched) { atomicAdd(&c->words_fetched, fetched); atomicAdd(&c->k_fetched[slot], fetched); }
        if (examined) { atomicAdd(&c->examined, examined); atomicAdd(&c->k_examined[slot], examined); }
        if (written) { atomicAdd(&c->words_written, written); atomicAdd(&c->k_written[slot], written); }
    }
}

template <int NW, int G, int U>
__global__ void __launch_bounds__(CTA) td_expand_kernel(TravArgs a, int blk_lo, int blk_hi) {
    constexpr int VPI = 32 / G; /* frontier vertices handled per pass of the warp */
    const int lane = threadIdx.x & 31;
    const int sub = lane % G, grp = lane / G;
    ull scanned = 0, fetched = 0, examined = 0, written = 0;
    for (int b = blk_lo + global_warp(); b < blk_hi; b += total_warps()) {
        const unsigned w = a.cur_nz[b];
        if (w == 0u) continue;
        const int nset = __popc(w);
        for (int k0 = 0; k0 < nset; k0 += VPI) {
            const int k = k0 + grp;
            if (k >= nset) continue;
            const int v = (b << 5) + (int)__fns(w, 0, k + 1);
            const LW<NW> f = lw_load<NW>(a.cur, v);
            const long long beg = a.row_ptr[v], end = a.row_ptr[v + 1];
            if (sub == 0) ++examined;
#ifdef MSBFS_TD_INT4
            if (G >= 8) td_push_vec<NW, G>(a, f, beg, end, sub, scanned, fetched, written);
            else
#endif
            td_push<NW, U>(a, f, beg + sub, end, G, scanned, fetched, written);
        }
    }
    flush_counters(a.ctrl, G == 32 ? SLOT_TD_G32 : G == 8 ? SLOT_TD_G8 : SLOT_TD_G2, scanned, fetched, examined, written);
}

/* chunk id -> (level k, vertex): level k holds vertices [0, base[k+1]-base[k]) */
__device__ __forceinline__ void chunk_lookup(const TravArgs& a, long long c, int& k_out, int& v_out) {
    int lo = 0, hi = a.n_chunk_levels; /* largest k with base[k] <= c */
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (a.chunk_level_base[mid] <= c) lo = mid; else hi = mid;
    }
    k_out = lo;
    v_out = (int)(c - a.chunk_level_base[lo]);
}

template <int NW, int U>
__global__ void __launch_bounds__(CTA) td_heavy_kernel(TravArgs a) {
    const int lane = threadIdx.x & 31;
    ull scanned = 0, fetched = 0, examined = 0, written = 0;
    for (long long c = global_warp(); c < a.n_chunks; c += total_warps()) {
        int k, v;
        chunk_lookup(a, c, k, v);
        if (!test_bit(a.cur_nz, v)) continue;
        const LW<NW> f = lw_load<NW>(a.cur, v);
        const long long beg = a.row_ptr[v] + (long long)k * MSBFS_CHUNK;
        const long long end = min(beg + (long long)MSBFS_CHUNK, a.row_ptr[v + 1]);
        if (lane == 0 && k == 0) ++examined;
#ifdef MSBFS_TD_INT4
        td_push_vec<NW, 32>(a, f, beg, end, lane, scanned, fetched, written);
#else
        td_push<NW, U>(a, f, beg + lane, end, 32, scanned, fetched, written);
#endif
    }
    flush_counters(a.ctrl, SLOT_TD_HEAVY, scanned, fetched, examined, written);
}

/* A warp takes a tile of 32 bitmap words (1024 vertices, one coalesced 128-byte read) and
 * hands the set bits out densely: round r gives lane j the (32 r + j)-th set bit of the
 * tile.  With a few frontier vertices per 32-vertex block this keeps all 32 lanes busy
 * and 32 independent load chains in flight instead of two or three. */
struct TileBits {
    unsigned w;      /* this lane's bitmap word */
    int excl, total; /* set bits in the lanes below / in the whole tile */
    __device__ __forceinline__ void load(unsigned word, int lane) {
        w = word;
        const int cnt = __popc(w);
        int incl = cnt;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += t;
        }
        excl = incl - cnt;
        total = __shfl_sync(0xffffffffu, incl, 31);
    }
    /* vertex offset inside the tile (0..1023) of set bit number `rank`; all lanes must call */
    __device__ __forceinline__ int select(int rank) const {
        int lo = 0;
#pragma unroll
        for (int step = 16; step > 0; step >>= 1) {
            const int cand = lo + step;
            const int e = __shfl_sync(0xffffffffu, excl, cand & 31);
            if (cand < 32 && e <= rank) lo = cand;
        }
        const unsigned wi = __shfl_sync(0xffffffffu, w, lo);
        const int ei = __shfl_sync(0xffffffffu, excl, lo);
        const int r = rank - ei;
        const int bit = (r >= 0 && r < __popc(wi)) ? (int)__fns(wi, 0, r + 1) : 0;
        return (lo << 5) + bit;
    }
};

/* acc -> new frontier.  acc already excludes visited lanes (td_push masks with
 * ~visited, and visited does not change while a level is expanded).  A warp takes a tile of 1024
 * vertices of the touched bitmap: where few are touched (the usual top-down level) the set bits
 * are dealt out densely to the 32 lanes, so every lane has a vertex and 32 independent load
 * chains are in flight; a densely touched tile is walked block by block, lane = vertex. */
template <int NW>
__device__ __forceinline__ void td_finalize_vertex(const TravArgs& a, int v, const LW<NW>& active, ull& full_acc) {
    /* both loads first: a store in between would keep the second load from being issued early */
    const LW<NW> x = lw_load<NW>(a.acc, v);
    const LW<NW> vis = lw_load<NW>(a.visited, v);
    const LW<NW> nv = lw_or(vis, x);
    lw_store<NW>(a.acc, v, lw_zero<NW>());
    lw_store<NW>(a.visited, v, nv);
    lw_store<NW>(a.nxt, v, x);
    if (lw_covers(nv, active)) {
        note_full(full_acc, a.row_ptr[v + 1] - a.row_ptr[v]);
        atomicOr(&a.vis_full[v >> 5], 1u << (v & 31));
    }
}

template <int NW>
__global__ void __launch_bounds__(CTA) td_finalize_kernel(TravArgs a, int nblk) {
    const int lane = threadIdx.x & 31;
    const LW<NW> active = lw_from<NW>(a.ctrl->active);
    const int ntiles = (nblk + 31) >> 5;
    ull full_acc = 0;
    if (ntiles < total_warps()) { /* small graph: a 32-vertex block per warp keeps more warps busy than tiles would */
        for (int b = global_warp(); b < nblk; b += total_warps()) {
            const unsigned w = a.nxt_nz[b];
            if ((w >> lane) & 1u) td_finalize_vertex<NW>(a, (b << 5) + lane, active, full_acc);
        }
        flush_full(a.ctrl, full_acc, lane);
        return;
    }
    for (int t = global_warp(); t < ntiles; t += total_warps()) {
        const int b = (t << 5) + lane;
        TileBits tb;
        tb.load(b < nblk ? a.nxt_nz[b] : 0u, lane);
        if (tb.total == 0) continue;
        if (tb.total > 256) {
            for (int k = 0; k < 32; ++k) {
                const unsigned w = __shfl_sync(0xffffffffu, tb.w, k);
                if ((w >> lane) & 1u) td_finalize_vertex<NW>(a, (t << 10) + (k << 5) + lane, active, full_acc);
            }
        } else {
            for (int r0 = 0; r0 < tb.total; r0 += 32) {
                const int off = tb.select(r0 + lane); /* warp-collective */
                if (r0 + lane < tb.total) td_finalize_vertex<NW>(a, (t << 10) + off, active, full_acc);
            }
        }
    }
    flush_full(a.ctrl, full_acc, lane);
}

/* ------------------------------------------------------------ bottom-up --- */

/* Bottom-up scan of one row segment [beg, end) by one thread, software-pipelined.
 * col is read four entries at a time with one 128-bit load (the group index is a multiple
 * of 4, so the address is 16-byte aligned; entries outside [beg, end) are masked) — a
 * thread walking its own row touches each 128-byte line of col with 8 requests instead of
 * 32.  While the lane words of group i are in flight, the col entries and the
 * frontier-bitmap words of group i+1 are already being fetched, so an iteration costs one
 * memory latency (the words) instead of three chained ones (col -> bitmap -> word): these
 * kernels are bound by latency x occupancy, not by bytes (ncu r1f).  Stops as soon as
 * every lane in `need` has been found. */
template <int W> struct alignas(4 * W) ColVec { int v[W]; };

/* W = col entries per load (4: one 128-bit load, 8: one 256-bit load). */
template <int NW, int W>
__device__ __forceinline__ LW<NW> bu_scan_row(const TravArgs& a, long long beg, long long end,
                                              const LW<NW>& need, unsigned& scanned, unsigned& fetched) {
    typedef ColVec<W> CV;
    auto probe = [&](int v) -> unsigned { return a.cur_nz[v >> 5]; };
    LW<NW> acc = lw_zero<NW>();
    long long e = beg & ~(long long)(W - 1);
    CV c = *reinterpret_cast<const CV*>(a.col + e);
    unsigned m = 0u; /* bit k: entry e + k belongs to the segment */
#pragma unroll
    for (int k = 0; k < W; ++k) m |= (unsigned)((e + k >= beg) && (e + k < end)) << k;
    unsigned bw[W];
#pragma unroll
    for (int k = 0; k < W; ++k) bw[k] = ((m >> k) & 1u) ? probe(c.v[k]) : 0u;
    for (;;) {
        const long long en = e + W;
        const bool more = en < end;
        CV cn;
#pragma unroll
        for (int k = 0; k < W; ++k) cn.v[k] = 0;
        if (more) cn = *reinterpret_cast<const CV*>(a.col + en);
        LW<NW> t = lw_zero<NW>();
        scanned += __popc(m);
#pragma unroll
        for (int k = 0; k < W; ++k) {
            const bool hit = ((m >> k) & 1u) && ((bw[k] >> (c.v[k] & 31)) & 1u);
            fetched += hit;
            if (hit) t = lw_or(t, lw_load<NW>(a.cur, c.v[k]));
        }
        unsigned mn = 0u, bwn[W];
#pragma unroll
        for (int k = 0; k < W; ++k) bwn[k] = 0u;
        if (more) {
#pragma unroll
            for (int k = 0; k < W; ++k) {
                const bool in = en + k < end;
                mn |= (unsigned)in << k;
                bwn[k] = in ? probe(cn.v[k]) : 0u;
            }
        }
        acc = lw_or(acc, t);
        if (!more || lw_covers(acc, need)) break;
        c = cn; m = mn; e = en;
#pragma unroll
        for (int k = 0; k < W; ++k) bw[k] = bwn[k];
    }
    return acc;
}

/* The block's result bits: nxt_nz is fully rewritten every bottom-up level. */
__device__ __forceinline__ void bu_store_block_bits(const TravArgs& a, int b, int lane, unsigned fullw, unsigned nzbits,
                                                    unsigned fullbits) {
    if (lane == 0) {
        a.nxt_nz[b] = nzbits;
        if (fullbits) a.vis_full[b] = fullw | fullbits;
    }
}

/* Rows of at most MSBFS_DEG_G8 entries: one thread per vertex, no cross-thread
 * traffic at all.  The 32 visited words of a block are one contiguous 1 KB read, and
 * each thread leaves its loop as soon as its own need is covered. */
template <int NW, int W>
__global__ void __launch_bounds__(CTA, (NW == 1) ? 6 : (NW == 8) ? MSBFS_NW8_BLOCKS : 4) bu_thread_kernel(TravArgs a, int blk_lo, int blk_hi) {
    const int lane = threadIdx.x & 31;
    unsigned scanned = 0, fetched = 0, examined = 0, written = 0;
    ull full_acc = 0, nf = 0, mf = 0;
    unsigned heavy = 0u;
    for (int b = blk_lo + global_warp(); b < blk_hi; b += total_warps()) {
        const unsigned fullw = a.vis_full[b];
        if (fullw == 0xffffffffu) {
            if (lane == 0) a.nxt_nz[b] = 0u;
            continue;
        }
        const int u = (b << 5) + lane;
        /* the 33 row offsets of the block staged in shared memory: one coalesced read instead of two
         * overlapping reads per thread (measured neutral to slightly better, the second read hit L1:
         * profiles/r2/sweep_f_*; kept as the tidier access pattern) */
        __shared__ long long s_row[CTA / 32][33];
        long long* tile = s_row[threadIdx.x >> 5];
        __syncwarp();
        tile[lane] = a.row_ptr[u];
        if (lane == 0) tile[32] = a.row_ptr[u + 32];
        __syncwarp();
        const long long beg = tile[lane], end = tile[lane + 1];
        bool has_new = false, full_now = false;
        if (!((fullw >> lane) & 1u) && end > beg) {
            const LW<NW> need = lw_andn(lw_from<NW>(a.ctrl->live), lw_load<NW>(a.visited, u));
            if (lw_any(need)) {
                ++examined;
                const LW<NW> acc = bu_scan_row<NW, W>(a, beg, end, need, scanned, fetched);
                const LW<NW> found = lw_and(acc, need);
                if (lw_any(found)) {
                    const LW<NW> nv = lw_or(lw_load<NW>(a.visited, u), found);
                    lw_store<NW>(a.visited, u, nv);
                    lw_store<NW>(a.nxt, u, found);
                    ++written;
                    has_new = true;
                    ++nf; mf += (ull)(end - beg);
                    if (b < a.blk_heavy) heavy = 1u;
                    if (lw_covers(nv, lw_from<NW>(a.ctrl->active))) { full_now = true; note_full(full_acc, end - beg); }
                }
            }
        }
        const unsigned nz = __ballot_sync(0xffffffffu, has_new);
        const unsigned fb = __ballot_sync(0xffffffffu, full_now);
        bu_store_block_bits(a, b, lane, fullw, nz, fb);
    }
    flush_counters(a.ctrl, SLOT_BU_G2, scanned, fetched, examined, written);
    flush_full(a.ctrl, full_acc, lane);
    flush_frontier_totals(a.ctrl, nf, mf, heavy, lane);
}

/* Rows longer than a.seg entries: one thread per segment of a.seg entries, work items
 * level-major (all first segments, then all second segments, ...), so consecutive
 * threads work on consecutive vertices (coalesced visited / hub_acc / row_ptr reads)
 * and the early segments of a row have usually run before its later ones start.
 * Segments of a row co-operate through hub_acc: a segment first subtracts what earlier
 * segments already found and skips the row when nothing is left to find. */
template <int NW, int W>
__global__ void __launch_bounds__(CTA, (NW == 1) ? 6 : (NW == 8) ? MSBFS_NW8_BLOCKS : 4) bu_segment_kernel(TravArgs a) {
    const int lane = threadIdx.x & 31;
    unsigned scanned = 0, fetched = 0, examined = 0, written = 0;
    for (long long c0 = 32ll * global_warp(); c0 < a.n_seg_items; c0 += 32ll * total_warps()) {
        const long long c = c0 + lane;
        if (c >= a.n_seg_items) continue;
        int k;
        {   /* largest k with base[k] <= c0, then walk forward to this thread's item */
            int lo = 0, hi = a.n_seg_levels;
            while (hi - lo > 1) {
                const int mid = (lo + hi) >> 1;
                if (a.seg_level_base[mid] <= c0) lo = mid; else hi = mid;
            }
            k = lo;
            while (c >= a.seg_level_base[k + 1]) ++k;
        }
        const int u = (int)(c - a.seg_level_base[k]);
        if (test_bit(a.vis_full, u)) continue;
        LW<NW> need = lw_andn(lw_from<NW>(a.ctrl->live), lw_load<NW>(a.visited, u));
        if (!lw_any(need)) continue;
        if (k == 0) ++examined;
        need = lw_andn(need, lw_load_volatile<NW>(a.hub_acc, u));
        if (!lw_any(need)) continue;
        const long long beg = a.row_ptr[u] + (long long)k * a.seg;
        const long long end = min(beg + (long long)a.seg, a.row_ptr[u + 1]);
        const LW<NW> acc = bu_scan_row<NW, W>(a, beg, end, need, scanned, fetched);
        const LW<NW> found = lw_and(acc, need);
        if (lw_any(found)) {
#pragma unroll
            for (int i = 0; i < NW; ++i)
                if (found.w[i]) atomicOr(&a.hub_acc[(long long)u * NW + i], found.w[i]);
            ++written;
        }
    }
    flush_counters(a.ctrl, SLOT_BU_HEAVY, scanned, fetched, examined, written);
}

/* hub_acc -> new frontier for the blocks [0, nblk_hub) handled by bu_segment_kernel */
template <int NW>
__global__ void __launch_bounds__(CTA) bu_heavy_finalize_kernel(TravArgs a, int nblk_hub) {
    const int lane = threadIdx.x & 31;
    const LW<NW> live = lw_from<NW>(a.ctrl->live);
    const LW<NW> active = lw_from<NW>(a.ctrl->active);
    ull full_acc = 0, nf = 0, mf = 0;
    unsigned heavy = 0u;
    for (int b = global_warp(); b < nblk_hub; b += total_warps()) {
        const int u = (b << 5) + lane;
        const unsigned fullw = a.vis_full[b];
        bool has_new = false, full_now = false;
        if (!((fullw >> lane) & 1u)) {
            const LW<NW> x = lw_load<NW>(a.hub_acc, u);
            if (lw_any(x)) {
                lw_store<NW>(a.hub_acc, u, lw_zero<NW>());
                const LW<NW> vis = lw_load<NW>(a.visited, u);
                const LW<NW> found = lw_andn(lw_and(x, live), vis);
                if (lw_any(found)) {
                    const LW<NW> nv = lw_or(vis, found);
                    lw_store<NW>(a.visited, u, nv);
                    lw_store<NW>(a.nxt, u, found);
                    has_new = true;
                    const long long deg = a.row_ptr[u + 1] - a.row_ptr[u];
                    ++nf; mf += (ull)deg;
                    if (b < a.blk_heavy) heavy = 1u;
                    if (lw_covers(nv, active)) { full_now = true; note_full(full_acc, deg); }
                }
            }
        }
        const unsigned nz = __ballot_sync(0xffffffffu, has_new);
        const unsigned fb = __ballot_sync(0xffffffffu, full_now);
        if (lane == 0) {
            a.nxt_nz[b] = nz;
            if (fb) a.vis_full[b] = fullw | fb;
        }
    }
    flush_full(a.ctrl, full_acc, lane);
    flush_frontier_totals(a.ctrl, nf, mf, heavy, lane);
}

/* ------------------------------------------- per-lane counts (popcount) --- */

/* 32 x 32 bit-matrix transpose across a warp: on return, bit i of lane j's word is
 * bit j of lane i's input. */
__device__ __forceinline__ unsigned warp_transpose32(unsigned x, int lane) {
#pragma unroll
    for (int s = 0; s < 5; ++s) {
        const int d = 16 >> s;
        const unsigned m = (s == 0) ? 0x0000ffffu : (s == 1) ? 0x00ff00ffu : (s == 2) ? 0x0f0f0f0fu
                         : (s == 3) ? 0x33333333u : 0x55555555u;
        const unsigned p = __shfl_xor_sync(0xffffffffu, x, d);
        x = (lane & d) ? ((x & ~m) | ((p >> d) & m)) : ((x & m) | ((p & m) << d));
    }
    return x;
}

/* Bit-sliced vertical counters: slice s holds bit s of a per-bit-position count.
 * Adding a word is a ripple of (AND, XOR) pairs; 5 slices take 31 additions. */
template <int NC> struct Csa {
    static constexpr int K = 5;
    unsigned sl[K][NC];
    unsigned cnt[NC]; /* after flush(): this lane's bit position, chunk c */
    __device__ __forceinline__ void init() {
#pragma unroll
        for (int s = 0; s < K; ++s)
#pragma unroll
            for (int c = 0; c < NC; ++c) sl[s][c] = 0u;
#pragma unroll
        for (int c = 0; c < NC; ++c) cnt[c] = 0u;
    }
    __device__ __forceinline__ void add(const ull* w) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            unsigned carry = (unsigned)(w[c >> 1] >> ((c & 1) * 32));
#pragma unroll
            for (int s = 0; s < K; ++s) {
                const unsigned t = sl[s][c] & carry;
                sl[s][c] ^= carry;
                carry = t;
            }
        }
    }
    /* two words at once: a full adder (3-input XOR + majority, one LOP3 each) folds both
     * into slice 0, then a single carry ripples up — half the work of two add() calls */
    __device__ __forceinline__ void add2(const ull* w0, const ull* w1) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const unsigned x = (unsigned)(w0[c >> 1] >> ((c & 1) * 32));
            const unsigned y = (unsigned)(w1[c >> 1] >> ((c & 1) * 32));
            const unsigned t0 = sl[0][c];
            sl[0][c] = t0 ^ x ^ y;
            unsigned carry = (t0 & x) | (t0 & y) | (x & y);
#pragma unroll
            for (int s = 1; s < K; ++s) {
                const unsigned t = sl[s][c] & carry;
                sl[s][c] ^= carry;
                carry = t;
            }
        }
    }
    /* four words at once: two full adders into slice 0, one into slice 1, then a single
     * carry ripples from slice 2 — 12 logic ops per 32 lanes x 4 words */
    __device__ __forceinline__ void add4(const ull* w0, const ull* w1, const ull* w2, const ull* w3) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
            const int sh = (c & 1) * 32;
            const unsigned x0 = (unsigned)(w0[c >> 1] >> sh), x1 = (unsigned)(w1[c >> 1] >> sh);
            const unsigned x2 = (unsigned)(w2[c >> 1] >> sh), x3 = (unsigned)(w3[c >> 1] >> sh);
            const unsigned t1 = x0 ^ x1 ^ x2;
            const unsigned c1 = (x0 & x1) | (x0 & x2) | (x1 & x2);
            const unsigned s0 = sl[0][c];
            sl[0][c] = s0 ^ t1 ^ x3;
            const unsigned c2 = (s0 & t1) | (s0 & x3) | (t1 & x3);
            const unsigned s1 = sl[1][c];
            sl[1][c] = s1 ^ c1 ^ c2;
            unsigned carry = (s1 & c1) | (s1 & c2) | (c1 & c2);
#pragma unroll
            for (int s = 2; s < K; ++s) {
                const unsigned t = sl[s][c] & carry;
                sl[s][c] ^= carry;
                carry = t;
            }
        }
    }
    /* warp-collective */
    __device__ __forceinline__ void flush(int lane) {
#pragma unroll
        for (int c = 0; c < NC; ++c) {
#pragma unroll
            for (int s = 0; s < K; ++s) {
                cnt[c] += (unsigned)__popc(warp_transpose32(sl[s][c], lane)) << s;
                sl[s][c] = 0u;
            }
        }
    }
};

/* mode 0: count the new frontier words (nxt, where nxt_nz) into ctrl->cnt, plus the
 *         frontier's vertex and entry totals for the direction heuristic;
 * mode 1: count visited words of vertices whose degree has bit `degbit` set into
 *         out[] scaled by 2^degbit (reached CSR entries per lane, for TEPS). */
/* A lane word wider than 256 bits is counted in 256-bit parts by different warps (a warp
 * keeps the bit-sliced counters of one part in registers), four vertices' words per pass. */
template <int NW> struct CountShape {
    static constexpr int PARTS = (NW > 4) ? NW / 4 : 1;  /* 256-bit parts per lane word */
    static constexpr int PW = NW / PARTS;                 /* 64-bit words per part        */
};
template <int PW> struct alignas(8 * PW) PartWord { ull w[PW]; };

/* MODE 0 has two ways to get the per-lane counts of a level, chosen on the device:
 *   - count the NEW frontier words (nxt where nxt_nz): cost ~ frontier vertices;
 *   - after a bottom-up level (`after_bu`), when fewer vertices remain not-full than entered the
 *     frontier, count instead the lanes each not-yet-full vertex still MISSES (~visited & active):
 *     newly visited (j) = unvisited before (j) - unvisited after (j), and the epilogue keeps the
 *     running `unvisited` per lane.  On the dense levels of a power-law graph (32 M new frontier
 *     words, 3 M vertices left) that is a tenth of the words.
 * After a bottom-up level the frontier's vertex / entry totals come from the kernels that wrote
 * it, so neither way needs them from here. */
template <int NW, int MODE>
__global__ void __launch_bounds__(CTA) count_kernel(TravArgs a, int nblk, int degbit, ull* __restrict__ out, int after_bu, int allow_missing) {
    constexpr int PARTS = CountShape<NW>::PARTS, PW = CountShape<NW>::PW, NC = 2 * PW;
    typedef PartWord<PW> Part;
    __shared__ unsigned s_cnt[64 * NW];
    for (int i = threadIdx.x; i < 64 * NW; i += CTA) s_cnt[i] = 0u;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    const int part = global_warp() % PARTS;          /* this warp's slice of every word */
    const int wslot = global_warp() / PARTS, nslots = total_warps() / PARTS;
    const bool missing = (MODE == 0) && after_bu && allow_missing && a.ctrl->n_u >= 0 && (ull)a.ctrl->n_u < a.ctrl->n_f; /* uniform: both final before this launch */
    const ull* words = (MODE == 0 && !missing) ? a.nxt : a.visited;
    Csa<NC> csa;
    csa.init();
    int pending = 0;
    ull nf = 0, mf = 0;
    unsigned heavy = 0u;
    Part zero, act;
#pragma unroll
    for (int i = 0; i < PW; ++i) { zero.w[i] = 0ull; act.w[i] = missing ? a.ctrl->active[part * PW + i] : 0ull; }
    if (wslot < nslots) {
        for (int b0 = 4 * wslot; b0 < nblk; b0 += 4 * nslots) {
            bool on[4];
            int v[4];
            unsigned any = 0u;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int b = b0 + q;
                v[q] = (b << 5) + lane;
                on[q] = false;
                if (b < nblk) {
                    if (MODE == 0 && missing) {
                        const unsigned w = ~a.vis_full[b];
                        on[q] = ((w >> lane) & 1u) && v[q] < a.n_active;
                        any |= w;
                    } else if (MODE == 0) {
                        const unsigned w = a.nxt_nz[b];
                        any |= w;
                        on[q] = (w >> lane) & 1u;
                        if (w && b < a.blk_heavy) heavy = 1u;
                    } else {
                        const ull d = (ull)(a.row_ptr[v[q] + 1] - a.row_ptr[v[q]]);
                        on[q] = (d >> degbit) & 1ull;
                        any |= on[q];
                    }
                }
            }
            if (MODE == 0 && any == 0u) continue; /* warp-uniform: the bitmap words are */
            Part x[4];
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                x[q] = on[q] ? *reinterpret_cast<const Part*>(words + (long long)v[q] * NW + part * PW) : zero;
                if (MODE == 0 && missing) {
#pragma unroll
                    for (int i = 0; i < PW; ++i) x[q].w[i] = on[q] ? (~x[q].w[i] & act.w[i]) : 0ull;
                }
            }
            csa.add4(x[0].w, x[1].w, x[2].w, x[3].w); /* a word that is off is zero and adds nothing */
            if (MODE == 0 && !after_bu && part == 0) {
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    if (on[q]) { ++nf; mf += (ull)(a.row_ptr[v[q] + 1] - a.row_ptr[v[q]]); }
            }
            pending += 4;
            if (pending >= 28) { csa.flush(lane); pending = 0; }
        }
    }
    csa.flush(lane);
#pragma unroll
    for (int c = 0; c < NC; ++c)
        if (csa.cnt[c]) atomicAdd(&s_cnt[part * 64 * PW + c * 32 + lane], csa.cnt[c]);
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * NW; i += CTA) {
        const unsigned cv = s_cnt[i];
        if (cv) {
            if (MODE == 0) atomicAdd(&a.ctrl->cnt[i], (ull)cv);
            else atomicAdd(&out[i], (ull)cv << degbit);
        }
    }
    if (MODE == 0) {
        if (missing && blockIdx.x == 0 && threadIdx.x == 0) a.ctrl->cnt_mode = 2;
        if (!after_bu) {
            nf = warp_sum(nf);
            mf = warp_sum(mf);
            if (lane == 0 && nf) { atomicAdd(&a.ctrl->n_f, nf); atomicAdd(&a.ctrl->m_f, mf); }
            if (__any_sync(0xffffffffu, heavy != 0u) && lane == 0) a.ctrl->heavy_frontier = 1u;
        }
    }
}

/* ------------------------------------------------------ level bookkeeping --- */

__global__ void init_batch_kernel(LevelCtrl* c, int lanes, long long nnz, long long n_active) {
    const int j = threadIdx.x; /* MSBFS_MAX_LANES threads */
    c->cnt[j] = 0ull;
    c->sums[j] = 0ull;
    c->unvisited[j] = (j < lanes) ? (ull)n_active : 0ull;
    if (j < MSBFS_MAX_NW) {
        c->live[j] = 0ull;
        ull m = 0ull;
        const int lo = j * 64;
        if (lanes >= lo + 64) m = ~0ull;
        else if (lanes > lo) m = (1ull << (lanes - lo)) - 1ull;
        c->active[j] = m;
    }
    if (j == 0) {
        c->n_f = 0; c->m_f = 0; c->m_u = nnz; c->n_u = n_active; c->heavy_frontier = 0u; c->cnt_mode = 0;
        c->scanned = c->words_fetched = c->examined = c->words_written = 0;
        c->level = -1;
    }
}

/* F(j) += (level + 1) * count(j)  (the reference's sum of distances, main.cu:81-87,
 * taken level by level), live lanes, and the direction of the next level.
 * `direction` is what the level just finished used (0 = injection). */
__device__ __forceinline__ void epilogue_body(LevelCtrl* c, LevelLog* log, int log_cap, LevelVerdict* verdict,
                                              int direction, long long alpha, int force_direction) {
    /* any block size that is a multiple of 32 */
    const int cnt_mode = c->cnt_mode; /* read by everyone before thread 0 resets it below */
    for (int j = threadIdx.x; j < MSBFS_MAX_LANES; j += blockDim.x) {
        ull cnt = c->cnt[j];
        if (cnt_mode == 2) { /* cnt[] holds what lane j still misses: newly visited = the difference */
            const ull after = cnt;
            cnt = c->unvisited[j] - after;
            c->unvisited[j] = after;
        } else {
            c->unvisited[j] -= cnt;
        }
        c->sums[j] += (ull)(c->level + 1) * cnt;
        c->cnt[j] = 0ull;
        const unsigned bal = __ballot_sync(0xffffffffu, cnt != 0ull);
        if ((j & 31) == 0) reinterpret_cast<unsigned*>(c->live)[j >> 5] = bal;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        const ull nf = c->n_f, mf = c->m_f;
        const long long mu = c->m_u;
        /* Top-down costs about alpha x more per CSR entry than bottom-up (random visited read
         * + atomic OR vs a streamed col entry and a bitmap probe); both pay per vertex they
         * touch: the frontier's vertices top-down, every not-yet-full vertex bottom-up.
         * Fitted on Kronecker-26, ER-24 and the 4096^2 lattice. */
        const long long nu = c->n_u;
        const long long td_cost = alpha * (long long)mf + (alpha * 4 / 5) * (long long)nf;
        const long long bu_cost = mu + alpha * nu;
        int next = (force_direction == 1 || force_direction == 2) ? force_direction : ((td_cost < bu_cost) ? 1 : 2);
        if (c->log_pos < log_cap) {
            LevelLog& l = log[c->log_pos];
            l.level = c->level; l.direction = direction;
            l.frontier_vertices = nf; l.frontier_entries = mf; l.unexplored_entries = mu;
            l.scanned = c->scanned; l.words_fetched = c->words_fetched;
            l.vertices_examined = c->examined; l.words_written = c->words_written;
        }
        c->log_pos += 1;
        verdict->done = (nf == 0ull);
        verdict->next_direction = next;
        verdict->n_f = nf; verdict->m_f = mf; verdict->m_u = mu; verdict->n_u = nu;
        verdict->heavy_frontier = (int)c->heavy_frontier;
        c->last_done = (nf == 0ull);
        c->last_next = next;
        c->heavy_frontier = 0u;
        c->cnt_mode = 0;
        c->n_f = 0; c->m_f = 0;
        c->scanned = c->words_fetched = c->examined = c->words_written = 0;
        c->level += 1;
        __threadfence_system();
    }
}

__global__ void epilogue_kernel(LevelCtrl* c, LevelLog* log, int log_cap, LevelVerdict* verdict,
                                int direction, long long alpha, int force_direction) {
    epilogue_body(c, log, log_cap, verdict, direction, alpha, force_direction);
}

/* -------------------------------------------- top-down streak (persistent) --- */

/* High-diameter, low-degree graphs (road-like lattices) run thousands of levels whose
 * frontiers are small: the cost is then launches and host round trips, not bytes.  This
 * cooperative kernel keeps one grid resident and runs consecutive top-down levels —
 * expand, finalize + per-lane count, epilogue — separated by grid-wide barriers, until
 * the BFS is done or the direction heuristic asks for a bottom-up level.  Used only when
 * every row is short (max degree <= option "streak_max_deg", default 256), so a frontier vertex is one thread.
 * On return the host finds ctrl / verdict exactly as after a normal epilogue, plus the
 * number of levels run (the frontier buffers have swapped that many times). */

template <int NW>
__global__ void __launch_bounds__(CTA, (NW == 8) ? 2 : 4) td_streak_kernel(TravArgs a, LevelLog* log, int log_cap, LevelVerdict* verdict,
                                                           long long alpha, int force_direction, int max_levels) {
    cg::grid_group grid = cg::this_grid();
    __shared__ unsigned s_cnt[64 * NW];
    __shared__ ull s_ctr[6]; /* scanned, fetched, examined, written, nf, mf */
    __shared__ ull s_mu; /* packed newly-full totals, see note_full */
    __shared__ unsigned s_flags; /* bit 0: this block saw new frontier words */
    const int lane = threadIdx.x & 31;
    const int nblk = a.n_pad >> 5;
    const int ntiles = (nblk + 31) >> 5;
    int levels_run = 0;

    for (;;) {
        for (int i = threadIdx.x; i < 64 * NW; i += CTA) s_cnt[i] = 0u;
        if (threadIdx.x < 6) s_ctr[threadIdx.x] = 0ull;
        if (threadIdx.x == 0) { s_mu = 0; s_flags = 0u; }
        __syncthreads();

        /* ---- expand: one thread per frontier vertex; consumed bitmap words are cleared so
         *      that this buffer is a clean nxt_nz two levels from now */
        {
            ull scanned = 0, fetched = 0, examined = 0, written = 0;
            for (int t = global_warp(); t < ntiles; t += total_warps()) {
                const int b = (t << 5) + lane;
                unsigned word = 0u;
                if (b < nblk) {
                    word = a.cur_nz[b];
                    if (word) a.cur_nz[b] = 0u;
                }
                TileBits tb;
                tb.load(word, lane);
                for (int r0 = 0; r0 < tb.total; r0 += 32) {
                    const int off = tb.select(r0 + lane);
                    if (r0 + lane < tb.total) {
                        const int v = (t << 10) + off;
                        const LW<NW> f = lw_load<NW>(a.cur, v);
                        ++examined;
                        td_push<NW, 4>(a, f, a.row_ptr[v], a.row_ptr[v + 1], 1, scanned, fetched, written);
                    }
                }
            }
            scanned = warp_sum(scanned); fetched = warp_sum(fetched);
            examined = warp_sum(examined); written = warp_sum(written);
            if (lane == 0 && (scanned | examined)) {
                atomicAdd(&s_ctr[0], scanned); atomicAdd(&s_ctr[1], fetched);
                atomicAdd(&s_ctr[2], examined); atomicAdd(&s_ctr[3], written);
            }
        }
        grid.sync();

        /* ---- finalize + count.  Two rounds (64 touched vertices) per pass so that four
         *      independent word loads per lane are in flight before the first store; lanes are
         *      counted bit by bit in shared memory (words of a road-like frontier carry few
         *      lanes), which keeps the kernel at 64 registers. */
        {
            ull nf = 0, mf = 0;
            ull full_acc = 0;
            bool warp_any = false;
            const LW<NW> active = lw_from<NW>(a.ctrl->active);
            for (int t = global_warp(); t < ntiles; t += total_warps()) {
                const int b = (t << 5) + lane;
                TileBits tb;
                tb.load(b < nblk ? a.nxt_nz[b] : 0u, lane);
                if (tb.total == 0) continue;
                warp_any = true;
                for (int r0 = 0; r0 < tb.total; r0 += 64) {
                    int v[2];
                    bool ok[2];
                    LW<NW> x[2], vis[2];
                    long long deg[2];
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        const int rank = r0 + 32 * q + lane;
                        const int off = tb.select(rank);
                        ok[q] = rank < tb.total;
                        v[q] = (t << 10) + off;
                    }
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        x[q] = ok[q] ? lw_load<NW>(a.acc, v[q]) : lw_zero<NW>();
                        vis[q] = ok[q] ? lw_load<NW>(a.visited, v[q]) : lw_zero<NW>();
                        deg[q] = ok[q] ? (a.row_ptr[v[q] + 1] - a.row_ptr[v[q]]) : 0;
                    }
#pragma unroll
                    for (int q = 0; q < 2; ++q) {
                        if (!ok[q]) continue;
                        lw_store<NW>(a.acc, v[q], lw_zero<NW>());
                        const LW<NW> nv = lw_or(vis[q], x[q]);
                        lw_store<NW>(a.visited, v[q], nv);
                        lw_store<NW>(a.nxt, v[q], x[q]);
#pragma unroll
                        for (int i = 0; i < NW; ++i) {
                            ull bits = x[q].w[i];
                            while (bits) {
                                const int j = __ffsll((long long)bits) - 1;
                                bits &= bits - 1ull;
                                atomicAdd(&s_cnt[i * 64 + j], 1u);
                            }
                        }
                        ++nf;
                        mf += (ull)deg[q];
                        if (lw_covers(nv, active)) {
                            note_full(full_acc, deg[q]);
                            atomicOr(&a.vis_full[v[q] >> 5], 1u << (v[q] & 31));
                        }
                    }
                }
            }
            if (warp_any) {
                nf = warp_sum(nf); mf = warp_sum(mf);
                full_acc = warp_sum(full_acc);
                if (lane == 0) {
                    atomicAdd(&s_ctr[4], nf); atomicAdd(&s_ctr[5], mf);
                    if (full_acc) atomicAdd(&s_mu, full_acc);
                    atomicOr(&s_flags, 1u);
                }
            }
        }
        __syncthreads();
        if (s_flags & 1u)
            for (int i = threadIdx.x; i < 64 * NW; i += CTA)
                if (s_cnt[i]) atomicAdd(&a.ctrl->cnt[i], (ull)s_cnt[i]);
        if (threadIdx.x == 0) {
            if (s_ctr[0]) { atomicAdd(&a.ctrl->scanned, s_ctr[0]); atomicAdd(&a.ctrl->k_scanned[SLOT_TD_G2], s_ctr[0]); }
            if (s_ctr[1]) { atomicAdd(&a.ctrl->words_fetched, s_ctr[1]); atomicAdd(&a.ctrl->k_fetched[SLOT_TD_G2], s_ctr[1]); }
            if (s_ctr[2]) { atomicAdd(&a.ctrl->examined, s_ctr[2]); atomicAdd(&a.ctrl->k_examined[SLOT_TD_G2], s_ctr[2]); }
            if (s_ctr[3]) { atomicAdd(&a.ctrl->words_written, s_ctr[3]); atomicAdd(&a.ctrl->k_written[SLOT_TD_G2], s_ctr[3]); }
            if (s_ctr[4]) { atomicAdd(&a.ctrl->n_f, s_ctr[4]); atomicAdd(&a.ctrl->m_f, s_ctr[5]); }
            commit_full(a.ctrl, s_mu);
        }
        __threadfence();
        grid.sync();

        /* ---- epilogue by block 0, verdict read by everyone */
        if (blockIdx.x == 0) epilogue_body(a.ctrl, log, log_cap, verdict, 1, alpha, force_direction);
        __threadfence();
        grid.sync();
        ++levels_run;
        const volatile LevelCtrl* vc = a.ctrl;
        const bool stop = vc->last_done || vc->last_next != 1 || levels_run >= max_levels;
        /* the frontier buffers trade places */
        ull* t = a.cur; a.cur = a.nxt; a.nxt = t;
        unsigned* tb = a.cur_nz; a.cur_nz = a.nxt_nz; a.nxt_nz = tb;
        if (stop) break;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        verdict->streak_levels = levels_run;
        __threadfence_system();
    }
}

/* ------------------------------------------------ the last levels (tail) --- */

/* After the dense levels a batch is left with a few ten thousand vertices that some lane still
 * misses (those outside the giant component, mostly) and three or four more levels to run, each of
 * which would sweep every bitmap and cost launches, a count pass and a host round trip for a
 * handful of frontier vertices.  Instead the unfinished vertices are listed once and ONE
 * cooperative kernel runs the remaining levels bottom-up over that list: scan, per-lane counts in
 * shared memory, epilogue by block 0, two grid barriers per level. */
constexpr int TAIL_CAP = 1 << 18;           /* most vertices the list may hold */
constexpr long long TAIL_MAX_ENTRIES = 1ll << 23; /* most CSR entries they may have between them */
constexpr int TAIL_MAX_ROW = 4096;          /* longest row a single thread is asked to scan */

/* list[0] = count, list[1] = longest row among them, list[4 ...] = the vertices */
__global__ void __launch_bounds__(CTA) tail_list_kernel(TravArgs a, int nblk, int* __restrict__ list) {
    const int lane = threadIdx.x & 31;
    for (int b = global_warp(); b < nblk; b += total_warps()) {
        const unsigned w = ~a.vis_full[b];
        const int v = (b << 5) + lane;
        const bool on = ((w >> lane) & 1u) && v < a.n_active;
        const unsigned m = __ballot_sync(0xffffffffu, on);
        if (m == 0u) continue;
        int base = 0;
        if (lane == 0) base = atomicAdd(&list[0], __popc(m));
        base = __shfl_sync(0xffffffffu, base, 0);
        if (on) {
            const int pos = base + __popc(m & ((1u << lane) - 1u));
            if (pos < TAIL_CAP) list[4 + pos] = v;
            const long long deg = a.row_ptr[v + 1] - a.row_ptr[v];
            atomicMax(&list[1], (int)min(deg, (long long)0x7fffffff));
        }
    }
}

template <int NW>
__global__ void __launch_bounds__(CTA, (NW == 8) ? 2 : 4) tail_levels_kernel(TravArgs a, const int* __restrict__ list, int nlist, LevelLog* log,
                                                                             int log_cap, LevelVerdict* verdict, long long alpha, int max_levels) {
    cg::grid_group grid = cg::this_grid();
    __shared__ unsigned s_cnt[64 * NW];
    const int lane = threadIdx.x & 31;
    const int nblk = a.n_pad >> 5;
    const int gtid = blockIdx.x * blockDim.x + threadIdx.x, gsize = gridDim.x * blockDim.x;
    const LW<NW> active = lw_from<NW>(a.ctrl->active);
    int levels_run = 0;
    for (;;) {
        for (int i = threadIdx.x; i < 64 * NW; i += CTA) s_cnt[i] = 0u;
        __syncthreads();
        const LW<NW> live = lw_load_volatile<NW>(a.ctrl->live, 0); /* rewritten by the epilogue every level */
        unsigned scanned = 0, fetched = 0, examined = 0, written = 0;
        ull full_acc = 0, nf = 0, mf = 0;
        for (int i = gtid; i < nlist; i += gsize) {
            const int u = list[i];
            if (test_bit(a.vis_full, u)) continue;
            const LW<NW> vis = lw_load<NW>(a.visited, u);
            const LW<NW> need = lw_andn(live, vis);
            if (!lw_any(need)) continue;
            ++examined;
            const long long beg = a.row_ptr[u], end = a.row_ptr[u + 1];
            const LW<NW> acc = bu_scan_row<NW, 4>(a, beg, end, need, scanned, fetched);
            const LW<NW> found = lw_and(acc, need);
            if (!lw_any(found)) continue;
            const LW<NW> nv = lw_or(vis, found);
            lw_store<NW>(a.visited, u, nv);
            lw_store<NW>(a.nxt, u, found);
            atomicOr(&a.nxt_nz[u >> 5], 1u << (u & 31));
            ++written; ++nf; mf += (ull)(end - beg);
#pragma unroll
            for (int w = 0; w < NW; ++w) {
                ull bits = found.w[w];
                while (bits) {
                    const int j = __ffsll((long long)bits) - 1;
                    bits &= bits - 1ull;
                    atomicAdd(&s_cnt[w * 64 + j], 1u);
                }
            }
            if (lw_covers(nv, active)) {
                note_full(full_acc, end - beg);
                atomicOr(&a.vis_full[u >> 5], 1u << (u & 31));
            }
        }
        flush_counters(a.ctrl, SLOT_BU_G2, scanned, fetched, examined, written);
        flush_full(a.ctrl, full_acc, lane);
        flush_frontier_totals(a.ctrl, nf, mf, 0u, lane);
        __syncthreads();
        for (int i = threadIdx.x; i < 64 * NW; i += CTA)
            if (s_cnt[i]) atomicAdd(&a.ctrl->cnt[i], (ull)s_cnt[i]);
        __threadfence();
        grid.sync();
        /* the frontier just read becomes the next level's output bitmap: wipe it (everyone is done probing it) */
        for (int b = gtid; b < nblk; b += gsize) a.cur_nz[b] = 0u;
        if (blockIdx.x == 0) epilogue_body(a.ctrl, log, log_cap, verdict, 2, alpha, 2);
        __threadfence();
        grid.sync();
        ++levels_run;
        const volatile LevelCtrl* vc = a.ctrl;
        const bool stop = vc->last_done || levels_run >= max_levels;
        ull* t = a.cur; a.cur = a.nxt; a.nxt = t;
        unsigned* tb = a.cur_nz; a.cur_nz = a.nxt_nz; a.nxt_nz = tb;
        if (stop) break;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        verdict->streak_levels = levels_run;
        __threadfence_system();
    }
}

__global__ void store_sums_kernel(const LevelCtrl* c, long long* __restrict__ sums, int group_base, int lanes) {
    const int j = threadIdx.x;
    if (j < lanes) sums[group_base + j] = (long long)c->sums[j];
}

/* ------------------------------------------------------------- host side --- */

constexpr size_t MAX_EVENT_PAIRS = 1u << 16;

struct Launcher {
    msbfs_handle* h;
    TravArgs a;
    long long launches = 0;
    int grid_cap;
    std::vector<int> pair_slot; /* slot of event pair i (events 2i, 2i+1 of h->ev_pool) */
    std::vector<int> pair_batch, pair_level;
    int cur_batch = 0, cur_level = -1;
    std::vector<cudaEvent_t> batch_ev; /* start / stop of every batch */

    /* bracket one launch with CUDA events on the launching stream (option "profile") */
    template <class F> void timed(int slot, F&& launch) {
        ++launches;
        ++h->slot_profile[slot].launches;
        const bool rec = h->opt_profile && pair_slot.size() < MAX_EVENT_PAIRS;
        size_t i = pair_slot.size();
        if (rec) {
            while (h->ev_pool.size() < 2 * (i + 1)) {
                cudaEvent_t e;
                if (cudaEventCreate(&e) != cudaSuccess) { launch(); return; }
                h->ev_pool.push_back(e);
            }
            cudaEventRecord(h->ev_pool[2 * i], h->stream);
        }
        launch();
        if (rec) {
            cudaEventRecord(h->ev_pool[2 * i + 1], h->stream);
            pair_slot.push_back(slot);
            pair_batch.push_back(cur_batch);
            pair_level.push_back(cur_level);
        }
    }
    void mark_batch() { /* called at the start and at the end of every batch */
        cudaEvent_t e;
        if (cudaEventCreate(&e) != cudaSuccess) return;
        cudaEventRecord(e, h->stream);
        batch_ev.push_back(e);
    }
    void collect() {
        h->launch_log.clear();
        for (size_t i = 0; i < pair_slot.size(); ++i) {
            float ms = 0.f;
            if (cudaEventElapsedTime(&ms, h->ev_pool[2 * i], h->ev_pool[2 * i + 1]) == cudaSuccess)
                h->slot_profile[pair_slot[i]].ms += ms;
            msbfs_launch_record r;
            r.batch = pair_batch[i]; r.level = pair_level[i]; r.slot = pair_slot[i]; r.ms = ms;
            h->launch_log.push_back(r);
        }
        pair_slot.clear(); pair_batch.clear(); pair_level.clear();
        h->batch_ms.clear();
        for (size_t i = 0; i + 1 < batch_ev.size(); i += 2) {
            float ms = 0.f;
            cudaEventElapsedTime(&ms, batch_ev[i], batch_ev[i + 1]);
            h->batch_ms.push_back((double)ms);
        }
        for (cudaEvent_t e : batch_ev) cudaEventDestroy(e);
        batch_ev.clear();
    }

    int grid_for_warps(long long warps) const {
        long long blocks = (warps + (CTA / 32) - 1) / (CTA / 32);
        return (int)std::max<long long>(1, std::min<long long>(blocks, grid_cap));
    }
};

template <int NW>
int run_batches(msbfs_handle* h, bool want_reached, std::vector<ull>* reached_out) {
    cudaStream_t s = h->stream;
    const int lanes_per_batch = 64 * NW;
    const int K = h->q_groups;
    const int nblk = h->n_pad / 32;
    const size_t word_bytes = (size_t)h->n_pad * NW * sizeof(ull);

    Launcher L;
    L.h = h;
    L.grid_cap = h->sm_count * 8;
    TravArgs& a = L.a;
    a.row_ptr = h->d_row_ptr; a.col = h->d_col;
    a.visited = h->d_visited; a.cur = h->d_cur; a.nxt = h->d_nxt; a.acc = h->d_acc; a.hub_acc = h->d_hub_acc;
    a.cur_nz = h->d_bitmaps; a.nxt_nz = h->d_bitmaps + nblk; a.vis_full = h->d_bitmaps + 2 * (size_t)nblk;
    a.ctrl = h->d_ctrl;
    a.chunk_level_base = h->d_chunk_level_base; a.n_chunk_levels = h->n_chunk_levels; a.n_chunks = h->n_chunks;
    a.n_pad = h->n_pad; a.n_active = h->n_active;
    a.blk_heavy = h->blk_heavy; a.blk_warp = h->blk_warp; a.blk_g8 = h->blk_g8; a.blk_end = h->blk_end;
    a.seg_level_base = h->d_seg_level_base; a.n_seg_levels = h->n_seg_levels; a.n_seg_items = h->n_seg_items;
    a.seg = h->seg; a.blk_multi = h->blk_multi;

    /* col entries per load: the long-row segments gain from 256-bit loads at 512 lanes (kron26
     * bu_segment 26.2 -> 23.9 ms per step), the short rows of the thread kernel lose (85.2 -> 89.8:
     * more masked work per row); "col_width" = 4 or 8 forces both */
    const bool wide_seg = h->opt_col_width == 8 || (h->opt_col_width == 0 && NW == 8);
    const bool wide = h->opt_col_width == 8;
    void (*k_thread)(TravArgs, int, int) = wide ? bu_thread_kernel<NW, 8> : bu_thread_kernel<NW, 4>;
    void (*k_segment)(TravArgs) = wide_seg ? bu_segment_kernel<NW, 8> : bu_segment_kernel<NW, 4>;

    /* persistent top-down kernel: only when every row is short and the device can co-schedule the grid */
    int streak_grid = 0;
    if (h->opt_streak && h->max_degree <= (unsigned)h->opt_streak_max_deg) {
        int coop = 0, per_sm = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device);
        if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, td_streak_kernel<NW>, CTA, 0) == cudaSuccess && per_sm > 0)
            streak_grid = per_sm * h->sm_count;
    }

    /* road-like graph: the sparse per-group engine (lane_major.cu) takes the batches it can hold */
    bool lm_on = (h->opt_lane_major == 1 || (h->opt_lane_major < 0 && h->max_degree <= 16u)) && h->opt_direction != 2 &&
                 h->max_degree <= (unsigned)h->opt_streak_max_deg;
    if (lm_on) {
        int rc = msbfs_lm_ensure(h, lanes_per_batch);
        if (rc) return rc;
        lm_on = h->lm_grid > 0;
    }
    h->lm_expanded = h->lm_scanned = h->lm_fresh = 0;

    /* the tail kernel: cooperative, one block per SM is plenty for a list of at most 2^18 vertices */
    int tail_grid = 0;
    /* (worth it from about a million vertices on: below that a level's sweeps and launches cost less than
     * listing the vertices — RMAT-16 0.81 -> 0.93 ms with it, Kronecker-26 133.4 -> 129.3) */
    if (h->opt_tail && h->opt_direction != 1 && (h->opt_tail > 1 || h->n_pad >= (1 << 20))) {
        int coop = 0, per_sm = 0;
        cudaDeviceGetAttribute(&coop, cudaDevAttrCooperativeLaunch, h->device);
        if (coop && cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, tail_levels_kernel<NW>, CTA, 0) == cudaSuccess && per_sm > 0)
            tail_grid = h->sm_count;
        else cudaGetLastError();
        if (tail_grid > 0 && !h->d_tail_list) MSBFS_CUDA(cudaMalloc(&h->d_tail_list, sizeof(int) * (size_t)(TAIL_CAP + 4)));
    }

    ull* d_reached = nullptr;
    if (want_reached) {
        MSBFS_CUDA(cudaMalloc(&d_reached, sizeof(ull) * MSBFS_MAX_LANES));
        reached_out->assign((size_t)K, 0ull);
    }

    MSBFS_CUDA(cudaMemsetAsync(h->d_ctrl, 0, sizeof(LevelCtrl), s)); /* log_pos and the per-kernel accumulators */
    int batches = 0;
    int tail_n = 0;
    long long levels = 0, levels_td = 0, levels_bu = 0;

    for (int group_base = 0; group_base < K; group_base += lanes_per_batch, ++batches) {
        const int lanes = std::min(lanes_per_batch, K - group_base);
        L.cur_batch = batches; L.cur_level = -1;
        MsbfsNvtxRange nvtx_batch("msbfs_batch");
        L.mark_batch();
        if (lm_on) {
            msbfs_lm_result r;
            r.want_reach = want_reached ? 1 : 0;
            MSBFS_CUDA(cudaMemcpyAsync(&r.log_pos_in, &h->d_ctrl->log_pos, sizeof(int), cudaMemcpyDeviceToHost, s));
            MSBFS_CUDA(cudaStreamSynchronize(s));
            int rc = MSBFS_OK;
            L.timed(SLOT_LANE_MAJOR, [&] { rc = msbfs_lm_run_batch(h, group_base, lanes, &r); });
            if (rc) return rc;
            if (!r.overflow) {
                L.launches += 2; /* inject + store beside the cooperative kernel */
                MSBFS_CUDA(cudaMemcpyAsync(&h->d_ctrl->log_pos, &r.log_pos_out, sizeof(int), cudaMemcpyHostToDevice, s));
                MSBFS_CUDA(cudaStreamSynchronize(s));
                levels += r.levels; levels_td += r.levels;
                h->lm_expanded += r.expanded; h->lm_scanned += r.scanned; h->lm_fresh += r.fresh;
                if (want_reached)
                    for (int j = 0; j < lanes; ++j) (*reached_out)[(size_t)group_base + j] = r.reach[j];
                L.mark_batch();
                continue;
            }
            /* a level outgrew the queue: this is not a graph for the sparse engine; the batch starts over below */
            lm_on = false;
        }
        L.timed(SLOT_OTHER, [&] { init_batch_kernel<<<1, MSBFS_MAX_LANES, 0, s>>>(h->d_ctrl, lanes, (long long)h->nnz, (long long)h->n_active); });
        MSBFS_CUDA(cudaMemsetAsync(h->d_visited, 0, word_bytes, s));
        MSBFS_CUDA(cudaMemsetAsync(h->d_bitmaps, 0, sizeof(unsigned) * 3 * (size_t)nblk, s));
        a.cur_nz = h->d_bitmaps; a.nxt_nz = h->d_bitmaps + nblk;
        a.cur = h->d_cur; a.nxt = h->d_nxt;
        L.timed(SLOT_OTHER, [&] {
            inject_kernel<NW><<<L.grid_for_warps(lanes), CTA, 0, s>>>(a, h->d_perm, h->n, h->n_active, h->d_q_offsets,
                                                                      h->d_q_sources, group_base, lanes);
        });
        L.timed(SLOT_TD_FINALIZE, [&] { td_finalize_kernel<NW><<<L.grid_for_warps(nblk), CTA, 0, s>>>(a, nblk); });
        L.timed(SLOT_COUNT, [&] { count_kernel<NW, 0><<<L.grid_for_warps(nblk), CTA, 0, s>>>(a, nblk, 0, nullptr, 0, 0); });
        L.timed(SLOT_OTHER, [&] {
            epilogue_kernel<<<1, MSBFS_MAX_LANES, 0, s>>>(h->d_ctrl, h->d_log, h->log_cap, h->h_verdict, 0, h->opt_alpha,
                                                           h->opt_direction);
        });
        MSBFS_CUDA(cudaGetLastError());

        for (;;) {
            std::swap(a.cur, a.nxt);
            std::swap(a.cur_nz, a.nxt_nz);
            MSBFS_CUDA(cudaStreamSynchronize(s));
            if (h->h_verdict->done) break;
            const int dir = h->h_verdict->next_direction;
            const bool heavy_frontier = h->h_verdict->heavy_frontier != 0;
            ++levels;
            ++L.cur_level;
            if (dir == 1 && streak_grid > 0) {
                /* short-row graph: stay on the device for as many top-down levels as it takes */
                MSBFS_CUDA(cudaMemsetAsync(a.nxt_nz, 0, sizeof(unsigned) * (size_t)nblk, s));
                long long alpha = h->opt_alpha;
                int force = h->opt_direction, max_levels = h->opt_streak_levels;
                void* args[] = {&a, &h->d_log, &h->log_cap, &h->h_verdict, &alpha, &force, &max_levels};
                cudaError_t ce = cudaSuccess;
                L.timed(SLOT_TD_G2, [&] {
                    ce = cudaLaunchCooperativeKernel((const void*)td_streak_kernel<NW>, dim3(streak_grid), dim3(CTA), args, 0, s);
                });
                if (ce != cudaSuccess) return msbfs_fail(MSBFS_ERR_CUDA, "cooperative launch failed: %s", cudaGetErrorString(ce));
                MSBFS_CUDA(cudaStreamSynchronize(s));
                const int run = h->h_verdict->streak_levels;
                levels += run - 1;
                L.cur_level += run - 1;
                levels_td += run;
                /* the kernel swapped the frontier buffers `run` times; the loop head swaps once more */
                if (run % 2 == 0) { std::swap(a.cur, a.nxt); std::swap(a.cur_nz, a.nxt_nz); }
                continue;
            }
            if (dir == 1) {
                ++levels_td;
                MSBFS_CUDA(cudaMemsetAsync(a.nxt_nz, 0, sizeof(unsigned) * (size_t)nblk, s));
                if (a.n_chunks > 0 && heavy_frontier)
                    L.timed(SLOT_TD_HEAVY, [&] { td_heavy_kernel<NW, 4><<<L.grid_for_warps(a.n_chunks), CTA, 0, s>>>(a); });
                if (a.blk_warp > a.blk_heavy)
                    L.timed(SLOT_TD_G32, [&] {
                        td_expand_kernel<NW, 32, 4><<<L.grid_for_warps(a.blk_warp - a.blk_heavy), CTA, 0, s>>>(a, a.blk_heavy, a.blk_warp);
                    });
                if (a.blk_g8 > a.blk_warp)
                    L.timed(SLOT_TD_G8, [&] {
                        td_expand_kernel<NW, 8, 4><<<L.grid_for_warps(a.blk_g8 - a.blk_warp), CTA, 0, s>>>(a, a.blk_warp, a.blk_g8);
                    });
                if (a.blk_end > a.blk_g8)
                    L.timed(SLOT_TD_G2, [&] {
                        td_expand_kernel<NW, 2, 4><<<L.grid_for_warps(a.blk_end - a.blk_g8), CTA, 0, s>>>(a, a.blk_g8, a.blk_end);
                    });
                L.timed(SLOT_TD_FINALIZE, [&] { td_finalize_kernel<NW><<<L.grid_for_warps(nblk), CTA, 0, s>>>(a, nblk); });
            } else if (tail_grid > 0 && h->h_verdict->n_u > 0 && h->h_verdict->n_u <= TAIL_CAP && h->h_verdict->m_u <= TAIL_MAX_ENTRIES &&
                       [&] { /* list the unfinished vertices; usable when they fit and none of them has a long row */
                           int head[4] = {0, 0, 0, 0};
                           if (cudaMemsetAsync(h->d_tail_list, 0, sizeof head, s) != cudaSuccess) return false;
                           L.timed(SLOT_OTHER, [&] { tail_list_kernel<<<L.grid_for_warps(nblk), CTA, 0, s>>>(a, nblk, h->d_tail_list); });
                           if (cudaMemcpyAsync(head, h->d_tail_list, sizeof head, cudaMemcpyDeviceToHost, s) != cudaSuccess) return false;
                           if (cudaStreamSynchronize(s) != cudaSuccess) return false;
                           tail_n = head[0];
                           return head[0] > 0 && head[0] <= TAIL_CAP && head[1] <= TAIL_MAX_ROW;
                       }()) {
                /* every remaining level in one launch */
                MSBFS_CUDA(cudaMemsetAsync(a.nxt_nz, 0, sizeof(unsigned) * (size_t)nblk, s));
                long long alpha = h->opt_alpha;
                int max_levels = 1 << 20;
                const int* list = h->d_tail_list + 4;
                void* args[] = {&a, &list, &tail_n, &h->d_log, &h->log_cap, &h->h_verdict, &alpha, &max_levels};
                cudaError_t ce = cudaSuccess;
                L.timed(SLOT_BU_G2, [&] {
                    ce = cudaLaunchCooperativeKernel((const void*)tail_levels_kernel<NW>, dim3(tail_grid), dim3(CTA), args, 0, s);
                });
                if (ce != cudaSuccess) return msbfs_fail(MSBFS_ERR_CUDA, "cooperative launch failed: %s", cudaGetErrorString(ce));
                MSBFS_CUDA(cudaStreamSynchronize(s));
                const int run = h->h_verdict->streak_levels;
                levels += run - 1;
                L.cur_level += run - 1;
                levels_bu += run;
                if (run % 2 == 0) { std::swap(a.cur, a.nxt); std::swap(a.cur_nz, a.nxt_nz); }
                continue;
            } else {
                ++levels_bu;
                if (a.blk_multi > 0) {
                    if (a.n_seg_items > 0)
                        L.timed(SLOT_BU_HEAVY, [&] {
                            k_segment<<<L.grid_for_warps((a.n_seg_items + 31) / 32), CTA, 0, s>>>(a);
                        });
                    L.timed(SLOT_BU_HEAVY_FINALIZE, [&] {
                        bu_heavy_finalize_kernel<NW><<<L.grid_for_warps(a.blk_multi), CTA, 0, s>>>(a, a.blk_multi);
                    });
                }
                if (a.blk_end > a.blk_multi)
                    L.timed(SLOT_BU_G2, [&] {
                        k_thread<<<L.grid_for_warps(a.blk_end - a.blk_multi), CTA, 0, s>>>(a, a.blk_multi, a.blk_end);
                    });
            }
            L.timed(SLOT_COUNT, [&] {
                count_kernel<NW, 0><<<L.grid_for_warps(nblk), CTA, 0, s>>>(a, nblk, 0, nullptr, dir == 2 ? 1 : 0, h->opt_count_missing);
            });
            L.timed(SLOT_OTHER, [&] {
                epilogue_kernel<<<1, MSBFS_MAX_LANES, 0, s>>>(h->d_ctrl, h->d_log, h->log_cap, h->h_verdict, dir,
                                                               h->opt_alpha, h->opt_direction);
            });
            MSBFS_CUDA(cudaGetLastError());
        }
        L.timed(SLOT_OTHER, [&] { store_sums_kernel<<<1, MSBFS_MAX_LANES, 0, s>>>(h->d_ctrl, h->d_sums, group_base, lanes); });
        L.mark_batch();

        if (want_reached) {
            MSBFS_CUDA(cudaMemsetAsync(d_reached, 0, sizeof(ull) * MSBFS_MAX_LANES, s));
            for (int bit = 0; bit < 40; ++bit)
                count_kernel<NW, 1><<<L.grid_for_warps(nblk), CTA, 0, s>>>(a, nblk, bit, d_reached, 0, 0);
            ull tmp[MSBFS_MAX_LANES];
            MSBFS_CUDA(cudaMemcpyAsync(tmp, d_reached, sizeof tmp, cudaMemcpyDeviceToHost, s));
            MSBFS_CUDA(cudaStreamSynchronize(s));
            for (int j = 0; j < lanes; ++j) (*reached_out)[(size_t)group_base + j] = tmp[j];
        }
    }
    MSBFS_CUDA(cudaStreamSynchronize(s));
    L.collect();
    MSBFS_CUDA(cudaGetLastError());
    if (d_reached) cudaFree(d_reached);

    h->stats.lanes = lanes_per_batch;
    h->stats.batches = batches;
    h->stats.levels = levels;
    h->stats.levels_top_down = levels_td;
    h->stats.levels_bottom_up = levels_bu;
    h->stats.kernel_launches = L.launches;
    return MSBFS_OK;
}

/* Buffers are sized for `nw` 64-bit words per lane word and serve any narrower width too
 * (every kernel strides by its own NW; acc / hub_acc are all-zero whatever the width). */
/* The runtime loads a kernel's code at its first launch; asking for its attributes loads it now.
 * Called when the traversal state is sized (graph load, the preprocessing window), so that the
 * first run does not pay a few milliseconds of module loading per kernel in its level loop. */
template <int NW> void preload_kernels() {
    cudaFuncAttributes at;
    const void* fns[] = {(const void*)inject_kernel<NW>, (const void*)td_expand_kernel<NW, 32, 4>, (const void*)td_expand_kernel<NW, 8, 4>,
                         (const void*)td_expand_kernel<NW, 2, 4>, (const void*)td_heavy_kernel<NW, 4>, (const void*)td_finalize_kernel<NW>,
                         (const void*)bu_thread_kernel<NW, 4>, (const void*)bu_thread_kernel<NW, 8>, (const void*)bu_segment_kernel<NW, 4>,
                         (const void*)bu_segment_kernel<NW, 8>, (const void*)bu_heavy_finalize_kernel<NW>, (const void*)count_kernel<NW, 0>,
                         (const void*)td_streak_kernel<NW>, (const void*)tail_levels_kernel<NW>, (const void*)tail_list_kernel, (const void*)init_batch_kernel, (const void*)epilogue_kernel,
                         (const void*)store_sums_kernel};
    for (const void* f : fns) cudaFuncGetAttributes(&at, f);
    cudaGetLastError();
}

int ensure_traversal_buffers(msbfs_handle* h, int nw) {
    if (h->nw_alloc >= nw && h->d_visited) return MSBFS_OK;
    msbfs_release_traversal(h);
    cudaStream_t s = h->stream;
    const size_t word_bytes = (size_t)std::max(h->n_pad, 32) * nw * sizeof(ull);
    const size_t nblk = (size_t)std::max(h->n_pad, 32) / 32;
    MSBFS_CUDA(cudaMalloc(&h->d_visited, word_bytes));
    MSBFS_CUDA(cudaMalloc(&h->d_cur, word_bytes));
    MSBFS_CUDA(cudaMalloc(&h->d_nxt, word_bytes));
    MSBFS_CUDA(cudaMalloc(&h->d_acc, word_bytes));
    const size_t hub_bytes = (size_t)std::max(h->blk_multi * 32, 32) * nw * sizeof(ull);
    MSBFS_CUDA(cudaMalloc(&h->d_hub_acc, hub_bytes));
    MSBFS_CUDA(cudaMalloc(&h->d_bitmaps, sizeof(unsigned) * 3 * nblk));
    MSBFS_CUDA(cudaMalloc(&h->d_ctrl, sizeof(LevelCtrl)));
    h->log_cap = 1 << 16;
    MSBFS_CUDA(cudaMalloc(&h->d_log, sizeof(LevelLog) * (size_t)h->log_cap));
    MSBFS_CUDA(cudaHostAlloc(&h->h_verdict, sizeof(LevelVerdict), cudaHostAllocMapped));
    MSBFS_CUDA(cudaMemsetAsync(h->d_acc, 0, word_bytes, s));      /* invariant: zero between kernels */
    MSBFS_CUDA(cudaMemsetAsync(h->d_hub_acc, 0, hub_bytes, s));   /* invariant: zero between kernels */
    MSBFS_CUDA(cudaMemsetAsync(h->d_ctrl, 0, sizeof(LevelCtrl), s));
    h->nw_alloc = nw;
    if (nw == 1) preload_kernels<1>(); else if (nw == 2) preload_kernels<2>(); else if (nw == 8) preload_kernels<8>(); else preload_kernels<4>();

    /* Blackwell's 126 MB L2: keep the three bitmaps (n/8 bytes each — every random
     * probe of the traversal hits them first) resident with an access-policy window. */
    if (h->opt_l2_persist) {
        int max_win = 0, max_persist = 0;
        cudaDeviceGetAttribute(&max_win, cudaDevAttrMaxAccessPolicyWindowSize, h->device);
        cudaDeviceGetAttribute(&max_persist, cudaDevAttrMaxPersistingL2CacheSize, h->device);
        const size_t want = sizeof(unsigned) * 3 * nblk;
        if (max_win > 0 && max_persist > 0) {
            const size_t win = std::min(want, (size_t)max_win);
            cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, std::min((size_t)max_persist, win));
            cudaStreamAttrValue attr;
            std::memset(&attr, 0, sizeof attr);
            attr.accessPolicyWindow.base_ptr = h->d_bitmaps;
            attr.accessPolicyWindow.num_bytes = win;
            attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)max_persist / (double)win);
            attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            cudaStreamSetAttribute(s, cudaStreamAttributeAccessPolicyWindow, &attr);
            cudaGetLastError(); /* best effort: a refusal only costs the hint */
        }
    }
    return MSBFS_OK;
}

int pick_nw(const msbfs_handle* h) {
    if (h->opt_lanes == 64) return 1;
    if (h->opt_lanes == 128) return 2;
    if (h->opt_lanes == 256) return 4;
    if (h->opt_lanes == 512) return 8;
    if (h->q_groups <= 64) return 1;
    if (h->q_groups <= 128) return 2;
    if (h->q_groups <= 256) return 4;
    return 8; /* 512 lanes: one 64-byte word per vertex (two sectors per gather) */
}

int traverse_impl(msbfs_handle* h, bool want_reached, std::vector<ull>* reached_out) {
    if (!h->has_graph) return msbfs_fail(MSBFS_ERR_INVALID, "no graph loaded");
    cudaStream_t s = h->stream;
    const int K = h->q_groups;
    h->stats.lanes = 0; h->stats.batches = 0; h->stats.levels = 0;
    h->stats.levels_top_down = h->stats.levels_bottom_up = 0;
    h->stats.scanned_top_down = h->stats.scanned_bottom_up = 0;
    h->stats.frontier_vertices = h->stats.candidate_vertices = 0;
    h->stats.alg_bytes = 0; h->stats.kernel_launches = 0; h->stats.run_ms = 0.0; h->stats.bottom_up_ms = 0.0;
    h->level_log.clear();
    static const char* kSlotNames[MSBFS_NUM_SLOTS] = {"td_heavy", "td_expand_g32", "td_expand_g8", "td_expand_g2",
                                                      "bu_segment", "lane_major", "(unused)", "bu_thread",
                                                      "td_finalize", "bu_heavy_finalize", "count", "bookkeeping"};
    for (int i = 0; i < MSBFS_NUM_SLOTS; ++i) {
        std::memset(&h->slot_profile[i], 0, sizeof h->slot_profile[i]);
        std::strncpy(h->slot_profile[i].name, kSlotNames[i], sizeof h->slot_profile[i].name - 1);
    }
    if (K == 0) return MSBFS_OK;
    if (h->n_active == 0) { /* no edges: every F is 0 */
        MSBFS_CUDA(cudaMemsetAsync(h->d_sums, 0, sizeof(long long) * (size_t)K, s));
        MSBFS_CUDA(cudaStreamSynchronize(s));
        if (want_reached) reached_out->assign((size_t)K, 0ull);
        return MSBFS_OK;
    }
    const int nw = pick_nw(h);
    int rc = ensure_traversal_buffers(h, nw);
    if (rc) return rc;

    struct EventPair { /* destroyed on every return path */
        cudaEvent_t a = nullptr, b = nullptr;
        ~EventPair() { if (a) cudaEventDestroy(a); if (b) cudaEventDestroy(b); }
    } evp;
    MSBFS_CUDA(cudaEventCreate(&evp.a));
    MSBFS_CUDA(cudaEventCreate(&evp.b));
    MSBFS_CUDA(cudaEventRecord(evp.a, s));
    if (nw == 1) rc = run_batches<1>(h, want_reached, reached_out);
    else if (nw == 2) rc = run_batches<2>(h, want_reached, reached_out);
    else if (nw == 8) rc = run_batches<8>(h, want_reached, reached_out);
    else rc = run_batches<4>(h, want_reached, reached_out);
    if (rc) return rc;
    MSBFS_CUDA(cudaEventRecord(evp.b, s));
    MSBFS_CUDA(cudaEventSynchronize(evp.b));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, evp.a, evp.b);
    h->stats.run_ms = ms;

    /* per-level log -> totals and algorithmic bytes (DESIGN.md "algorithmic bytes") */
    int log_pos = 0;
    MSBFS_CUDA(cudaMemcpy(&log_pos, &h->d_ctrl->log_pos, sizeof(int), cudaMemcpyDeviceToHost));
    const int nlog = std::min(log_pos, h->log_cap);
    h->level_log.resize((size_t)nlog);
    if (nlog > 0)
        MSBFS_CUDA(cudaMemcpy(h->level_log.data(), h->d_log, sizeof(LevelLog) * (size_t)nlog, cudaMemcpyDeviceToHost));
    const long long w = 8ll * nw;
    {
        LevelCtrl hc;
        MSBFS_CUDA(cudaMemcpy(&hc, h->d_ctrl, sizeof hc, cudaMemcpyDeviceToHost));
        for (int i = 0; i < MSBFS_NUM_SLOTS; ++i) {
            msbfs_kernel_profile& p = h->slot_profile[i];
            p.scanned = (int64_t)hc.k_scanned[i]; p.words_fetched = (int64_t)hc.k_fetched[i];
            p.vertices_examined = (int64_t)hc.k_examined[i]; p.words_written = (int64_t)hc.k_written[i];
            p.alg_bytes = 4ll * p.scanned + w * p.words_fetched + (16ll + w) * p.vertices_examined + 2ll * w * p.words_written;
        }
        {   /* the lane-major batches keep their own counters */
            msbfs_kernel_profile& p = h->slot_profile[SLOT_LANE_MAJOR];
            p.scanned = h->lm_scanned; p.words_fetched = h->lm_scanned; p.vertices_examined = h->lm_expanded; p.words_written = h->lm_fresh;
            p.alg_bytes = 24ll * p.vertices_examined + 8ll * p.scanned + 16ll * p.words_written;
        }
        /* filled only when every launch was timed (option "profile") */
        h->stats.bottom_up_ms = h->slot_profile[SLOT_BU_HEAVY].ms + h->slot_profile[SLOT_BU_G8].ms +
                                h->slot_profile[SLOT_BU_G2].ms + h->slot_profile[SLOT_BU_HEAVY_FINALIZE].ms;
    }
    long long bytes = 0;
    for (const LevelLog& l : h->level_log) {
        h->stats.frontier_vertices += (int64_t)l.frontier_vertices;
        if (l.direction == 3) { /* lane-major level: queue entry 8 B + two row offsets 16 B per visit, col 4 B + bitmap word
                                   4 B per entry, bitmap update 8 B + queue entry 8 B per newly visited vertex */
            h->stats.scanned_top_down += (int64_t)l.scanned;
            bytes += 24ll * (long long)l.vertices_examined + 8ll * (long long)l.scanned + 16ll * (long long)l.words_written;
            continue;
        }
        if (l.direction == 1) h->stats.scanned_top_down += (int64_t)l.scanned;
        if (l.direction == 2) { h->stats.scanned_bottom_up += (int64_t)l.scanned; h->stats.candidate_vertices += (int64_t)l.vertices_examined; }
        /* col entry 4 B; every lane word moved (fetched through col, or stored) w B;
         * every examined vertex: two row offsets 16 B + its own word w B;
         * the new frontier is read once more by the count kernel: w B per vertex */
        bytes += 4ll * (long long)l.scanned + w * (long long)l.words_fetched + (16ll + w) * (long long)l.vertices_examined
               + 2ll * w * (long long)l.words_written + w * (long long)l.frontier_vertices;
    }
    /* per batch: clearing visited (w per vertex) and the bitmaps */
    bytes += (long long)h->stats.batches * ((long long)h->n_pad * w + 3ll * (h->n_pad / 8));
    h->stats.alg_bytes = bytes;
    return MSBFS_OK;
}

} // namespace

void msbfs_release_traversal(msbfs_handle* h) {
    msbfs_lm_release(h);
    if (h->d_bitmaps && h->stream && h->opt_l2_persist) {
        /* the access-policy window points at memory that is about to be freed */
        cudaStreamAttrValue attr;
        std::memset(&attr, 0, sizeof attr);
        attr.accessPolicyWindow.num_bytes = 0;
        cudaStreamSetAttribute(h->stream, cudaStreamAttributeAccessPolicyWindow, &attr);
        cudaGetLastError();
    }
    cudaFree(h->d_visited); h->d_visited = nullptr;
    cudaFree(h->d_cur); h->d_cur = nullptr;
    cudaFree(h->d_nxt); h->d_nxt = nullptr;
    cudaFree(h->d_acc); h->d_acc = nullptr;
    cudaFree(h->d_hub_acc); h->d_hub_acc = nullptr;
    cudaFree(h->d_bitmaps); h->d_bitmaps = nullptr;
    cudaFree(h->d_ctrl); h->d_ctrl = nullptr;
    cudaFree(h->d_log); h->d_log = nullptr;
    if (h->h_verdict) { cudaFreeHost(h->h_verdict); h->h_verdict = nullptr; }
    cudaFree(h->d_tail_list); h->d_tail_list = nullptr;
    h->nw_alloc = 0;
}

int msbfs_traverse(msbfs_handle* h) { return traverse_impl(h, false, nullptr); }

/* Called at the end of a graph build: size the traversal state for the widest lane word
 * while the clock is still in the preprocessing window (the reference allocates its
 * distance array there too, main.cu:295).  Best effort — when memory is short the state
 * is sized on the first run instead, for the width that run needs. */
void msbfs_prealloc_traversal(msbfs_handle* h) {
    if (!h->has_graph || h->n_active == 0) return;
    size_t free_b = 0, total_b = 0;
    if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) { cudaGetLastError(); return; }
    /* the widest word, unless the caller has said how many groups its runs will carry */
    int nw = MSBFS_MAX_NW;
    if (h->opt_lanes) nw = h->opt_lanes / 64;
    else if (h->opt_expect_groups > 0) nw = h->opt_expect_groups <= 64 ? 1 : h->opt_expect_groups <= 128 ? 2 : h->opt_expect_groups <= 256 ? 4 : 8;
    const size_t need = (size_t)h->n_pad * nw * sizeof(ull) * 4 + (size_t)h->blk_multi * 32 * nw * sizeof(ull);
    if (need + (need >> 1) > free_b) return;
    if (ensure_traversal_buffers(h, nw) != MSBFS_OK) {
        cudaGetLastError();
        msbfs_release_traversal(h);
        msbfs_clear_error();
    }
}

int msbfs_reached_entries_impl(msbfs_handle* h, int64_t* out_entries, int32_t count) {
    std::vector<ull> reached;
    int rc = traverse_impl(h, true, &reached);
    if (rc) return rc;
    for (int32_t i = 0; i < count; ++i) out_entries[i] = (i < (int32_t)reached.size()) ? (int64_t)reached[(size_t)i] : 0;
    return MSBFS_OK;
}
