/*
 * engine.h — internal layout of an msbfs handle and the launch interfaces between
 * the translation units of libmsbfs.so.  Not part of the C ABI.
 *
 * Data layout in HBM (all owned by the handle):
 *   perm      int32 [n]            file vertex id -> internal id.  Internal ids are
 *                                  sorted by degree, descending; isolated vertices
 *                                  get ids >= n_active and take no further part.
 *   row_ptr   int64 [n_pad + 1]    CSR offsets over internal ids (64-bit: 2m can
 *                                  exceed 2^31, unlike main.cu:119-124)
 *   col       int32 [2m]           neighbours (internal ids), each row sorted
 *                                  ascending = highest-degree neighbour first
 *   visited / cur / nxt / acc      lane words, n_pad x NW x uint64.  Bit j of a
 *                                  vertex's word belongs to query group (lane) j of
 *                                  the running batch.  NW = 1, 2, 4 or 8 (64 ... 512
 *                                  lanes); at NW = 4 a word is one 32-byte sector
 *                                  and moves as a single LDG/STG.256, at NW = 8 two.
 *   cur_nz / nxt_nz / vis_full     bitmaps, one bit per vertex: frontier word is
 *                                  non-zero / all batch lanes have visited it.  A
 *                                  cur/nxt word is only meaningful when its bit is
 *                                  set, so the word arrays are never cleared.
 *   hub_acc                        lane words for the heavy (chunked) vertices,
 *                                  all-zero between kernels.
 *   ctrl                           per-batch control block (counts, sums, live
 *                                  lanes, direction decision) — the level loop's
 *                                  state lives on the device; the host only reads
 *                                  a 32-byte verdict per level.
 */
#ifndef MSBFS_ENGINE_H
#define MSBFS_ENGINE_H

#include "internal.h"

#include <cstdint>
#include <vector>

constexpr long long MSBFS_MAX_RECORDS = 1ll << 39; /* edge records per graph: 2 m and 8 m stay far from 2^63 */
constexpr int MSBFS_MAX_NW      = 8;     /* 512 lanes: the automatic choice above 256 groups */
constexpr int MSBFS_MAX_LANES   = 64 * MSBFS_MAX_NW;
constexpr int MSBFS_CHUNK       = 1024;  /* CSR entries per heavy-row chunk (one warp) */
constexpr int MSBFS_DEG_HEAVY   = 1024;  /* deg >= this: row is split into chunks      */
constexpr int MSBFS_DEG_WARP    = 64;    /* deg >  this: one warp per vertex           */
constexpr int MSBFS_DEG_G8      = 8;     /* deg >  this: 8 threads per vertex, else 2  */

/* kernels that are timed / accounted separately */
enum KernelSlot {
    SLOT_TD_HEAVY = 0, SLOT_TD_G32, SLOT_TD_G8, SLOT_TD_G2,
    SLOT_BU_HEAVY, SLOT_LANE_MAJOR, SLOT_BU_G8, SLOT_BU_G2,
    SLOT_TD_FINALIZE, SLOT_BU_HEAVY_FINALIZE, SLOT_COUNT, SLOT_OTHER,
    MSBFS_NUM_SLOTS
};

/* per-level record kept on the device and copied back after a run */
struct LevelLog {
    int32_t  level;        /* BFS depth that was expanded                */
    int32_t  direction;    /* 1 top-down, 2 bottom-up, 0 source injection */
    uint64_t frontier_vertices;   /* vertices with a non-empty NEW frontier word */
    uint64_t frontier_entries;    /* their CSR entries                           */
    int64_t  unexplored_entries;  /* CSR entries of vertices not yet full         */
    uint64_t scanned;             /* CSR entries scanned by this level            */
    uint64_t words_fetched;       /* lane words fetched through col (random)      */
    uint64_t vertices_examined;   /* candidates (BU) / expanded vertices (TD)     */
    uint64_t words_written;       /* lane words stored or OR-ed                   */
};

struct LevelCtrl {
    unsigned long long cnt[MSBFS_MAX_LANES];   /* newly visited vertices per lane, this level */
    unsigned long long sums[MSBFS_MAX_LANES];  /* running distance sums per lane              */
    unsigned long long unvisited[MSBFS_MAX_LANES]; /* active vertices lane j has not reached yet     */
    unsigned long long live[MSBFS_MAX_NW];     /* lanes whose frontier is non-empty           */
    unsigned long long active[MSBFS_MAX_NW];   /* lanes that carry a query group              */
    unsigned long long n_f, m_f;               /* new frontier: vertices, CSR entries         */
    long long          m_u;                    /* CSR entries of not-yet-full vertices        */
    long long          n_u;                    /* not-yet-full vertices (degree > 0)          */
    unsigned long long scanned, words_fetched, examined, words_written; /* this level        */
    int level;                                 /* depth being expanded (-1 = injection)       */
    int log_pos;
    unsigned heavy_frontier;                   /* new frontier touches a chunked (hub) row    */
    int last_done, last_next;                  /* device-side copy of the verdict (streak kernel) */
    int cnt_mode;                              /* what cnt[] holds: 0 = newly visited per lane, 2 = still unvisited per lane */
    /* per-kernel accumulators over a whole run (roofline accounting), see KernelSlot */
    unsigned long long k_scanned[MSBFS_NUM_SLOTS], k_fetched[MSBFS_NUM_SLOTS], k_examined[MSBFS_NUM_SLOTS],
        k_written[MSBFS_NUM_SLOTS];
};

/* what the host reads back after each level (pinned, written by the epilogue) */
struct LevelVerdict {
    int32_t done;
    int32_t next_direction;
    int32_t heavy_frontier;
    int32_t streak_levels;  /* levels run by the last persistent top-down kernel */
    uint64_t n_f, m_f;
    int64_t m_u;
    int64_t n_u;            /* vertices some lane still misses */
};

struct TravArgs {
    const long long* row_ptr;
    const int*       col;
    unsigned long long* visited;
    unsigned long long* cur;
    unsigned long long* nxt;
    unsigned long long* acc;
    unsigned long long* hub_acc;
    unsigned int* cur_nz;
    unsigned int* nxt_nz;
    unsigned int* vis_full;
    LevelCtrl* ctrl;
    const long long* chunk_level_base; /* [n_chunk_levels + 1] */
    int n_chunk_levels;
    long long n_chunks;
    int n_pad;      /* vertices incl. padding to a multiple of 32 */
    int n_active;   /* vertices with at least one CSR entry (internal ids [0, n_active)) */
    int blk_heavy;  /* 32-vertex blocks [0, blk_heavy)        : chunked rows    */
    int blk_warp;   /*                 [blk_heavy, blk_warp)  : warp per vertex */
    int blk_g8;     /*                 [blk_warp, blk_g8)     : 8 threads/vertex */
    int blk_end;    /*                 [blk_g8, blk_end)      : 2 threads/vertex */
    /* bottom-up: rows longer than `seg` entries are cut into segments of `seg`; segment
     * k of vertex u is work item seg_level_base[k] + u (level-major: the first segments
     * of all long rows come first).  Blocks [0, blk_multi) hold the long rows. */
    const long long* seg_level_base; /* [n_seg_levels + 1] */
    int n_seg_levels;
    long long n_seg_items;
    int seg;
    int blk_multi;
};

struct msbfs_handle {
    int device = 0;
    int sm_count = 148;
    cudaStream_t stream = nullptr;

    /* graph */
    int32_t n = 0, n_active = 0, n_pad = 0;
    int64_t m = 0, nnz = 0;
    int32_t* d_perm = nullptr;
    long long* d_row_ptr = nullptr;
    int* d_col = nullptr;
    int blk_heavy = 0, blk_warp = 0, blk_g8 = 0, blk_end = 0;
    long long* d_chunk_level_base = nullptr;
    int n_chunk_levels = 0;
    long long n_chunks = 0;
    long long* d_seg_level_base = nullptr;
    int n_seg_levels = 0, blk_multi = 0, seg = 256;
    long long n_seg_items = 0;
    bool has_graph = false;
    bool graph_imported = false;  /* perm / row_ptr / col belong to another process (share.cu): close, do not free */
    unsigned max_degree = 0;
    double build_ms = 0.0;

    /* traversal state */
    int nw_alloc = 0;
    unsigned long long *d_visited = nullptr, *d_cur = nullptr, *d_nxt = nullptr, *d_acc = nullptr, *d_hub_acc = nullptr;
    unsigned int *d_bitmaps = nullptr; /* cur_nz | nxt_nz | vis_full, contiguous (one L2 window) */
    LevelCtrl* d_ctrl = nullptr;
    LevelLog* d_log = nullptr;
    int log_cap = 0;
    LevelVerdict* h_verdict = nullptr; /* pinned */
    int* d_tail_list = nullptr;        /* unfinished vertices for the tail kernel + 4 ints of control in front */

    /* lane-major engine for road-like graphs (lane_major.cu): a visited bitmap per group, two
     * queues of (group, vertex) pairs, a control block */
    unsigned int* d_lm_bits = nullptr;
    unsigned long long* d_lm_q[2] = {nullptr, nullptr};
    void* d_lm_ctrl = nullptr;
    int* d_lm_ell = nullptr;      /* rows as fixed-width records when every row has <= 8 entries */
    int lm_ell_width = 0;
    int lm_lanes = 0, lm_grid = 0;
    long long lm_cap = 0;
    int64_t lm_expanded = 0, lm_scanned = 0, lm_fresh = 0; /* totals of the last run's lane-major batches */

    /* resident queries / results */
    int32_t q_groups = 0;
    int64_t q_sources = 0;
    long long* d_q_offsets = nullptr;
    int* d_q_sources = nullptr;
    long long* d_sums = nullptr;
    int64_t q_cap_groups = 0, q_cap_sources = 0;
    long long* h_pinned_io = nullptr; /* pinned staging for queries in / sums out */
    size_t h_pinned_io_bytes = 0;
    /* groups run in the order q_order (run position -> caller's index), heaviest first, so that
     * groups whose sources include hubs share batches; empty = caller's order */
    std::vector<int32_t> q_order;
    long long* d_deg = nullptr;
    int64_t deg_cap = 0;
    unsigned long long* d_scalar = nullptr; /* 8-byte scratch for the packed min all-reduce */

    /* pinned staging buffers of the loader (api.cu), allocated on first use */
    char* stage_buf[32] = {};   /* two per filler thread */
    size_t stage_cap[32] = {};
    cudaEvent_t stage_ev[32] = {};

    /* options */
    int opt_lanes = 0, opt_direction = 0, opt_l2_persist = 1;
    int opt_deg_thread = MSBFS_DEG_G8, opt_deg_warp = MSBFS_DEG_WARP, opt_deg_heavy = MSBFS_DEG_HEAVY; /* bin limits (next graph load) */
    int opt_streak_levels = 1 << 20;           /* cap on levels per persistent launch (profiling) */
    int opt_streak_max_deg = 256;              /* longest row for which the persistent kernel is used */
    int opt_reorder = 1;                       /* run groups heaviest-first (by their sources' degrees) */
    int opt_tail = 1;                          /* last levels (few unfinished vertices) in one persistent bottom-up kernel over their list */
    int opt_streak = 1;                        /* persistent top-down kernel on short-row graphs  */
    int opt_expect_groups = 0;                 /* most groups a run will carry (sizes the traversal state at graph load); 0 = any */
    int opt_order_key = 0;                     /* run-order key: 0 = two-hop weight of the sources, 1 = smallest source id (spatial proxy) */
    int opt_lane_major = -1;                   /* sparse per-group engine: -1 = when every row has <= 16 entries, 0 = never, 1 = whenever the graph allows */
    int opt_lm_ell = 1;                        /* lane-major engine: fixed-width row records when every row has <= 8 entries */
    int opt_stage_threads = 0;                 /* host threads filling the loader's pinned staging buffers; 0 = by size of the load */
    int opt_count_missing = 1;                 /* after dense bottom-up levels count what lanes still miss instead of the new frontier */
    int opt_seg = 0;                           /* bottom-up segment length (next graph load); 0 = by graph size */
    int opt_col_width = 0;                     /* col entries per load in the bottom-up scan: 4 (128-bit), 8 (256-bit), 0 = 8 for the long-row kernel at 512 lanes, else 4 */
    long long opt_alpha = 40;

    /* stats of the last run */
    msbfs_stats stats{};
    std::vector<LevelLog> level_log;
    int opt_profile = 0;                       /* time every kernel launch with CUDA events */
    std::vector<cudaEvent_t> ev_pool;          /* start/stop pairs, grown on demand         */
    msbfs_kernel_profile slot_profile[MSBFS_NUM_SLOTS]{};
    std::vector<msbfs_launch_record> launch_log; /* one record per timed launch of the last run (option "profile") */
    std::vector<double> batch_ms;              /* device time of every batch of the last run */
};

/* graph_build.cu */
int msbfs_build_from_device_edges(msbfs_handle* h, int32_t n, int64_t m, const int32_t* d_uv);
void msbfs_release_graph(msbfs_handle* h);

/* api.cu: pieces of the loader that comm.cu's sharded load reuses */
int msbfs_open_graph_file(const char* path, int* fd_out, int32_t* n_out, int64_t* m_out);
int msbfs_stage_records(msbfs_handle* h, int fd, int64_t first, int64_t count, char* d_dst, const char* what);

/* lane_major.cu */
struct msbfs_lm_result {
    int overflow = 0, levels = 0, log_pos_in = 0, log_pos_out = 0, want_reach = 0;
    int64_t expanded = 0, scanned = 0, fresh = 0;
    unsigned long long reach[MSBFS_MAX_LANES];
};
int msbfs_lm_ensure(msbfs_handle* h, int lanes);
int msbfs_lm_run_batch(msbfs_handle* h, int group_base, int lanes, msbfs_lm_result* out);
void msbfs_lm_release(msbfs_handle* h);

/* traverse.cu */
int msbfs_traverse(msbfs_handle* h);
int msbfs_reached_entries_impl(msbfs_handle* h, int64_t* out_entries, int32_t count);
void msbfs_release_traversal(msbfs_handle* h);
void msbfs_prealloc_traversal(msbfs_handle* h);

#endif
