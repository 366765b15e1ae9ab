/*
 * msbfs.h — C ABI of the B200-native multi-source-BFS engine (libmsbfs.so).
 *
 * The reference (/root/reference/main.cu) is a single executable with no plugin
 * or FFI surface; this ABI is cut at the function-level seam its main() already
 * has: device set-up (main.cu:227-228, 283-295), the loaders (92-130, 134-164),
 * the per-query BFS + sum pair called from the scheduler loop (312-322 ->
 * GPUMultiSourceBFS 40-73, ComputeFofU 75-89) and the result reduction
 * (324-397).  Each entry point names the reference lines it replaces.
 *
 * Conventions
 *   - plain C types only; no C++ objects or exceptions cross the boundary;
 *   - every function returns an int status (MSBFS_OK == 0); the message for the
 *     calling thread's last failure is msbfs_last_error();
 *   - the opaque handle owns all device memory and streams; the caller owns every
 *     host buffer and may free it as soon as the call returns;
 *   - one handle per (process, device); a handle is not thread-safe, distinct
 *     handles are independent;
 *   - there is NO CPU fallback: with no usable CUDA device msbfs_create fails.
 */
#ifndef MSBFS_H
#define MSBFS_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define MSBFS_OK          0
#define MSBFS_ERR_INVALID 1 /* bad argument / call order                      */
#define MSBFS_ERR_IO      2 /* file cannot be opened or is truncated          */
#define MSBFS_ERR_CUDA    3 /* CUDA runtime failure (no device, OOM, fault)   */
#define MSBFS_ERR_FORMAT  4 /* vertex id outside [0, n) in an edge record     */
#define MSBFS_ERR_NOMEM   5 /* host allocation failed                         */
#define MSBFS_ERR_RANGE   6 /* value does not fit the packed representation   */

typedef struct msbfs_handle msbfs_t;

/* Message describing the last failure on the calling thread ("" if none). */
const char* msbfs_last_error(void);

/* ------------------------------------------------------------ life cycle -- */

/* Replaces cudaSetDevice(rank % numGPU) and the device allocations,
 * main.cu:227-228, 283-295.  Binds the handle to `device`, creates its stream. */
int msbfs_create(int device, msbfs_t** out);

/* Replaces the cudaFree calls, main.cu:416-418. */
int msbfs_destroy(msbfs_t* h);

/* Tuning / test knobs; unknown keys -> MSBFS_ERR_INVALID.
 *   "lanes"      0 = auto (64/128/256/512 by group count), or 64 / 128 / 256 / 512
 *   "direction"  0 = direction-optimizing, 1 = top-down only, 2 = bottom-up only
 *   "alpha"      cost of a top-down CSR entry relative to a bottom-up one (default 40): the next
 *                level runs top-down while alpha * (frontier_entries + 0.8 frontier_vertices)
 *                < unexplored_entries + alpha * unexplored_vertices
 *   "l2_persist" 1 = pin frontier bitmaps in L2 with an access-policy window (default), 0 = off
 *   "reorder"    1 = run the groups heaviest-first (sum of their sources' degrees) so that groups
 *                with hubs among their sources share batches (default), 0 = caller's order
 *   "profile"    1 = time every kernel launch with CUDA events (msbfs_get_kernel_profile)
 *   "expect_groups"  most groups a run on this handle will carry (0 = unknown): the traversal
 *                state allocated at graph load is sized for it instead of for 512 lanes
 *   "order_key"  what "reorder" sorts the groups by: 0 = two-hop weight of their sources (default),
 *                1 = smallest source id, a locality proxy for graphs whose ids are spatial
 *                (lattices, road networks); measured without effect on the 4096^2 lattice
 *   "lane_major"  the sparse per-group engine for road-like graphs (a visited bitmap per group and a
 *                queue of (group, vertex) pairs, one persistent kernel per batch): -1 = when every
 *                row has at most 16 entries (default), 0 = never, 1 = on any graph whose longest
 *                row fits "streak_max_deg"; a batch whose level outgrows the queue (2 x the
 *                non-isolated vertices) is redone by the bit-parallel engine
 *   "lm_ell"     1 = the lane-major engine reads rows from fixed-width records (4 or 8 ids per vertex,
 *                one 16- or 32-byte load) when every row has at most 8 entries (default), 0 = from the CSR
 *   "stage_threads"  host threads that fill the loader's pinned staging buffers, 1 ... 16; 0 = 4 below
 *                1 GiB, 8 from there on, never more than half the cores (default)
 *   "tail"       1 = once at most 2^18 vertices (2^23 CSR entries, rows <= 4096) are left unfinished, list them
 *                and run every remaining level bottom-up over the list in one persistent kernel — on graphs of
 *                a million vertices or more (default); 2 = on any graph; 0 = never
 *   "count_missing"  1 = after a bottom-up level that leaves fewer vertices unfinished than it put
 *                into the frontier, take the per-lane counts from what the unfinished vertices
 *                still miss instead of from the new frontier words (default), 0 = always the latter
 *   "col_width"  col entries per load in the bottom-up scan: 4 (128-bit), 8 (256-bit), 0 = 8 for the
 *                long-row kernel of 512-lane batches and 4 elsewhere (default)
 * Layout knobs read at the next graph load (defaults chosen by sweeps):
 *   "seg"        bottom-up segment length, 0 = by graph size (256 ... 1024)
 *   "deg_heavy" / "deg_warp" / "deg_thread"   degree limits of the top-down bins
 * Persistent top-down kernel for short-row (road-like) graphs:
 *   "streak" 0/1, "streak_max_deg" longest row it accepts (256), "streak_levels" levels per launch */
int msbfs_set_option(msbfs_t* h, const char* key, int64_t value);

/* ----------------------------------------------------------------- graph -- */

/* Replaces LoadGraphBin + the CSR upload, main.cu:92-130 and 283-291.
 * File layout (main.cu:101-116): int32 n, int64 m, m x (int32 u, int32 v); every
 * record is an undirected edge stored in both directions, duplicates and
 * self-loops kept.  The file is streamed through pinned staging buffers with
 * asynchronous H2D copies and the CSR (64-bit offsets) is built on the device.
 * Unlike the reference (UB, main.cu:114-115) ids outside [0, n) are rejected
 * with MSBFS_ERR_FORMAT.  A file that cannot be opened gives MSBFS_ERR_IO. */
int msbfs_load_graph_file(msbfs_t* h, const char* path, int32_t* n_out, int64_t* m_out);

/* Same, from `m` (u, v) records already in host memory. */
int msbfs_set_graph_edges(msbfs_t* h, int32_t n, int64_t m, const int32_t* uv_pairs);

/* Same, from records already resident in this handle's device memory
 * (e.g. produced by msbfs_gen_edges_device). The buffer is only read. */
int msbfs_set_graph_edges_device(msbfs_t* h, int32_t n, int64_t m, const int32_t* d_uv_pairs);

/* One CSR per GPU for ranks that share it (replaces the per-rank cudaMalloc + upload of
 * main.cu:283-291 when `mpirun -np R ... -gn G` puts R / G ranks on a device, main.cu:227).
 * The rank that loaded the graph exports a blob — sizes, layout tables and CUDA IPC handles
 * of the CSR arrays; call with capacity 0 to learn its size — the host program ships the
 * bytes to the other ranks on that device (MPI_Bcast in host/main.cpp), and each of them
 * imports it into its own handle, which then reads the exporter's arrays.  The exporter's
 * handle must outlive every importer's.  Import fails with MSBFS_ERR_CUDA where CUDA IPC is
 * not available (load the file instead) and refuses a blob made for another device. */
int msbfs_export_graph(msbfs_t* h, void* blob, int64_t capacity, int64_t* size_out);
int msbfs_import_graph(msbfs_t* h, const void* blob, int64_t size);

/* --------------------------------------------------------------- queries -- */

/* Replaces the scheduler loop body, main.cu:312-322 (GPUMultiSourceBFS 40-73 +
 * ComputeFofU 75-89), for `num_groups` groups at once.  Group i's sources are
 * sources[group_offsets[i] .. group_offsets[i+1]) (host arrays; offsets need not
 * start at 0).  out_sums[i] = F(group i) = sum over reachable vertices of the hop
 * distance to the nearest valid source; ids outside [0, n) are ignored
 * (main.cu:49); unreachable vertices add nothing (main.cu:84-85); a group with no
 * valid source gives 0.  Copies queries H2D and sums D2H inside the call. */
int msbfs_run(msbfs_t* h, int32_t num_groups, const int64_t* group_offsets,
              const int32_t* sources, int64_t* out_sums);

/* The same work split in three so that the traversal can be timed with inputs
 * and outputs resident in HBM: upload (H2D), run (device only), download (D2H). */
int msbfs_upload_queries(msbfs_t* h, int32_t num_groups, const int64_t* group_offsets,
                         const int32_t* sources);
int msbfs_run_resident(msbfs_t* h);
int msbfs_download_sums(msbfs_t* h, int64_t* out_sums, int32_t count);

/* Degrees (CSR row lengths, both directions counted, main.cu:114-115) of `count` vertices
 * given by their file ids; ids outside [0, n) give 0.  Host arrays.  Lets a scheduler
 * estimate how heavy a group is (a source with a million neighbours floods the frontier
 * of its whole batch) before it deals groups out to ranks; the reference's round-robin,
 * main.cu:303-307, has no such notion. */
int msbfs_vertex_degrees(msbfs_t* h, int64_t count, const int32_t* vertex_ids, int64_t* out_degrees);

/* The weight the engine itself orders groups by ("reorder"): per group, the sum over its
 * sources of (degree + degrees of the source's first 256 neighbours) — how fast the group's
 * BFS reaches the hubs.  A multi-rank scheduler sorts by it and hands out batches of the
 * order (host/main.cpp); groups as in msbfs_run, host arrays.  No counterpart in the
 * reference's round-robin (main.cu:303-307). */
int msbfs_group_weights(msbfs_t* h, int32_t num_groups, const int64_t* group_offsets, const int32_t* sources,
                        int64_t* out_weights);

/* Measurement metadata, not part of the reference's output: for the groups of the
 * last run, the number of CSR entries of reached vertices (Graph500 traversed
 * edges = half of it).  Runs an extra device pass when called. */
int msbfs_reached_entries(msbfs_t* h, int64_t* out_entries, int32_t count);

/* No counterpart in the reference.  F of up to 64 groups on a graph of at most 65,536
 * vertices, computed without allocating: for callers that may not call cudaMalloc or
 * cudaFree where this runs (a latency-critical loop that cannot afford the device-wide
 * stall an allocation may cause, a host application whose memory pool owns the whole
 * GPU, a process held to a memory budget fixed at start-up).  The traversal state lives
 * in a 1.5 MiB __device__ array inside the library (every GPU that loads the library
 * holds it until exit); larger inputs are refused with MSBFS_ERR_RANGE — use msbfs_run.
 *   inputs   CSR of the graph in the caller's own vertex ids, both directions stored
 *            (d_row_offsets int64[n + 1], d_col_indices int32), groups as in msbfs_run
 *            (d_group_offsets int64[num_groups + 1] starting at 0, d_sources int32);
 *   output   d_out_sums int64[num_groups];
 *   none of the pointers may be NULL (pass any valid device pointer for an array that has
 *   no elements); column entries outside [0, n) are skipped, as out-of-range sources are;
 *   all five are device pointers on the CURRENT device; the work is queued on
 *   `cuda_stream` (a cudaStream_t, NULL = default stream) and the sums are complete when
 *   that stream reaches the end of it — the call itself does not wait.
 * The first call on each device also looks up the array's device address; make that
 * first call before any stream capture.  The address is kept for the life of the process:
 * a cudaDeviceReset() in between invalidates it, so do not reset a device this entry
 * point has been used on.  All calls share the one array (each call
 * clears the part it uses before reading it), so calls that overlap on the same GPU —
 * different streams or threads — race with each other: serialise them. */
int msbfs_small_run_noalloc(int32_t n, const int64_t* d_row_offsets, const int32_t* d_col_indices,
                            int32_t num_groups, const int64_t* d_group_offsets, const int32_t* d_sources,
                            int64_t* d_out_sums, void* cuda_stream);

/* ------------------------------------------------------------- reduction -- */

/* Replaces the argmin, main.cu:379-397, over a full array: first F >= 0 seeds,
 * strict '<' keeps the lowest index among ties; K = 0 gives (-1, -1).  The report
 * prints min_k + 1 (main.cu:409).  Host-only. */
int msbfs_argmin(int32_t num_groups, const int64_t* sums, int32_t* min_k_out, int64_t* min_f_out);

/* Local part of the distributed argmin that replaces the (q, F) Pair gather,
 * main.cu:166-193, 324-375: packed = (F << 24) | global_query_id, minimised over
 * sums[0..count) with ids first_qid, first_qid + 1, ...; UINT64_MAX when count = 0.
 * A plain unsigned min over ranks then reproduces the lowest-index tie-break.
 * MSBFS_ERR_RANGE when an id needs more than 24 bits or a sum more than 39. */
int msbfs_argmin_packed(const int64_t* sums, int32_t first_qid, int32_t count, uint64_t* packed_out);
int msbfs_unpack_argmin(uint64_t packed, int32_t* min_k_out, int64_t* min_f_out);
/* Same packing for a share whose groups are not contiguous: group i of the share has the
 * global query id query_ids[i]. */
int msbfs_argmin_packed_ids(const int64_t* sums, const int32_t* query_ids, int32_t count, uint64_t* packed_out);

/* One 8-byte min all-reduce over NCCL / NVLink (replaces MPI_Gather + MPI_Gatherv of the
 * Pair struct and the scan on rank 0, main.cu:166-193, 340-397) when every rank owns its
 * own GPU.  `nccl_comm` is an ncclComm_t — the caller's own, or one made by
 * msbfs_nccl_comm_create.  NCCL is resolved at run time (dlopen), so the library has no NCCL
 * link dependency.  With several ranks on one device (mpirun -np 2 ... -gn 1) NCCL cannot
 * form a communicator: reduce the packed word with MPI_Allreduce instead. */
int msbfs_allreduce_min_u64(msbfs_t* h, void* nccl_comm, uint64_t* inout);

/* Communicator bootstrap for a host program that has MPI (or any launcher) but does not
 * want to link NCCL itself — no counterpart in the reference, which has no NCCL.  Rank 0
 * calls msbfs_nccl_unique_id and broadcasts the MSBFS_NCCL_ID_BYTES bytes (MPI_Bcast of
 * MPI_BYTE); every rank then calls msbfs_nccl_comm_create with the same id, the world size
 * and its rank (collective: all ranks must call it), on the device its handle is bound to. */
#define MSBFS_NCCL_ID_BYTES 128
int msbfs_nccl_unique_id(void* id_out);
int msbfs_nccl_comm_create(msbfs_t* h, const void* id, int32_t world, int32_t rank, void** comm_out);
int msbfs_nccl_comm_destroy(void* comm);

/* Replaces LoadGraphBin on rank 0 + MPI_Bcast of the CSR + the per-rank upload (main.cu:237-255,
 * 283-291) for ranks that own a GPU each: rank r reads only records [r * ceil(m / world), ...)
 * of the file, one ncclAllGather over `nccl_comm` (NVLink) completes the edge list on every
 * GPU, and every rank builds its CSR there.  Collective: all `world` ranks of the communicator
 * call it with the same file.  Errors as msbfs_load_graph_file. */
int msbfs_load_graph_file_sharded(msbfs_t* h, const char* path, void* nccl_comm, int32_t rank, int32_t world,
                                  int32_t* n_out, int64_t* m_out);

/* -------------------------------------------------------------- scheduler -- */

/* Replaces the round-robin ownership of main.cu:303-307.  Orders the groups heaviest first by
 * `weights` (msbfs_group_weights), cuts the order into batches of `lanes` groups and assigns
 * the batches to `world` ranks: by measured cost (longest first onto the least-loaded rank)
 * when batch_cost[0 .. num_costs) holds one figure per batch (msbfs_get_batch_ms of an earlier
 * pass over the same plan, concatenated over ranks in batch order), else in snake order
 * 0..R-1, R-1..0, ...  Outputs: order_out[i] = group at run position i (all groups);
 * owner_out[g] = rank that runs group g; *lanes_out = batch width the ranks should set with
 * msbfs_set_option("lanes") (0 = the engine's own choice, when world == 1).  A rank runs its
 * groups in the relative order of order_out.  lanes = 0 chooses: 512; 256 when that leaves a
 * rank fewer than two batches and the 512-group batches' weights differ by more than 3 x
 * (power-law graphs: a rank then mixes a fast and a slow batch; hub-less graphs keep 512);
 * halved further while ranks would stay idle.  msbfs_plan_lanes is the same rule without the
 * weights (a sizing hint before the graph is loaded).  Host-only; every rank computes the same plan from the same inputs. */
int msbfs_plan_lanes(int32_t num_groups, int32_t world);
int msbfs_plan_shares(int32_t num_groups, const int64_t* weights, int32_t world, int32_t lanes,
                      const double* batch_cost, int32_t num_costs, int32_t* order_out, int32_t* owner_out,
                      int32_t* lanes_out);

/* ------------------------------------------------------------ statistics -- */

typedef struct msbfs_stats {
    int32_t n;                 /* vertices in the file                              */
    int32_t n_active;          /* vertices with degree > 0 (relabelled to the front) */
    int64_t m;                 /* edge records                                      */
    int64_t csr_entries;       /* 2 m                                               */
    int32_t lanes;             /* query groups advanced together in the last run    */
    int32_t batches;           /* lane batches in the last run                      */
    int64_t levels;            /* BFS levels over all batches                       */
    int64_t levels_top_down;
    int64_t levels_bottom_up;
    int64_t scanned_top_down;  /* CSR entries scanned by top-down kernels           */
    int64_t scanned_bottom_up; /* CSR entries scanned by bottom-up kernels          */
    int64_t frontier_vertices; /* sum over levels of vertices with a non-empty frontier word */
    int64_t candidate_vertices;/* sum over bottom-up levels of vertices examined    */
    int64_t alg_bytes;         /* algorithmic bytes of the last run     */
    int64_t kernel_launches;   /* kernels of this library launched by the last run  */
    double  build_ms;          /* device time of the last graph build               */
    double  run_ms;            /* device time of the last run (CUDA events)         */
    double  bottom_up_ms;      /* of which bottom-up kernels (if profiling enabled) */
} msbfs_stats;

int msbfs_get_stats(msbfs_t* h, msbfs_stats* out);

/* Per-kernel accounting of the last run, for roofline reporting.  With option
 * "profile" = 1 every launch is bracketed by CUDA events on the handle's stream and
 * `ms` is their sum; the counters are always kept.  alg_bytes = 4 B per scanned CSR
 * entry + one lane word per word fetched + (16 B + one word) per vertex examined +
 * two words per word written.  Slots: 0-3 top-down (chunked rows, 32, 8,
 * 2 threads per vertex; 3 is also the persistent short-row kernel), 4 bottom-up long rows
 * (one thread per segment), 5-6 unused,
 * 7 bottom-up short rows (one thread per vertex), 8 top-down finalize, 9 bottom-up long-row
 * finalize, 10 per-lane count, 11 bookkeeping.  Measurement only: nothing in
 * the reference corresponds to it. */
typedef struct msbfs_kernel_profile {
    char    name[32];
    int64_t launches;
    double  ms;
    int64_t alg_bytes;
    int64_t scanned, words_fetched, vertices_examined, words_written;
} msbfs_kernel_profile;

/* One record per BFS level of the last run (all batches, in execution order); the
 * counters behind MSBFS_STATS and the roofline accounting.  direction: 0 = source
 * injection, 1 = top-down, 2 = bottom-up, 3 = a level of the lane-major engine (scanned = CSR entries read,
 * vertices_examined = (group, vertex) visits, words_written = newly visited).  Returns the number of records through
 * count_out (at most `capacity` are copied). */
typedef struct msbfs_level_record {
    int32_t  level;
    int32_t  direction;
    uint64_t frontier_vertices;   /* vertices that got a non-empty new frontier word */
    uint64_t frontier_entries;    /* their CSR entries                               */
    int64_t  unexplored_entries;  /* CSR entries of vertices some lane still misses  */
    uint64_t scanned;             /* CSR entries scanned                             */
    uint64_t words_fetched;       /* lane words fetched through col                  */
    uint64_t vertices_examined;
    uint64_t words_written;
} msbfs_level_record;

int msbfs_get_level_log(msbfs_t* h, msbfs_level_record* out, int32_t capacity, int32_t* count_out);

/* With option "profile" = 1: every kernel launch of the last run, in launch order, with the
 * batch and BFS level it belongs to (level -1 = source injection) and its CUDA-event time.
 * Measurement only (per-level cost of each kernel); nothing in the reference corresponds. */
typedef struct msbfs_launch_record {
    int32_t batch;
    int32_t level;
    int32_t slot;   /* as in msbfs_get_kernel_profile */
    float   ms;
} msbfs_launch_record;
int msbfs_get_launch_log(msbfs_t* h, msbfs_launch_record* out, int32_t capacity, int32_t* count_out);

/* Device time (CUDA events) of every lane batch of the last run, in run order: what a
 * multi-rank scheduler needs to plan shares from measured cost (msbfs_plan_shares). */
int msbfs_get_batch_ms(msbfs_t* h, double* out, int32_t capacity, int32_t* count_out);

int msbfs_num_kernel_slots(void);
int msbfs_get_kernel_profile(msbfs_t* h, int slot, msbfs_kernel_profile* out);

/* ----------------------------------------------- host-side file helpers -- */

/* Replaces LoadQueryBin, main.cu:134-164: uint8 K, then per group uint8 S + S x
 * int32.  Also accepts the extended container (first byte 0, "MSBFSQ2", uint32 K,
 * per group uint32 S + ids) needed for K or S > 255, which the reference format
 * cannot express.  Arrays are malloc'ed; release with msbfs_free. */
int msbfs_load_query_file(const char* path, int32_t* num_groups_out,
                          int64_t** group_offsets_out, int32_t** sources_out);
int msbfs_write_query_file(const char* path, int32_t num_groups, const int64_t* group_offsets,
                           const int32_t* sources, int force_extended);
int msbfs_write_graph_file(const char* path, int32_t n, int64_t m, const int32_t* uv_pairs);
void msbfs_free(void* p);

/* ------------------------------------------------- synthetic benchmarks -- */

/* kind: 0 R-MAT (p0 = scale), 1 Kronecker = R-MAT with scrambled labels
 * (p0 = scale), 2 lattice (p0 = width, p1 = height; m is implied), 3 Erdos-Renyi
 * (p0 = n).  Host and device versions write identical records. */
int msbfs_gen_num_vertices(int kind, int32_t p0, int32_t p1, int32_t* n_out);
int msbfs_gen_edges_host(int kind, int32_t p0, int32_t p1, int64_t m, uint64_t seed, int32_t* uv_out);
int msbfs_gen_edges_device(int device, int kind, int32_t p0, int32_t p1, int64_t m, uint64_t seed,
                           int32_t* d_uv_out);
int msbfs_gen_queries_host(int32_t n, int32_t num_groups, int32_t sources_per_group, uint64_t seed,
                           int64_t* group_offsets_out, int32_t* sources_out);

#ifdef __cplusplus
}
#endif
#endif /* MSBFS_H */
