/*
 * oracle.c — TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C CPU restatement of the reference's multi-source-BFS path
 * (/root/reference/main.cu).  Only tests/, __graft_entry__.smoke() and
 * bench.py's cpu_baseline / --impl reference legs may load this library; the
 * product (libmsbfs.so, ./main) never links or calls it and has no CPU fallback.
 *
 * Parity pin: the reference ships no tests or golden vectors, so
 * this restatement is pinned against the reference's OWN source executed on the
 * host (oracle/_ref/main_host, built by oracle/Makefile from main.cu where it
 * lies, through stub headers) — see tests/test_oracle_pin.py and tests/golden/.
 *
 * Each function cites the reference lines it restates.  Deliberate widenings
 * (documented, result-neutral on inputs the reference can load): CSR offsets are
 * int64 instead of int32 (main.cu:119-124 overflows at 2m >= 2^31), and the
 * query reader also accepts the extended "MSBFSQ2" container for K or S > 255.
 */
#define _GNU_SOURCE
#include "oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#ifdef _OPENMP
#include <omp.h>
#endif

/* ------------------------------------------------------------------ CSR --- */

/* main.cu:106-129 — adjacency in record order, both directions, duplicates and
 * self-loops kept (a self-loop (u,u) lands twice in row u), then flattened.
 * A counting sort that walks the records in file order reproduces the
 * vector<vector<int>> push_back order exactly. */
int oracle_build_csr(int32_t n, int64_t m, const int32_t* uv, int64_t** row_out, int32_t** col_out) {
    if (n < 0 || m < 0) return -1;
    int64_t* row = (int64_t*)calloc((size_t)n + 2, sizeof(int64_t));
    int32_t* col = (int32_t*)malloc(sizeof(int32_t) * (size_t)(2 * m > 0 ? 2 * m : 1));
    if (!row || !col) { free(row); free(col); return -2; }
    for (int64_t i = 0; i < m; ++i) {
        int32_t u = uv[2 * i], v = uv[2 * i + 1];
        if (u < 0 || u >= n || v < 0 || v >= n) { free(row); free(col); return -3; } /* UB in the reference (main.cu:114-115) */
        row[u + 2]++; row[v + 2]++;
    }
    /* row[i+1] becomes the write cursor of vertex i, row[i] its final start */
    for (int32_t i = 0; i < n; ++i) row[i + 2] += row[i + 1];
    for (int64_t i = 0; i < m; ++i) {
        int32_t u = uv[2 * i], v = uv[2 * i + 1];
        col[row[u + 1]++] = v;   /* adj[u].push_back(v)  main.cu:114 */
        col[row[v + 1]++] = u;   /* adj[v].push_back(u)  main.cu:115 */
    }
    *row_out = row; /* row[0..n] are the offsets (main.cu:119-121) */
    *col_out = col;
    return 0;
}

/* Same adjacency multiset built by all host threads (atomic counting + atomic fill), for
 * the multi-GiB bench inputs where the serial walk above takes minutes.  The order
 * inside a row is then arbitrary; no hop distance or distance sum depends on it. */
int oracle_build_csr_parallel(int32_t n, int64_t m, const int32_t* uv, int64_t** row_out, int32_t** col_out) {
    if (n < 0 || m < 0) return -1;
    int64_t* row = (int64_t*)calloc((size_t)n + 2, sizeof(int64_t));
    int32_t* col = (int32_t*)malloc(sizeof(int32_t) * (size_t)(2 * m > 0 ? 2 * m : 1));
    int64_t* cur = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    if (!row || !col || !cur) { free(row); free(col); free(cur); return -2; }
    int bad = 0;
#pragma omp parallel for schedule(static) reduction(|:bad)
    for (int64_t i = 0; i < m; ++i) {
        int32_t u = uv[2 * i], v = uv[2 * i + 1];
        if (u < 0 || u >= n || v < 0 || v >= n) { bad = 1; continue; }
#pragma omp atomic
        row[u + 1]++;
#pragma omp atomic
        row[v + 1]++;
    }
    if (bad) { free(row); free(col); free(cur); return -3; }
    for (int32_t i = 0; i < n; ++i) row[i + 1] += row[i];
#pragma omp parallel for schedule(static)
    for (int32_t i = 0; i < n; ++i) cur[i] = row[i];
#pragma omp parallel for schedule(static)
    for (int64_t i = 0; i < m; ++i) {
        int32_t u = uv[2 * i], v = uv[2 * i + 1];
        int64_t pu, pv;
#pragma omp atomic capture
        pu = cur[u]++;
#pragma omp atomic capture
        pv = cur[v]++;
        col[pu] = v;
        col[pv] = u;
    }
    free(cur);
    *row_out = row;
    *col_out = col;
    return 0;
}

/* main.cu:92-117 — header int32 n, int64 m, then m x (int32 u, int32 v). */
int oracle_load_graph_bin(const char* path, int32_t* n_out, int64_t* m_out, int32_t** uv_out) {
    FILE* f = fopen(path, "rb");
    if (!f) return -1; /* reference: "Could not open graph file", exit(1) (main.cu:95-99) */
    int32_t n = 0; int64_t m = 0;
    if (fread(&n, sizeof n, 1, f) != 1 || fread(&m, sizeof m, 1, f) != 1 || n < 0 || m < 0) { fclose(f); return -2; }
    int32_t* uv = (int32_t*)malloc(sizeof(int32_t) * (size_t)(2 * m > 0 ? 2 * m : 1));
    if (!uv) { fclose(f); return -3; }
    if ((int64_t)fread(uv, 2 * sizeof(int32_t), (size_t)m, f) != m) { free(uv); fclose(f); return -2; }
    fclose(f);
    *n_out = n; *m_out = m; *uv_out = uv;
    return 0;
}

/* main.cu:134-164 — uint8 K, then per group uint8 S + S x int32.
 * Extension: first byte 0 followed by "MSBFSQ2" switches to
 * uint32 K, then per group uint32 S + S x int32.  A lone 0x00 byte stays K = 0. */
int oracle_load_query_bin(const char* path, int32_t* k_out, int64_t** offsets_out, int32_t** sources_out) {
    FILE* f = fopen(path, "rb");
    if (!f) return -1; /* reference: "Could not open query file", exit(1) (main.cu:137-141) */
    unsigned char k8 = 0;
    if (fread(&k8, 1, 1, f) != 1) { fclose(f); return -2; }
    int64_t K = k8;
    int extended = 0;
    if (k8 == 0) {
        char magic[7];
        if (fread(magic, 1, 7, f) == 7 && memcmp(magic, "MSBFSQ2", 7) == 0) {
            uint32_t k32 = 0;
            if (fread(&k32, 4, 1, f) != 1) { fclose(f); return -2; }
            K = k32; extended = 1;
        }
    }
    int64_t* offs = (int64_t*)calloc((size_t)K + 1, sizeof(int64_t));
    size_t cap = 1024, cnt = 0;
    int32_t* src = (int32_t*)malloc(cap * sizeof(int32_t));
    if (!offs || !src) { fclose(f); free(offs); free(src); return -3; }
    for (int64_t i = 0; i < K; ++i) {
        uint32_t S = 0;
        if (extended) { if (fread(&S, 4, 1, f) != 1) S = 0; }
        else { unsigned char s8 = 0; if (fread(&s8, 1, 1, f) != 1) s8 = 0; S = s8; }
        if (cnt + S > cap) {
            while (cnt + S > cap) cap *= 2;
            int32_t* bigger = (int32_t*)realloc(src, cap * sizeof(int32_t));
            if (!bigger) { fclose(f); free(offs); free(src); return -3; }
            src = bigger;
        }
        size_t got = fread(src + cnt, sizeof(int32_t), S, f);
        for (size_t j = got; j < S; ++j) src[cnt + j] = -1; /* short file: reference reads garbage; we ignore */
        cnt += S;
        offs[i + 1] = (int64_t)cnt;
    }
    fclose(f);
    *k_out = (int32_t)K; *offsets_out = offs; *sources_out = src;
    return 0;
}

/* ------------------------------------------------------------------ BFS --- */

/* main.cu:40-73 + 16-38 — distances start at -1, every source id in [0,n) is set
 * to 0 (others silently ignored, main.cu:49), then level-synchronous relaxation
 * until a level changes nothing.  Level-synchronous relaxation of an unweighted
 * graph yields exact hop distances, so a frontier queue gives the same array.
 * `queue` is caller scratch of n int32. */
static void bfs_group(int32_t n, const int64_t* row, const int32_t* col,
                      const int32_t* sources, int64_t nsrc, int32_t* dist, int32_t* queue) {
    memset(dist, 0xFF, sizeof(int32_t) * (size_t)n);           /* -1 everywhere (main.cu:44) */
    int64_t head = 0, tail = 0;
    for (int64_t i = 0; i < nsrc; ++i) {
        int32_t s = sources[i];
        if (s >= 0 && s < n && dist[s] != 0) { dist[s] = 0; queue[tail++] = s; } /* main.cu:46-51 */
    }
    while (head < tail) {
        int32_t v = queue[head++];
        int32_t next_level = dist[v] + 1;
        for (int64_t e = row[v]; e < row[v + 1]; ++e) {         /* main.cu:26-35 */
            int32_t u = col[e];
            if (dist[u] == -1) { dist[u] = next_level; queue[tail++] = u; }
        }
    }
}

/* main.cu:75-89 — sum of the non-negative distances (unreachable = -1 skipped).
 * Also returns the Graph500-style traversed-edge count: half the CSR entries of
 * reached vertices; that is measurement metadata, not part of
 * the reference's output. */
static int64_t sum_distances(int32_t n, const int32_t* dist, const int64_t* row, int64_t* entries_out) {
    int64_t sum = 0, entries = 0;
    for (int32_t i = 0; i < n; ++i) {
        if (dist[i] < 0) continue;
        sum += dist[i];
        entries += row[i + 1] - row[i];
    }
    if (entries_out) *entries_out = entries;
    return sum;
}

int oracle_group_distances(int32_t n, const int64_t* row, const int32_t* col,
                           const int32_t* sources, int64_t nsrc, int32_t* dist_out) {
    int32_t* queue = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
    if (!queue) return -2;
    bfs_group(n, row, col, sources, nsrc, dist_out, queue);
    free(queue);
    return 0;
}

/* main.cu:312-322 — F(U_k) for every group.  The reference walks groups serially
 * per rank; groups are independent, so here they run one per OpenMP thread. */
int oracle_run(int32_t n, const int64_t* row, const int32_t* col,
               int32_t K, const int64_t* offsets, const int32_t* sources,
               int64_t* sums_out, int64_t* reached_entries_out, int32_t nthreads) {
    int err = 0;
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
#pragma omp parallel num_threads(nthreads)
    {
        int32_t* dist = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
        int32_t* queue = (int32_t*)malloc(sizeof(int32_t) * (size_t)(n > 0 ? n : 1));
        if (!dist || !queue) {
#pragma omp atomic write
            err = -2;
        } else {
#pragma omp for schedule(dynamic, 1)
            for (int32_t k = 0; k < K; ++k) {
                bfs_group(n, row, col, sources + offsets[k], offsets[k + 1] - offsets[k], dist, queue);
                int64_t entries = 0;
                sums_out[k] = sum_distances(n, dist, row, &entries);
                if (reached_entries_out) reached_entries_out[k] = entries;
            }
        }
        free(dist); free(queue);
    }
    return err;
}

/* ------------------------------------------------- one BFS on all threads --- */

/* The same distances as bfs_group, computed the way the reference's kernel does it
 * (main.cu:16-38, 61-71): level by level, every frontier vertex relaxing its row, a
 * vertex taken by whichever thread reaches it first (all writers store the same level,
 * main.cu:30-33).  Host threads take the place of the kernel's threads; a compare-and-swap
 * claims the vertex so that it enters the next frontier once.  Used where one BFS per
 * thread (oracle_run) would make a single timing step last half a minute: Kronecker-26
 * under bench.py --impl reference.  Returns F and the reached CSR entries. */
static int64_t bfs_group_wide(int32_t n, const int64_t* row, const int32_t* col, const int32_t* sources, int64_t nsrc,
                              int32_t* dist, int32_t* cur, int32_t* nxt, int nthreads, int64_t* entries_out) {
    int64_t ncur = 0;
#pragma omp parallel for schedule(static) num_threads(nthreads)
    for (int32_t i = 0; i < n; ++i) dist[i] = -1;                /* main.cu:44 */
    for (int64_t i = 0; i < nsrc; ++i) {
        int32_t s = sources[i];
        if (s >= 0 && s < n && dist[s] != 0) { dist[s] = 0; cur[ncur++] = s; } /* main.cu:46-51 */
    }
    for (int32_t level = 0; ncur > 0; ++level) {                 /* main.cu:61-71 */
        int64_t nnext = 0;
#pragma omp parallel num_threads(nthreads)
        {
            enum { LOCAL = 4096 };
            int32_t buf[LOCAL];
            int cnt = 0;
#pragma omp for schedule(dynamic, 64) nowait
            for (int64_t i = 0; i < ncur; ++i) {
                const int32_t v = cur[i];
                for (int64_t e = row[v]; e < row[v + 1]; ++e) {  /* main.cu:26-35 */
                    const int32_t u = col[e];
                    if (dist[u] == -1 && __sync_bool_compare_and_swap(&dist[u], -1, level + 1)) {
                        buf[cnt++] = u;
                        if (cnt == LOCAL) {
                            const int64_t at = __sync_fetch_and_add(&nnext, (int64_t)cnt);
                            memcpy(nxt + at, buf, sizeof(int32_t) * (size_t)cnt);
                            cnt = 0;
                        }
                    }
                }
            }
            if (cnt) {
                const int64_t at = __sync_fetch_and_add(&nnext, (int64_t)cnt);
                memcpy(nxt + at, buf, sizeof(int32_t) * (size_t)cnt);
            }
        }
        int32_t* t = cur; cur = nxt; nxt = t;
        ncur = nnext;
    }
    int64_t sum = 0, entries = 0;                                /* main.cu:81-87 */
#pragma omp parallel for schedule(static) reduction(+:sum, entries) num_threads(nthreads)
    for (int32_t i = 0; i < n; ++i) {
        if (dist[i] < 0) continue;
        sum += dist[i];
        entries += row[i + 1] - row[i];
    }
    if (entries_out) *entries_out = entries;
    return sum;
}

/* main.cu:312-322 with the groups walked one after the other (as a rank of the reference
 * does) and all host threads inside each BFS. */
int oracle_run_wide(int32_t n, const int64_t* row, const int32_t* col,
                    int32_t K, const int64_t* offsets, const int32_t* sources,
                    int64_t* sums_out, int64_t* reached_entries_out, int32_t nthreads) {
#ifdef _OPENMP
    if (nthreads <= 0) nthreads = omp_get_max_threads();
#else
    nthreads = 1;
#endif
    const size_t cap = (size_t)(n > 0 ? n : 1);
    int32_t* dist = (int32_t*)malloc(sizeof(int32_t) * cap);
    int32_t* cur = (int32_t*)malloc(sizeof(int32_t) * cap);
    int32_t* nxt = (int32_t*)malloc(sizeof(int32_t) * cap);
    if (!dist || !cur || !nxt) { free(dist); free(cur); free(nxt); return -2; }
    for (int32_t k = 0; k < K; ++k) {
        int64_t entries = 0;
        sums_out[k] = bfs_group_wide(n, row, col, sources + offsets[k], offsets[k + 1] - offsets[k], dist, cur, nxt,
                                     nthreads, &entries);
        if (reached_entries_out) reached_entries_out[k] = entries;
    }
    free(dist); free(cur); free(nxt);
    return 0;
}

/* main.cu:379-397 — minF/minK start at -1; first F >= 0 seeds them, then a strict
 * '<' scan keeps the lowest index among equal minima.  K = 0 leaves (-1, -1); the
 * report prints minK + 1 (main.cu:409). */
void oracle_argmin(int32_t K, const int64_t* sums, int32_t* min_k_out, int64_t* min_f_out) {
    int64_t minF = -1; int32_t minK = -1;
    for (int32_t i = 0; i < K; ++i) {
        if (sums[i] >= 0) { minF = sums[i]; minK = i; break; }
    }
    for (int32_t i = 0; i < K; ++i) {
        if (sums[i] < minF && sums[i] >= 0) { minF = sums[i]; minK = i; }
    }
    *min_k_out = minK; *min_f_out = minF;
}

/* Launchers such as torchrun export OMP_NUM_THREADS=1; the timing legs of bench.py state
 * their own thread count instead of inheriting that. */
void oracle_set_threads(int nthreads) {
#ifdef _OPENMP
    if (nthreads > 0) omp_set_num_threads(nthreads);
#else
    (void)nthreads;
#endif
}

int oracle_num_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

void oracle_free(void* p) { free(p); }
