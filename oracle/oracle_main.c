/*
 * oracle_main.c — TEST INFRASTRUCTURE.  Command-line face of the CPU oracle:
 * same flags and the same seven report lines as the reference
 * (/root/reference/main.cu:203-224, 403-414) so outputs can be diffed textually.
 * Extra flag --dump-sums <file> writes one "k F" line per group.
 */
#include "oracle.h"

#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

static double now_s(void) {
    struct timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int main(int argc, char** argv) {
    if (argc < 5) { /* main.cu:204-212 */
        fprintf(stderr, "Usage: mpirun -np <ranks> %s -g <graph.bin> -q <query.bin> -gn <numGPU>\n", argv[0]);
        return -1;
    }
    const char *g = "", *q = "", *dump = NULL; int gn = 1;
    for (int i = 1; i < argc; ++i) { /* main.cu:216-224 */
        if (!strcmp(argv[i], "-g") && i + 1 < argc) g = argv[++i];
        else if (!strcmp(argv[i], "-q") && i + 1 < argc) q = argv[++i];
        else if (!strcmp(argv[i], "-gn") && i + 1 < argc) gn = atoi(argv[++i]);
        else if (!strcmp(argv[i], "--dump-sums") && i + 1 < argc) dump = argv[++i];
    }
    double t0 = now_s();
    int32_t n; int64_t m; int32_t* uv;
    if (oracle_load_graph_bin(g, &n, &m, &uv) != 0) { fprintf(stderr, "Could not open graph file %s\n", g); return 1; }
    int64_t* row; int32_t* col;
    if (oracle_build_csr(n, m, uv, &row, &col) != 0) { fprintf(stderr, "bad graph file %s\n", g); return 1; }
    free(uv);
    int32_t K; int64_t* offs; int32_t* src;
    if (oracle_load_query_bin(q, &K, &offs, &src) != 0) { fprintf(stderr, "Could not open query file %s\n", q); return 1; }
    double t1 = now_s();
    int64_t* sums = (int64_t*)malloc(sizeof(int64_t) * (size_t)(K > 0 ? K : 1));
    oracle_run(n, row, col, K, offs, src, sums, NULL, 0);
    int32_t minK; int64_t minF;
    oracle_argmin(K, sums, &minK, &minF);
    double t2 = now_s();
    printf("Graph: %s\n", g);
    printf("Query: %s\n", q);
    printf("Query number (k) with minimum F value: %d\n", minK + 1);
    printf("Minimum F value: %lld\n", (long long)minF);
    printf("GPU # : %d GPU\n", gn);
    printf("Preprocessing time: %.9f s\n", t1 - t0);
    printf("Computation time: %.9f s\n", t2 - t1);
    if (dump) {
        FILE* f = fopen(dump, "w");
        if (f) { for (int32_t k = 0; k < K; ++k) fprintf(f, "%d %lld\n", k, (long long)sums[k]); fclose(f); }
    }
    free(sums); free(row); free(col); free(offs); free(src);
    return 0;
}
