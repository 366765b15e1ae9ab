/*
 * oracle.h — TEST INFRASTRUCTURE (see oracle.c).  C interface of the CPU
 * restatement of /root/reference/main.cu used as the parity checker.
 */
#ifndef MSBFS_ORACLE_H
#define MSBFS_ORACLE_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* main.cu:106-129 */
int oracle_build_csr(int32_t n, int64_t m, const int32_t* uv, int64_t** row_out, int32_t** col_out);
/* same multiset, all host threads, row order arbitrary (large bench inputs only) */
int oracle_build_csr_parallel(int32_t n, int64_t m, const int32_t* uv, int64_t** row_out, int32_t** col_out);
/* main.cu:92-117 (returns the raw records; pair with oracle_build_csr) */
int oracle_load_graph_bin(const char* path, int32_t* n_out, int64_t* m_out, int32_t** uv_out);
/* main.cu:134-164 (+ extended container) */
int oracle_load_query_bin(const char* path, int32_t* k_out, int64_t** offsets_out, int32_t** sources_out);
/* main.cu:40-73: distances of one group (dist_out: n int32, -1 = unreachable) */
int oracle_group_distances(int32_t n, const int64_t* row, const int32_t* col,
                           const int32_t* sources, int64_t nsrc, int32_t* dist_out);
/* main.cu:312-322: F of every group; reached_entries_out (optional) = CSR entries of reached vertices */
int oracle_run(int32_t n, const int64_t* row, const int32_t* col,
               int32_t K, const int64_t* offsets, const int32_t* sources,
               int64_t* sums_out, int64_t* reached_entries_out, int32_t nthreads);
/* same results, groups one after the other, all host threads inside each BFS (level-synchronous,
 * as the reference's kernel, main.cu:16-38) */
int oracle_run_wide(int32_t n, const int64_t* row, const int32_t* col,
                    int32_t K, const int64_t* offsets, const int32_t* sources,
                    int64_t* sums_out, int64_t* reached_entries_out, int32_t nthreads);
/* main.cu:379-397 */
void oracle_argmin(int32_t K, const int64_t* sums, int32_t* min_k_out, int64_t* min_f_out);
void oracle_set_threads(int nthreads); /* OpenMP team size for the calls that take no count */
int oracle_num_threads(void);
void oracle_free(void* p);

#ifdef __cplusplus
}
#endif
#endif
