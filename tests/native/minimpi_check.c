/* Test program for the bundled MPI subset (host/minimpi): every collective the new host
 * program and the unmodified reference use, with payloads larger than the shared window
 * (run with MINIMPI_WINDOW_MB=1) and with ranks that contribute nothing. */
#include <mpi.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define CHECK(cond) do { if (!(cond)) { fprintf(stderr, "rank %d: FAILED %s (line %d)\n", rank, #cond, __LINE__); MPI_Abort(MPI_COMM_WORLD, 3); } } while (0)

typedef struct { int q; long long fv; } Pair; /* main.cu:166-170 */

int main(int argc, char** argv) {
    int rank = 0, size = 1;
    MPI_Init(&argc, &argv);
    MPI_Comm_size(MPI_COMM_WORLD, &size);
    MPI_Comm_rank(MPI_COMM_WORLD, &rank);

    /* Bcast of 3 MiB through a 1 MiB window, from a non-zero root when there is one */
    const int n = 3 * 262144 + 17, root = size > 1 ? 1 : 0;
    int* buf = (int*)malloc(sizeof(int) * (size_t)n);
    for (int i = 0; i < n; ++i) buf[i] = rank == root ? i * 7 + 1 : -1;
    MPI_Bcast(buf, n, MPI_INT, root, MPI_COMM_WORLD);
    for (int i = 0; i < n; i += 4099) CHECK(buf[i] == i * 7 + 1);
    CHECK(buf[n - 1] == (n - 1) * 7 + 1);
    int zero_count = 0;
    MPI_Bcast(buf, zero_count, MPI_INT, 0, MPI_COMM_WORLD); /* empty groups are broadcast with count 0 (main.cu:273) */

    /* Gather + Gatherv of a struct datatype, rank r sends r % 3 pairs (some send none) */
    MPI_Datatype pair_t;
    {
        int bl[2] = {1, 1};
        MPI_Aint disp[2], base;
        Pair tmp;
        MPI_Get_address(&tmp, &base);
        MPI_Get_address(&tmp.q, &disp[0]);
        MPI_Get_address(&tmp.fv, &disp[1]);
        disp[0] -= base; disp[1] -= base;
        MPI_Datatype types[2] = {MPI_INT, MPI_LONG_LONG};
        MPI_Type_create_struct(2, bl, disp, types, &pair_t);
        MPI_Type_commit(&pair_t);
    }
    int mine = rank % 3;
    Pair send[3];
    for (int i = 0; i < mine; ++i) { send[i].q = rank * 10 + i; send[i].fv = 1000000000000LL + rank * 100 + i; }
    int* counts = (int*)calloc((size_t)size, sizeof(int));
    int* displs = (int*)calloc((size_t)size, sizeof(int));
    MPI_Gather(&mine, 1, MPI_INT, counts, 1, MPI_INT, 0, MPI_COMM_WORLD);
    int total = 0;
    if (rank == 0) for (int r = 0; r < size; ++r) { CHECK(counts[r] == r % 3); displs[r] = total; total += counts[r]; }
    Pair* all = (Pair*)calloc((size_t)(total > 0 ? total : 1), sizeof(Pair));
    MPI_Gatherv(send, mine, pair_t, all, rank == 0 ? counts : NULL, rank == 0 ? displs : NULL, pair_t, 0, MPI_COMM_WORLD);
    if (rank == 0)
        for (int r = 0; r < size; ++r)
            for (int i = 0; i < counts[r]; ++i) {
                CHECK(all[displs[r] + i].q == r * 10 + i);
                CHECK(all[displs[r] + i].fv == 1000000000000LL + r * 100 + i);
            }
    MPI_Type_free(&pair_t);

    /* Allreduce: packed unsigned min (the argmin word), long long sum, int max */
    unsigned long long packed = ((unsigned long long)(50 - (rank % 4)) << 24) | (unsigned)rank;
    MPI_Allreduce(MPI_IN_PLACE, &packed, 1, MPI_UNSIGNED_LONG_LONG, MPI_MIN, MPI_COMM_WORLD);
    {
        unsigned long long want = ~0ull;
        for (int r = 0; r < size; ++r) { unsigned long long w = ((unsigned long long)(50 - (r % 4)) << 24) | (unsigned)r; if (w < want) want = w; }
        CHECK(packed == want);
    }
    long long s_in[2] = {rank + 1, -rank}, s_out[2];
    MPI_Allreduce(s_in, s_out, 2, MPI_LONG_LONG, MPI_SUM, MPI_COMM_WORLD);
    CHECK(s_out[0] == (long long)size * (size + 1) / 2 && s_out[1] == -(long long)size * (size - 1) / 2);
    int mx = rank;
    MPI_Allreduce(MPI_IN_PLACE, &mx, 1, MPI_INT, MPI_MAX, MPI_COMM_WORLD);
    CHECK(mx == size - 1);

    /* Allgather + Barrier */
    int* ag = (int*)malloc(sizeof(int) * 2 * (size_t)size);
    int me2[2] = {rank, rank * rank};
    MPI_Allgather(me2, 2, MPI_INT, ag, 2, MPI_INT, MPI_COMM_WORLD);
    for (int r = 0; r < size; ++r) CHECK(ag[2 * r] == r && ag[2 * r + 1] == r * r);
    for (int i = 0; i < 20; ++i) MPI_Barrier(MPI_COMM_WORLD);

    if (rank == 0) printf("minimpi OK with %d ranks\n", size);
    free(buf); free(counts); free(displs); free(all); free(ag);
    MPI_Finalize();
    return 0;
}
