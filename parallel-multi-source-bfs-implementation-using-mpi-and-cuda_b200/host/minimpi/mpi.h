/*
 * minimpi — the subset of the MPI C API that the multi-source-BFS host program
 * (and the unmodified reference, /root/reference/main.cu:197-420) uses, for
 * boxes where no MPI is installed.  Ranks are processes started by the bundled
 * `mpirun` launcher on ONE node and talk through a POSIX shared-memory window.
 *
 * Written against the standard API: building with a real MPI (mpicxx) needs no
 * source change in the callers — just drop this directory from the include path.
 */
#ifndef MINIMPI_MPI_H
#define MINIMPI_MPI_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef int       MPI_Comm;
typedef int       MPI_Datatype;
typedef int       MPI_Op;
typedef ptrdiff_t MPI_Aint;

#define MPI_COMM_WORLD 0
#define MPI_SUCCESS    0
#define MPI_ERR_OTHER  15

/* predefined datatypes (handle = small int; size table in minimpi.c) */
#define MPI_DATATYPE_NULL      0
#define MPI_CHAR               1
#define MPI_BYTE               2
#define MPI_UNSIGNED_CHAR      3
#define MPI_INT                4
#define MPI_UNSIGNED           5
#define MPI_LONG_LONG          6
#define MPI_LONG_LONG_INT      6
#define MPI_UNSIGNED_LONG_LONG 7
#define MPI_UINT64_T           7
#define MPI_INT64_T            6
#define MPI_INT32_T            4
#define MPI_FLOAT              8
#define MPI_DOUBLE             9
#define MPI_LONG               10
#define MINIMPI_FIRST_DERIVED  32

#define MPI_MAX 1
#define MPI_MIN 2
#define MPI_SUM 3

#define MPI_IN_PLACE ((void*)1)

int MPI_Init(int* argc, char*** argv);
int MPI_Finalize(void);
int MPI_Abort(MPI_Comm comm, int errorcode);
int MPI_Comm_size(MPI_Comm comm, int* size);
int MPI_Comm_rank(MPI_Comm comm, int* rank);
int MPI_Barrier(MPI_Comm comm);
double MPI_Wtime(void);

int MPI_Bcast(void* buffer, int count, MPI_Datatype datatype, int root, MPI_Comm comm);
int MPI_Gather(const void* sendbuf, int sendcount, MPI_Datatype sendtype,
               void* recvbuf, int recvcount, MPI_Datatype recvtype, int root, MPI_Comm comm);
int MPI_Gatherv(const void* sendbuf, int sendcount, MPI_Datatype sendtype,
                void* recvbuf, const int* recvcounts, const int* displs,
                MPI_Datatype recvtype, int root, MPI_Comm comm);
int MPI_Allgather(const void* sendbuf, int sendcount, MPI_Datatype sendtype,
                  void* recvbuf, int recvcount, MPI_Datatype recvtype, MPI_Comm comm);
int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype datatype,
                  MPI_Op op, MPI_Comm comm);

int MPI_Get_address(const void* location, MPI_Aint* address);
int MPI_Type_create_struct(int count, const int* blocklengths, const MPI_Aint* displacements,
                           const MPI_Datatype* types, MPI_Datatype* newtype);
int MPI_Type_commit(MPI_Datatype* datatype);
int MPI_Type_free(MPI_Datatype* datatype);
int MPI_Type_size(MPI_Datatype datatype, int* size);

#ifdef __cplusplus
}
#endif

#endif /* MINIMPI_MPI_H */
