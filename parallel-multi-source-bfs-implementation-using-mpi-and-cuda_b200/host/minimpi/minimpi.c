/*
 * minimpi.c — single-node MPI subset over a POSIX shared-memory window.
 * See mpi.h.  Collectives used by the reference (/root/reference/main.cu:242-278
 * Bcast, 340-365 Gather/Gatherv, 181-191 struct datatype) plus Allreduce /
 * Allgather / Barrier for the new host program.
 *
 * Layout of the segment created by `mpirun`:
 *   [ header page(s) | data window ]
 * The window is cut into `size` equal slots for gathers; broadcasts use all of it.
 * Synchronisation is a sense-reversing barrier on C11 atomics in the header.
 * Without the launcher's environment (MINIMPI_SHM unset) the process is a
 * singleton world: size 1, rank 0, every collective is a local copy.
 */
#define _GNU_SOURCE
#include "mpi.h"

#include <errno.h>
#include <fcntl.h>
#include <sched.h>
#include <stdatomic.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#define MINIMPI_MAX_RANKS 1024
#define MINIMPI_HDR_BYTES 16384

typedef struct {
    atomic_int barrier_count;
    atomic_int barrier_sense;
    atomic_int abort_flag;
    int        world_size;
    uint64_t   window_bytes;
    uint64_t   lens[MINIMPI_MAX_RANKS];
} minimpi_hdr;

static minimpi_hdr* g_hdr  = NULL;
static char*        g_win  = NULL;
static size_t       g_win_bytes = 0;
static int          g_rank = 0, g_size = 1, g_sense = 0, g_inited = 0;

/* derived datatypes: only the extent matters (data moves as raw bytes) */
typedef struct { int used; size_t size; size_t extent; } minimpi_dtype;
static minimpi_dtype g_types[64];

static size_t type_extent(MPI_Datatype t) {
    switch (t) {
        case MPI_CHAR: case MPI_BYTE: case MPI_UNSIGNED_CHAR: return 1;
        case MPI_INT: case MPI_UNSIGNED: case MPI_FLOAT: return 4;
        case MPI_LONG_LONG: case MPI_UNSIGNED_LONG_LONG: case MPI_DOUBLE: case MPI_LONG: return 8;
        default: break;
    }
    if (t >= MINIMPI_FIRST_DERIVED && t < MINIMPI_FIRST_DERIVED + 64 && g_types[t - MINIMPI_FIRST_DERIVED].used)
        return g_types[t - MINIMPI_FIRST_DERIVED].extent;
    fprintf(stderr, "minimpi: unknown datatype %d\n", t);
    abort();
}

static void die_if_aborted(void) {
    if (g_hdr && atomic_load(&g_hdr->abort_flag)) _exit(1);
}

static void barrier(void) {
    if (g_size == 1) return;
    g_sense = !g_sense;
    if (atomic_fetch_add(&g_hdr->barrier_count, 1) == g_size - 1) {
        atomic_store(&g_hdr->barrier_count, 0);
        atomic_store(&g_hdr->barrier_sense, g_sense);
    } else {
        unsigned spins = 0;
        while (atomic_load(&g_hdr->barrier_sense) != g_sense) {
            ++spins;
            if (spins < 2000) {
#if defined(__x86_64__)
                __builtin_ia32_pause();
#endif
            } else if (spins < 20000) {
                sched_yield();
            } else {
                struct timespec ts = {0, 50000};
                nanosleep(&ts, NULL);
                die_if_aborted();
                if (getppid() == 1) _exit(1); /* launcher died */
            }
        }
    }
}

int MPI_Init(int* argc, char*** argv) {
    (void)argc; (void)argv;
    const char* shm = getenv("MINIMPI_SHM");
    const char* r = getenv("MINIMPI_RANK");
    const char* s = getenv("MINIMPI_SIZE");
    g_inited = 1;
    if (!shm || !r || !s) { g_rank = 0; g_size = 1; return MPI_SUCCESS; }
    g_rank = atoi(r);
    g_size = atoi(s);
    if (g_size <= 1) { g_size = 1; g_rank = 0; return MPI_SUCCESS; }
    int fd = shm_open(shm, O_RDWR, 0600);
    if (fd < 0) { perror("minimpi: shm_open"); exit(1); }
    struct stat st;
    if (fstat(fd, &st) != 0) { perror("minimpi: fstat"); exit(1); }
    void* p = mmap(NULL, (size_t)st.st_size, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (p == MAP_FAILED) { perror("minimpi: mmap"); exit(1); }
    close(fd);
    g_hdr = (minimpi_hdr*)p;
    g_win = (char*)p + MINIMPI_HDR_BYTES;
    g_win_bytes = (size_t)g_hdr->window_bytes;
    return MPI_SUCCESS;
}

int MPI_Finalize(void) { g_inited = 0; return MPI_SUCCESS; }

int MPI_Abort(MPI_Comm comm, int errorcode) {
    (void)comm;
    if (g_hdr) atomic_store(&g_hdr->abort_flag, 1);
    fflush(NULL);
    _exit(errorcode ? errorcode : 1);
}

int MPI_Comm_size(MPI_Comm comm, int* size) { (void)comm; *size = g_size; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm comm, int* rank) { (void)comm; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm comm) { (void)comm; barrier(); return MPI_SUCCESS; }

double MPI_Wtime(void) {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int MPI_Bcast(void* buffer, int count, MPI_Datatype datatype, int root, MPI_Comm comm) {
    (void)comm;
    if (g_size == 1 || count <= 0) { if (g_size > 1) { barrier(); } return MPI_SUCCESS; }
    size_t bytes = (size_t)count * type_extent(datatype);
    char* buf = (char*)buffer;
    for (size_t off = 0; off < bytes; off += g_win_bytes) {
        size_t len = bytes - off < g_win_bytes ? bytes - off : g_win_bytes;
        if (g_rank == root) memcpy(g_win, buf + off, len);
        barrier();
        if (g_rank != root) memcpy(buf + off, g_win, len);
        barrier();
    }
    return MPI_SUCCESS;
}

/* every rank contributes `sbytes`; receivers (root, or all when root < 0) place
 * rank r's bytes at recv + roffs[r]. */
static void gather_bytes(const void* send, size_t sbytes, char* recv, const size_t* roffs, int root) {
    if (g_size == 1) {
        if (sbytes && recv + roffs[0] != (const char*)send) memmove(recv + roffs[0], send, sbytes);
        return;
    }
    uint64_t lens[MINIMPI_MAX_RANKS];
    g_hdr->lens[g_rank] = sbytes;
    barrier();
    size_t maxlen = 0;
    for (int r = 0; r < g_size; ++r) { lens[r] = g_hdr->lens[r]; if (lens[r] > maxlen) maxlen = lens[r]; }
    barrier();
    size_t slot = (g_win_bytes / (size_t)g_size) & ~(size_t)63;
    const int receiver = (root < 0 || g_rank == root);
    for (size_t off = 0; off < maxlen; off += slot) {
        if (off < sbytes) {
            size_t len = sbytes - off < slot ? sbytes - off : slot;
            memcpy(g_win + (size_t)g_rank * slot, (const char*)send + off, len);
        }
        barrier();
        if (receiver) {
            for (int r = 0; r < g_size; ++r) {
                if (off < lens[r]) {
                    size_t len = lens[r] - off < slot ? lens[r] - off : slot;
                    memcpy(recv + roffs[r] + off, g_win + (size_t)r * slot, len);
                }
            }
        }
        barrier();
    }
}

int MPI_Gather(const void* sendbuf, int sendcount, MPI_Datatype sendtype,
               void* recvbuf, int recvcount, MPI_Datatype recvtype, int root, MPI_Comm comm) {
    (void)comm;
    size_t roffs[MINIMPI_MAX_RANKS];
    size_t sbytes = (size_t)sendcount * type_extent(sendtype);
    if (g_rank == root) {
        size_t rb = (size_t)recvcount * type_extent(recvtype);
        for (int r = 0; r < g_size; ++r) roffs[r] = (size_t)r * rb;
    }
    gather_bytes(sendbuf, sbytes, (char*)recvbuf, roffs, root);
    return MPI_SUCCESS;
}

int MPI_Gatherv(const void* sendbuf, int sendcount, MPI_Datatype sendtype,
                void* recvbuf, const int* recvcounts, const int* displs,
                MPI_Datatype recvtype, int root, MPI_Comm comm) {
    (void)comm; (void)recvcounts;
    size_t roffs[MINIMPI_MAX_RANKS];
    size_t sbytes = (size_t)sendcount * type_extent(sendtype);
    if (g_rank == root) {
        size_t ext = type_extent(recvtype);
        for (int r = 0; r < g_size; ++r) roffs[r] = (size_t)displs[r] * ext;
    }
    gather_bytes(sendbuf, sbytes, (char*)recvbuf, roffs, root);
    return MPI_SUCCESS;
}

int MPI_Allgather(const void* sendbuf, int sendcount, MPI_Datatype sendtype,
                  void* recvbuf, int recvcount, MPI_Datatype recvtype, MPI_Comm comm) {
    (void)comm;
    size_t roffs[MINIMPI_MAX_RANKS];
    size_t rb = (size_t)recvcount * type_extent(recvtype);
    for (int r = 0; r < g_size; ++r) roffs[r] = (size_t)r * rb;
    const void* src = sendbuf;
    size_t sbytes = (size_t)sendcount * type_extent(sendtype);
    if (sendbuf == MPI_IN_PLACE) { src = (char*)recvbuf + roffs[g_rank]; sbytes = rb; }
    gather_bytes(src, sbytes, (char*)recvbuf, roffs, -1);
    return MPI_SUCCESS;
}

#define REDUCE_LOOP(T)                                                            \
    do {                                                                          \
        T* out = (T*)recvbuf;                                                     \
        const T* all = (const T*)tmp;                                             \
        for (int i = 0; i < count; ++i) {                                         \
            T acc = all[i];                                                       \
            for (int r = 1; r < g_size; ++r) {                                    \
                T x = all[(size_t)r * count + i];                                 \
                if (op == MPI_SUM) acc = (T)(acc + x);                            \
                else if (op == MPI_MIN) acc = x < acc ? x : acc;                  \
                else acc = x > acc ? x : acc;                                     \
            }                                                                     \
            out[i] = acc;                                                         \
        }                                                                         \
    } while (0)

int MPI_Allreduce(const void* sendbuf, void* recvbuf, int count, MPI_Datatype datatype,
                  MPI_Op op, MPI_Comm comm) {
    (void)comm;
    size_t ext = type_extent(datatype);
    size_t bytes = (size_t)count * ext;
    const void* src = (sendbuf == MPI_IN_PLACE) ? recvbuf : sendbuf;
    if (g_size == 1) { if (src != recvbuf) memcpy(recvbuf, src, bytes); return MPI_SUCCESS; }
    char* tmp = (char*)malloc(bytes * (size_t)g_size + 8);
    char* mine = (char*)malloc(bytes + 8);
    if (!tmp || !mine) { fprintf(stderr, "minimpi: out of memory\n"); MPI_Abort(comm, 1); }
    memcpy(mine, src, bytes);
    size_t roffs[MINIMPI_MAX_RANKS];
    for (int r = 0; r < g_size; ++r) roffs[r] = (size_t)r * bytes;
    gather_bytes(mine, bytes, tmp, roffs, -1);
    switch (datatype) {
        case MPI_INT:                REDUCE_LOOP(int); break;
        case MPI_UNSIGNED:           REDUCE_LOOP(unsigned); break;
        case MPI_LONG_LONG:          REDUCE_LOOP(long long); break;
        case MPI_LONG:               REDUCE_LOOP(long); break;
        case MPI_UNSIGNED_LONG_LONG: REDUCE_LOOP(unsigned long long); break;
        case MPI_DOUBLE:             REDUCE_LOOP(double); break;
        case MPI_FLOAT:              REDUCE_LOOP(float); break;
        default:
            fprintf(stderr, "minimpi: Allreduce on unsupported datatype %d\n", datatype);
            MPI_Abort(comm, 1);
    }
    free(tmp);
    free(mine);
    return MPI_SUCCESS;
}

int MPI_Get_address(const void* location, MPI_Aint* address) {
    *address = (MPI_Aint)(intptr_t)location;
    return MPI_SUCCESS;
}

int MPI_Type_create_struct(int count, const int* blocklengths, const MPI_Aint* displacements,
                           const MPI_Datatype* types, MPI_Datatype* newtype) {
    size_t ub = 0, payload = 0, align = 1;
    for (int i = 0; i < count; ++i) {
        size_t e = type_extent(types[i]);
        size_t end = (size_t)displacements[i] + e * (size_t)blocklengths[i];
        if (end > ub) ub = end;
        payload += e * (size_t)blocklengths[i];
        size_t a = e > 8 ? 8 : e;
        if (a > align) align = a;
    }
    size_t extent = (ub + align - 1) / align * align; /* trailing padding as a C struct */
    for (int k = 0; k < 64; ++k) {
        if (!g_types[k].used) {
            g_types[k].used = 1; g_types[k].size = payload; g_types[k].extent = extent;
            *newtype = MINIMPI_FIRST_DERIVED + k;
            return MPI_SUCCESS;
        }
    }
    return MPI_ERR_OTHER;
}

int MPI_Type_commit(MPI_Datatype* datatype) { (void)datatype; return MPI_SUCCESS; }

int MPI_Type_free(MPI_Datatype* datatype) {
    int k = *datatype - MINIMPI_FIRST_DERIVED;
    if (k >= 0 && k < 64) g_types[k].used = 0;
    *datatype = MPI_DATATYPE_NULL;
    return MPI_SUCCESS;
}

int MPI_Type_size(MPI_Datatype datatype, int* size) {
    int k = datatype - MINIMPI_FIRST_DERIVED;
    if (k >= 0 && k < 64 && g_types[k].used) *size = (int)g_types[k].size;
    else *size = (int)type_extent(datatype);
    return MPI_SUCCESS;
}
