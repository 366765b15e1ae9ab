/*
 * mpirun.c — launcher for minimpi: `mpirun -np <ranks> <program> [args...]`.
 * Creates the shared-memory segment, starts <ranks> copies of <program> with
 * MINIMPI_RANK / MINIMPI_SIZE / MINIMPI_SHM set, waits for them, and — like a
 * real mpirun — tears the job down and returns the failing status as soon as a
 * rank exits non-zero (the reference relies on that: a rank that cannot open its
 * input calls exit(1) while its peers sit in MPI_Bcast, main.cu:97-98, 242).
 */
#define _GNU_SOURCE
#include <errno.h>
#include <fcntl.h>
#include <signal.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#define MINIMPI_HDR_BYTES 16384

static pid_t* g_pids = NULL;
static int g_np = 0;
static char g_shm[64];

static void kill_all(int sig) {
    for (int i = 0; i < g_np; ++i)
        if (g_pids[i] > 0) kill(g_pids[i], sig);
}

static void on_signal(int sig) {
    kill_all(SIGTERM);
    shm_unlink(g_shm);
    _exit(128 + sig);
}

int main(int argc, char** argv) {
    int np = 1, i = 1;
    for (; i < argc; ++i) {
        if (!strcmp(argv[i], "-np") || !strcmp(argv[i], "-n") || !strcmp(argv[i], "--np")) {
            if (i + 1 >= argc) { fprintf(stderr, "mpirun: %s needs a value\n", argv[i]); return 2; }
            np = atoi(argv[++i]);
        } else if (!strcmp(argv[i], "--oversubscribe") || !strcmp(argv[i], "--allow-run-as-root")) {
            /* accepted for command-line compatibility; nothing to do */
        } else if (!strcmp(argv[i], "--bind-to") || !strcmp(argv[i], "--map-by")) {
            ++i;
        } else {
            break;
        }
    }
    if (i >= argc || np < 1 || np > 1024) {
        fprintf(stderr, "usage: mpirun -np <ranks> <program> [args...]\n");
        return 2;
    }

    size_t window_mb = 64;
    const char* w = getenv("MINIMPI_WINDOW_MB");
    if (w && atoi(w) > 0) window_mb = (size_t)atoi(w);
    size_t window = window_mb << 20;
    size_t total = MINIMPI_HDR_BYTES + window;

    snprintf(g_shm, sizeof g_shm, "/minimpi_%d", (int)getpid());
    int fd = shm_open(g_shm, O_CREAT | O_EXCL | O_RDWR, 0600);
    if (fd < 0) { perror("mpirun: shm_open"); return 1; }
    if (ftruncate(fd, (off_t)total) != 0) { perror("mpirun: ftruncate"); shm_unlink(g_shm); return 1; }
    char* p = (char*)mmap(NULL, MINIMPI_HDR_BYTES, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (p == MAP_FAILED) { perror("mpirun: mmap"); shm_unlink(g_shm); return 1; }
    memset(p, 0, MINIMPI_HDR_BYTES);
    /* header: 3 x int32 atomics, int32 world_size, uint64 window_bytes (see minimpi.c) */
    ((int32_t*)p)[3] = np;
    ((uint64_t*)p)[2] = (uint64_t)window;
    munmap(p, MINIMPI_HDR_BYTES);
    close(fd);

    g_np = np;
    g_pids = (pid_t*)calloc((size_t)np, sizeof(pid_t));
    signal(SIGINT, on_signal);
    signal(SIGTERM, on_signal);

    for (int r = 0; r < np; ++r) {
        pid_t pid = fork();
        if (pid < 0) { perror("mpirun: fork"); kill_all(SIGTERM); shm_unlink(g_shm); return 1; }
        if (pid == 0) {
            char buf[32];
            snprintf(buf, sizeof buf, "%d", r);  setenv("MINIMPI_RANK", buf, 1);
            snprintf(buf, sizeof buf, "%d", np); setenv("MINIMPI_SIZE", buf, 1);
            setenv("MINIMPI_SHM", g_shm, 1);
            execvp(argv[i], &argv[i]);
            fprintf(stderr, "mpirun: cannot execute %s: %s\n", argv[i], strerror(errno));
            _exit(127);
        }
        g_pids[r] = pid;
    }

    /* Like a real mpirun, a rank that exits non-zero ends the job with that status.  The
     * peers get a short grace period first: when every rank is leaving on its own (the usage
     * path: all return -1, rank 0 prints) nobody must be cut off mid-message; ranks that are
     * stuck in a collective waiting for the dead rank are terminated after it. */
    int rc = 0, alive = np;
    double deadline = 0.0;
    while (alive > 0) {
        int status = 0;
        pid_t pid = waitpid(-1, &status, rc ? WNOHANG : 0);
        if (pid < 0) { if (errno == EINTR) continue; break; }
        if (pid == 0) { /* after a failure: poll until the grace period is over */
            struct timespec ts;
            clock_gettime(CLOCK_MONOTONIC, &ts);
            const double now = (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
            if (now >= deadline) { kill_all(SIGTERM); deadline = now + 3600.0; }
            struct timespec nap = {0, 2000000};
            nanosleep(&nap, NULL);
            continue;
        }
        for (int r = 0; r < np; ++r) if (g_pids[r] == pid) g_pids[r] = 0;
        --alive;
        int code = WIFEXITED(status) ? WEXITSTATUS(status) : 128 + WTERMSIG(status);
        if (code != 0 && rc == 0) {
            rc = code;
            struct timespec ts;
            clock_gettime(CLOCK_MONOTONIC, &ts);
            deadline = (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec + 0.5;
        }
    }
    shm_unlink(g_shm);
    free(g_pids);
    return rc;
}
