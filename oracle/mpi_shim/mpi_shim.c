/* oracle/mpi_shim/mpi_shim.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 * See mpi.h.  Ranks = forked processes; one MAP_SHARED|MAP_ANONYMOUS region
 * holds a process-shared barrier and a bounce buffer that every collective is
 * chunked through.  Deterministic: MPI_Allreduce sums contributions in rank
 * order 0..P-1, left to right. */
#define _GNU_SOURCE
#include "mpi.h"
#include <pthread.h>
#include <signal.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/types.h>
#include <sys/wait.h>
#include <time.h>
#include <unistd.h>

#define SHIM_BUF_BYTES ((size_t)64 << 20)

typedef struct {
    pthread_barrier_t bar;
    int np;
    char pad[64];
    char buf[];
} shim_region;

static shim_region *g_shm = NULL;
static int g_rank = 0, g_np = 1;
static pid_t *g_kids = NULL;

static void shim_setup(int np)
{
    g_np = np < 1 ? 1 : np;
    g_shm = mmap(NULL, sizeof(shim_region) + SHIM_BUF_BYTES, PROT_READ | PROT_WRITE,
                 MAP_SHARED | MAP_ANONYMOUS, -1, 0);
    if (g_shm == MAP_FAILED) { perror("mpi_shim mmap"); exit(1); }
    pthread_barrierattr_t at;
    pthread_barrierattr_init(&at);
    pthread_barrierattr_setpshared(&at, PTHREAD_PROCESS_SHARED);
    pthread_barrier_init(&g_shm->bar, &at, (unsigned)g_np);
    g_shm->np = g_np;
    g_kids = calloc((size_t)g_np, sizeof(pid_t));
    g_rank = 0;
    fflush(NULL);
    for (int r = 1; r < g_np; ++r) {
        pid_t c = fork();
        if (c < 0) { perror("mpi_shim fork"); exit(1); }
        if (c == 0) { g_rank = r; free(g_kids); g_kids = NULL; return; }
        g_kids[r] = c;
    }
}

static void shim_reap(void)
{
    if (g_rank != 0 || !g_kids) return;
    for (int r = 1; r < g_np; ++r) {
        int status;
        if (g_kids[r] > 0) waitpid(g_kids[r], &status, 0);
    }
    free(g_kids); g_kids = NULL;
}

static void shim_teardown(void)
{
    if (g_shm) {
        if (g_rank == 0) pthread_barrier_destroy(&g_shm->bar);
        munmap(g_shm, sizeof(shim_region) + SHIM_BUF_BYTES);
        g_shm = NULL;
    }
    g_np = 1; g_rank = 0;
}

static inline void bar(void) { if (g_np > 1) pthread_barrier_wait(&g_shm->bar); }

int MPI_Init(int *argc, char ***argv)
{
    (void)argc; (void)argv;
    const char *e = getenv("CGSHIM_NP");
    shim_setup(e ? atoi(e) : 1);
    return MPI_SUCCESS;
}

int MPI_Finalize(void)
{
    bar();
    fflush(NULL);
    if (g_rank != 0) _exit(0);
    shim_reap();
    shim_teardown();
    return MPI_SUCCESS;
}

int MPI_Abort(MPI_Comm comm, int code)
{
    (void)comm;
    fflush(NULL);
    if (g_rank == 0 && g_kids)
        for (int r = 1; r < g_np; ++r) if (g_kids[r] > 0) kill(g_kids[r], SIGKILL);
    if (g_rank != 0) kill(getppid(), SIGTERM);
    _exit(code);
}

int MPI_Comm_size(MPI_Comm c, int *size) { (void)c; *size = g_np; return MPI_SUCCESS; }
int MPI_Comm_rank(MPI_Comm c, int *rank) { (void)c; *rank = g_rank; return MPI_SUCCESS; }
int MPI_Barrier(MPI_Comm c) { (void)c; bar(); return MPI_SUCCESS; }

double MPI_Wtime(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int MPI_Bcast(void *buf, int count, MPI_Datatype dt, int root, MPI_Comm c)
{
    (void)c;
    if (g_np == 1) return MPI_SUCCESS;
    size_t total = (size_t)count * (size_t)dt;
    for (size_t off = 0; off < total; off += SHIM_BUF_BYTES) {
        size_t len = total - off < SHIM_BUF_BYTES ? total - off : SHIM_BUF_BYTES;
        if (g_rank == root) memcpy(g_shm->buf, (char *)buf + off, len);
        bar();
        if (g_rank != root) memcpy((char *)buf + off, g_shm->buf, len);
        bar();
    }
    return MPI_SUCCESS;
}

int MPI_Scatter(const void *sendbuf, int sendcount, MPI_Datatype sdt, void *recvbuf,
                int recvcount, MPI_Datatype rdt, int root, MPI_Comm c)
{
    (void)c; (void)recvcount; (void)rdt;
    size_t per = (size_t)sendcount * (size_t)sdt;
    if (g_np == 1) { memcpy(recvbuf, sendbuf, per); return MPI_SUCCESS; }
    size_t cap = (SHIM_BUF_BYTES / (size_t)g_np) & ~(size_t)63;
    for (size_t off = 0; off < per; off += cap) {
        size_t len = per - off < cap ? per - off : cap;
        if (g_rank == root)
            for (int r = 0; r < g_np; ++r)
                memcpy(g_shm->buf + (size_t)r * cap, (const char *)sendbuf + (size_t)r * per + off, len);
        bar();
        memcpy((char *)recvbuf + off, g_shm->buf + (size_t)g_rank * cap, len);
        bar();
    }
    return MPI_SUCCESS;
}

int MPI_Allreduce(const void *sendbuf, void *recvbuf, int count, MPI_Datatype dt,
                  MPI_Op op, MPI_Comm c)
{
    (void)c; (void)op;
    size_t bytes = (size_t)count * (size_t)dt;
    if (g_np == 1) { memcpy(recvbuf, sendbuf, bytes); return MPI_SUCCESS; }
    size_t cap = (SHIM_BUF_BYTES / (size_t)g_np) & ~(size_t)63;
    if (bytes > cap) { fprintf(stderr, "mpi_shim: allreduce too large\n"); MPI_Abort(0, 1); }
    memcpy(g_shm->buf + (size_t)g_rank * cap, sendbuf, bytes);
    bar();
    for (int i = 0; i < count; ++i) {
        if (dt == MPI_FLOAT) {
            float s = ((const float *)(g_shm->buf))[i];
            for (int r = 1; r < g_np; ++r) s += ((const float *)(g_shm->buf + (size_t)r * cap))[i];
            ((float *)recvbuf)[i] = s;
        } else {
            double s = ((const double *)(g_shm->buf))[i];
            for (int r = 1; r < g_np; ++r) s += ((const double *)(g_shm->buf + (size_t)r * cap))[i];
            ((double *)recvbuf)[i] = s;
        }
    }
    bar();
    return MPI_SUCCESS;
}

int MPI_Allgather(const void *sendbuf, int sendcount, MPI_Datatype sdt, void *recvbuf,
                  int recvcount, MPI_Datatype rdt, MPI_Comm c)
{
    (void)c; (void)recvcount; (void)rdt;
    size_t per = (size_t)sendcount * (size_t)sdt;
    if (g_np == 1) { memcpy(recvbuf, sendbuf, per); return MPI_SUCCESS; }
    size_t cap = (SHIM_BUF_BYTES / (size_t)g_np) & ~(size_t)63;
    for (size_t off = 0; off < per; off += cap) {
        size_t len = per - off < cap ? per - off : cap;
        memcpy(g_shm->buf + (size_t)g_rank * cap, (const char *)sendbuf + off, len);
        bar();
        for (int r = 0; r < g_np; ++r)
            memcpy((char *)recvbuf + (size_t)r * per + off, g_shm->buf + (size_t)r * cap, len);
        bar();
    }
    return MPI_SUCCESS;
}

int cgshim_run_conjugrad(void (*fn)(void *, void *, void *, int, int, int),
                         const void *A, const void *b, void *x,
                         int n, int np, int elem_size)
{
    if (np < 1 || n % np != 0) return -1;
    const size_t es = (size_t)elem_size;
    const int m = n / np;
    shim_setup(np);                       /* children return from here too */
    char *la = malloc((size_t)m * (size_t)n * es);
    char *lb = malloc((size_t)m * es);
    char *lx = malloc((size_t)n * es);
    if (!la || !lb || !lx) { fprintf(stderr, "mpi_shim: out of memory\n"); _exit(1); }
    /* what MPI_Bcast / MPI_Scatter deliver at parallel_cg.c:111-115 */
    memcpy(la, (const char *)A + (size_t)g_rank * (size_t)m * (size_t)n * es, (size_t)m * (size_t)n * es);
    memcpy(lb, (const char *)b + (size_t)g_rank * (size_t)m * es, (size_t)m * es);
    memcpy(lx, x, (size_t)n * es);
    fn(la, lb, lx, m, g_rank, g_np);
    fflush(NULL);
    if (g_rank != 0) _exit(0);
    memcpy(x, lx, (size_t)n * es);
    free(la); free(lb); free(lx);
    shim_reap();
    shim_teardown();
    return 0;
}
