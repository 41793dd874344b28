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
#define SHIM_MAX_NP 16
#define SHIM_CHAN_BYTES ((size_t)256 << 10)

/* One channel per ordered (source, destination) pair for MPI_Send / MPI_Recv: a single slot that
 * the sender fills and the receiver drains, so messages of a pair arrive in the order sent
 * (MPI's non-overtaking rule); longer messages travel as a sequence of slot-sized pieces. */
typedef struct {
    volatile int full;
    int tag;
    size_t len;
    char data[SHIM_CHAN_BYTES];
} shim_chan;

typedef struct {
    pthread_barrier_t bar;
    int np;
    char pad[64];
    shim_chan chan[SHIM_MAX_NP][SHIM_MAX_NP];     /* [source][destination]; untouched pages cost nothing */
    char buf[];
} shim_region;

static shim_region *g_shm = NULL;
static int g_rank = 0, g_np = 1;
static pid_t *g_kids = NULL;

static void shim_setup(int np)
{
    g_np = np < 1 ? 1 : np;
    if (g_np > SHIM_MAX_NP) { fprintf(stderr, "mpi_shim: at most %d ranks\n", SHIM_MAX_NP); exit(1); }
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

/* Blocking point-to-point, as point-to-point_cg.c uses it (:183-256 distribution, :339-394 the
 * emulated reduce / gather): the receiver names its source and tag, counts match exactly. */
#include <sched.h>
static void shim_wait(volatile int *flag, int until)
{
    for (unsigned spins = 0; __atomic_load_n(flag, __ATOMIC_ACQUIRE) != until; ++spins) {
        if (spins < 2000) continue;
        if (spins < 4000) { sched_yield(); continue; }
        struct timespec ts = {0, 20000};
        nanosleep(&ts, NULL);
    }
}

int MPI_Send(const void *buf, int count, MPI_Datatype dt, int dest, int tag, MPI_Comm c)
{
    (void)c;
    if (dest < 0 || dest >= g_np || dest == g_rank) { fprintf(stderr, "mpi_shim: bad destination %d\n", dest); MPI_Abort(0, 1); }
    shim_chan *ch = &g_shm->chan[g_rank][dest];
    size_t total = (size_t)count * (size_t)dt, off = 0;
    do {
        const size_t len = total - off < SHIM_CHAN_BYTES ? total - off : SHIM_CHAN_BYTES;
        shim_wait(&ch->full, 0);
        memcpy(ch->data, (const char *)buf + off, len);
        ch->tag = tag; ch->len = len;
        __atomic_store_n(&ch->full, 1, __ATOMIC_RELEASE);
        off += len;
    } while (off < total);
    return MPI_SUCCESS;
}

int MPI_Recv(void *buf, int count, MPI_Datatype dt, int source, int tag, MPI_Comm c, MPI_Status *status)
{
    (void)c;
    if (source < 0 || source >= g_np || source == g_rank) { fprintf(stderr, "mpi_shim: bad source %d\n", source); MPI_Abort(0, 1); }
    shim_chan *ch = &g_shm->chan[source][g_rank];
    size_t total = (size_t)count * (size_t)dt, off = 0;
    do {
        shim_wait(&ch->full, 1);
        if (ch->tag != tag || ch->len > total - off) {
            fprintf(stderr, "mpi_shim: rank %d expected tag %d (%zu bytes left) from %d, got tag %d (%zu bytes)\n",
                    g_rank, tag, total - off, source, ch->tag, ch->len);
            MPI_Abort(0, 1);
        }
        memcpy((char *)buf + off, ch->data, ch->len);
        off += ch->len;
        __atomic_store_n(&ch->full, 0, __ATOMIC_RELEASE);
    } while (off < total);
    if (status) { status->MPI_SOURCE = source; status->MPI_TAG = tag; status->MPI_ERROR = MPI_SUCCESS; }
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
