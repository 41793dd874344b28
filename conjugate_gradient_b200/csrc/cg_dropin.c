/* cg_dropin.c -- the reference's own entry points, re-hosted on the GPU solver.
 *
 *   conjugrad(A, b, x)                          serialConjugate.c:40,180
 *   conjugrad_mpi(lA, lb, lx, m, rank, np)      parallel_cg.c:46-48,248-250
 *   cgb_conjugrad_host(A, b, x, n, ..., gpus)   runtime-n form of the above
 *
 * The reference's conjugrad() allocates and frees its temporaries on every call
 * (serialConjugate.c:199-204,253-258) and is re-entrant.  Here the expensive
 * state (CUDA context, NCCL communicator, device buffers, captured graphs) is
 * kept in explicit context objects that are created on first use, cached per
 * thread / per GPU set, and released by cgb_shutdown().
 */
#define _GNU_SOURCE
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <time.h>
#include <unistd.h>

#include "../../include/cgb200.h"
#include "cg_internal.h"

#define MAX_GPUS 16

/* ------------------------------------------------------------ multi-GPU team (one process) */
typedef struct {
    int gpus;
    cgb_ctx *ctx[MAX_GPUS];
} team_t;

static pthread_mutex_t g_team_mu = PTHREAD_MUTEX_INITIALIZER;
static team_t g_team;                 /* cached across calls; guarded by g_team_mu */

typedef struct {
    int g, gpus;
    const unsigned char *id;
    cgb_ctx **slot;
    /* solve */
    const double *A, *b;
    double *x;
    int64_t n;
    const cgb_options *opt;
    cgb_stats stats;
    int rc;
    char err[512];
} member_t;

static void *team_create_main(void *arg)
{
    member_t *me = arg;
    me->rc = cgb_create_rank(me->slot, me->g, me->g, me->gpus, me->id);
    if (me->rc != CGB_OK) snprintf(me->err, sizeof me->err, "%s", cgb_last_error());
    return NULL;
}

static void *team_solve_main(void *arg)
{
    member_t *me = arg;
    const int64_t m = me->n / me->gpus;
    double *xloc = me->x;
    if (me->g != 0) {                         /* every rank holds the full x (parallel_cg.c:98,111) */
        xloc = malloc((size_t)me->n * sizeof(double));
        if (!xloc) { me->rc = CGB_ERR_NOMEM; snprintf(me->err, sizeof me->err, "out of host memory"); return NULL; }
        memcpy(xloc, me->x, (size_t)me->n * sizeof(double));
    }
    me->rc = cgb_conjugrad_rank(*me->slot, me->A + (size_t)me->g * (size_t)m * (size_t)me->n,
                                me->b + (size_t)me->g * (size_t)m, xloc, m, me->opt, &me->stats);
    if (me->rc != CGB_OK) snprintf(me->err, sizeof me->err, "%s", cgb_last_error());
    if (me->g != 0) free(xloc);
    return NULL;
}

/* Every member unmaps its peers' windows before any member frees its own (cg_solver.c: free_problem). */
static void team_release(void)
{
    for (int g = 0; g < g_team.gpus; ++g) cgb_unmap_peers(g_team.ctx[g]);
    for (int g = 0; g < g_team.gpus; ++g) { cgb_destroy(g_team.ctx[g]); g_team.ctx[g] = NULL; }
    g_team.gpus = 0;
}

/* One host thread per member, all or none: the members meet in collectives, so a team with a thread
 * missing would hang.  The threads wait at a gate until every pthread_create has succeeded; if one
 * fails the gate opens with "abort", nobody calls fn, and the caller gets CGB_ERR_NOMEM. */
typedef struct { pthread_mutex_t mu; pthread_cond_t cv; int state; } team_gate;      /* 0 wait, 1 go, -1 abort */
typedef struct { void *(*fn)(void *); void *arg; team_gate *gate; } team_thunk;

static void *team_thunk_main(void *arg)
{
    team_thunk *t = arg;
    pthread_mutex_lock(&t->gate->mu);
    while (t->gate->state == 0) pthread_cond_wait(&t->gate->cv, &t->gate->mu);
    const int go = t->gate->state == 1;
    pthread_mutex_unlock(&t->gate->mu);
    if (go) t->fn(t->arg);
    return NULL;
}

static int team_run(void *(*fn)(void *), void *members, size_t stride, int gpus)
{
    pthread_t th[MAX_GPUS];
    team_thunk tk[MAX_GPUS];
    team_gate gate;
    pthread_mutex_init(&gate.mu, NULL);
    pthread_cond_init(&gate.cv, NULL);
    gate.state = 0;
    int started = 0;
    for (int g = 1; g < gpus; ++g) {                 /* member 0 runs on the calling thread */
        tk[g].fn = fn; tk[g].arg = (char *)members + (size_t)g * stride; tk[g].gate = &gate;
        if (pthread_create(&th[g], NULL, team_thunk_main, &tk[g]) != 0) break;
        started = g;
    }
    const int all = started == gpus - 1;
    pthread_mutex_lock(&gate.mu);
    gate.state = all ? 1 : -1;
    pthread_cond_broadcast(&gate.cv);
    pthread_mutex_unlock(&gate.mu);
    if (all) fn(members);
    for (int g = 1; g <= started; ++g) pthread_join(th[g], NULL);
    pthread_cond_destroy(&gate.cv);
    pthread_mutex_destroy(&gate.mu);
    if (!all) {
        cgb_set_error("%s%s (cg_dropin.c:%d)", "could not start one host thread per GPU (pthread_create failed)", "", __LINE__);
        return CGB_ERR_NOMEM;
    }
    return CGB_OK;
}

/* rc and message of a team phase: the member that failed by itself names the cause; members that only
 * learnt of it through the status agreement ("another rank failed") speak only if nobody else does */
#define TEAM_COLLECT(mb, gpus, rc, prefix)                                                              \
    do {                                                                                                \
        for (int pass_ = 0; pass_ < 2 && (rc) == CGB_OK; ++pass_)                                       \
            for (int g_ = 0; g_ < (gpus) && (rc) == CGB_OK; ++g_)                                       \
                if ((mb)[g_].rc != CGB_OK && (pass_ == 1 || !strstr((mb)[g_].err, "another rank failed"))) { \
                    (rc) = (mb)[g_].rc;                                                                 \
                    cgb_set_error("%s%s (cg_dropin.c:%d)", prefix, (mb)[g_].err, __LINE__);             \
                }                                                                                       \
    } while (0)

static void team_release_keep_error(void)
{
    char keep[512];
    snprintf(keep, sizeof keep, "%s", cgb_last_error());
    team_release();
    cgb_set_error("%s%s", keep, "", 0);
}

static int team_ensure(int gpus)
{
    if (g_team.gpus == gpus) return CGB_OK;
    team_release();
    if (gpus == 1) {
        int rc = cgb_create(&g_team.ctx[0], 0);
        if (rc == CGB_OK) g_team.gpus = 1;
        return rc;
    }
    unsigned char id[128];
    int rc = cgb_unique_id(id);
    if (rc != CGB_OK) return rc;
    member_t mb[MAX_GPUS];
    memset(mb, 0, sizeof mb);
    for (int g = 0; g < gpus; ++g) { mb[g].g = g; mb[g].gpus = gpus; mb[g].id = id; mb[g].slot = &g_team.ctx[g]; }
    rc = team_run(team_create_main, mb, sizeof mb[0], gpus);
    if (rc != CGB_OK) return rc;
    for (int g = 0; g < gpus; ++g)
        if (mb[g].rc != CGB_OK) {
            cgb_set_error("%s%s (cg_dropin.c:%d)", "team create failed: ", mb[g].err, __LINE__);
            int bad = mb[g].rc;
            for (int q = 0; q < gpus; ++q) { cgb_destroy(g_team.ctx[q]); g_team.ctx[q] = NULL; }
            return bad;
        }
    g_team.gpus = gpus;
    return CGB_OK;
}

int cgb_shutdown(void)
{
    pthread_mutex_lock(&g_team_mu);
    team_release();
    pthread_mutex_unlock(&g_team_mu);
    return CGB_OK;
}

int cgb_conjugrad_host_opt(const double *A, const double *b, double *x, int64_t n, int gpus,
                           const cgb_options *opt, cgb_stats *stats)
{
    if (!A || !b || !x || n < 1 || n > CGB_MAX_ORDER) { cgb_set_error("%s%s (cg_dropin.c:%d)", "cgb_conjugrad_host: bad argument (null pointer or matrix order out of range)", "", __LINE__); return CGB_ERR_ARG; }
    if (gpus < 1) gpus = 1;
    if (gpus > MAX_GPUS) { cgb_set_error("%s%s (cg_dropin.c:%d)", "cgb_conjugrad_host: too many GPUs", "", __LINE__); return CGB_ERR_ARG; }
    if (n % gpus != 0) {
        char msg[96];
        snprintf(msg, sizeof msg, "%lld is not divisible by %d", (long long)n, gpus);   /* parallel_cg.c:88 */
        cgb_set_error("%s%s (cg_dropin.c:%d)", msg, "", __LINE__);
        return CGB_ERR_NOT_DIVISIBLE;
    }
    int ndev = cgb_device_count();
    if (ndev < gpus) { cgb_set_error("%s%s (cg_dropin.c:%d)", "not enough CUDA devices for the requested --gpus", "", __LINE__); return CGB_ERR_CUDA; }
    pthread_mutex_lock(&g_team_mu);
    int rc = team_ensure(gpus);
    if (rc != CGB_OK) { pthread_mutex_unlock(&g_team_mu); return rc; }
    if (gpus == 1) {
        rc = cgb_conjugrad_rank(g_team.ctx[0], A, b, x, n, opt, stats);
        pthread_mutex_unlock(&g_team_mu);
        return rc;
    }
    member_t mb[MAX_GPUS];
    memset(mb, 0, sizeof mb);
    for (int g = 0; g < gpus; ++g) {
        mb[g].g = g; mb[g].gpus = gpus; mb[g].slot = &g_team.ctx[g];
        mb[g].A = A; mb[g].b = b; mb[g].x = x; mb[g].n = n; mb[g].opt = opt;
    }
    rc = team_run(team_solve_main, mb, sizeof mb[0], gpus);
    TEAM_COLLECT(mb, gpus, rc, "rank failed: ");
    if (rc != CGB_OK) team_release_keep_error();       /* contexts and communicators of a failed call are not reused */
    if (rc == CGB_OK && stats) {
        *stats = mb[0].stats;
        for (int g = 1; g < gpus; ++g) {                  /* MPI_Barrier-bracketed timers: slowest rank */
            if (mb[g].stats.solve_seconds > stats->solve_seconds) stats->solve_seconds = mb[g].stats.solve_seconds;
            if (mb[g].stats.h2d_seconds > stats->h2d_seconds) stats->h2d_seconds = mb[g].stats.h2d_seconds;
            if (mb[g].stats.loop_ms > stats->loop_ms) stats->loop_ms = mb[g].stats.loop_ms;
        }
    }
    pthread_mutex_unlock(&g_team_mu);
    return rc;
}

/* ---- generated problems: every GPU writes its own row block, nothing is uploaded */
typedef struct {
    int g, gpus;
    cgb_ctx **slot;
    int64_t n;
    uint64_t seed;
    unsigned flags;
    const cgb_options *opt;
    double *x;
    cgb_stats stats;
    int rc;
    char err[512];
} genmember_t;

static void *team_generated_main(void *arg)
{
    genmember_t *me = arg;
    me->rc = cgb_generate(*me->slot, me->n, me->seed, me->flags);
    if (me->rc == CGB_OK) me->rc = cgb_solve(*me->slot, me->opt, &me->stats);
    if (me->rc == CGB_OK && me->g == 0 && me->x) me->rc = cgb_read_x(*me->slot, me->x);
    if (me->rc != CGB_OK) snprintf(me->err, sizeof me->err, "%s", cgb_last_error());
    return NULL;
}

int cgb_solve_generated(int64_t n, uint64_t seed, unsigned flags, int gpus,
                        const cgb_options *opt, double *x, cgb_stats *stats)
{
    if (n < 1 || n > CGB_MAX_ORDER) { cgb_set_error("%s%s (cg_dropin.c:%d)", "cgb_solve_generated: matrix order out of range", "", __LINE__); return CGB_ERR_ARG; }
    if (gpus < 1) gpus = 1;
    if (gpus > MAX_GPUS) { cgb_set_error("%s%s (cg_dropin.c:%d)", "cgb_solve_generated: too many GPUs", "", __LINE__); return CGB_ERR_ARG; }
    if (n % gpus != 0) {
        char msg[96];
        snprintf(msg, sizeof msg, "%lld is not divisible by %d", (long long)n, gpus);
        cgb_set_error("%s%s (cg_dropin.c:%d)", msg, "", __LINE__);
        return CGB_ERR_NOT_DIVISIBLE;
    }
    if (cgb_device_count() < gpus) { cgb_set_error("%s%s (cg_dropin.c:%d)", "not enough CUDA devices", "", __LINE__); return CGB_ERR_CUDA; }
    pthread_mutex_lock(&g_team_mu);
    int rc = team_ensure(gpus);
    if (rc != CGB_OK) { pthread_mutex_unlock(&g_team_mu); return rc; }
    genmember_t mb[MAX_GPUS];
    memset(mb, 0, sizeof mb);
    for (int g = 0; g < gpus; ++g) {
        mb[g].g = g; mb[g].gpus = gpus; mb[g].slot = &g_team.ctx[g];
        mb[g].n = n; mb[g].seed = seed; mb[g].flags = flags; mb[g].opt = opt; mb[g].x = x;
    }
    rc = team_run(team_generated_main, mb, sizeof mb[0], gpus);
    TEAM_COLLECT(mb, gpus, rc, "rank failed: ");
    if (rc != CGB_OK) team_release_keep_error();
    if (rc == CGB_OK && stats) {
        *stats = mb[0].stats;
        for (int g = 1; g < gpus; ++g) {
            if (mb[g].stats.solve_seconds > stats->solve_seconds) stats->solve_seconds = mb[g].stats.solve_seconds;
            if (mb[g].stats.loop_ms > stats->loop_ms) stats->loop_ms = mb[g].stats.loop_ms;
        }
    }
    pthread_mutex_unlock(&g_team_mu);
    return rc;
}

/* ---- the three text files: every GPU ingests its own rows (cg_ingest.c) */
typedef struct {
    int g, gpus;
    cgb_ctx **slot;
    int64_t n;
    const char *fileA, *fileB, *fileX;
    const cgb_options *opt;
    double *x;
    cgb_stats stats;
    int rc;
    char err[512];
} filemember_t;

static void *team_files_main(void *arg)
{
    filemember_t *me = arg;
    me->rc = cgb_load_text(*me->slot, me->n, me->fileA, me->fileB, me->fileX);
    if (me->rc != CGB_OK) snprintf(me->err, sizeof me->err, "%s", cgb_last_error());
    return NULL;
}

static void *team_files_solve_main(void *arg)
{
    filemember_t *me = arg;
    me->rc = cgb_solve(*me->slot, me->opt, &me->stats);
    if (me->rc == CGB_OK && me->g == 0 && me->x) me->rc = cgb_read_x(*me->slot, me->x);
    if (me->rc != CGB_OK) snprintf(me->err, sizeof me->err, "%s", cgb_last_error());
    return NULL;
}

int cgb_conjugrad_files(const char *fileA, const char *fileB, const char *fileX, int64_t n, int gpus,
                        const cgb_options *opt, double *x, cgb_stats *stats)
{
    if (!fileA || !fileB || !fileX || n < 1 || n > CGB_MAX_ORDER) { cgb_set_error("%s%s (cg_dropin.c:%d)", "cgb_conjugrad_files: bad argument (null pointer or matrix order out of range)", "", __LINE__); return CGB_ERR_ARG; }
    if (gpus < 1) gpus = 1;
    if (gpus > MAX_GPUS) { cgb_set_error("%s%s (cg_dropin.c:%d)", "cgb_conjugrad_files: too many GPUs", "", __LINE__); return CGB_ERR_ARG; }
    if (n % gpus != 0) {
        char msg[96];
        snprintf(msg, sizeof msg, "%lld is not divisible by %d", (long long)n, gpus);   /* parallel_cg.c:88 */
        cgb_set_error("%s%s (cg_dropin.c:%d)", msg, "", __LINE__);
        return CGB_ERR_NOT_DIVISIBLE;
    }
    if (cgb_device_count() < gpus) { cgb_set_error("%s%s (cg_dropin.c:%d)", "not enough CUDA devices for the requested --gpus", "", __LINE__); return CGB_ERR_CUDA; }
    pthread_mutex_lock(&g_team_mu);
    int rc = team_ensure(gpus);
    if (rc != CGB_OK) { pthread_mutex_unlock(&g_team_mu); return rc; }
    filemember_t mb[MAX_GPUS];
    memset(mb, 0, sizeof mb);
    /* two phases (cgb_load_text lets the ranks agree on its outcome, so one phase would do; two keep the
     * first failing rank's message -- the name of the bad file -- rather than "another rank failed") */
    for (int phase = 0; phase < 2 && rc == CGB_OK; ++phase) {
        for (int g = 0; g < gpus; ++g) {
            mb[g].g = g; mb[g].gpus = gpus; mb[g].slot = &g_team.ctx[g]; mb[g].n = n;
            mb[g].fileA = fileA; mb[g].fileB = fileB; mb[g].fileX = fileX; mb[g].opt = opt; mb[g].x = x;
        }
        rc = team_run(phase == 0 ? team_files_main : team_files_solve_main, mb, sizeof mb[0], gpus);
        TEAM_COLLECT(mb, gpus, rc, "");
    }
    if (rc != CGB_OK && rc != CGB_ERR_IO) team_release_keep_error();
    if (rc == CGB_OK && stats) {
        *stats = mb[0].stats;
        for (int g = 1; g < gpus; ++g) {
            if (mb[g].stats.solve_seconds > stats->solve_seconds) stats->solve_seconds = mb[g].stats.solve_seconds;
            if (mb[g].stats.h2d_seconds > stats->h2d_seconds) stats->h2d_seconds = mb[g].stats.h2d_seconds;
            if (mb[g].stats.loop_ms > stats->loop_ms) stats->loop_ms = mb[g].stats.loop_ms;
        }
    }
    pthread_mutex_unlock(&g_team_mu);
    return rc;
}

int cgb_conjugrad_host(const double *A, const double *b, double *x, int64_t n, double eps,
                       int64_t max_iter, int gpus, cgb_stats *stats)
{
    cgb_options o;
    cgb_default_options(&o);
    if (eps > 0.0) o.eps = eps;
    o.max_iter = max_iter;
    return cgb_conjugrad_host_opt(A, b, x, n, gpus, &o, stats);
}

/* ------------------------------------------------------------ reference-named wrappers */
/* The run-time ROWS.  The reference's conjugrad() reads a compile-time macro and is re-entrant
 * (serialConjugate.c:180, no globals); here the order set by a thread belongs to that thread, so two
 * threads can run systems of different order side by side.  A thread that never set one uses the
 * last value any thread set (a main() that sets the order and calls conjugrad() from a worker). */
static int64_t g_dimension = 0;
static __thread int64_t tl_dimension = 0;
static __thread cgb_stats tl_stats;
static __thread int tl_have_stats = 0;

void cgb_set_dimension(int64_t n)
{
    tl_dimension = n;
    __atomic_store_n(&g_dimension, n, __ATOMIC_RELAXED);
}
static int64_t current_dimension(void)
{
    return tl_dimension > 0 ? tl_dimension : __atomic_load_n(&g_dimension, __ATOMIC_RELAXED);
}

int cgb_last_stats(cgb_stats *out)
{
    if (!out || !tl_have_stats) return CGB_ERR_STATE;
    *out = tl_stats;
    return CGB_OK;
}

static int env_int(const char *name, int dflt)
{
    const char *e = getenv(name);
    return (e && *e) ? atoi(e) : dflt;
}

/* serialConjugate.c:180: x in/out, nothing returned; the CG-only time is
 * printed from inside (serialConjugate.c:249-251). */
void conjugrad(double *matrixA, double *vectorB, double *vectorX)
{
    tl_have_stats = 0;
    const int64_t dim = current_dimension();
    if (dim < 1) {
        fprintf(stderr, "conjugrad: matrix order not set (call cgb_set_dimension(n): the run-time ROWS)\n");
        return;
    }
    int rc = cgb_conjugrad_host(matrixA, vectorB, vectorX, dim, CGB_EPSILON, 0,
                                env_int("CGB_GPUS", 1), &tl_stats);
    if (rc != CGB_OK) { fprintf(stderr, "conjugrad: %s\n", cgb_last_error()); return; }
    tl_have_stats = 1;
    printf("average clock execution time in seconds: %f\n", tl_stats.solve_seconds);
}

/* Rendezvous for the NCCL id between the SPMD callers of conjugrad_mpi():
 * ranks that are threads of one process meet in a process-wide slot; ranks
 * that are processes (mpiexec, torchrun) meet through the file named by
 * CGB_ID_FILE (rank 0 writes it atomically, the others wait for it). */
static pthread_mutex_t g_rdv_mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t g_rdv_cv = PTHREAD_COND_INITIALIZER;
static unsigned char g_rdv_id[128];
static int g_rdv_have = 0, g_rdv_taken = 0;

static void launch_tag(char tag[64])
{
    static const char *const vars[] = {"CGB_RUN_ID", "TORCHELASTIC_RUN_ID", "SLURM_JOB_ID", "OMPI_MCA_ess_base_jobid", "PMIX_NAMESPACE", "PMI_JOBID"};
    memset(tag, 0, 64);
    for (size_t i = 0; i < sizeof vars / sizeof vars[0]; ++i) {
        const char *v = getenv(vars[i]);
        if (v && *v) { snprintf(tag, 64, "%.15s=%.46s", vars[i], v); return; }
    }
}

static int rendezvous_id(int rank, int np, unsigned char *id)
{
    const char *path = getenv("CGB_ID_FILE");
    if (path && *path) {
        /* File layout: 128 bytes of id, then a 64-byte launch tag.  The tag is whatever the launcher gives
         * every rank of ONE launch (CGB_RUN_ID, torchrun's TORCHELASTIC_RUN_ID, SLURM_JOB_ID, OpenMPI's
         * OMPI_MCA_ess_base_jobid, PMI_JOBID / PMIX_NAMESPACE); readers ignore a file whose tag is not theirs,
         * so the id a crashed run left behind is never taken for this run's.  Without any such variable a
         * reader accepts only a file written after it started (less two minutes of launch skew).  Rank 0
         * also removes a stale file before it makes the new id. */
        char tag[64];
        launch_tag(tag);
        if (rank == 0) {
            (void)unlink(path);
            int rc = cgb_unique_id(id);
            if (rc != CGB_OK) return rc;
            char tmp[4096];
            snprintf(tmp, sizeof tmp, "%s.tmp.%d", path, (int)getpid());
            FILE *f = fopen(tmp, "wb");
            if (!f || fwrite(id, 1, 128, f) != 128 || fwrite(tag, 1, sizeof tag, f) != sizeof tag) { if (f) fclose(f); return CGB_ERR_IO; }
            fclose(f);
            if (rename(tmp, path) != 0) return CGB_ERR_IO;
            return CGB_OK;
        }
        const time_t started = time(NULL);
        for (int tries = 0; tries < 600000; ++tries) {
            FILE *f = fopen(path, "rb");
            if (f) {
                char got_tag[64];
                size_t got = fread(id, 1, 128, f);
                size_t gt = fread(got_tag, 1, sizeof got_tag, f);
                struct stat sb;
                int fresh = fstat(fileno(f), &sb) == 0 && sb.st_mtime >= started - 120;
                fclose(f);
                if (got == 128 && gt == sizeof got_tag && memcmp(got_tag, tag, sizeof tag) == 0 && (tag[0] != '\0' || fresh)) return CGB_OK;
            }
            usleep(100);
        }
        return CGB_ERR_IO;
    }
    pthread_mutex_lock(&g_rdv_mu);
    int rc = CGB_OK;
    if (rank == 0) {
        while (g_rdv_have) pthread_cond_wait(&g_rdv_cv, &g_rdv_mu);     /* previous round draining */
        rc = cgb_unique_id(g_rdv_id);
        g_rdv_have = 1; g_rdv_taken = 0;
        pthread_cond_broadcast(&g_rdv_cv);
    } else {
        while (!g_rdv_have) pthread_cond_wait(&g_rdv_cv, &g_rdv_mu);
    }
    memcpy(id, g_rdv_id, 128);
    if (++g_rdv_taken == np) { g_rdv_have = 0; pthread_cond_broadcast(&g_rdv_cv); }
    pthread_mutex_unlock(&g_rdv_mu);
    return rc;
}

/* test hook: the rendezvous alone (CPU tests of the stale-file rules) */
int cgb_debug_rendezvous_id(int rank, int np, void *id128)
{
    if (!id128 || np < 1 || rank < 0 || rank >= np) return CGB_ERR_ARG;
    return rendezvous_id(rank, np, id128);
}

static __thread cgb_ctx *tl_rank_ctx = NULL;
static __thread int tl_rank = -1, tl_np = -1;

/* parallel_cg.c:248: SPMD, every rank calls it; local_vectorX is the FULL
 * vector; rank 0 prints the CG-only time (parallel_cg.c:331-336). */
void conjugrad_mpi(double *local_matrixA, double *local_vectorB, double *local_vectorX,
                   int local_row, int myrank, int procsnum)
{
    tl_have_stats = 0;
    if (procsnum < 1 || myrank < 0 || myrank >= procsnum || local_row < 1) {
        fprintf(stderr, "conjugrad_mpi: bad rank/size arguments\n");
        return;
    }
    if (!tl_rank_ctx || tl_rank != myrank || tl_np != procsnum) {
        if (tl_rank_ctx) { cgb_destroy(tl_rank_ctx); tl_rank_ctx = NULL; }
        int ndev = cgb_device_count();
        if (ndev < 1) { fprintf(stderr, "conjugrad_mpi: no CUDA device\n"); return; }
        int dev = env_int("CGB_DEVICE", env_int("LOCAL_RANK", myrank)) % ndev;
        unsigned char id[128];
        int rc = CGB_OK;
        if (procsnum > 1) rc = rendezvous_id(myrank, procsnum, id);
        if (rc == CGB_OK) rc = cgb_create_rank(&tl_rank_ctx, dev, myrank, procsnum, id);
        if (rc != CGB_OK) { fprintf(stderr, "conjugrad_mpi: %s\n", cgb_last_error()); tl_rank_ctx = NULL; return; }
        tl_rank = myrank; tl_np = procsnum;
        if (myrank == 0 && procsnum > 1) {
            /* ncclCommInitRank is collective: once it has returned here every rank has read the id,
             * so the rendezvous file can go -- a later run never meets a stale one */
            const char *path = getenv("CGB_ID_FILE");
            if (path && *path) (void)unlink(path);
        }
    }
    int rc = cgb_conjugrad_rank(tl_rank_ctx, local_matrixA, local_vectorB, local_vectorX,
                                local_row, NULL, &tl_stats);
    if (rc != CGB_OK) { fprintf(stderr, "conjugrad_mpi: %s\n", cgb_last_error()); return; }
    tl_have_stats = 1;
    if (myrank == 0) printf("cg method execution time in seconds: %f\n", tl_stats.solve_seconds);
}

/* Releases the calling thread's conjugrad_mpi() context (collective). */
int cgb_release_rank(void)
{
    if (tl_rank_ctx) { cgb_destroy(tl_rank_ctx); tl_rank_ctx = NULL; tl_rank = tl_np = -1; }
    return CGB_OK;
}
