/* oracle/cg_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU oracle for the dense conjugate-gradient hot path of
 * mawunyega/conjugate_gradient (serialConjugate.c, parallel_cg.c).  Only
 * tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
 * legs may load this library; the product (libcgb200.so, cg_b200) never does.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement
 *   - against the reference's shipped fixtures (matrixA.txt/vectorb.txt/
 *     initialguess.txt -> x = (2/3, 1/3); matrixA1.txt/vectorb1.txt/X0.txt ->
 *     x = (-1, 1, -1, 1)), committed under tests/golden/;
 *   - bit for bit against the real reference sources compiled into oracle/_ref
 *     (oracle/build_ref.sh) at n in {2, 4, 512, 2048, 16384}, FP32 as shipped and the
 *     mechanical float->double build, serial and both MPI programs (parallel_cg.c
 *     and point-to-point_cg.c, through oracle/mpi_shim);
 *   - against golden vectors the real reference produced on long runs (cond ~ 2^7
 *     and 2^12: 149 and 506 iterations, FP32 leaving through the loop bound) and
 *     on the rho = 0 breakdown (NaNs to the loop bound), tests/golden/golden.json.
 *
 * Build: see oracle/Makefile (gcc -O2 -ffp-contract=off, no -march).
 */
#define _GNU_SOURCE
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <math.h>
#include <time.h>
#include <pthread.h>

#include "spd_gen_ref.h"

typedef struct {
    int64_t iterations;   /* loop bodies executed (x and r updated that many times) */
    int64_t break_k;      /* 0-based loop index at `break`, -1 if the loop ran out   */
    int64_t first_hit_k;  /* first k at which sqrt(beta) < eps held, -1 if never     */
    double  final_beta;   /* last r.r                                                */
    double  rsold0;       /* initial r.r                                             */
    double  seconds;      /* wall time of the solver region, where the reference     */
                          /* times it (serialConjugate.c:208,249; parallel_cg.c:278-279,328-329) */
    double  loop_seconds; /* the iteration loop alone (after the initial A*x0 / r.r) */
} orc_stats;

static void *orc_alloc(size_t bytes)
{
    void *p = NULL;
    if (bytes == 0) bytes = 64;
    if (posix_memalign(&p, 64, bytes) != 0 || !p) {
        fprintf(stderr, "oracle: cannot allocate %zu bytes\n", bytes);
        abort();
    }
    return p;
}

static double orc_now(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

#define REAL double
#define SFX f64
#include "cg_oracle_impl.h"
#undef REAL
#undef SFX

#define REAL float
#define SFX f32
#include "cg_oracle_impl.h"
#undef REAL
#undef SFX

/* ------------------------------------------------------------------ generator */

/* Rows [row0, row0+m) of the seeded SPD matrix, row-major, into A (m x n). */
void orc_gen_fill_A(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                    unsigned flags)
{
    const uint64_t key = orc_gen_key(seed);
    if (flags & ORC_GEN_SPECTRUM) {
        const orc_spectrum sp = orc_gen_spectrum_setup(key, (uint64_t)n, flags);
        for (int64_t i = 0; i < m; ++i)
            for (int64_t j = 0; j < n; ++j)
                A[i * n + j] = orc_gen_spectrum_a(key, (uint64_t)n, (uint64_t)(row0 + i), (uint64_t)j, flags, &sp);
        return;
    }
    if (flags & ORC_GEN_GRAM) {
        /* R in full (n x n), then the m x n block of R R' by rows: plain loops, k ascending */
        double *R = malloc((size_t)n * (size_t)n * sizeof(double));
        if (!R) { for (int64_t e = 0; e < m * n; ++e) A[e] = NAN; return; }
        for (int64_t i = 0; i < n; ++i)
            for (int64_t k = 0; k < n; ++k) R[i * n + k] = orc_gen_u(key, (uint64_t)(i * n + k));
        for (int64_t i = 0; i < m; ++i) {
            const double *ri = R + (row0 + i) * n;
            for (int64_t j = 0; j < n; ++j) {
                const double *rj = R + j * n;
                double acc = 0.0;
                for (int64_t k = 0; k < n; ++k) acc += ri[k] * rj[k];
                if (row0 + i == j) acc = acc + (double)n * orc_gen_d((uint64_t)n, (uint64_t)j, flags);
                if (flags & ORC_GEN_ROUND4) acc = orc_gen_round4(acc);
                A[i * n + j] = acc;
            }
        }
        free(R);
        return;
    }
    for (int64_t i = 0; i < m; ++i)
        for (int64_t j = 0; j < n; ++j)
            A[i * n + j] = orc_gen_a(key, (uint64_t)n, (uint64_t)(row0 + i), (uint64_t)j, flags);
}

void orc_gen_fill_b(double *b, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                    unsigned flags)
{
    const uint64_t key = orc_gen_key(seed);
    for (int64_t i = 0; i < m; ++i)
        b[i] = orc_gen_b(key, (uint64_t)n, (uint64_t)(row0 + i), flags);
}

double orc_gen_sample_u(uint64_t seed, uint64_t k)
{
    return orc_gen_u(orc_gen_key(seed), k);
}

/* ------------------------------------------------- threaded SPMD (timing leg) */
/* The same per-rank arithmetic as orc_conjugrad_spmd_f64, with the P ranks run
 * as P host threads (the reference runs them as P MPI processes,
 * parallel_cg.c:81-83).  Each thread generates and first-touches its own row
 * block, so the matrix never exists twice (the reference keeps a second full
 * copy on rank 0, parallel_cg.c:103) and lands on the thread's NUMA node.
 * All-gather and all-reduce are shared arrays + barriers with rank-ordered
 * sums, i.e. the same results as the sequential restatement.  Used by
 * bench.py's cpu_baseline and --impl reference legs at sizes the real
 * reference cannot index (n > 46340). */

typedef struct {
    int64_t n, P, m, max_iter;
    uint64_t seed; unsigned flags;
    double eps; int ignore_stop;
    double *b, *x, *r, *ploc, *pall, *Ap, *tx, *tr, *tp;
    double *slots;
    pthread_barrier_t bar;
    /* outputs (written by rank 0) */
    orc_stats st;
    double gen_seconds;
} orc_team;

typedef struct { orc_team *tm; int64_t g; } orc_member;

static double team_allreduce(orc_team *tm, int64_t g, double part)
{
    tm->slots[g * 8] = part;                  /* one cache line per rank */
    pthread_barrier_wait(&tm->bar);
    double s = tm->slots[0];
    for (int64_t q = 1; q < tm->P; ++q) s += tm->slots[q * 8];
    pthread_barrier_wait(&tm->bar);
    return s;
}

static void *team_rank_main(void *arg)
{
    orc_member *me = (orc_member *)arg;
    orc_team *tm = me->tm;
    const int64_t g = me->g, n = tm->n, m = tm->m, o = g * m;
    double t0 = orc_now();
    double *Ablk = orc_alloc((size_t)m * (size_t)n * sizeof(double));
    orc_gen_fill_A(Ablk, n, o, m, tm->seed, tm->flags);
    orc_gen_fill_b(tm->b + o, n, o, m, tm->seed, tm->flags);
    memset(tm->x + o, 0, (size_t)m * sizeof(double));
    pthread_barrier_wait(&tm->bar);
    if (g == 0) tm->gen_seconds = orc_now() - t0;

    double rsold, alpha, beta = 0.0;
    int64_t k, done_iters = 0, break_k = -1, first_hit = -1;
    pthread_barrier_wait(&tm->bar);
    t0 = orc_now();
    orc_matvec_f64(tm->Ap + o, Ablk, tm->x, m, n);
    orc_residual_f64(tm->r + o, tm->b + o, tm->Ap + o, m);
    orc_residual_f64(tm->ploc + o, tm->b + o, tm->Ap + o, m);
    rsold = team_allreduce(tm, g, orc_vecvec_f64(tm->r + o, tm->r + o, m));
    double rs0 = rsold;
    const double tl0 = orc_now();
    for (k = 0; k < tm->max_iter; ++k) {
        memcpy(tm->pall + o, tm->ploc + o, (size_t)m * sizeof(double));
        pthread_barrier_wait(&tm->bar);                      /* allgather done */
        orc_matvec_f64(tm->Ap + o, Ablk, tm->pall, m, n);
        alpha = team_allreduce(tm, g, orc_vecvec_f64(tm->ploc + o, tm->Ap + o, m));
        alpha = rsold / alpha;
        /* x is replicated in the reference; each thread updates its slice,
         * element-wise identical arithmetic (parallel_cg.c:299,303). */
        orc_scalarvec_f64(tm->tx + o, tm->pall + o, alpha, m);
        orc_vecadd_f64(tm->x + o, tm->x + o, tm->tx + o, m);
        orc_scalarvec_f64(tm->tr + o, tm->Ap + o, alpha, m);
        orc_vecsub_f64(tm->r + o, tm->r + o, tm->tr + o, m);
        beta = team_allreduce(tm, g, orc_vecvec_f64(tm->r + o, tm->r + o, m));
        done_iters = k + 1;
        if (sqrt(beta) < tm->eps) {
            if (first_hit < 0) first_hit = k;
            if (!tm->ignore_stop) { break_k = k; break; }
        }
        orc_scalarvec_f64(tm->tp + o, tm->ploc + o, beta / rsold, m);
        orc_vecadd_f64(tm->ploc + o, tm->r + o, tm->tp + o, m);
        rsold = beta;
    }
    pthread_barrier_wait(&tm->bar);
    if (g == 0) {
        tm->st.seconds = orc_now() - t0;
        tm->st.loop_seconds = orc_now() - tl0;
        tm->st.iterations = done_iters;
        tm->st.break_k = break_k;
        tm->st.first_hit_k = first_hit;
        tm->st.final_beta = beta;
        tm->st.rsold0 = rs0;
    }
    free(Ablk);
    return NULL;
}

/* Generated-problem SPMD solve on P host threads.  x_out (n doubles) may be NULL.
 * Returns 0, or -1 if n % P != 0 (parallel_cg.c:86-90). */
int orc_team_solve_generated(int64_t n, uint64_t seed, unsigned flags, int64_t P,
                             double eps, int64_t max_iter, int ignore_stop,
                             double *x_out, orc_stats *st, double *gen_seconds)
{
    if (P < 1 || n % P != 0) return -1;
    orc_team tm;
    memset(&tm, 0, sizeof tm);
    tm.n = n; tm.P = P; tm.m = n / P; tm.max_iter = max_iter;
    tm.seed = seed; tm.flags = flags; tm.eps = eps; tm.ignore_stop = ignore_stop;
    size_t vb = (size_t)n * sizeof(double);
    tm.b = orc_alloc(vb); tm.x = orc_alloc(vb); tm.r = orc_alloc(vb);
    tm.ploc = orc_alloc(vb); tm.pall = orc_alloc(vb); tm.Ap = orc_alloc(vb);
    tm.tx = orc_alloc(vb); tm.tr = orc_alloc(vb); tm.tp = orc_alloc(vb);
    tm.slots = orc_alloc((size_t)P * 8 * sizeof(double));
    pthread_barrier_init(&tm.bar, NULL, (unsigned)P);
    pthread_t *th = malloc((size_t)P * sizeof *th);
    orc_member *mb = malloc((size_t)P * sizeof *mb);
    for (int64_t g = 0; g < P; ++g) {
        mb[g].tm = &tm; mb[g].g = g;
        pthread_create(&th[g], NULL, team_rank_main, &mb[g]);
    }
    for (int64_t g = 0; g < P; ++g) pthread_join(th[g], NULL);
    if (x_out) memcpy(x_out, tm.x, vb);
    if (st) *st = tm.st;
    if (gen_seconds) *gen_seconds = tm.gen_seconds;
    pthread_barrier_destroy(&tm.bar);
    free(th); free(mb); free(tm.slots);
    free(tm.b); free(tm.x); free(tm.r); free(tm.ploc); free(tm.pall);
    free(tm.Ap); free(tm.tx); free(tm.tr); free(tm.tp);
    return 0;
}

/* Bounded matVec sample for the cpu_baseline leg: `rows` rows of the generated
 * n-column matrix times a generated vector, `reps` passes, one thread; returns
 * seconds per pass.  checksum defeats dead-code elimination. */
double orc_time_matvec_sample(int64_t n, int64_t rows, uint64_t seed, int reps,
                              double *checksum)
{
    double *A = orc_alloc((size_t)rows * (size_t)n * sizeof(double));
    double *v = orc_alloc((size_t)n * sizeof(double));
    double *y = orc_alloc((size_t)rows * sizeof(double));
    orc_gen_fill_A(A, n, 0, rows, seed, 0);
    orc_gen_fill_b(v, n, 0, n, seed, 0);
    double t0 = orc_now(), acc = 0.0;
    for (int r = 0; r < reps; ++r) {
        orc_matvec_f64(y, A, v, rows, n);
        acc += y[(r * 7919) % rows];
    }
    double dt = (orc_now() - t0) / (double)reps;
    if (checksum) *checksum = acc;
    free(A); free(v); free(y);
    return dt;
}

/* ------------------------------------------------------------------ fused-reductions CG
 * CPU statement of the Chronopoulos-Gear recurrence the product offers as
 * cgb_options.cg_variant = CGB_CG_FUSED_REDUCTIONS.  The REFERENCE has no such
 * variant, so this function is not pinned against reference code: it follows the published
 * algorithm (Chronopoulos & Gear, "s-step iterative methods for symmetric linear systems",
 * J. Comput. Appl. Math. 25 (1989), the s = 1 case) with the reference's stop rule and exit
 * state (serialConjugate.c:234-238: test after the x, r update, p stale on exit).  It is
 * validated in tests/test_oracle.py against the pinned classic oracle above: the two produce
 * the same iterates in exact arithmetic and agree to ~1e-15 on the reference's matrix family. */
void orc_cg_fused_f64(const double *A, const double *b, double *x, int64_t n, double eps,
                      int64_t max_iter, int ignore_stop, orc_stats *st)
{
    double *r = orc_alloc((size_t)n * sizeof(double)), *w = orc_alloc((size_t)n * sizeof(double));
    double *p = orc_alloc((size_t)n * sizeof(double)), *s = orc_alloc((size_t)n * sizeof(double));
    const double t0 = orc_now();
    double gamma = 0.0, delta = 0.0, gamma_prev = 0.0, alpha_prev = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        double ax = 0.0;
        for (int64_t j = 0; j < n; ++j) ax += A[i * n + j] * x[j];
        r[i] = b[i] - ax;
        p[i] = 0.0; s[i] = 0.0;
    }
    for (int64_t i = 0; i < n; ++i) gamma += r[i] * r[i];
    st->rsold0 = gamma; st->final_beta = gamma;
    st->iterations = 0; st->break_k = -1; st->first_hit_k = -1;
    const double t1 = orc_now();
    for (int64_t k = 0; k < max_iter; ++k) {
        delta = 0.0;
        for (int64_t i = 0; i < n; ++i) {
            double wi = 0.0;
            for (int64_t j = 0; j < n; ++j) wi += A[i * n + j] * r[j];
            w[i] = wi;
        }
        for (int64_t i = 0; i < n; ++i) delta += r[i] * w[i];
        const double beta = (k == 0) ? 0.0 : gamma / gamma_prev;
        const double alpha = (k == 0) ? gamma / delta : gamma / (delta - beta * gamma / alpha_prev);
        double gnew = 0.0;
        for (int64_t i = 0; i < n; ++i) {
            p[i] = r[i] + p[i] * beta;
            s[i] = w[i] + s[i] * beta;
            x[i] = x[i] + p[i] * alpha;
            r[i] = r[i] - s[i] * alpha;
            gnew += r[i] * r[i];
        }
        gamma_prev = gamma; alpha_prev = alpha; gamma = gnew;
        st->iterations = k + 1; st->final_beta = gamma;
        if (sqrt(gamma) < eps) {
            if (st->first_hit_k < 0) st->first_hit_k = k;
            if (!ignore_stop) { st->break_k = k; break; }
        }
    }
    const double t2 = orc_now();
    st->seconds = t2 - t0; st->loop_seconds = t2 - t1;
    free(r); free(w); free(p); free(s);
}
