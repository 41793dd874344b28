/* cg_main.c -- the reference's command line, re-hosted on the B200 solver.
 *
 *   cg_b200 matrixA.txt vectorb.txt initialguess.txt [options]
 *
 * Same three positional files in the same order as the reference
 * (serialConjugate.c:48-52,65-67; parallel_cg.c:105-107) and the same stdout
 * lines.  The problem size, a compile-time macro in the
 * reference (#define ROWS/COLS, serialConjugate.c:29-30), is read at run time
 * from dimensions.txt (rows(A) cols(A) rows(b) cols(b), one per line).
 *
 * Options (none of them disturbs the positional form):
 *   --dims FILE       dimensions file (default: ./dimensions.txt, else next to A,
 *                     else inferred from the number of values in vectorb)
 *   --gpus G          row-shard A over G GPUs (the reference's mpiexec -np G)
 *   --eps E           stop tolerance (default 1e-6 = EPSILON)
 *   --max-iter K      loop bound (default n = ROWS)
 *   --print-x         print the solution in printer()'s format (the call is
 *                     commented out in the reference: serialConjugate.c:252)
 *   --write-x FILE    write x one value per line ("%.17g"): usable as initialguess
 *   --gen N           no input files: seeded on-device SPD system of order N
 *   --seed S          generator seed (default 20170209)   --round4  "%.4f" rounding
 *   --kappa-log2 K    generator: spread the diagonal so that cond(A) ~ 2^K (K = 0: the
 *                     reference's kappa ~ 1.5 family, ~5 iterations)
 *   --family F        generator: shift (0.5 (R + R') + n I, generateSPDmatrix.m:11,17: the default),
 *                     gram (R R' + n I, the commented alternative of :12), spectrum (dense matrix
 *                     with the prescribed eigenvalues n (1 + (2^K - 1) t^2))
 *   --bench-iters K   fixed-iteration timing mode (stop test recorded, not acted on)
 *   --gemv V          1 = streaming loads, 2 = bulk-async ring
 *   --comm nccl|peer  per-iteration exchange with --gpus > 1 (default: peer when the GPUs
 *                     can map each other's memory: fused NVLink pushes, else NCCL)
 *   --stats           extra lines: iterations, residual, GEMV GB/s
 *   --emit-files P    with --gen: also write the system in generateSPDmatrix.m's file format
 *                     (generateSPDmatrix.m:23-43: one value per line, A and b "%.4f", x0 "%.1f")
 *                     as P_matrix.txt, P_vector.txt, P_x0.txt and P_dimensions.txt; implies
 *                     --round4 so that the files hold exactly the system that is solved
 *   --symmetric       matVec from the upper tiles only (half the bytes); the matrix must be exactly
 *                     symmetric (checked on the GPU; with --gpus > 1 only for --gen systems)
 *   --cg classic|fused  recurrence: the reference's (default) or Chronopoulos-Gear CG with one
 *                     exposed reduction per iteration (same iterates in exact arithmetic)
 *   --guard           end the solve on p.Ap = 0 / exact x0 instead of iterating NaNs to the
 *                     loop bound as the reference does (parallel_cg.c:298,318)
 *   --run-record FILE machine-readable record of the run (JSON: iterations, sqrt(r.r), exit
 *                     reason, timings, GB/s) -- the reference reports only a time
 */
#define _GNU_SOURCE
#include <errno.h>
#include <math.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <libgen.h>
#include <unistd.h>

#include "../../include/cgb200.h"
#include "cg_io.h"

typedef struct {
    const char *fileA, *fileB, *fileX, *dims, *write_x, *record, *emit;
    int gpus, print_x, stats, round4, gemv, comm, kappa_log2, guard, symmetric, cg_variant;
    unsigned family;             /* 0, CGB_GEN_SPECTRUM or CGB_GEN_GRAM */
    int64_t gen_n, max_iter, bench_iters;
    uint64_t seed;
    double eps;
} cli_t;

/* Whole-token numeric options: "--gpus 0", "--gpus x" or "--eps -1" are usage errors, not
 * silently a default (and never a division by zero in the divisibility test). */
static int parse_ll(const char *opt, const char *v, long long lo, long long *out)
{
    char *end = NULL;
    errno = 0;
    const long long x = strtoll(v, &end, 10);
    if (end == v || *end != '\0' || x < lo || errno == ERANGE) {
        fprintf(stderr, "option %s takes an integer >= %lld (got \"%s\")\n", opt, lo, v);
        return -1;
    }
    *out = x;
    return 0;
}

static int parse_cli(int argc, char **argv, cli_t *c)
{
    const char *pos[3] = {0, 0, 0};
    int npos = 0;
    memset(c, 0, sizeof *c);
    c->gpus = 1; c->seed = 20170209ULL; c->eps = CGB_EPSILON;
    for (int i = 1; i < argc; ++i) {
        const char *a = argv[i];
        const char *v = (i + 1 < argc) ? argv[i + 1] : NULL;
        if (strncmp(a, "--", 2) != 0) { if (npos < 3) pos[npos] = a; ++npos; continue; }
        if (!strcmp(a, "--print-x")) c->print_x = 1;
        else if (!strcmp(a, "--stats")) c->stats = 1;
        else if (!strcmp(a, "--round4")) c->round4 = 1;
        else if (!strcmp(a, "--guard")) c->guard = 1;
        else if (!strcmp(a, "--symmetric")) c->symmetric = 1;
        else if (!v) { fprintf(stderr, "option %s needs a value\n", a); return -1; }
        else if (!strcmp(a, "--dims")) { c->dims = v; ++i; }
        else if (!strcmp(a, "--write-x")) { c->write_x = v; ++i; }
        else if (!strcmp(a, "--run-record")) { c->record = v; ++i; }
        else if (!strcmp(a, "--emit-files")) { c->emit = v; c->round4 = 1; ++i; }
        else if (!strcmp(a, "--gpus")) {
            long long g = 0;
            if (parse_ll(a, v, 1, &g)) return -1;
            if (g > 1024) { fprintf(stderr, "option --gpus: too many\n"); return -1; }
            c->gpus = (int)g; ++i;
        }
        else if (!strcmp(a, "--eps")) {
            char *end = NULL;
            c->eps = strtod(v, &end);
            if (end == v || *end != '\0' || !(c->eps > 0.0)) { fprintf(stderr, "option --eps takes a positive number (got \"%s\")\n", v); return -1; }
            ++i;
        }
        else if (!strcmp(a, "--max-iter")) { long long k; if (parse_ll(a, v, 1, &k)) return -1; c->max_iter = k; ++i; }
        else if (!strcmp(a, "--gen")) { long long k; if (parse_ll(a, v, 1, &k)) return -1; c->gen_n = k; ++i; }
        else if (!strcmp(a, "--seed")) {
            char *end = NULL;
            c->seed = strtoull(v, &end, 10);
            if (end == v || *end != '\0' || *v == '-') { fprintf(stderr, "option --seed takes an unsigned integer (got \"%s\")\n", v); return -1; }
            ++i;
        }
        else if (!strcmp(a, "--bench-iters")) { long long k; if (parse_ll(a, v, 1, &k)) return -1; c->bench_iters = k; ++i; }
        else if (!strcmp(a, "--gemv")) {
            long long k = 0;
            if (parse_ll(a, v, 0, &k)) return -1;
            if (k > 2) { fprintf(stderr, "option --gemv takes 0 (default), 1 (streaming loads) or 2 (bulk-async copies)\n"); return -1; }
            c->gemv = (int)k; ++i;
        }
        else if (!strcmp(a, "--cg")) {
            if (!strcmp(v, "classic")) c->cg_variant = CGB_CG_CLASSIC;
            else if (!strcmp(v, "fused")) c->cg_variant = CGB_CG_FUSED_REDUCTIONS;
            else { fprintf(stderr, "--cg takes classic or fused\n"); return -1; }
            ++i;
        }
        else if (!strcmp(a, "--family")) {
            if (!strcmp(v, "shift")) c->family = 0u;                      /* generateSPDmatrix.m:11,17 */
            else if (!strcmp(v, "gram")) c->family = CGB_GEN_GRAM;         /* its commented alternative, :12 */
            else if (!strcmp(v, "spectrum")) c->family = CGB_GEN_SPECTRUM; /* prescribed eigenvalues */
            else { fprintf(stderr, "--family takes shift, gram or spectrum\n"); return -1; }
            ++i;
        }
        else if (!strcmp(a, "--kappa-log2")) {
            long long k = 0;
            if (parse_ll(a, v, 0, &k)) return -1;
            if (k > 63) { fprintf(stderr, "option --kappa-log2: at most 63\n"); return -1; }
            c->kappa_log2 = (int)k; ++i;
        }
        else if (!strcmp(a, "--comm")) {
            if (!strcmp(v, "nccl")) c->comm = CGB_COMM_NCCL;
            else if (!strcmp(v, "peer")) c->comm = CGB_COMM_PEER;
            else if (!strcmp(v, "auto")) c->comm = CGB_COMM_AUTO;
            else { fprintf(stderr, "--comm takes nccl, peer or auto\n"); return -1; }
            ++i;
        }
        else { fprintf(stderr, "unknown option %s\n", a); return -1; }
    }
    c->fileA = pos[0]; c->fileB = pos[1]; c->fileX = pos[2];
    if (c->gen_n > 0) {
        if (npos != 0) { fprintf(stderr, "--gen N solves a generated system: it takes no input files (got %d)\n", npos); return -1; }
        return 0;
    }
    return npos == 3 ? 0 : -2;
}

static const char *exit_reason_name(int r)
{
    switch (r) {
    case CGB_EXIT_STOP_RULE: return "stop_rule";
    case CGB_EXIT_MAX_ITER: return "max_iter";
    case CGB_EXIT_BREAKDOWN: return "breakdown";
    case CGB_EXIT_EXACT_X0: return "exact_initial_guess";
    default: return "none";
    }
}

static int write_run_record(const char *path, const cli_t *cli, const cgb_stats *st, double total_s)
{
    FILE *f = fopen(path, "w");
    if (!f) return -1;
    const double it_s = (st->iterations > 0 && st->loop_ms > 0.0) ? 1.0e3 * (double)st->iterations / st->loop_ms : 0.0;
    fprintf(f, "{\"n\": %lld, \"gpus\": %d, \"rows_per_gpu\": %lld, \"source\": \"%s\", \"seed\": %llu,\n",
            (long long)st->n, st->ranks, (long long)st->m, cli->gen_n > 0 ? "generated" : "files",
            (unsigned long long)cli->seed);
    fprintf(f, " \"eps\": %.17g, \"iterations\": %lld, \"break_k\": %lld, \"exit\": \"%s\",\n",
            cli->eps, (long long)st->iterations, (long long)st->break_k, exit_reason_name(st->exit_reason));
    /* NaN / inf are not JSON: a broken-down unguarded run reports null */
    if (isfinite(st->final_residual)) fprintf(f, " \"final_residual\": %.17g,", st->final_residual);
    else fprintf(f, " \"final_residual\": null,");
    if (isfinite(st->rsold0)) fprintf(f, " \"initial_rr\": %.17g,\n", st->rsold0);
    else fprintf(f, " \"initial_rr\": null,\n");
    fprintf(f, " \"solve_seconds\": %.9f, \"loop_ms\": %.6f, \"init_ms\": %.6f, \"h2d_seconds\": %.9f, \"total_cpu_seconds\": %.6f,\n",
            st->solve_seconds, st->loop_ms, st->init_ms, st->h2d_seconds, total_s);
    fprintf(f, " \"iterations_per_second\": %.6f, \"gemv_bytes_per_gpu\": %.0f, \"iteration_gbs_per_gpu\": %.3f,\n",
            it_s, st->gemv_bytes, st->gemv_bytes * it_s * 1.0e-9);
    fprintf(f, " \"exchange\": \"%s\", \"graph_chunk\": %d, \"kernel_launches\": %lld,\n",
            st->comm == CGB_COMM_PEER ? "peer" : st->comm == CGB_COMM_NCCL ? "nccl" : "none", st->graph_chunk,
            (long long)st->kernel_launches);
    fprintf(f, " \"gemv\": {\"variant\": %d, \"rows\": %d, \"unroll\": %d, \"threads\": %d, \"vec\": %d, \"grid\": %d}}\n",
            st->gemv_variant, st->gemv_rows, st->gemv_unroll, st->gemv_threads, st->gemv_vec, st->gemv_grid);
    return fclose(f) == 0 ? 0 : -1;
}

/* The generated system as the text files generateSPDmatrix.m:23-43 writes (the row blocks come
 * from the on-device generator, a slab of rows at a time). */
static int emit_files(const cli_t *cli, int64_t n, unsigned flags)
{
    char path[4200];
    cgb_ctx *ctx = NULL;
    if (cgb_create(&ctx, 0) != CGB_OK) return -1;
    int64_t slab = ((int64_t)1 << 25) / n;                 /* <= 256 MiB of doubles per slab */
    if (slab < 1) slab = 1;
    if (slab > n) slab = n;
    double *rows = malloc((size_t)slab * (size_t)n * sizeof(double));
    double *vec = malloc((size_t)n * sizeof(double));
    int rc = (rows && vec) ? 0 : -1;
    snprintf(path, sizeof path, "%s_matrix.txt", cli->emit);
    for (int64_t r0 = 0; r0 < n && rc == 0; r0 += slab) {
        const int64_t cnt = n - r0 < slab ? n - r0 : slab;
        if (cgb_op_generate(ctx, n, r0, cnt, cli->seed, flags, rows, NULL) != CGB_OK ||
            cgb_write_values_mode(path, rows, cnt * n, 4, r0 != 0) != CGB_OK) rc = -1;
    }
    if (rc == 0) {
        snprintf(path, sizeof path, "%s_vector.txt", cli->emit);
        if (cgb_op_generate(ctx, n, 0, n, cli->seed, flags, NULL, vec) != CGB_OK ||
            cgb_write_values(path, vec, n, 4) != CGB_OK) rc = -1;
    }
    if (rc == 0) {
        snprintf(path, sizeof path, "%s_x0.txt", cli->emit);
        for (int64_t i = 0; i < n; ++i) vec[i] = 0.0;      /* generateSPDmatrix.m:8 */
        if (cgb_write_values(path, vec, n, 1) != CGB_OK) rc = -1;
    }
    if (rc == 0) {
        snprintf(path, sizeof path, "%s_dimensions.txt", cli->emit);
        FILE *f = fopen(path, "w");
        if (!f) rc = -1;
        else { fprintf(f, "%lld\n%lld\n%lld\n1\n", (long long)n, (long long)n, (long long)n); fclose(f); }
    }
    free(rows); free(vec);
    cgb_destroy(ctx);
    return rc;
}

/* dimensions.txt lookup: --dims, ./dimensions.txt, <dir of A>/dimensions.txt */
static int find_dimensions(const cli_t *c, int64_t dims[4], int *inferred)
{
    *inferred = 0;
    if (c->dims) return cgb_read_dimensions(c->dims, dims);
    if (cgb_read_dimensions("dimensions.txt", dims) == CGB_OK) return CGB_OK;
    char tmp[4096], path[4200];
    snprintf(tmp, sizeof tmp, "%s", c->fileA);
    snprintf(path, sizeof path, "%s/dimensions.txt", dirname(tmp));
    if (cgb_read_dimensions(path, dims) == CGB_OK) return CGB_OK;
    int64_t nb = cgb_count_values(c->fileB);
    if (nb < 1) return CGB_ERR_IO;
    dims[0] = dims[1] = dims[2] = nb; dims[3] = 1;
    *inferred = 1;
    return CGB_OK;
}

int main(int argc, char **argv)
{
    const clock_t c0 = clock();                      /* parallel_cg.c:75 */
    /* stdout carries the reference's lines only: NCCL's banner (NCCL_DEBUG=VERSION/INFO) goes to stderr */
    setenv("NCCL_DEBUG_FILE", "/dev/stderr", 0);
    cli_t cli;
    int pc = parse_cli(argc, argv, &cli);
    if (pc == -2) {
        /* serialConjugate.c:48-52 (message kept, exit status made non-zero) */
        printf("%s requires four (4) files \n", "cg_b200");
        return 2;
    }
    if (pc != 0) return 2;

    cgb_options opt;
    cgb_default_options(&opt);
    opt.eps = cli.eps;
    opt.max_iter = cli.max_iter;
    opt.gemv_variant = cli.gemv;
    opt.comm = cli.comm;
    opt.guard = cli.guard;
    opt.symmetric = cli.symmetric;
    opt.cg_variant = cli.cg_variant;
    if (cli.bench_iters > 0) { opt.ignore_stop = 1; opt.max_iter = cli.bench_iters; }

    cgb_stats st;
    memset(&st, 0, sizeof st);
    double *x = NULL;
    int64_t n = 0;
    int rc;
    int saved_stdout = -1;
    /* While the solver runs, fd 1 points at stderr: whatever a library underneath prints
     * (NCCL's version banner with NCCL_DEBUG=VERSION) cannot land between the reference's lines. */
#define SOLVER_STDOUT_OFF() do { fflush(stdout); saved_stdout = dup(1); if (saved_stdout >= 0) dup2(2, 1); } while (0)
#define SOLVER_STDOUT_ON()  do { fflush(stdout); if (saved_stdout >= 0) { dup2(saved_stdout, 1); close(saved_stdout); saved_stdout = -1; } } while (0)

    /* n * n is printed and sized below: keep it inside 63 bits (no GPU set holds such a matrix anyway) */
#define MAX_ORDER 2147483647LL
    if (cli.gen_n > 0) {
        n = cli.gen_n;
        if (n > MAX_ORDER) { fprintf(stderr, "cg_b200: matrix order %lld is out of range\n", (long long)n); return 1; }
        if (n % cli.gpus != 0) {                     /* parallel_cg.c:86-90 */
            printf("%lld is not divisible by %d\n", (long long)n, cli.gpus);
            return 1;
        }
        printf("Computing cg of matrix size : %lld\n", (long long)n * (long long)n);
        x = malloc((size_t)n * sizeof(double));
        if (!x) { printf("can't allocate memory for vector\n"); return 1; }
        const unsigned gflags = (cli.round4 ? CGB_GEN_ROUND4 : 0u) | CGB_GEN_KAPPA_LOG2(cli.kappa_log2) | cli.family;
        if (cli.emit && emit_files(&cli, n, gflags) != 0) {
            fprintf(stderr, "cg_b200: cannot write the %s_* files: %s\n", cli.emit, cgb_last_error());
            return 1;
        }
        SOLVER_STDOUT_OFF();
        rc = cgb_solve_generated(n, cli.seed, gflags, cli.gpus, &opt, x, &st);
        SOLVER_STDOUT_ON();
    } else {
        int64_t dims[4];
        int inferred = 0;
        if (find_dimensions(&cli, dims, &inferred) != CGB_OK) {
            printf("Could not open file\n");          /* serialConjugate.c:103 */
            fprintf(stderr, "cg_b200: no usable dimensions.txt and cannot size %s\n", cli.fileB);
            return 1;
        }
        if (inferred) fprintf(stderr, "cg_b200: dimensions.txt not found; n = %lld from %s\n", (long long)dims[0], cli.fileB);
        if (dims[0] != dims[1]) {                     /* serialConjugate.c:53-57 */
            printf("%lld and %lld must be same size\n", (long long)dims[0], (long long)dims[1]);
            return 1;
        }
        n = dims[0];
        if (n > MAX_ORDER) { fprintf(stderr, "cg_b200: matrix order %lld is out of range\n", (long long)n); return 1; }
        if (dims[2] != n || dims[3] != 1) {
            fprintf(stderr, "cg_b200: dimensions of b (%lld x %lld) do not match A\n", (long long)dims[2], (long long)dims[3]);
            return 1;
        }
        if (n % cli.gpus != 0) {
            printf("%lld is not divisible by %d\n", (long long)n, cli.gpus);
            return 1;
        }
        printf("Computing cg of matrix size : %lld\n", (long long)n * (long long)n);   /* :58, 64-bit */
        x = malloc((size_t)n * sizeof(double));
        if (!x) { printf("can't allocate memory for vector\n"); return 1; }             /* :80 */
        /* initialize() x3 + MPI_Scatter (parallel_cg.c:105-115): every GPU ingests the bytes of
         * its own rows and parses them on the device; the matrix never exists on the host */
        SOLVER_STDOUT_OFF();
        rc = cgb_conjugrad_files(cli.fileA, cli.fileB, cli.fileX, n, cli.gpus, &opt, x, &st);
        SOLVER_STDOUT_ON();
        if (rc == CGB_ERR_IO) {
            if (strstr(cgb_last_error(), "Could not open file")) printf("Could not open file\n");   /* :103 */
            fprintf(stderr, "cg_b200: %s\n", cgb_last_error());
            return 1;
        }
    }
    if (rc != CGB_OK) {
        fprintf(stderr, "cg_b200: %s\n", cgb_last_error());
        return 1;
    }
    if (cli.print_x) cgb_print_vector(stdout, x, n, 1);
    if (cli.gpus == 1) {
        printf("average clock execution time in seconds: %f\n", st.solve_seconds);       /* serialConjugate.c:249-251 */
    } else {
        printf("cg method execution time in seconds: %f\n", st.solve_seconds);           /* parallel_cg.c:334-335 */
        printf("collective data distribution time in seconds: %f\n", st.h2d_seconds);    /* :123-124 */
        printf("clock execution time in seconds: %f\n", (double)(clock() - c0) / CLOCKS_PER_SEC); /* :125-126 */
    }
    if (cli.stats || cli.bench_iters > 0) {
        printf("iterations: %lld\n", (long long)st.iterations);
        printf("loop index at break: %lld\n", (long long)st.break_k);
        printf("final residual sqrt(r.r): %.6e\n", st.final_residual);
        printf("exit: %s\n", exit_reason_name(st.exit_reason));
        if (cli.gpus > 1) printf("exchange: %s\n", st.comm == CGB_COMM_PEER ? "peer (fused NVLink pushes)" : "nccl");
        if (st.iterations > 0 && st.loop_ms > 0.0) {
            double it_s = 1.0e3 * (double)st.iterations / st.loop_ms;
            printf("iterations per second: %.3f\n", it_s);
            printf("iteration-level GEMV GB/s per GPU (8*(m*n+n+m) bytes / iteration time): %.1f\n",
                   st.gemv_bytes * it_s * 1.0e-9);
        }
    }
    if (cli.record && write_run_record(cli.record, &cli, &st, (double)(clock() - c0) / CLOCKS_PER_SEC) != 0) {
        fprintf(stderr, "cg_b200: cannot write %s\n", cli.record);
        return 1;
    }
    if (cli.write_x && cgb_write_values(cli.write_x, x, n, -1) != CGB_OK) {
        fprintf(stderr, "cg_b200: cannot write %s\n", cli.write_x);
        return 1;
    }
    free(x);
    cgb_shutdown();
    return 0;
}
