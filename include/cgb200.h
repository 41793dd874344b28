/* cgb200.h -- C ABI of libcgb200.so: a B200-native (sm_100a) dense FP64
 * conjugate-gradient solver behind the drop-in surface of
 * mawunyega/conjugate_gradient.
 *
 * The reference has no plugin/FFI layer: its boundary is a C function,
 * conjugrad(), compiled into each program, plus a three-file CLI.  Every entry
 * point below states the reference interface it replaces (file:line are into
 * the reference repository).  Plain C: pointers, sizes, POD structs; no C++ or
 * torch types cross this boundary.  There is no CPU fallback: every call that
 * computes needs a CUDA device and fails with CGB_ERR_CUDA otherwise.
 *
 * Element type: the reference is `float` throughout (serialConjugate.c:33-41,
 * parallel_cg.c:34-52); this library is `double` (the FP64 target of the
 * north star).  Problem size: the reference bakes n in as `#define ROWS/COLS`
 * (serialConjugate.c:29-30, parallel_cg.c:30-31); here n is a run-time
 * argument (or dimensions.txt for the CLI).
 */
#ifndef CGB200_H
#define CGB200_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CGB_VERSION 100

/* status codes (the reference returns nothing and exit(0)s / MPI_Aborts on
 * failure, serialConjugate.c:48-57,78-82, parallel_cg.c:76-95; a library must not) */
#define CGB_OK                 0
#define CGB_ERR_ARG           -1   /* bad argument / null pointer / n < 1 or > 2^24  */
#define CGB_ERR_CUDA          -2   /* CUDA runtime error or no device                 */
#define CGB_ERR_NCCL          -3   /* NCCL error                                      */
#define CGB_ERR_NOT_DIVISIBLE -4   /* n % ranks != 0   (parallel_cg.c:86-90)          */
#define CGB_ERR_NOMEM         -5   /* host or device allocation failed (:140-144)     */
#define CGB_ERR_STATE         -6   /* no problem loaded / wrong call order            */
#define CGB_ERR_IO            -7   /* text input missing, short or malformed          */
#define CGB_ERR_COMM          -8   /* a peer rank stopped answering during a solve    */

#define CGB_EPSILON 1.0e-6         /* serialConjugate.c:28, parallel_cg.c:29          */

/* generator flags */
#define CGB_GEN_ROUND4 1u          /* mimic generateSPDmatrix.m's "%.4f" round trip   */
/* Harder SPD family for iteration-bound runs: the n*I shift of generateSPDmatrix.m:17
 * becomes n*diag(1 + (2^k - 1) t_i^2), t_i = i/(n-1), k in 1..63 (condition number ~2^k:
 * k = 4 -> ~35 CG iterations, 7 -> ~100, 10 -> ~270 at eps = 1e-6).  k = 0: the reference's. */
#define CGB_GEN_KAPPA_LOG2(k) (((unsigned)(k) & 0x3fu) << 8)
/* A dense matrix with a PRESCRIBED spectrum: A = H diag(lambda) H, H = I - 2 v v' a Householder
 * reflector from the seeded stream, lambda_i = n (1 + (2^k - 1) t_i^2) with k from
 * CGB_GEN_KAPPA_LOG2 -- the eigenvalues are exactly the lambda_i, so the iteration count of CG is a
 * property of k alone (k = 10: ~230 iterations).  Exactly symmetric, bit-identical to the CPU twin. */
#define CGB_GEN_SPECTRUM 2u
/* The reference's commented alternative `A = A*A'` (generateSPDmatrix.m:12) followed by its
 * `A = A + n*eye(n)` (:17): A = R R' + n diag(d), R = rand(n, n) from the seeded stream, d the
 * spread of CGB_GEN_KAPPA_LOG2 (1 for k = 0).  An FP64 contraction of 2 n^3 flops, done on the
 * FP64 tensor pipe; exactly symmetric; equal to the CPU twin to ~1e-15 relative per entry (the
 * summation order of the tensor instruction is the hardware's). */
#define CGB_GEN_GRAM 4u

/* GEMV kernel selection (0 = library default) */
#define CGB_GEMV_AUTO 0
#define CGB_GEMV_LDG  1            /* 128/256-bit streaming loads, registers          */
#define CGB_GEMV_BULK 2            /* cp.async.bulk (TMA engine) shared-memory ring   */
#define CGB_GEMV_SYMV 3            /* reported in cgb_stats when cgb_options.symmetric ran */

/* How the ranks exchange p and the two dot scalars each iteration (what MPI_Allgather /
 * MPI_Allreduce do in parallel_cg.c:290-294,313); every rank must pass the same value */
#define CGB_COMM_AUTO 0            /* PEER when every rank could map every window, else NCCL */
#define CGB_COMM_NCCL 1            /* ncclAllGather + 2 x ncclAllReduce between the kernels   */
#define CGB_COMM_PEER 2            /* fused: the kernels push over NVLink into peer memory    */

/* Recurrence.  CLASSIC is the reference's (serialConjugate.c:213-245): three
 * synchronisation points per iteration (p, p.Ap, r.r).  FUSED_REDUCTIONS is Chronopoulos-Gear
 * CG -- the same iterates in exact arithmetic, computed as w = A r, gamma = r.r, delta = r.w,
 * beta = gamma/gamma_prev, alpha = gamma/(delta - beta gamma/alpha_prev), p = r + beta p,
 * s = w + beta s, x += alpha p, r -= alpha s -- with two kernels and ONE exposed reduction per
 * iteration; same stop rule, same exit state (x, r updated).  One extra matVec per solve. */
#define CGB_CG_CLASSIC 0
#define CGB_CG_FUSED_REDUCTIONS 1

typedef struct cgb_ctx cgb_ctx;    /* one per rank (= per GPU); owns all device state */

typedef struct cgb_options {
    double  eps;           /* stop when sqrt(r.r) < eps; <= 0 -> CGB_EPSILON (parallel_cg.c:314)  */
    int64_t max_iter;      /* <= 0 -> n, the reference's loop bound (parallel_cg.c:288)            */
    int     ignore_stop;   /* 1: fixed-iteration timing mode, the test is evaluated and recorded   */
                           /*    but not acted on (not in the reference)         */
    int     use_graph;     /* 1 (default): iterations are replayed from a captured CUDA graph      */
    int     check_every;   /* iterations per graph launch between host polls of the stop flag;     */
                           /* 0 -> chosen from the estimated iteration time; rounded up to even    */
    int     time_gemv;     /* 1: stream mode with a CUDA-event pair around every GEMV launch       */
    int     gemv_variant;  /* CGB_GEMV_*                                                           */
    int     gemv_rows;     /* rows per CTA (0 auto)                                                */
    int     gemv_unroll;   /* LDG: load steps in flight per thread (0 auto)                        */
    int     gemv_threads;  /* threads per CTA (0 auto)                                             */
    int     gemv_vec;      /* LDG: doubles per load 1/2/4 (0 auto)                                 */
    int     gemv_stages;   /* BULK: ring depth (0 auto)                                            */
    int     gemv_tile_cols;/* BULK: columns per stage (0 auto)                                     */
    int     comm;          /* CGB_COMM_* (ranks > 1)                                               */
    int     guard;         /* 1: end the solve instead of iterating NaNs when p.Ap = 0 (alpha not  */
                           /*    finite) or the initial residual is exactly 0.  Default 0 mirrors  */
                           /*    the reference, which has no such test (parallel_cg.c:298,318)     */
    int     gemv_l2_keep_pct; /* small n (A within ~0.8..3 L2 sizes): % of L2 in which the first rows  */
                           /* of A are kept (evict_last) from one iteration to the next; 0 = the   */
                           /* library default, < 0 = plain streaming                               */
    int     symmetric;     /* 1: matVec reads each pair a_ij, a_ji once (half the bytes, SURVEY.md  */
                           /*    8f-4).  Needs a_ij == a_ji bit for bit -- verified on the device   */
                           /*    once per loaded matrix on one rank (CGB_ERR_ARG otherwise);        */
                           /*    several ranks (peer exchange only) accept generated systems, or    */
                           /*    2 = the caller asserts symmetry (cgb_conjugrad* then upload only   */
                           /*    the tiles the product reads, see cgb_load_host_symmetric).          */
                           /*    n / ranks a multiple of 512.                                        */
                           /*    0 (default): the reference's full row products                     */
    int     cg_variant;    /* CGB_CG_*                                                              */
} cgb_options;

#define CGB_EXIT_STOP_RULE 1       /* sqrt(r.r) < eps                 (parallel_cg.c:314-317)  */
#define CGB_EXIT_MAX_ITER  2       /* loop bound reached              (parallel_cg.c:288)      */
#define CGB_EXIT_BREAKDOWN 3       /* guard: alpha = rho/p.Ap not finite                       */
#define CGB_EXIT_EXACT_X0  4       /* guard: initial residual exactly zero                     */

typedef struct cgb_stats {
    int64_t n, m;              /* order; rows owned by this rank (n / ranks, parallel_cg.c:83)     */
    int     rank, ranks;
    int64_t iterations;        /* x/r updates applied = loop index at break + 1                    */
    int64_t break_k;           /* 0-based loop index at `break` (parallel_cg.c:314-317), -1 if the */
                               /* loop ran out                                                     */
    int64_t first_hit_k;       /* first k at which sqrt(beta) < eps held, -1 if never              */
    double  final_beta;        /* last r.r                                                         */
    double  final_residual;    /* sqrt(final_beta): what the stop rule compares with eps           */
    double  rsold0;            /* initial r.r                                                      */
    double  solve_seconds;     /* host wall clock around the solver region, where the reference    */
                               /* times it (serialConjugate.c:208,249; parallel_cg.c:278-279,328-329) */
    double  loop_ms;           /* CUDA-event time of the iteration loop on the solver stream       */
    double  init_ms;           /* CUDA-event time of A*x0, residual, initial r.r                   */
    double  h2d_seconds;       /* last cgb_load_host                                               */
    double  d2h_seconds;       /* last cgb_read_x                                                  */
    double  gemv_ms_avg;       /* time_gemv=1: mean / min CUDA-event duration of the loop's GEMVs  */
    double  gemv_ms_min;
    int64_t gemv_timed;        /* number of GEMV launches in that mean                             */
    double  gemv_bytes;        /* algorithmic bytes per GEMV launch on this rank: 8*(m*n + n + m)  */
    int64_t kernel_launches;   /* kernels of this library enqueued by the last solve (incl. those  */
                               /* replayed from graphs; NCCL kernels not counted)                  */
    int     gemv_variant, gemv_rows, gemv_unroll, gemv_threads, gemv_vec, gemv_stages,
            gemv_tile_cols, gemv_grid;
    int     graph_chunk;       /* iterations per captured graph (0: stream mode)                   */
    int     comm;              /* CGB_COMM_NCCL / CGB_COMM_PEER as used (0 with one rank)          */
    int     exit_reason;       /* CGB_EXIT_*                                                       */
    int     cg_variant;        /* CGB_CG_* as run                                                  */
    int64_t gemv_keep_rows;    /* rows of A held in L2 across iterations (0: plain streaming)      */
    int64_t h2d_bytes;         /* bytes the last cgb_load_host / cgb_load_host_symmetric copied    */
    int     pdl;               /* 1: the iteration kernels overlapped their launches (programmatic */
                               /* dependent launch; one rank, CGB_PDL=0 in the environment: off)   */
    int     h2d_path;          /* CGB_H2D_*: how the last upload of A reached the GPU              */
} cgb_stats;

/* cgb_stats.h2d_path: the caller's A may be pinned (cudaHostAlloc / cudaHostRegister) or plain
 * malloc memory as in the reference's main() (memAllocate(), serialConjugate.c:62-64,76-84) */
#define CGB_H2D_NONE       0       /* nothing uploaded (generated on the device, or text ingest)  */
#define CGB_H2D_DIRECT     1       /* pinned source: DMA straight out of the caller's buffer      */
#define CGB_H2D_STAGED     2       /* pageable source: host threads pack it into a ring of pinned */
                                   /* slots, one async copy per slot (cg_stage.c)                 */
#define CGB_H2D_DRIVER     3       /* pageable source left to the driver (small, or CGB_STAGE=0)  */
#define CGB_H2D_REGISTERED 4       /* pageable source pinned in place piece by piece (CGB_STAGE=register) */

/* ---- debugging -------------------------------------------------------------- */
/* With CGB_DEBUG_GUARDS=1 in the environment at cgb_create*(), every device array of the
 * context lies between two 256-byte guard bands; this reads them back after the stream has
 * drained.  CGB_OK: untouched.  CGB_ERR_STATE: a kernel wrote outside an array (the message
 * names it), or the bands are off. */
int cgb_debug_check_guards(cgb_ctx *ctx);
/* Self test of the above (refused without CGB_DEBUG_GUARDS=1): stores one double at element
 * `offset` of array `what` (1 b, 2 x, 3 r, 5 Ap), -32 <= offset < length + 32. */
int cgb_debug_poke(cgb_ctx *ctx, int what, int64_t offset, double value);
/* The NCCL-id rendezvous of conjugrad_mpi() alone (tests): rank 0 makes a 128-byte id, every rank of
 * the np callers receives it -- through the file named by CGB_ID_FILE when set (ranks = processes; a file
 * left by another launch is ignored), else through a process-wide slot (ranks = threads). */
int cgb_debug_rendezvous_id(int rank, int np, void *id128);
/* With CGB_PROBE=1 in the environment of a solve on several ranks (peer-memory exchange), the
 * exchange-fused kernels leave per-CTA time stamps of their last launch: when a CTA started,
 * finished its own column block, saw the peers' flags, ended, and how long it polled.  Copies up
 * to max_words 64-bit words (layout: 8 words per CTA, see tools/probe_exchange.py) and returns
 * the number copied, or a negative status.  Measurement aid for the part of the path ncu cannot
 * replay (a replayed pass would wait for pushes the peers never repeat). */
int64_t cgb_debug_probe_read(cgb_ctx *ctx, unsigned long long *dst, int64_t max_words);

/* ---- library / device ------------------------------------------------------ */
int         cgb_version(void);
int         cgb_device_count(void);               /* < 0: CUDA unusable                */
const char *cgb_last_error(void);                 /* message of this thread's last failure */
void        cgb_default_options(cgb_options *o);

/* ---- contexts -------------------------------------------------------------- */
/* One rank on one GPU, no communicator: the serial program (serialConjugate.c). */
int cgb_create(cgb_ctx **out, int device);
/* SPMD rank `rank` of `ranks` on GPU `device` (parallel_cg.c:81-83: MPI_Comm_rank /
 * MPI_Comm_size).  `id` is the 128-byte NCCL unique id made by cgb_unique_id() on
 * one rank and handed to all others by the caller's launcher (torchrun, MPI, a
 * file, threads).  Collective: every rank must call it. */
int cgb_unique_id(void *id128);
int cgb_create_rank(cgb_ctx **out, int device, int rank, int ranks, const void *id128);
int cgb_destroy(cgb_ctx *ctx);

/* ---- problem set-up -------------------------------------------------------- */
/* What MPI_Bcast(x0) + MPI_Scatter(A) + MPI_Scatter(b) deliver to a rank
 * (parallel_cg.c:111-115): the rank's m = n/ranks rows of A (row-major, m x n),
 * its m entries of b, and the FULL initial guess x0 (n).  Host buffers,
 * caller-owned, copied to the GPU.  Pinned buffers are read in place by the DMA engine; plain
 * malloc memory (what the reference's main() passes) goes through a multi-threaded ring of
 * pinned staging slots so that it, too, moves at PCIe rate (cgb_stats.h2d_path says which). */
int cgb_load_host(cgb_ctx *ctx, int64_t n, const double *A_rows, const double *b_rows,
                  const double *x0_full);
/* cgb_load_host for a matrix the caller asserts to be symmetric (a_ij == a_ji bit for bit, as
 * cgb_options.symmetric = 2): only the 512 x 512 tiles the symmetric product reads are copied
 * -- on one rank those on and above the tile diagonal, n (n + 512) / 2 values; entries below
 * are never referenced (BLAS dsymv's uplo = 'U', at tile granularity).  Same buffers and
 * layout as cgb_load_host (the full row block; the unread part is simply skipped).  The
 * problem must then be solved with cgb_options.symmetric set (CGB_ERR_STATE otherwise).
 * CGB_ERR_ARG unless n / ranks is a multiple of 512 (several ranks: peer-memory exchange). */
int cgb_load_host_symmetric(cgb_ctx *ctx, int64_t n, const double *A_rows, const double *b_rows,
                            const double *x0_full);
/* The same three inputs as the reference's TEXT files (initialize(), parallel_cg.c:147-168,
 * "%f%*c" token rules): the rank locates the bytes of its own m rows in matrixA.txt, streams
 * them through pinned memory and the GPU tokenises and converts them (bit-identical to
 * strtod; tokens outside the exact fast path are resolved by strtod on the host).  The full
 * matrix is never held on the host.  vectorb.txt / initialguess.txt: n values each.
 * Short or missing files are CGB_ERR_IO. */
int cgb_load_text(cgb_ctx *ctx, int64_t n, const char *fileA, const char *fileB, const char *fileX);
/* North-star item 4: the rank's rows of the seeded SPD system generated in place
 * on the GPU (generateSPDmatrix.m:4-17): A = 0.5(R+R') + nI, b = rand, x0 = 0. */
int cgb_generate(cgb_ctx *ctx, int64_t n, uint64_t seed, unsigned flags);
/* Replace only the initial guess (n doubles, host) of a loaded problem, e.g. to
 * restart from a previous solution (the reference's x0 file is its restart input). */
int cgb_set_x0(cgb_ctx *ctx, const double *x0_full);

/* ---- the solver ------------------------------------------------------------ */
/* conjugrad() on the resident problem (serialConjugate.c:180-259 /
 * parallel_cg.c:248-345): r = p = b - A x0; loop { Ap; alpha = rho/(p.Ap);
 * x += alpha p; r -= alpha Ap; beta = r.r; if sqrt(beta) < eps break;
 * p = r + (beta/rho) p; rho = beta }.  Collective over the ranks.  x stays on
 * the device (replicated on every rank on return, as in the reference). */
int cgb_solve(cgb_ctx *ctx, const cgb_options *opt, cgb_stats *stats);
/* Full solution vector (n doubles) to a host buffer. */
int cgb_read_x(cgb_ctx *ctx, double *x_full);

/* ---- one-call drop-ins ----------------------------------------------------- */
/* Runtime-n equivalent of `void conjugrad(float *matrixA, float *vectorB, float
 * *vectorX)` (serialConjugate.c:40,180): host A (n x n), b, x; x is the initial
 * guess on entry and the solution on exit.  `gpus` > 1 row-shards A over that
 * many GPUs of this node from one process (one host thread per GPU). */
int cgb_conjugrad_host(const double *A, const double *b, double *x, int64_t n,
                       double eps, int64_t max_iter, int gpus, cgb_stats *stats);
/* Same, with the full option block (fixed-iteration mode, kernel selection). */
int cgb_conjugrad_host_opt(const double *A, const double *b, double *x, int64_t n, int gpus,
                           const cgb_options *opt, cgb_stats *stats);
/* North-star item 4 end to end: generate the seeded SPD system of order n on
 * `gpus` GPUs (each its own row block, generateSPDmatrix.m:4-17), solve, and
 * return x (n doubles, may be NULL).  No text files, no host copy of A. */
int cgb_solve_generated(int64_t n, uint64_t seed, unsigned flags, int gpus,
                        const cgb_options *opt, double *x, cgb_stats *stats);
/* The reference's whole program minus its printing: `prog matrixA.txt vectorb.txt
 * initialguess.txt` (serialConjugate.c:62-68; rank-0 read + MPI_Scatter of parallel_cg.c:
 * 100-118) on `gpus` GPUs through cgb_load_text(); x (n doubles, may be NULL) receives the
 * solution.  stats->h2d_seconds is the text ingest ("data distribution") time. */
int cgb_conjugrad_files(const char *fileA, const char *fileB, const char *fileX, int64_t n, int gpus,
                        const cgb_options *opt, double *x, cgb_stats *stats);
/* Releases the contexts cached by the one-call entry points (GPU memory, NCCL
 * communicators, graphs).  The reference frees its temporaries at the end of
 * every conjugrad() (serialConjugate.c:253-258); here it is explicit. */
int cgb_shutdown(void);
/* Same for a rank of an SPMD program, replacing `void conjugrad(float
 * *local_matrixA, float *local_vectorB, float *local_vectorX, int local_row, int
 * myrank, int procsnum)` (parallel_cg.c:46-48,248-250) on an existing rank
 * context: local_x is the FULL vector (n = local_row * ranks), in/out. */
int cgb_conjugrad_rank(cgb_ctx *ctx, const double *local_A, const double *local_b,
                       double *local_x, int64_t local_row, const cgb_options *opt,
                       cgb_stats *stats);

/* Source-compatible wrappers with the reference's exact names and argument
 * lists (element type double).  n comes from cgb_set_dimension(), the run-time
 * replacement of `#define ROWS` (or from dimensions.txt through the CLI).
 * They print what the reference prints from inside conjugrad():
 *   "average clock execution time in seconds: %f\n"  (serialConjugate.c:249-251)
 *   "cg method execution time in seconds: %f\n"      (parallel_cg.c:334-335, rank 0)
 * and report failures on stderr (the reference would exit/abort). */
void cgb_set_dimension(int64_t n);
void conjugrad(double *matrixA, double *vectorB, double *vectorX);
void conjugrad_mpi(double *local_matrixA, double *local_vectorB, double *local_vectorX,
                   int local_row, int myrank, int procsnum);
/* Stats of the calling thread's last conjugrad()/conjugrad_mpi() call. */
int  cgb_last_stats(cgb_stats *out);
/* Releases the calling thread's conjugrad_mpi() rank context (collective). */
int  cgb_release_rank(void);

/* ---- text inputs and output of the reference CLI --------------------------- */
/* dimensions.txt: shipped by the reference (dimensions.txt:1-4) but opened by
 * none of its programs; here it is the run-time ROWS/COLS: four integers, one
 * per line: rows(A) cols(A) rows(b) cols(b). */
int     cgb_read_dimensions(const char *path, int64_t dims[4]);
/* initialize() (serialConjugate.c:85-105, parallel_cg.c:147-168): `count`
 * numbers, row-major, "%f%*c" token rules (any whitespace or one separator
 * character between numbers; a UTF-8 BOM is skipped).  Unlike the reference,
 * a missing file or a short file is an error (CGB_ERR_IO), not garbage. */
int     cgb_read_values(const char *path, double *dst, int64_t count, int64_t *got);
int64_t cgb_count_values(const char *path);      /* numeric tokens in a file, < 0 on error */
/* generateSPDmatrix.m:28-43 file format: one value per line, "%.<d>f"
 * (4 for A and b, 1 for x0); decimals < 0 writes round-trippable "%.17g". */
int     cgb_write_values(const char *path, const double *v, int64_t count, int decimals);

/* ---- kernel-level operators (parity tests, profiling) ---------------------- */
/* Each runs ONE product kernel on host data: upload, launch, download. */
/* matVec (parallel_cg.c:172-184) with the fused vecVec(w, y) (:293):
 * y = A v for an m x n block; if w != NULL *dot = sum_i w[i]*y[i]. */
int cgb_op_matvec(cgb_ctx *ctx, const double *A, const double *v, const double *w,
                  double *y, double *dot, int64_t m, int64_t n, const cgb_options *opt);
/* residual twice + vecVec (parallel_cg.c:284-286): r = p = b - Ax, *rr = r.r */
int cgb_op_init_residual(cgb_ctx *ctx, const double *b, const double *Ax, double *r,
                         double *p, double *rr, int64_t m);
/* scalarVec/vecAdd/vecSub/vecVec (parallel_cg.c:298-309): alpha = rho/pAp,
 * x += alpha p, r -= alpha Ap, *rr = r.r (x, r in/out). */
int cgb_op_xr_update(cgb_ctx *ctx, double rho, double pAp, double *x, double *r,
                     const double *p, const double *Ap, double *rr, int64_t m);
/* stop rule + scalarVec/vecAdd (parallel_cg.c:314-322): returns *stopped = 1 and
 * leaves p untouched if sqrt(beta) < eps, else p = r + (beta/rho) p (p in/out). */
int cgb_op_p_update(cgb_ctx *ctx, double rho, double beta, double eps, double *p,
                    const double *r, int *stopped, int64_t m);
/* Rows [row0,row0+m) of the generated A (m x n) and b to host buffers (either may be NULL). */
int cgb_op_generate(cgb_ctx *ctx, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                    unsigned flags, double *A_rows, double *b_rows);

/* ---- no-allocation entry point ---------------------------------------------- */
/* No counterpart in the reference: added for callers that
 * cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid
 * it).  The matVec with the fused dot of cgb_op_matvec -- y = A v for an m x n row-major block, and
 * *d_dot = sum_i w[i] y[i] when d_w / d_dot are given (both, or neither) -- on DEVICE pointers of
 * the current device, enqueued on `stream` (asynchronous: results are ready when the stream is).
 * It calls no cudaMalloc / cudaFree / allocating library routine and copies nothing: the kernel's
 * scratch (one dot partial per CTA, ticket, flags) is a static device array of
 * 512 + 8 * CGB_NOALLOC_MAX_CTAS bytes (1 MiB) that every GPU loading the library holds until
 * exit.  Limit: m <= CGB_NOALLOC_MAX_ROWS, else CGB_ERR_ARG with a message.  All calls share the
 * array: the part that must start at zero is cleared on `stream` before each launch, and two calls
 * that overlap on one GPU (different streams or threads) race -- serialise them. */
#define CGB_NOALLOC_MAX_CTAS 131072
#define CGB_NOALLOC_MAX_ROWS ((int64_t)262144)      /* at least 2 rows per CTA for every shape the plan picks */
int cgb_matvec_noalloc(const double *d_A, const double *d_v, const double *d_w, double *d_y,
                       double *d_dot, int64_t m, int64_t n, void *stream);

/* ---- measurement helpers --------------------------------------------------- */
/* Times `reps` back-to-back launches of the GEMV on the resident problem
 * (v = current direction vector) with CUDA events on the solver stream, after
 * `warmup` launches; A must exceed L2 or flush_l2 != 0.  Outputs ms per launch. */
int cgb_time_gemv(cgb_ctx *ctx, const cgb_options *opt, int warmup, int reps, int flush_l2,
                  double *ms_avg, double *ms_min);
/* Copies rows/entries of resident device arrays to the host (tests).
 * what: 0 A rows, 1 b, 2 x (full), 3 r, 4 p (full), 5 Ap. */
int cgb_debug_read(cgb_ctx *ctx, int what, int64_t offset, int64_t count, double *dst);
int cgb_device_info(cgb_ctx *ctx, int *device, int *sm_count, int64_t *free_bytes,
                    int64_t *total_bytes, char *name64);

#ifdef __cplusplus
}
#endif
#endif
