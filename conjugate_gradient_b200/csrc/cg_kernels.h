/* cg_kernels.h -- the thin C layer between the C host solver (cg_solver.c) and
 * the hand-written sm_100a kernels (cg_kernels.cu).  Plain C: device pointers,
 * sizes and a stream; every launcher returns a cudaError_t as int.
 *
 * Reference functions each kernel replaces:
 *   cgk_gemv          matVec   parallel_cg.c:172-184  (+ vecVec(p,Ap) :293 fused)
 *   cgk_init_residual residual parallel_cg.c:188-196 twice (:284-285) + vecVec :286
 *   cgk_xr_update     scalarVec+vecAdd (:299,303), scalarVec+vecSub (:304,308), vecVec :309
 *   cgk_p_update      stop test :314-317, scalarVec+vecAdd :318,322, rsold=beta :323
 *   cgk_generate      generateSPDmatrix.m:4-17
 * Not in the reference: cgk_symv* (symmetric upper-tile product), cgk_cg2_step*
 * (fused-reductions recurrence), cgk_gemv_comm / the <COMM> vector kernels (exchange over peer
 * memory), cgk_ingest_* (text parser).  Launch-latency machinery shared by all of them: deferred
 * reductions and programmatic dependent launch (cgk_set_pdl), both described where declared.
 */
#ifndef CG_KERNELS_H
#define CG_KERNELS_H
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* Device-resident solver scalars: nothing in an iteration returns to the host. */
typedef struct cgk_scalars {
    double rr[2];          /* rr[cur] = rho of the running iteration; rr[cur^1] receives beta */
    double pAp;            /* all-reduced p.Ap                                              */
    double pAp_local;      /* this rank's partial (all-reduce send buffer)                  */
    double rr_local;       /* this rank's r.r partial (all-reduce send buffer)              */
    double final_beta;     /* r.r after the last applied x/r update                         */
    double rsold0;         /* initial r.r                                                   */
    long long iters;       /* x/r updates applied                                           */
    long long first_hit;   /* first loop index k with sqrt(beta) < eps, -1 if none          */
    long long break_k;     /* loop index at which the stop rule fired, -1                   */
    int stop_req;          /* written by ONE thread of cgk_p_update: 1 stop rule fired,      */
                           /* 2 max_iter reached; read by gemv / xr_update of later iterations */
    int stopped;           /* stop_req latched by the next gemv launch; read by p_update, so  */
                           /* no kernel ever reads a flag that is written during its own run  */
    unsigned int ticket_gemv;
    unsigned int ticket_vec;
    /* peer-memory exchange (cgk_peers): per-kernel epoch counters, advanced by the last CTA of
     * each non-stopped launch, so a replayed graph needs no per-iteration launch argument */
    unsigned int gen;      /* solve generation: high bits of every tag, set by cgk_reset      */
    unsigned int ep_gemv, ep_xr, ep_p;
    int comm_error;        /* a spin on a peer's flag ran into CGK_SPIN_TIMEOUT_NS            */
    double cg2_gamma_prev; /* CGB_CG_FUSED_REDUCTIONS: gamma and alpha of the previous iteration      */
    double cg2_alpha_prev;
    int breakdown;         /* guard mode only: 3 = alpha not finite (p.Ap = 0), 4 = r0.r0 = 0;   */
                           /* written by cgk_xr_update / cgk_after_init, read by LATER kernels  */
    /* CGB_CG_FUSED_REDUCTIONS on one rank (deferred reductions, no last-CTA phase): the step of
     * parity c reads slot [c] and CTA 0 writes slot [c ^ 1], so no launch reads what it writes */
    unsigned int cg2_k[2];
    double cg2_gprev[2], cg2_aprev[2];
    unsigned int ticket_p; /* second ticket of cgk_xrp_update (its p phase)                      */
    unsigned int pad_;
} cgk_scalars;

/* ---- peer-memory exchange over NVLink (replaces MPI_Allgather(p) parallel_cg.c:290-291 and the
 * MPI_Allreduce of the two dot scalars :294,313 inside the solver kernels themselves).
 * Every rank owns a window [cgk_mailbox | p_full(n)] that all ranks of the node map (CUDA IPC or
 * same-process peer access).  Producers PUSH: the p-update kernel stores its slice of the new
 * direction straight into every peer's p_full and then posts an epoch tag into the peers'
 * p_flag[rank]; the last CTA of the GEMV / xr-update kernels stores the rank's dot partial into
 * every peer's dot[][][rank] slot in NCCL-LL style (two 8-byte words, each {32 data bits, 32-bit
 * tag}, so data and flag arrive atomically without a fence).  Consumers only ever read LOCAL
 * memory: the next kernel spins on its own mailbox and adds the G partials in rank order, so all
 * ranks get bit-identical sums and take the same `break` (as parallel_cg.c:313-317 relies on). */
#define CGK_MAX_RANKS 8
#define CGK_SPIN_TIMEOUT_NS 20000000000ull      /* 20 s: a dead peer ends the solve with an error */
#define CGK_WINDOW_HEADER_BYTES 4096            /* mailbox at offset 0, p_full at this offset     */

typedef struct cgk_mailbox {
    unsigned long long p_flag[CGK_MAX_RANKS];             /* tag of the last p slice rank g pushed */
    unsigned long long dot[2][2][CGK_MAX_RANKS][2];       /* [0 p.Ap | 1 r.r][parity][src][LL word] */
    unsigned long long c_flag[CGK_MAX_RANKS];             /* symmetric product: tag of the column sums rank g pushed */
} cgk_mailbox;

typedef struct cgk_peers {
    int rank, ranks;
    double *p_full[CGK_MAX_RANKS];      /* every rank's gathered-direction vector, peer-mapped */
    cgk_mailbox *mbox[CGK_MAX_RANKS];   /* every rank's mailbox; mbox[rank] is local           */
    double *crecv[CGK_MAX_RANKS];       /* every rank's [ranks][m] receive area for column sums */
    unsigned long long *probe;          /* NULL, or CGK_PROBE_WORDS words of device memory into which the
                                         * exchange kernels of the classic iteration write per-CTA time stamps
                                         * (CGB_PROBE=1; tools/probe_exchange.py reads them) */
} cgk_peers;
/* probe layout: 8 words per CTA; product CTAs first, then the CTAs of xr_update, then those of p_update.
 * product:   [0] globaltimer at entry, [1] after the own column block, [2] after the flag wait, [3] before the
 *            epilogue, [4] SM clocks spent in the wait (warp 0), [5] %smid, [6] epoch, [7] 1
 * xr / p:    [0] entry, [1] after the mailbox gather, [2] end, [4] SM clocks in the gather, [5] %smid, [6] epoch, [7] 1 */
#define CGK_PROBE_MAX_PRODUCT_CTAS 16384
#define CGK_PROBE_VEC_CTAS 592
#define CGK_PROBE_WORDS (8 * (CGK_PROBE_MAX_PRODUCT_CTAS + 2 * CGK_PROBE_VEC_CTAS))

/* Row-sharded symmetric product (cg_symv.cu): which tiles of its m x n row block a rank reads.
 * Block pairs {g, h} are split so that every rank reads exactly half of its block: its own
 * diagonal block's upper tiles, all of the blocks at cyclic distance 1 .. G/2 - 1 (G even;
 * .. (G-1)/2 for odd G), and half of the block at distance G/2 (the lower-numbered rank of the
 * pair takes tile rows < row_end, the other one the tile columns >= col_begin). */
typedef struct cgk_symplan {
    int nblk;                           /* foreign column blocks this rank reads                */
    int blk_rank[CGK_MAX_RANKS];        /* their owners, in reading order (distance ascending)  */
    int row_end[CGK_MAX_RANKS];         /* local tile rows [0, row_end) contribute              */
    int col_begin[CGK_MAX_RANKS];       /* tile columns [col_begin, mt) of the block are read   */
    int nsrc;                           /* ranks that send column sums to this rank             */
    int src_rank[CGK_MAX_RANKS];
    int src_col_begin[CGK_MAX_RANKS];   /* they cover my local tile columns >= this             */
} cgk_symplan;

typedef struct cgk_gemv_plan {
    int variant;           /* CGK_GEMV_*                                  */
    int rows_per_cta;      /* R                                           */
    int unroll;            /* U column chunks in flight per thread        */
    int threads;           /* NT                                          */
    int vec;               /* doubles per load: 1, 2 (128-bit), 4 (256-bit) */
    int stages;            /* bulk-copy ring depth (variant BULK)         */
    int tile_cols;         /* C: columns per ring stage (variant BULK)    */
    int grid;              /* CTAs (resolved)                             */
    int smem_bytes;        /* dynamic shared memory (resolved)            */
    int keep_pct;          /* request: % of L2 to keep rows of A in across iterations;
                              0 = default (60), < 0 = never                  */
    long long l2_bytes;    /* request: the device's L2 size (0: no L2 policy) */
    long long keep_rows;   /* resolved: rows [0, keep_rows) are loaded evict_last */
} cgk_gemv_plan;

enum { CGK_GEMV_AUTO = 0, CGK_GEMV_LDG = 1, CGK_GEMV_BULK = 2, CGK_GEMV_SYMV = 3 };

/* Upper bound on gemv CTAs for any plan (sizes the partials workspace). */
int64_t cgk_gemv_max_ctas(int64_t m);
int64_t cgk_vec_max_ctas(int64_t m);

/* Resolves a launch plan for an m x n row block.  On entry the fields of
 * *plan are the request (0 = choose for me); on return they are the resolved
 * configuration.  Shapes the requested variant cannot tile (odd n, n not a
 * multiple of the stage width, unaligned bases) fall back to the streaming-load
 * kernel with the widest legal vector width -- always a CUDA kernel. */
int cgk_gemv_make_plan(cgk_gemv_plan *plan, int64_t m, int64_t n, const void *A,
                       const void *v, int sm_count);
/* Plan for cgk_gemv_comm: streaming-load kernel only, vector width limited by the alignment of
 * the column-block boundaries (multiples of m). */
int cgk_gemv_make_plan_comm(cgk_gemv_plan *plan, int64_t m, int64_t n, const void *A,
                            const void *v, int sm_count);

/* y = A v (m x n row block, row-major).  If w != NULL the CTA partials of
 * sum_i w[i]*y[i] go to `partials` and the last CTA to finish sums them in a
 * fixed order into *dot_out.  Returns immediately on device once sc->stop_req is set. */
int cgk_gemv(const cgk_gemv_plan *plan, const double *A, const double *v, const double *w,
             double *y, double *partials, double *dot_out, cgk_scalars *sc,
             int64_t m, int64_t n, void *stream);
/* Same product with the exchange fused in (ranks > 1, peer windows): the CTA first walks the
 * rank's own column block [rank*m, (rank+1)*m) -- which needs only the slice of p this GPU
 * wrote itself -- then waits for the peers' p_flag tags and walks the remote columns, so the
 * p exchange overlaps the local-block product; the last CTA pushes the p.Ap partial to all
 * peers.  v must be peers->p_full[rank], w the rank's own slice of it. */
int cgk_gemv_comm(const cgk_gemv_plan *plan, const double *A, const double *v, const double *w,
                  double *y, double *partials, double *dot_out, cgk_scalars *sc,
                  int64_t m, int64_t n, const cgk_peers *peers, void *stream);

/* r = b - Ax0, p_loc = b - Ax0, rr partials -> *rr_out (fixed-order sum). */
int cgk_init_residual(const double *b, const double *Ax, double *r, double *p_loc,
                      double *partials, double *rr_out, cgk_scalars *sc, int64_t m,
                      const cgk_peers *peers, void *stream);

/* alpha = rr[cur]/pAp; x += alpha p; r -= alpha Ap; r.r partials -> *rr_out.
 * guard != 0 (not in the reference, which iterates NaNs to the loop bound):
 * a non-finite alpha leaves x and r untouched and ends the solve (stop_req = breakdown = 3). */
/* Deferred reductions (one rank; takes the store -> fence -> ticket -> reload chain off the end
 * of every producer, worth ~4 us per iteration where an iteration is ~20 us): a producer
 * called with a NULL result pointer (cgk_gemv: dot_out, cgk_xr_update: rr_out) leaves its
 * per-CTA partials only, and the consumer adds them itself in a fixed order --
 * cgk_xr_update takes the n_dot GEMV partials (n_dot = the GEMV's grid; 0: read sc->pAp),
 * cgk_p_update the n_rr = cgk_vec_grid(m) partials of cgk_xr_update (0: read sc->rr). */
int cgk_xr_update(double *x, double *r, const double *p_loc, const double *Ap,
                  double *partials, double *rr_out, cgk_scalars *sc, int cur, int guard,
                  int64_t m, const cgk_peers *peers, const double *dot_partials, int n_dot,
                  void *stream);
/* cgk_xr_update + cgk_p_update as ONE launch for the peer exchange (ranks > 1): the grid (at most
 * cgk_vec_max_ctas CTAs, all co-resident) updates x and r, its last CTA pushes the rank's r.r
 * partial, every CTA then waits on the local mailbox for the G partials, applies the stop rule
 * and updates / pushes its part of p with r and p still in registers.  One launch ramp, one
 * drain and one reload of r less per iteration than the two-kernel form; same arithmetic, same
 * order, same tags and epochs (the two forms can alternate between solves). */
int cgk_xrp_update(double *x, double *r, double *p_loc, const double *Ap, double *partials,
                   cgk_scalars *sc, int cur, int guard, double eps, int ignore_stop,
                   int64_t max_iter, int64_t m, const cgk_peers *peers, void *stream);
/* 1 when the merged kernel's grid for m elements is co-resident on this device. */
int cgk_xrp_supported(int64_t m, int sm_count);

/* Programmatic dependent launch for the calling thread's following launches of the one-rank
 * iteration kernels (cg_device.cuh explains the rule the kernels keep); 0 = plain launches. */
void cgk_set_pdl(int on);
int cgk_pdl_enabled(void);
int cgk_vec_grid(int64_t m);      /* CTAs (= partials) of the vector kernels for m elements */

/* beta = rr[cur^1]: bookkeeping + stop rule on thread 0; p = r + (beta/rho) p. */
int cgk_p_update(double *p_loc, const double *r, cgk_scalars *sc, int cur, double eps,
                 int ignore_stop, int64_t max_iter, int64_t m, const cgk_peers *peers,
                 const double *rr_partials, int n_rr, void *stream);

/* One iteration of the Chronopoulos-Gear form of CG (the two per-iteration
 * reductions of parallel_cg.c:294,313 merged into one exposed synchronisation).  With gamma =
 * r.r and delta = r.(A r) of the CURRENT residual (both complete when the product w = A r ends):
 *   stop rule on gamma (as parallel_cg.c:314-317, after iteration k-1's update);
 *   beta = gamma/gamma_prev, alpha = gamma / (delta - beta gamma / alpha_prev)   (k = 0: 0, gamma/delta)
 *   p = r + beta p;  s = w + beta s;  x += alpha p;  r -= alpha s;  partial of the new r.r
 * r lives in the gathered vector (r_loc = the rank's slice of it) because the NEXT product is
 * A r; with peers the new slice is pushed to every rank and tagged like p in cgk_p_update. */
int cgk_cg2_step(double *x, double *r_loc, double *pvec, double *svec, const double *w, double *partials,
                 cgk_scalars *sc, double eps, int ignore_stop, int guard, int64_t max_iter, int64_t m,
                 const cgk_peers *peers, void *stream);
/* The same step on one rank with deferred reductions: `cur` = parity of the launch; gamma is
 * the sum of the previous step's partials (partials[cur][...], this launch writes
 * partials[cur ^ 1][...]: `partials` holds 2 x cgk_vec_max_ctas doubles), delta the sum of the
 * product's n_dot partials (0: read sc->pAp). */
int cgk_cg2_step_one(double *x, double *r_loc, double *pvec, double *svec, const double *w, double *partials,
                     cgk_scalars *sc, int cur, double eps, int ignore_stop, int guard, int64_t max_iter,
                     int64_t m, const double *dot_partials, int n_dot, void *stream);

/* Resets the scalars for a new solve (first kernel of the init phase). */
int cgk_reset(cgk_scalars *sc, unsigned int gen, void *stream);
/* rr[0] bookkeeping after the initial all-reduce (peers != NULL: it IS the all-reduce --
 * gathers the ranks' r.r partials from the local mailbox). */
int cgk_after_init(cgk_scalars *sc, int guard, const cgk_peers *peers, void *stream);

/* Rows [row0,row0+m) of the seeded SPD matrix into A (m x n); b rows likewise. */
int cgk_generate_A(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                   unsigned flags, void *stream);
int cgk_generate_b(double *b, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                   unsigned flags, void *stream);
/* The other two families of spd_gen.cuh: CGB_GEN_SPECTRUM (inv, s4 from cgb_gen_spectrum_scalars on the
 * host) and CGB_GEN_GRAM (cg_gram.cu: R R' + n diag(d) on the FP64 tensor pipe). */
int cgk_generate_spectrum_A(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                            unsigned flags, double inv, double s4, void *stream);
int cgk_generate_gram_A(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed, unsigned flags, void *stream);

/* Writes `bytes` of zeros-then-ones traffic over buf: evicts L2 between timed launches. */
int cgk_l2_flush(void *buf, int64_t bytes, void *stream);

/* ---- symmetric product (cg_symv.cu): y = A v reading only the upper T x T tiles of an exactly
 * symmetric A (n a multiple of cgk_symv_tile(), one GPU).  tiles: device array of (I, J) pairs
 * made from cgk_symv_tile_list(); rowpart: (n / T) * n doubles, colpart: split * (n / T) * n. */
int     cgk_symv_tile(void);
int64_t cgk_symv_tile_list(int64_t n, int split, int *tiles_xy);   /* split: CTAs (row parts) per tile: 1, 2, 4 */
int     cgk_symv_pick_split(int64_t tiles, int sm_count);
int     cgk_symv_finish_grid(int64_t n);   /* CTAs (= dot partials) of the finishing kernel */
int cgk_symv(const double *A, const double *v, const double *w, double *y, const void *tiles,
             int64_t ntiles, double *rowpart, double *colpart, double *partials, double *dot_out,
             cgk_scalars *sc, int64_t n, int rows_in_flight, int split, void *stream);
/* Row-sharded form (ranks > 1, peer windows): local tile rows, p exchange flags checked per
 * tile, column sums pushed to their owners (cgk_symv_comm = tile kernel + push kernel + finish
 * kernel; the finish also pushes the p.Ap partial like cgk_gemv_comm).  tiles: (I_local, J)
 * pairs from cgk_symv_tile_list_rank(); rowpart: (n / T) * m doubles; colpart: split * (m / T) * n. */
void    cgk_symv_make_plan(cgk_symplan *sp, int rank, int ranks, int64_t m);
int64_t cgk_symv_tile_list_rank(const cgk_symplan *sp, int rank, int64_t m, int split, int *tiles_xy);
int cgk_symv_comm(const double *A, const double *v, const double *w, double *y, const void *tiles,
                  int64_t ntiles, double *rowpart, double *colpart, double *partials, double *dot_out,
                  cgk_scalars *sc, int64_t m, int64_t n, int rows_in_flight, int split,
                  const cgk_symplan *sp, const cgk_peers *peers, void *stream);
/* *d_flag = 1 unless a_ij and a_ji are bit-identical for all i, j (and no NaN). */
int cgk_symmetry_check(const double *A, int64_t n, int *d_flag, void *stream);
/* the same for an n x n block inside a wider matrix (leading dimension ld) */
int cgk_symmetry_check_ld(const double *A, int64_t n, int64_t ld, int *d_flag, void *stream);
/* Row-sharded A: 64-bit fingerprints of the rank's G column blocks (sums[0..ranks)) and a NaN flag
 * (sums[ranks]); blocks (g, h) and (h, g) have equal fingerprints iff a_ij == a_ji throughout them
 * (up to a 2^-64 chance).  cg_symv.cu. */
int cgk_symmetry_hash(const double *A, int64_t n, int64_t row0, int64_t m, int ranks, unsigned long long *sums, void *stream);

/* ---- text ingest (cg_ingest.cu): the rank's part of matrixA.txt parsed on the GPU */
typedef struct cgk_ingest_state {
    long long base;            /* values produced by the chunks so far (next chunk's first index) */
    unsigned int exc_count;    /* tokens left to the host (outside the exact fast path)           */
    unsigned int pad_;
} cgk_ingest_state;
/* CTAs (= entries of block_counts / block_offsets) a chunk of `padded_len` bytes needs. */
int64_t cgk_ingest_blocks(int64_t padded_len);
/* One chunk of text already in device memory: `txt` holds whole tokens, starts at a token or
 * separator, is padded with separator bytes to `padded_len` (a multiple of 16) and followed by
 * at least 64 more separator bytes.  Values land at dst[state->base ...] (indices >= cap are
 * counted but not stored); `abs_pos` is the chunk's byte offset in the file, reported back in
 * exc_pos[] for tokens the host has to convert. */
int cgk_ingest_chunk(const unsigned char *txt, int64_t padded_len, int64_t abs_pos, double *dst,
                     int64_t cap, unsigned int *block_counts, long long *block_offsets,
                     cgk_ingest_state *state, unsigned long long *exc_pos, long long *exc_idx,
                     unsigned int exc_cap, void *stream);
/* dst[idx[i]] = val[i] (the host-converted exceptions). */
int cgk_scatter(double *dst, const long long *idx, const double *val, int64_t count, void *stream);

const char *cgk_error_string(int err);

#ifdef __cplusplus
}
#endif
#endif
