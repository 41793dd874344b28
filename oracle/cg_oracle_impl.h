/* oracle/cg_oracle_impl.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU restatement of the reference's dense conjugate-gradient path with 64-bit
 * sizes and indices.  Included twice by cg_oracle.c with
 *     REAL = double, SFX = f64      (the FP64 oracle the GPU path is held to)
 *     REAL = float,  SFX = f32      (the reference's shipped precision)
 *
 * Every function states the reference lines whose arithmetic (operation order,
 * temporaries, rounding points) it follows.  The reference's `int` index math
 * (serialConjugate.c:78,117) overflows above n = 46340; here everything is
 * int64_t.  Build WITHOUT FMA contraction (-ffp-contract=off, no -march): the
 * reference built with plain `gcc -O2` on x86-64 rounds each product before the
 * add, and this file is checked bit for bit against that build (oracle/_ref).
 */

#define ORC_CAT2(a, b) a##_##b
#define ORC_CAT(a, b) ORC_CAT2(a, b)
#define ORC_FN(name) ORC_CAT(name, SFX)

/* y = A_blk * v for an m x n row block; the sum runs left to right and is
 * accumulated through y[i] itself, starting from 0.
 * Reference: serialConjugate.c:109-120, parallel_cg.c:172-184. */
void ORC_FN(orc_matvec)(REAL *y, const REAL *A, const REAL *v, int64_t m, int64_t n)
{
    for (int64_t i = 0; i < m; ++i) {
        const REAL *row = A + i * n;
        REAL acc = (REAL)0.0;
        for (int64_t j = 0; j < n; ++j)
            acc += row[j] * v[j];
        y[i] = acc;
    }
}

/* out = b - Av.  Reference: serialConjugate.c:124-131, parallel_cg.c:188-196. */
void ORC_FN(orc_residual)(REAL *out, const REAL *b, const REAL *Av, int64_t m)
{
    for (int64_t i = 0; i < m; ++i)
        out[i] = b[i] - Av[i];
}

/* out = s * v.  Reference: serialConjugate.c:135-142, parallel_cg.c:200-208. */
void ORC_FN(orc_scalarvec)(REAL *out, const REAL *v, REAL s, int64_t m)
{
    for (int64_t i = 0; i < m; ++i)
        out[i] = v[i] * s;
}

/* dot product, ascending index, REAL accumulator.
 * Reference: serialConjugate.c:145-155, parallel_cg.c:211-221. */
REAL ORC_FN(orc_vecvec)(const REAL *a, const REAL *b, int64_t m)
{
    REAL s = (REAL)0.0;
    for (int64_t i = 0; i < m; ++i)
        s += a[i] * b[i];
    return s;
}

/* out = a + b.  Reference: serialConjugate.c:159-166, parallel_cg.c:225-233. */
void ORC_FN(orc_vecadd)(REAL *out, const REAL *a, const REAL *b, int64_t m)
{
    for (int64_t i = 0; i < m; ++i)
        out[i] = a[i] + b[i];
}

/* out = a - b.  Reference: serialConjugate.c:170-177, parallel_cg.c:237-245. */
void ORC_FN(orc_vecsub)(REAL *out, const REAL *a, const REAL *b, int64_t m)
{
    for (int64_t i = 0; i < m; ++i)
        out[i] = a[i] - b[i];
}

/* Serial solver.  Operation order of serialConjugate.c:209-244:
 *   Ax0; r = b-Ax0; p = b-Ax0 (recomputed); rsold = r.r
 *   loop k < max_iter: Ap; alpha = rsold/(p.Ap); x += alpha p (via temp);
 *     r -= alpha Ap (via temp); beta = r.r; if sqrt(beta) < eps break;
 *     p = r + (beta/rsold) p (via temp); rsold = beta.
 * The stop test promotes beta to double before sqrt and compares with the
 * double eps (serialConjugate.c:28,235).  max_iter = n in the reference
 * (serialConjugate.c:213).  ignore_stop (not in the reference) keeps iterating
 * for the fixed-iteration timing mode; the test is still evaluated and the
 * first iteration at which it held is recorded. */
void ORC_FN(orc_conjugrad)(const REAL *A, const REAL *b, REAL *x, int64_t n,
                           double eps, int64_t max_iter, int ignore_stop,
                           orc_stats *st)
{
    REAL *Ap = orc_alloc(n * sizeof(REAL)), *r = orc_alloc(n * sizeof(REAL));
    REAL *p = orc_alloc(n * sizeof(REAL)), *tx = orc_alloc(n * sizeof(REAL));
    REAL *tr = orc_alloc(n * sizeof(REAL)), *tp = orc_alloc(n * sizeof(REAL));
    REAL rsold, alpha, beta = (REAL)0.0;
    int64_t k, done_iters = 0, break_k = -1, first_hit = -1;
    double t0 = orc_now();

    ORC_FN(orc_matvec)(Ap, A, x, n, n);
    ORC_FN(orc_residual)(r, b, Ap, n);
    ORC_FN(orc_residual)(p, b, Ap, n);
    rsold = ORC_FN(orc_vecvec)(r, r, n);
    if (st) st->rsold0 = (double)rsold;
    const double tl0 = orc_now();
    for (k = 0; k < max_iter; ++k) {
        ORC_FN(orc_matvec)(Ap, A, p, n, n);
        alpha = ORC_FN(orc_vecvec)(p, Ap, n);
        alpha = rsold / alpha;
        ORC_FN(orc_scalarvec)(tx, p, alpha, n);
        ORC_FN(orc_vecadd)(x, x, tx, n);
        ORC_FN(orc_scalarvec)(tr, Ap, alpha, n);
        ORC_FN(orc_vecsub)(r, r, tr, n);
        beta = ORC_FN(orc_vecvec)(r, r, n);
        done_iters = k + 1;
        if (sqrt((double)beta) < eps) {
            if (first_hit < 0) first_hit = k;
            if (!ignore_stop) { break_k = k; break; }
        }
        ORC_FN(orc_scalarvec)(tp, p, beta / rsold, n);
        ORC_FN(orc_vecadd)(p, r, tp, n);
        rsold = beta;
    }
    if (st) {
        st->seconds = orc_now() - t0;
        st->loop_seconds = orc_now() - tl0;
        st->iterations = done_iters;
        st->break_k = break_k;
        st->first_hit_k = first_hit;
        st->final_beta = (double)beta;
    }
    free(Ap); free(r); free(p); free(tx); free(tr); free(tp);
}

/* Row-block SPMD solver, the arithmetic of parallel_cg.c:283-323 for P ranks,
 * evaluated rank after rank in one thread (deterministic).  Rank g owns rows
 * [g*m, (g+1)*m), m = n/P (parallel_cg.c:83).  Differences from the serial
 * form that change rounding and are therefore mirrored:
 *   - every dot product is a per-rank partial (ascending local index) followed
 *     by an all-reduce (parallel_cg.c:286-287,293-294,309-313); the reduction
 *     order of MPI_Allreduce is the MPI library's -- here and in
 *     oracle/mpi_shim it is rank 0,1,...,P-1, left to right;
 *   - x is replicated and updated over all n entries from the gathered p
 *     (parallel_cg.c:299,303), r and p are updated on local rows only.
 * A is passed whole; rank g reads its block in place. */
void ORC_FN(orc_conjugrad_spmd)(const REAL *A, const REAL *b, REAL *x, int64_t n,
                                int64_t P, double eps, int64_t max_iter,
                                int ignore_stop, orc_stats *st)
{
    const int64_t m = n / P;
    REAL *Ap = orc_alloc(n * sizeof(REAL)), *r = orc_alloc(n * sizeof(REAL));
    REAL *ploc = orc_alloc(n * sizeof(REAL)), *pall = orc_alloc(n * sizeof(REAL));
    REAL *tx = orc_alloc(n * sizeof(REAL)), *tr = orc_alloc(n * sizeof(REAL));
    REAL *tp = orc_alloc(n * sizeof(REAL));
    REAL rsold, alpha, beta = (REAL)0.0, part;
    int64_t k, g, done_iters = 0, break_k = -1, first_hit = -1;
    double t0 = orc_now();

    rsold = (REAL)0.0;
    for (g = 0; g < P; ++g) {
        ORC_FN(orc_matvec)(Ap + g * m, A + g * m * n, x, m, n);
        ORC_FN(orc_residual)(r + g * m, b + g * m, Ap + g * m, m);
        ORC_FN(orc_residual)(ploc + g * m, b + g * m, Ap + g * m, m);
        part = ORC_FN(orc_vecvec)(r + g * m, r + g * m, m);
        rsold = (g == 0) ? part : rsold + part;
    }
    if (st) st->rsold0 = (double)rsold;
    const double tl0 = orc_now();
    for (k = 0; k < max_iter; ++k) {
        memcpy(pall, ploc, (size_t)n * sizeof(REAL));          /* allgather */
        alpha = (REAL)0.0;
        for (g = 0; g < P; ++g) {
            ORC_FN(orc_matvec)(Ap + g * m, A + g * m * n, pall, m, n);
            part = ORC_FN(orc_vecvec)(ploc + g * m, Ap + g * m, m);
            alpha = (g == 0) ? part : alpha + part;
        }
        alpha = rsold / alpha;
        ORC_FN(orc_scalarvec)(tx, pall, alpha, n);             /* replicated x */
        ORC_FN(orc_vecadd)(x, x, tx, n);
        beta = (REAL)0.0;
        for (g = 0; g < P; ++g) {
            ORC_FN(orc_scalarvec)(tr + g * m, Ap + g * m, alpha, m);
            ORC_FN(orc_vecsub)(r + g * m, r + g * m, tr + g * m, m);
            part = ORC_FN(orc_vecvec)(r + g * m, r + g * m, m);
            beta = (g == 0) ? part : beta + part;
        }
        done_iters = k + 1;
        if (sqrt((double)beta) < eps) {
            if (first_hit < 0) first_hit = k;
            if (!ignore_stop) { break_k = k; break; }
        }
        for (g = 0; g < P; ++g) {
            ORC_FN(orc_scalarvec)(tp + g * m, ploc + g * m, beta / rsold, m);
            ORC_FN(orc_vecadd)(ploc + g * m, r + g * m, tp + g * m, m);
        }
        rsold = beta;
    }
    if (st) {
        st->seconds = orc_now() - t0;
        st->loop_seconds = orc_now() - tl0;
        st->iterations = done_iters;
        st->break_k = break_k;
        st->first_hit_k = first_hit;
        st->final_beta = (double)beta;
    }
    free(Ap); free(r); free(ploc); free(pall); free(tx); free(tr); free(tp);
}

#undef ORC_FN
#undef ORC_CAT
#undef ORC_CAT2
