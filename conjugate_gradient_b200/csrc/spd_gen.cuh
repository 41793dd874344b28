/* spd_gen.cuh -- seeded counter-based SPD system generator (product side).
 *
 * On-device equivalent of the reference's generateSPDmatrix.m (north-star
 * item 4): R = rand(n,n) (generateSPDmatrix.m:4), b = rand(n,1) (:6),
 * x0 = zeros (:8), A = 0.5*(R+R') (:11), A = A + n*I (:17).  MATLAB's
 * Mersenne-Twister stream cannot be reproduced without MATLAB, so rand() is a
 * documented counter-based stream instead (splitmix64 in counter mode):
 *
 *     key     = mix(seed + GOLDEN)
 *     u(k)    = (mix(key + (k+1)*GOLDEN) >> 11) * 2^-53          in [0,1)
 *     R[i][j] = u(i*n + j)        b[i] = u(n*n + i)        x0 = 0
 *     A[i][j] = 0.5*(R[i][j] + R[j][i]) + (i == j ? n : 0)
 *
 * A is exactly symmetric (IEEE addition commutes) and every element is a pure
 * function of (seed, n, i, j), so each GPU writes its own row block in place
 * and the text files of the reference are never needed for large n.
 * CGB_GEN_ROUND4 mimics the generator's "%.4f" file round trip (:29,:35).
 */
#pragma once
#include <stdint.h>

#define CGB_GEN_ROUND4 1u

#if defined(__CUDACC__)
#define CGB_HD __host__ __device__ __forceinline__
#else
#define CGB_HD static inline
#endif

CGB_HD uint64_t cgb_mix64(uint64_t z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}

CGB_HD uint64_t cgb_gen_key(uint64_t seed) { return cgb_mix64(seed + 0x9E3779B97F4A7C15ULL); }

CGB_HD double cgb_gen_u(uint64_t key, uint64_t k)
{
    return (double)(cgb_mix64(key + (k + 1ULL) * 0x9E3779B97F4A7C15ULL) >> 11) * 0x1.0p-53;
}

/* ---- harder SPD families for iteration-bound runs ----------------------------
 * Diagonal spread d_i = 1 + (2^k - 1) t_i^2, t_i = i/(n-1), k = CGB_GEN_KAPPA_LOG2 bits of the flags
 * (every operation individually rounded: the CPU twin reproduces the bits; k = 0 gives exactly 1).
 *
 * CGB_GEN_SPECTRUM -- a dense matrix with a PRESCRIBED spectrum: A = H L H, H = I - 2 v v' (a
 * Householder reflector: orthogonal and symmetric), L = diag(lambda_i), lambda_i = n d_i, so the
 * eigenvalues of A are exactly the lambda_i (n .. n 2^k, quadratically spaced) up to rounding.
 * Element-wise, with s = sum_k lambda_k v_k^2:
 *     A[i][j] = v_i v_j (4 s - 2 (lambda_i + lambda_j)) + (i == j ? lambda_i : 0)
 * v = w / |w|, w_i = u(n^2 + n + i) - 0.5.  |w| and s are two O(n) sums done on the host in index
 * order (cgb_gen_spectrum_scalars); the element formula is symmetric in (i, j) operation by
 * operation (IEEE + and * commute), so A is exactly symmetric.
 *
 * CGB_GEN_GRAM -- the reference's commented alternative `A = A*A'` (generateSPDmatrix.m:12),
 * followed by its `A = A + n*eye(n)` (:17, here with the spread d_i):
 *     A[i][j] = sum_k R[i][k] R[j][k] + (i == j ? n d_i : 0),  R[i][k] = u(i n + k)
 * an FP64 contraction (2 m n^2 flops for an m-row block): cg_gram.cu, on the FP64 tensor pipe. */
#define CGB_GEN_SPECTRUM 2u
#define CGB_GEN_GRAM 4u

CGB_HD double cgb_gen_spread(uint64_t n, uint64_t i, unsigned flags)
{
    const unsigned k = (flags >> 8) & 0x3fu;
    if (k == 0u || n <= 1) return 1.0;
#if defined(__CUDA_ARCH__)
    const double t = __ddiv_rn((double)i, (double)(n - 1));
    return __dadd_rn(1.0, __dmul_rn(__dsub_rn((double)(1ULL << k), 1.0), __dmul_rn(t, t)));
#else
    volatile double t = (double)i / (double)(n - 1);
    volatile double tt = t * t;
    volatile double span = (double)(1ULL << k) - 1.0;
    volatile double st = span * tt;
    return 1.0 + st;
#endif
}

CGB_HD double cgb_gen_lambda(uint64_t n, uint64_t i, unsigned flags)
{
#if defined(__CUDA_ARCH__)
    return __dmul_rn((double)n, cgb_gen_spread(n, i, flags));
#else
    volatile double l = (double)n * cgb_gen_spread(n, i, flags);
    return l;
#endif
}

/* w_i of the reflector, before normalisation */
CGB_HD double cgb_gen_w(uint64_t key, uint64_t n, uint64_t i)
{
    return cgb_gen_u(key, n * n + n + i) - 0.5;            /* exact: both operands are multiples of 2^-53 below 1 */
}

/* element of the prescribed-spectrum matrix; inv = 1/|w|, s4 = 4 sum_k lambda_k v_k^2 */
CGB_HD double cgb_gen_spectrum_elem(uint64_t key, uint64_t n, uint64_t i, uint64_t j, unsigned flags, double inv, double s4)
{
    const double li = cgb_gen_lambda(n, i, flags), lj = cgb_gen_lambda(n, j, flags);
#if defined(__CUDA_ARCH__)
    const double vi = __dmul_rn(cgb_gen_w(key, n, i), inv), vj = __dmul_rn(cgb_gen_w(key, n, j), inv);
    const double c = __dsub_rn(s4, __dmul_rn(2.0, __dadd_rn(li, lj)));
    double a = __dmul_rn(__dmul_rn(vi, vj), c);
    if (i == j) a = __dadd_rn(a, li);
    return a;
#else
    volatile double vi = cgb_gen_w(key, n, i) * inv, vj = cgb_gen_w(key, n, j) * inv;
    volatile double sum = li + lj;
    volatile double two = 2.0 * sum;
    volatile double c = s4 - two;
    volatile double vv = vi * vj;
    volatile double a = vv * c;
    if (i == j) a = a + li;
    return a;
#endif
}

#if !defined(__CUDACC__)
#include <math.h>
/* the two O(n) sums of the reflector, in index order, every operation rounded by itself */
static inline void cgb_gen_spectrum_scalars(uint64_t key, uint64_t n, unsigned flags, double *inv, double *s4)
{
    volatile double w2 = 0.0;
    for (uint64_t i = 0; i < n; ++i) { volatile double w = cgb_gen_w(key, n, i); volatile double q = w * w; w2 = w2 + q; }
    volatile double nv = 1.0 / sqrt(w2);
    volatile double s = 0.0;
    for (uint64_t i = 0; i < n; ++i) {
        volatile double v = cgb_gen_w(key, n, i) * nv;
        volatile double q = v * v;
        volatile double t = cgb_gen_lambda(n, i, flags) * q;
        s = s + t;
    }
    *inv = nv;
    *s4 = 4.0 * s;
}
#endif
