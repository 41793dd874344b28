/* oracle/spd_gen_ref.h -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * CPU twin of the seeded on-device SPD generator (product side:
 * conjugate_gradient_b200/csrc/spd_gen.cuh).  It is written independently of
 * the product header on purpose: tests/ compare the two bit for bit.
 *
 * Construction follows the reference's generateSPDmatrix.m:
 *   R = rand(n,n)            generateSPDmatrix.m:4
 *   b = rand(n,1)            generateSPDmatrix.m:6
 *   x = zeros(n,1)           generateSPDmatrix.m:8
 *   A = 0.5*(R+R')           generateSPDmatrix.m:11
 *   A = A + n*eye(n)         generateSPDmatrix.m:17
 * MATLAB's Mersenne-Twister stream is not reproducible here (no MATLAB/Octave),
 * so rand() is replaced by a documented counter-based stream:
 *   key  = splitmix64_mix(seed)
 *   u(k) = top 53 bits of splitmix64_mix(key + (k+1)*GOLDEN) * 2^-53   in [0,1)
 *   R[i][j] = u(i*n + j),  b[i] = u(n*n + i),  x0 = 0.
 * The optional ROUND4 flag mimics the generator's "%.4f" text round trip
 * (generateSPDmatrix.m:29,35): v -> rint(v*1e4)/1e4.
 */
#ifndef ORACLE_SPD_GEN_REF_H
#define ORACLE_SPD_GEN_REF_H
#include <stdint.h>
#include <math.h>

#define ORC_GEN_ROUND4 1u
/* Harder family: bits 8..13 of the flags hold k; the n*I shift of
 * generateSPDmatrix.m:17 becomes n*diag(1 + (2^k - 1) * t_i^2), t_i = i/(n-1), which spreads
 * the spectrum over [~n, ~n*2^k] (still strictly diagonally dominant, hence SPD) so CG needs
 * tens to hundreds of iterations instead of ~5.  k = 0 is the reference's construction. */
#define ORC_GEN_KAPPA_LOG2(k) (((unsigned)(k) & 0x3fu) << 8)

static inline uint64_t orc_mix64(uint64_t z)
{
    z ^= z >> 30; z *= 0xBF58476D1CE4E5B9ULL;
    z ^= z >> 27; z *= 0x94D049BB133111EBULL;
    z ^= z >> 31;
    return z;
}

static inline uint64_t orc_gen_key(uint64_t seed)
{
    return orc_mix64(seed + 0x9E3779B97F4A7C15ULL);
}

static inline double orc_gen_u(uint64_t key, uint64_t k)
{
    uint64_t h = orc_mix64(key + (k + 1ULL) * 0x9E3779B97F4A7C15ULL);
    return (double)(h >> 11) * 0x1.0p-53;
}

static inline double orc_gen_round4(double v)
{
    return rint(v * 1.0e4) / 1.0e4;
}

static inline double orc_gen_a(uint64_t key, uint64_t n, uint64_t i, uint64_t j,
                               unsigned flags)
{
    double v = 0.5 * (orc_gen_u(key, i * n + j) + orc_gen_u(key, j * n + i));
    if (i == j) {
        unsigned k = (flags >> 8) & 0x3fu;
        double scale = 1.0;
        if (k && n > 1) {
            double t = (double)i / (double)(n - 1);
            double tt = t * t;
            double span = (double)(1ULL << k) - 1.0;
            double st = span * tt;
            scale = 1.0 + st;
        }
        double shift = (double)n * scale;
        v = v + shift;
    }
    if (flags & ORC_GEN_ROUND4) v = orc_gen_round4(v);
    return v;
}

/* ---- the two further families (product side: spd_gen.cuh, cg_gram.cu) -- written from their
 * definitions, not from the product's code.
 * ORC_GEN_SPECTRUM: A = H L H with H = I - 2 v v', L = diag(lambda), lambda_i = n d_i,
 *   d_i = 1 + (2^k - 1) (i/(n-1))^2, v = w/|w|, w_i = u(n^2 + n + i) - 1/2:
 *   (H L H)_ij = lambda_i [i==j] - 2 v_i v_j lambda_j - 2 lambda_i v_i v_j + 4 v_i v_j (v' L v)
 *              = v_i v_j (4 s - 2 (lambda_i + lambda_j)) + lambda_i [i==j],   s = sum_k lambda_k v_k^2.
 * ORC_GEN_GRAM: A = R R' + n diag(d), R_ik = u(i n + k)  (generateSPDmatrix.m:12 then :17). */
#define ORC_GEN_SPECTRUM 2u
#define ORC_GEN_GRAM 4u

static inline double orc_gen_d(uint64_t n, uint64_t i, unsigned flags)
{
    unsigned k = (flags >> 8) & 0x3fu;
    if (!k || n < 2) return 1.0;
    double t = (double)i / (double)(n - 1);
    double tt = t * t;
    double span = (double)(1ULL << k) - 1.0;
    double st = span * tt;
    return 1.0 + st;
}

typedef struct { double inv_norm, four_s; } orc_spectrum;

static inline orc_spectrum orc_gen_spectrum_setup(uint64_t key, uint64_t n, unsigned flags)
{
    orc_spectrum sp;
    double norm2 = 0.0;
    for (uint64_t i = 0; i < n; ++i) {
        double w = orc_gen_u(key, n * n + n + i) - 0.5;
        double ww = w * w;
        norm2 = norm2 + ww;
    }
    sp.inv_norm = 1.0 / sqrt(norm2);
    double s = 0.0;
    for (uint64_t i = 0; i < n; ++i) {
        double v = (orc_gen_u(key, n * n + n + i) - 0.5) * sp.inv_norm;
        double vv = v * v;
        double lam = (double)n * orc_gen_d(n, i, flags);
        double term = lam * vv;
        s = s + term;
    }
    sp.four_s = 4.0 * s;
    return sp;
}

static inline double orc_gen_spectrum_a(uint64_t key, uint64_t n, uint64_t i, uint64_t j, unsigned flags,
                                        const orc_spectrum *sp)
{
    double lam_i = (double)n * orc_gen_d(n, i, flags);
    double lam_j = (double)n * orc_gen_d(n, j, flags);
    double v_i = (orc_gen_u(key, n * n + n + i) - 0.5) * sp->inv_norm;
    double v_j = (orc_gen_u(key, n * n + n + j) - 0.5) * sp->inv_norm;
    double lsum = lam_i + lam_j;
    double twice = 2.0 * lsum;
    double coef = sp->four_s - twice;
    double vv = v_i * v_j;
    double a = vv * coef;
    if (i == j) a = a + lam_i;
    if (flags & ORC_GEN_ROUND4) a = orc_gen_round4(a);
    return a;
}

static inline double orc_gen_b(uint64_t key, uint64_t n, uint64_t i, unsigned flags)
{
    double v = orc_gen_u(key, n * n + i);
    if (flags & ORC_GEN_ROUND4) v = orc_gen_round4(v);
    return v;
}

#endif
