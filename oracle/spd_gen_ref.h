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

static inline double orc_gen_b(uint64_t key, uint64_t n, uint64_t i, unsigned flags)
{
    double v = orc_gen_u(key, n * n + i);
    if (flags & ORC_GEN_ROUND4) v = orc_gen_round4(v);
    return v;
}

#endif
