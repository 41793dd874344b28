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
