/* cg_gram.cu -- rows of A = R R' + n diag(d): the reference's commented alternative `A = A*A'`
 * (generateSPDmatrix.m:12) followed by its `A = A + n*eye(n)` (:17), generated on the device.
 *
 * Unlike everything else on the path this IS a dense contraction (2 m n^2 flops for an m-row block,
 * an FP64 SYRK), so it runs on the FP64 tensor pipe: mma.sync.m8n8k4.f64 (SASS DMMA; tcgen05 has no
 * FP64 kind).  R is never stored: element R[i][k] = u(i n + k) of the counter stream is a pure
 * function of its index, so the [64 x 16] panels of R a CTA needs are computed straight into
 * shared memory (what a load would cost in bandwidth costs ~15 integer operations here).
 *
 * CTA: 64 x 64 tile of A, 8 warps as 4 x 2, a warp owns 16 x 32 = 2 x 4 m8n8 accumulator tiles.
 * Fragment layout of mma.m8n8k4.f64 (lane = 4 g + t): a = A[g][t] (row), b = B[t][g] (col),
 * c0, c1 = C[g][2t], C[g][2t + 1].  With B = R' the col-major B operand is simply a row of R.
 *
 * Exact symmetry: A[i][j] and A[j][i] are sums of the same products (IEEE * commutes) added in the
 * same order -- k ascending, four per instruction, the instruction's internal order being the same
 * for every output element -- so they agree bit for bit whichever tile, warp or rank computes them
 * (checked on the device by cgb_symmetry_check_kernel in the tests).  The bits of an element do
 * depend on the instruction's internal rounding, so parity with the CPU twin (a plain loop) is by
 * tolerance: <= 1e-13 relative per entry. */
#include <cuda_runtime.h>
#include <stdint.h>
#include "cg_kernels.h"
#include "spd_gen.cuh"

#define GR_BM 64
#define GR_BN 64
#define GR_BK 16
#define GR_PAD 1             /* row stride 17 doubles: the 8 rows a fragment load touches fall in different banks */

__device__ __forceinline__ void dmma_m8n8k4(double &c0, double &c1, double a, double b)
{
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(256)
cgb_gram_rows_kernel(double *__restrict__ A, long long n, long long row0, long long m, uint64_t key, unsigned flags)
{
    __shared__ double sA[GR_BM][GR_BK + GR_PAD];
    __shared__ double sB[GR_BN][GR_BK + GR_PAD];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int g = lane >> 2, t = lane & 3;
    const int wr = (wid >> 1) * 16, wc = (wid & 1) * 32;
    const long long i0 = (long long)blockIdx.y * GR_BM;          /* local row of the block */
    const long long j0 = (long long)blockIdx.x * GR_BN;

    double acc[2][4][2];
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b) acc[a][b][0] = acc[a][b][1] = 0.0;

    for (long long k0 = 0; k0 < n; k0 += GR_BK) {
        /* the two panels: 64 x 16 elements each, 4 + 4 per thread; rows / columns beyond the matrix are zeros */
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const int e = tid + q * 256, r = e >> 4, kk = e & 15;
            const long long k = k0 + kk;
            const long long gi = row0 + i0 + r, gj = j0 + r;
            sA[r][kk] = (i0 + r < m && k < n) ? cgb_gen_u(key, (uint64_t)gi * (uint64_t)n + (uint64_t)k) : 0.0;
            sB[r][kk] = (gj < n && k < n) ? cgb_gen_u(key, (uint64_t)gj * (uint64_t)n + (uint64_t)k) : 0.0;
        }
        __syncthreads();
#pragma unroll
        for (int ks = 0; ks < GR_BK; ks += 4) {
            double af[2], bf[4];
#pragma unroll
            for (int a = 0; a < 2; ++a) af[a] = sA[wr + a * 8 + g][ks + t];
#pragma unroll
            for (int b = 0; b < 4; ++b) bf[b] = sB[wc + b * 8 + g][ks + t];
#pragma unroll
            for (int a = 0; a < 2; ++a)
#pragma unroll
                for (int b = 0; b < 4; ++b) dmma_m8n8k4(acc[a][b][0], acc[a][b][1], af[a], bf[b]);
        }
        __syncthreads();
    }
#pragma unroll
    for (int a = 0; a < 2; ++a)
#pragma unroll
        for (int b = 0; b < 4; ++b)
#pragma unroll
            for (int h = 0; h < 2; ++h) {
                const long long li = i0 + wr + a * 8 + g, gj = j0 + wc + b * 8 + 2 * t + h;
                if (li >= m || gj >= n) continue;
                const long long gi = row0 + li;
                double v = acc[a][b][h];
                if (gi == gj) v = __dadd_rn(v, __dmul_rn((double)n, cgb_gen_spread((uint64_t)n, (uint64_t)gi, flags)));
                if (flags & CGB_GEN_ROUND4) v = __ddiv_rn(rint(__dmul_rn(v, 1.0e4)), 1.0e4);
                A[li * n + gj] = v;
            }
}

extern "C" int cgk_generate_gram_A(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed, unsigned flags, void *stream)
{
    const dim3 grid((unsigned)((n + GR_BN - 1) / GR_BN), (unsigned)((m + GR_BM - 1) / GR_BM));
    if (grid.y > 65535u) return (int)cudaErrorInvalidValue;
    cgb_gram_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(A, (long long)n, (long long)row0, (long long)m, cgb_gen_key(seed), flags);
    return (int)cudaGetLastError();
}
