/* cg_symv.cu -- y = A p for an exactly symmetric A, reading only its upper tiles.
 *
 * The reference's matVec (parallel_cg.c:172-184) reads every element of A; conjugate gradient
 * is only defined for symmetric positive-definite A, and generateSPDmatrix.m:11 builds an
 * exactly symmetric one.  When a_ij == a_ji bit for bit (checked on the device before this
 * path is used, cgk_symmetry_check), every off-diagonal element read once can serve both
 * y_i += a_ij p_j and y_j += a_ij p_i: the bandwidth-bound product needs HALF the bytes
 * (the change of algorithmic bytes is declared in the roofline accounting:
 * 8 (n(n+1)/2 + 2n) per launch instead of 8 (n*n + 2n)).
 *
 * Tiling: A is cut into T x T tiles (T = 512: 2 MiB each); only tiles (I, J) with J >= I are
 * read.  One CTA (8 warps) owns one tile; warp w walks rows w, w+8, ... of the tile and one
 * row is 4 warp-wide 256-bit loads (4 KiB contiguous).  Per element: one FMA into the row sum
 * (with p_J, 16 values per lane kept in registers) and, off the diagonal, one FMA into the
 * lane's 16 column sums (with p_I[row], broadcast from shared memory).  A diagonal tile is
 * read whole and contributes row sums only (n*T/2 extra elements, 0.8 % at n = 65536).
 * Nothing is accumulated with atomics: every tile writes its T row sums to rowpart[J][rows of I]
 * and its T column sums to colpart[I][cols of J]; symv_finish adds the n/T partials of every y_i
 * in a fixed order (bitwise reproducible) and fuses the p.Ap partial like the plain GEMV.
 */
#include "cg_device.cuh"

#define SY_T 512
#define SY_NT 256
#define SY_NW (SY_NT / 32)

template <int RU>
__global__ void __launch_bounds__(SY_NT)
symv_tile_kernel(const double *__restrict__ A, const double *__restrict__ p,
                 const int2 *__restrict__ tiles, double *__restrict__ rowpart,
                 double *__restrict__ colpart, cgk_scalars *__restrict__ sc, long long n)
{
    const int stop = *(volatile int *)&sc->stop_req;
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->stopped = stop;
        return;
    }
    __shared__ double s_pI[SY_T];
    __shared__ double s_row[SY_T];
    __shared__ double s_col[SY_NW][SY_T];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int2 t = tiles[blockIdx.x];
    const long long I = t.x, J = t.y;
    const bool diag = (I == J);

    double pj[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) Ld<4>::keep(p + J * SY_T + k * 128 + lane * 4, pj[k]);
    for (int i = tid; i < SY_T; i += SY_NT) s_pI[i] = p[I * SY_T + i];
    __syncthreads();

    double cacc[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 4; ++e) cacc[k][e] = 0.0;

    const double *base = A + (I * SY_T) * n + J * SY_T + lane * 4;
    for (int r = wid; r < SY_T; r += SY_NW * RU) {
        double av[RU][4][4];
#pragma unroll
        for (int u = 0; u < RU; ++u)
#pragma unroll
            for (int k = 0; k < 4; ++k) Ld<4>::stream(base + (long long)(r + SY_NW * u) * n + k * 128, av[u][k]);
#pragma unroll
        for (int u = 0; u < RU; ++u) {
            double ra[4] = {0.0, 0.0, 0.0, 0.0};
#pragma unroll
            for (int k = 0; k < 4; ++k)
#pragma unroll
                for (int e = 0; e < 4; ++e) ra[k] = fma(av[u][k][e], pj[k][e], ra[k]);
            if (!diag) {
                const double pi = s_pI[r + SY_NW * u];
#pragma unroll
                for (int k = 0; k < 4; ++k)
#pragma unroll
                    for (int e = 0; e < 4; ++e) cacc[k][e] = fma(av[u][k][e], pi, cacc[k][e]);
            }
            const double rs = warp_sum((ra[0] + ra[1]) + (ra[2] + ra[3]));
            if (lane == 0) s_row[r + SY_NW * u] = rs;
        }
    }
    if (!diag) {
#pragma unroll
        for (int k = 0; k < 4; ++k)
#pragma unroll
            for (int e = 0; e < 4; ++e) s_col[wid][k * 128 + lane * 4 + e] = cacc[k][e];
    }
    __syncthreads();
    /* rowpart[J][i]: contiguous in i for a fixed J, so both this store and the finish are coalesced */
    for (int i = tid; i < SY_T; i += SY_NT) rowpart[J * n + I * SY_T + i] = s_row[i];
    if (!diag) {
        for (int c = tid; c < SY_T; c += SY_NT) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < SY_NW; ++w) s += s_col[w][c];
            colpart[I * n + J * SY_T + c] = s;
        }
    }
}

/* y_i = sum_{J >= I(i)} rowpart[J][i] + sum_{I' < I(i)} colpart[I'][i], fixed order; then the
 * fused dot partial of vecVec(p, Ap) (parallel_cg.c:293) and the last-CTA reduction. */
__global__ void __launch_bounds__(SY_NT)
symv_finish_kernel(const double *__restrict__ rowpart, const double *__restrict__ colpart,
                   const double *__restrict__ w, double *__restrict__ y,
                   double *__restrict__ partials, double *__restrict__ dot_out,
                   cgk_scalars *__restrict__ sc, long long n)
{
    if (*(volatile int *)&sc->stop_req) return;
    const long long i = (long long)blockIdx.x * SY_NT + threadIdx.x;     /* n % SY_T == 0: always < n */
    const long long I = i / SY_T, ntl = n / SY_T;
    double s = 0.0;
    for (long long J = I; J < ntl; ++J) s += __ldcs(rowpart + J * n + i);
    for (long long I2 = 0; I2 < I; ++I2) s += __ldcs(colpart + I2 * n + i);
    y[i] = s;
    if (w == nullptr) return;
    double contrib = s * w[i];
    __shared__ double s_w[SY_NW];
    contrib = warp_sum(contrib);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = contrib;
    __syncthreads();
    double mine = 0.0;
    if (threadIdx.x == 0)
        for (int q = 0; q < SY_NW; ++q) mine += s_w[q];
    cta_partials_finish<SY_NT>(mine, partials, dot_out, &sc->ticket_gemv);
}

/* Exact symmetry test: one CTA per pair of 32 x 32 blocks (bi <= bj); both blocks are read
 * row-wise (256-byte segments) and compared through a shared-memory transpose.  Any a_ij that
 * is not bit-identical to a_ji (NaNs included) raises *flag. */
__global__ void __launch_bounds__(256)
symmetry_check_kernel(const double *__restrict__ A, long long n, long long nb, int *flag)
{
    /* blockIdx.x enumerates (bi, bj), bi <= bj, row-major over the upper block triangle */
    long long t = blockIdx.x, bi = 0, rowlen = nb;
    while (t >= rowlen) { t -= rowlen; ++bi; --rowlen; }
    const long long bj = bi + t;
    __shared__ double s_a[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;            /* 32 x 8 */
    for (int r = ty; r < 32; r += 8) {
        const long long i = bi * 32 + r, j = bj * 32 + tx;
        s_a[r][tx] = (i < n && j < n) ? A[i * n + j] : 0.0;
    }
    __syncthreads();
    int bad = 0;
    for (int r = ty; r < 32; r += 8) {
        const long long i = bj * 32 + r, j = bi * 32 + tx;             /* element (i, j) of the mirror block */
        if (i < n && j < n) {
            const unsigned long long a = (unsigned long long)__double_as_longlong(A[i * n + j]);
            const unsigned long long b = (unsigned long long)__double_as_longlong(s_a[tx][r]);
            const double av = __longlong_as_double((long long)a);
            if (a != b || av != av) bad = 1;
        }
    }
    if (bad) atomicExch(flag, 1);
}

extern "C" int cgk_symv_tile(void) { return SY_T; }

/* Fills tiles[] (host) with the upper tile triangle, row-major; returns the count. */
extern "C" int64_t cgk_symv_tile_list(int64_t n, int *tiles_xy)
{
    const int64_t ntl = n / SY_T;
    int64_t k = 0;
    for (int64_t I = 0; I < ntl; ++I)
        for (int64_t J = I; J < ntl; ++J) {
            if (tiles_xy) { tiles_xy[2 * k] = (int)I; tiles_xy[2 * k + 1] = (int)J; }
            ++k;
        }
    return k;
}

extern "C" int cgk_symv(const double *A, const double *v, const double *w, double *y,
                        const void *tiles, int64_t ntiles, double *rowpart, double *colpart,
                        double *partials, double *dot_out, cgk_scalars *sc, int64_t n, int rows_in_flight,
                        void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (n % SY_T != 0 || ntiles < 1) return (int)cudaErrorInvalidValue;
    const int2 *tl = (const int2 *)tiles;
    switch (rows_in_flight) {
    case 1: symv_tile_kernel<1><<<(unsigned)ntiles, SY_NT, 0, st>>>(A, v, tl, rowpart, colpart, sc, (long long)n); break;
    case 4: symv_tile_kernel<4><<<(unsigned)ntiles, SY_NT, 0, st>>>(A, v, tl, rowpart, colpart, sc, (long long)n); break;
    default: symv_tile_kernel<2><<<(unsigned)ntiles, SY_NT, 0, st>>>(A, v, tl, rowpart, colpart, sc, (long long)n); break;
    }
    symv_finish_kernel<<<(unsigned)(n / SY_NT), SY_NT, 0, st>>>(rowpart, colpart, w, y, partials, dot_out, sc, (long long)n);
    return (int)cudaGetLastError();
}

extern "C" int cgk_symmetry_check(const double *A, int64_t n, int *d_flag, void *stream)
{
    const long long nb = (n + 31) / 32;
    const long long blocks = nb * (nb + 1) / 2;
    if (blocks > 2147483647LL) return (int)cudaErrorInvalidValue;
    cudaMemsetAsync(d_flag, 0, sizeof(int), (cudaStream_t)stream);
    symmetry_check_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(A, (long long)n, nb, d_flag);
    return (int)cudaGetLastError();
}
