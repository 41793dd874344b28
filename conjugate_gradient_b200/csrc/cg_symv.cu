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
cgb_halfmat_tile_kernel(const double *__restrict__ A, const double *__restrict__ p,
                 const int2 *__restrict__ tiles, double *__restrict__ rowpart,
                 double *__restrict__ colpart, cgk_scalars *__restrict__ sc, long long n, int split)
{
    pdl_wait();                       /* p and the flags come from the previous kernel */
    pdl_launch_dependents();
    const int stop = *(volatile int *)&sc->stop_req;
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->stopped = stop;
        return;
    }
    __shared__ double s_pI[SY_T];
    __shared__ double s_row[SY_T];
    __shared__ double s_col[SY_NW][SY_T];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    /* a tile is read by `split` CTAs, each taking SY_T / split of its rows (more, smaller work
     * items when a GPU has few tiles: the last wave of CTAs stays full) */
    const int2 t = tiles[blockIdx.x];
    const long long I = t.x & 0xffffff, J = t.y;
    const int part = t.x >> 24, rows_per = SY_T / split, r_lo = part * rows_per, r_hi = r_lo + rows_per;
    const bool diag = (I == J);

    double pj[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) Ld<4>::coherent(p + J * SY_T + k * 128 + lane * 4, pj[k]);   /* not .nc: see pdl_wait */
    for (int i = tid; i < SY_T; i += SY_NT) s_pI[i] = __ldcg(p + I * SY_T + i);
    __syncthreads();

    double cacc[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 4; ++e) cacc[k][e] = 0.0;

    const double *base = A + (I * SY_T) * n + J * SY_T + lane * 4;
    for (int r = r_lo + wid; r < r_hi; r += SY_NW * RU) {
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
    for (int i = r_lo + tid; i < r_hi; i += SY_NT) rowpart[J * n + I * SY_T + i] = s_row[i];
    if (!diag) {
        for (int c = tid; c < SY_T; c += SY_NT) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < SY_NW; ++w) s += s_col[w][c];
            colpart[(I * split + part) * n + J * SY_T + c] = s;
        }
    }
}

/* y_i = sum_{J >= I(i)} rowpart[J][i] + sum_{I' < I(i)} colpart[I'][i], fixed order; then the
 * fused dot partial of vecVec(p, Ap) (parallel_cg.c:293).  A CTA owns SY_FR = 32 consecutive
 * i (one 256-byte segment per term) and its 8 warps take the terms k = w, w + 8, ...: at small
 * n that is 8 x the CTAs and an 8 x shorter chain of dependent loads per thread than one
 * thread per i (9.1 -> measured below at n = 4096); the 8 sums are added in warp order. */
#define SY_FR 32
__global__ void __launch_bounds__(SY_NT)
cgb_halfmat_finish_kernel(const double *__restrict__ rowpart, const double *__restrict__ colpart,
                   const double *__restrict__ w, double *__restrict__ y,
                   double *__restrict__ partials, double *__restrict__ dot_out,
                   cgk_scalars *__restrict__ sc, long long n, int split)
{
    pdl_wait();
    pdl_launch_dependents();
    if (*(volatile int *)&sc->stop_req) return;
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    const long long i = (long long)blockIdx.x * SY_FR + lane;            /* n % SY_T == 0: always < n */
    const long long I = i / SY_T, ntl = n / SY_T;                        /* the same tile row for the whole CTA */
    const long long nrow = ntl - I, total = nrow + I * split;
    double s = 0.0;
#pragma unroll 4
    for (long long k = wid; k < total; k += SY_NW) {
        const double *src = k < nrow ? rowpart + (I + k) * n : colpart + (k - nrow) * n;
        s += __ldcs(src + i);
    }
    __shared__ double s_part[SY_NW][SY_FR];
    s_part[wid][lane] = s;
    __syncthreads();
    double mine = 0.0;
    if (wid == 0) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < SY_NW; ++q) t += s_part[q][lane];
        y[i] = t;
        if (w != nullptr) mine = warp_sum(t * __ldcg(w + i));
    }
    if (w == nullptr) return;
    cta_partials_finish<SY_NT>(mine, partials, dot_out, &sc->ticket_gemv);
}

/* Exact symmetry test: one CTA per pair of 32 x 32 blocks (bi <= bj); both blocks are read
 * row-wise (256-byte segments) and compared through a shared-memory transpose.  Any a_ij that
 * is not bit-identical to a_ji (NaNs included) raises *flag. */
__global__ void __launch_bounds__(256)
cgb_symmetry_check_kernel(const double *__restrict__ A, long long n, long long ld, long long nb, int *flag)
{
    /* blockIdx.x enumerates (bi, bj), bi <= bj, row-major over the upper block triangle */
    long long t = blockIdx.x, bi = 0, rowlen = nb;
    while (t >= rowlen) { t -= rowlen; ++bi; --rowlen; }
    const long long bj = bi + t;
    __shared__ double s_a[32][33];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;            /* 32 x 8 */
    for (int r = ty; r < 32; r += 8) {
        const long long i = bi * 32 + r, j = bj * 32 + tx;
        s_a[r][tx] = (i < n && j < n) ? A[i * ld + j] : 0.0;
    }
    __syncthreads();
    int bad = 0;
    for (int r = ty; r < 32; r += 8) {
        const long long i = bj * 32 + r, j = bi * 32 + tx;             /* element (i, j) of the mirror block */
        if (i < n && j < n) {
            const unsigned long long a = (unsigned long long)__double_as_longlong(A[i * ld + j]);
            const unsigned long long b = (unsigned long long)__double_as_longlong(s_a[tx][r]);
            const double av = __longlong_as_double((long long)a);
            if (a != b || av != av) bad = 1;
        }
    }
    if (bad) atomicExch(flag, 1);
}

extern "C" int cgk_symv_tile(void) { return SY_T; }

/* Fills tiles[] (host) with the upper tile triangle, row-major; returns the count. */
extern "C" int64_t cgk_symv_tile_list(int64_t n, int split, int *tiles_xy)
{
    const int64_t ntl = n / SY_T;
    int64_t k = 0;
    for (int64_t I = 0; I < ntl; ++I)
        for (int64_t J = I; J < ntl; ++J)
            for (int part = 0; part < split; ++part) {
                if (tiles_xy) { tiles_xy[2 * k] = (int)I | (part << 24); tiles_xy[2 * k + 1] = (int)J; }
                ++k;
            }
    return k;
}

/* Row parts per tile.  Measured on B200 (profiles/r01_symv_split_sweep.txt): splitting costs 2-5 %
 * (every part repeats the tile's prologue and column-sum epilogue) and never pays once there is a
 * tile per SM; it only serves to put every SM to work when a GPU has fewer tiles than SMs. */
extern "C" int cgk_symv_pick_split(int64_t tiles, int sm_count)
{
    const int64_t sms = sm_count > 0 ? sm_count : 148;
    if (tiles * 10 >= sms * 9) return 1;
    if (tiles * 2 >= sms) return 2;
    return 4;
}

extern "C" int cgk_symv(const double *A, const double *v, const double *w, double *y,
                        const void *tiles, int64_t ntiles, double *rowpart, double *colpart,
                        double *partials, double *dot_out, cgk_scalars *sc, int64_t n, int rows_in_flight,
                        int split, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (n % SY_T != 0 || ntiles < 1 || (split != 1 && split != 2 && split != 4)) return (int)cudaErrorInvalidValue;
    const int2 *tl = (const int2 *)tiles;
    switch (rows_in_flight) {
    case 1: cgk_launch(cgb_halfmat_tile_kernel<1>, dim3((unsigned)ntiles), dim3(SY_NT), 0, st, A, v, tl, rowpart, colpart, sc, (long long)n, split); break;
    case 2: cgk_launch(cgb_halfmat_tile_kernel<2>, dim3((unsigned)ntiles), dim3(SY_NT), 0, st, A, v, tl, rowpart, colpart, sc, (long long)n, split); break;
    default: cgk_launch(cgb_halfmat_tile_kernel<4>, dim3((unsigned)ntiles), dim3(SY_NT), 0, st, A, v, tl, rowpart, colpart, sc, (long long)n, split); break;
    }
    cgk_launch(cgb_halfmat_finish_kernel, dim3((unsigned)(n / SY_FR)), dim3(SY_NT), 0, st, rowpart, colpart, w, y, partials, dot_out, sc, (long long)n, split);
    return (int)cudaGetLastError();
}

extern "C" int cgk_symv_finish_grid(int64_t n) { return (int)(n / SY_FR); }

extern "C" int cgk_symmetry_check_ld(const double *A, int64_t n, int64_t ld, int *d_flag, void *stream)
{
    const long long nb = (n + 31) / 32;
    const long long blocks = nb * (nb + 1) / 2;
    if (blocks > 2147483647LL) return (int)cudaErrorInvalidValue;
    cudaMemsetAsync(d_flag, 0, sizeof(int), (cudaStream_t)stream);
    cgb_symmetry_check_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(A, (long long)n, (long long)ld, nb, d_flag);
    return (int)cudaGetLastError();
}
extern "C" int cgk_symmetry_check(const double *A, int64_t n, int *d_flag, void *stream)
{
    return cgk_symmetry_check_ld(A, n, n, d_flag, stream);
}

/* Symmetry ACROSS row blocks (ranks > 1): rank g cannot see a_ji for j in another rank's rows, but it can
 * fingerprint its part of every block pair.  For the m x m block (g, h) of its rows the rank adds up, modulo
 * 2^64, f(min(i,j), max(i,j), bits(a_ij)) over the block's elements, f a 64-bit mixing function: block
 * (h, g) on rank h gives the same sum when a_ij == a_ji bit for bit throughout, and a different one
 * otherwise except with probability ~2^-64.  The ranks exchange their G sums (one small all-gather) and every
 * rank compares S[g][h] with S[h][g]; the diagonal block (g, g) is compared exactly by the kernel above.
 * A NaN anywhere raises sums[ranks] (no a_ij == a_ji holds for NaNs).  One pass over the block at HBM rate. */
__device__ __forceinline__ unsigned long long sym_mix(unsigned long long z)
{
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ULL;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBULL;
    return z ^ (z >> 31);
}
__global__ void __launch_bounds__(256)
cgb_symmetry_hash_kernel(const double *__restrict__ A, long long n, long long row0, long long m, int ranks,
                         unsigned long long *__restrict__ sums)
{
    __shared__ unsigned long long s_sum[CGK_MAX_RANKS + 1];
    if (threadIdx.x <= CGK_MAX_RANKS) s_sum[threadIdx.x] = 0ull;
    __syncthreads();
    unsigned long long nan = 0ull;
    for (long long r = blockIdx.x; r < m; r += gridDim.x) {
        const unsigned long long i = (unsigned long long)(row0 + r);
        const double *row = A + r * n;
        for (int h = 0; h < ranks; ++h) {
            unsigned long long acc = 0ull;
            for (long long j = (long long)h * m + threadIdx.x; j < (long long)(h + 1) * m; j += 256) {
                const double v = row[j];
                const unsigned long long uj = (unsigned long long)j, lo = i < uj ? i : uj, hi = i < uj ? uj : i;
                acc += sym_mix(sym_mix(lo * (unsigned long long)n + hi) ^ (unsigned long long)__double_as_longlong(v));
                if (v != v) nan = 1ull;
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            if ((threadIdx.x & 31) == 0) atomicAdd(&s_sum[h], acc);
        }
    }
    if (nan) atomicOr(&s_sum[CGK_MAX_RANKS], 1ull);
    __syncthreads();
    if (threadIdx.x < ranks) atomicAdd(&sums[threadIdx.x], s_sum[threadIdx.x]);
    if (threadIdx.x == 0 && s_sum[CGK_MAX_RANKS]) atomicOr(&sums[ranks], 1ull);
}
/* sums: ranks + 1 words of device memory (zeroed here): the fingerprints of the rank's G blocks, then the NaN flag */
extern "C" int cgk_symmetry_hash(const double *A, int64_t n, int64_t row0, int64_t m, int ranks, unsigned long long *sums, void *stream)
{
    if (ranks < 1 || ranks > CGK_MAX_RANKS) return (int)cudaErrorInvalidValue;
    cudaMemsetAsync(sums, 0, (size_t)(ranks + 1) * sizeof(unsigned long long), (cudaStream_t)stream);
    long long g = m < 148 * 8 ? m : 148 * 8;
    cgb_symmetry_hash_kernel<<<(unsigned)g, 256, 0, (cudaStream_t)stream>>>(A, (long long)n, (long long)row0, (long long)m, ranks, sums);
    return (int)cudaGetLastError();
}

/* ================================================================== row-sharded form (ranks > 1)
 * Rank g stores rows [g m, (g+1) m) of A and reads HALF of that block (cgk_symplan).  A tile
 * (I_local, J) gives row sums for the rank's own rows and column sums for columns of block
 * h = J / mt; the latter belong to rank h and travel there over NVLink:
 *   symv_tile_comm   as symv_tile, but p_J of a foreign block is read only after that rank's
 *                    p_flag of this iteration has been seen (tiles of the own block come first
 *                    in the list, so the exchange of p hides behind them)
 *   symv_push        column sums of each foreign block, added over the local tile rows in a
 *                    fixed order, stored into the owner's crecv[rank][...]; last CTA posts c_flag
 *   symv_finish_comm waits for the c_flags of the ranks that read blocks of MY columns, adds
 *                    row partials, own-block column partials and the received sums in a fixed
 *                    order; p.Ap partial pushed to every mailbox as in cgb_rowblock_matvec_comm */
template <int RU>
__global__ void __launch_bounds__(SY_NT)
cgb_halfmat_tile_comm_kernel(const double *__restrict__ A, const double *p, const int2 *__restrict__ tiles,
                      double *__restrict__ rowpart, double *__restrict__ colpart, cgk_scalars *sc,
                      long long m, long long n, int split, const cgk_peers peers)
{
    pdl_wait();
    pdl_launch_dependents();
    const int stop = *(volatile int *)&sc->stop_req;
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->stopped = stop;
        return;
    }
    const unsigned int ep = *(volatile unsigned int *)&sc->ep_gemv;
    const unsigned int gen = *(volatile unsigned int *)&sc->gen;
    __shared__ double s_pI[SY_T];
    __shared__ double s_row[SY_T];
    __shared__ double s_col[SY_NW][SY_T];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int2 t = tiles[blockIdx.x];
    const long long mt = m / SY_T;
    const long long Il = t.x & 0xffffff, J = t.y;
    const int part = t.x >> 24, rows_per = SY_T / split, r_lo = part * rows_per, r_hi = r_lo + rows_per;
    const bool diag = ((long long)peers.rank * mt + Il == J);
    const int h = (int)(J / mt);

    /* the rank's own slice of p needs no flag */
    for (int i = tid; i < SY_T; i += SY_NT) s_pI[i] = p[(long long)peers.rank * m + Il * SY_T + i];
    if (h != peers.rank) {
        if (tid == 0) {
            const unsigned long long *f = &peers.mbox[peers.rank]->p_flag[h];
            const unsigned long long tag = (unsigned long long)cgk_tag(gen, ep);
            unsigned long long t0 = 0ull;
            unsigned int polls = 0;
            while (ld_relaxed_sys(f) != tag)
                if ((++polls & 1023u) == 0u && spin_give_up(&t0, sc)) break;
            (void)ld_acquire_sys(f);
        }
    }
    __syncthreads();
    double pj[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k) Ld<4>::coherent(p + J * SY_T + k * 128 + lane * 4, pj[k]);

    double cacc[4][4];
#pragma unroll
    for (int k = 0; k < 4; ++k)
#pragma unroll
        for (int e = 0; e < 4; ++e) cacc[k][e] = 0.0;

    const double *base = A + (Il * SY_T) * n + J * SY_T + lane * 4;
    for (int r = r_lo + wid; r < r_hi; r += SY_NW * RU) {
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
    for (int i = r_lo + tid; i < r_hi; i += SY_NT) rowpart[J * m + Il * SY_T + i] = s_row[i];
    if (!diag) {
        for (int c = tid; c < SY_T; c += SY_NT) {
            double s = 0.0;
#pragma unroll
            for (int w = 0; w < SY_NW; ++w) s += s_col[w][c];
            colpart[(Il * split + part) * n + J * SY_T + c] = s;
        }
    }
}

/* grid: (m / SY_NT, nblk).  Column j of foreign block b: sum over the contributing local tile
 * rows, pushed into the owner's receive area.  The last CTA posts c_flag to every owner. */
__global__ void __launch_bounds__(SY_NT)
cgb_halfmat_push_kernel(const double *__restrict__ colpart, cgk_scalars *sc, long long m, long long n, int split,
                 const cgk_symplan sp, const cgk_peers peers)
{
    pdl_wait();
    pdl_launch_dependents();
    if (*(volatile int *)&sc->stop_req) return;
    const unsigned int ep = *(volatile unsigned int *)&sc->ep_gemv;
    const unsigned int gen = *(volatile unsigned int *)&sc->gen;
    const int b = blockIdx.y;
    const int h = sp.blk_rank[b];
    const long long jl = (long long)blockIdx.x * SY_NT + threadIdx.x;        /* column inside block h */
    if (jl >= (long long)sp.col_begin[b] * SY_T) {
        double s = 0.0;
        for (int Il = 0; Il < sp.row_end[b] * split; ++Il) s += __ldcs(colpart + (long long)Il * n + (long long)h * m + jl);
        peers.crecv[h][(long long)peers.rank * m + jl] = s;
    }
    __shared__ int s_last;
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence_system();
        const unsigned int total = gridDim.x * gridDim.y;
        s_last = (atomicAdd(&sc->ticket_vec, 1u) == total - 1u);
        if (s_last) {
            __threadfence_system();
            const unsigned long long tag = (unsigned long long)cgk_tag(gen, ep);
            for (int q = 0; q < sp.nblk; ++q)
                st_relaxed_sys(&peers.mbox[sp.blk_rank[q]]->c_flag[peers.rank], tag);
            sc->ticket_vec = 0u;
        }
    }
}

__global__ void __launch_bounds__(SY_NT)
cgb_halfmat_finish_comm_kernel(const double *__restrict__ rowpart, const double *__restrict__ colpart,
                        const double *w, double *__restrict__ y, double *__restrict__ partials,
                        double *__restrict__ dot_out, cgk_scalars *sc, long long m, long long n, int split,
                        const cgk_symplan sp, const cgk_peers peers)
{
    pdl_wait();
    pdl_launch_dependents();
    if (*(volatile int *)&sc->stop_req) return;
    const unsigned int ep = *(volatile unsigned int *)&sc->ep_gemv;
    const unsigned int gen = *(volatile unsigned int *)&sc->gen;
    /* 32 local rows per CTA; the 8 warps take every 8th term of the row's list of partials
     * (as cgb_halfmat_finish_kernel: 8 x the CTAs, an 8 x shorter chain of loads per thread) */
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const long long mt = m / SY_T;
    const long long i = (long long)blockIdx.x * SY_FR + lane;                /* local row, < m */
    const long long Il = i / SY_T;                                           /* the same for the whole CTA */
    double s = 0.0;
    /* everything this rank computed itself first (hides the wait for the peers' sums): the flat
     * list [own block's row partials | each foreign block's | own column partials], stride 8 */
    long long k = wid;
    {
        const long long cnt = mt - Il, J0 = (long long)peers.rank * mt + Il;
#pragma unroll 4
        for (; k < cnt; k += SY_NW) s += __ldcs(rowpart + (J0 + k) * m + i);
        k -= cnt;
    }
    for (int b = 0; b < sp.nblk; ++b)
        if (Il < sp.row_end[b]) {
            const long long cnt = mt - sp.col_begin[b], J0 = (long long)sp.blk_rank[b] * mt + sp.col_begin[b];
#pragma unroll 4
            for (; k < cnt; k += SY_NW) s += __ldcs(rowpart + (J0 + k) * m + i);
            k -= cnt;
        }
    {
        const long long cnt = Il * split;
#pragma unroll 4
        for (; k < cnt; k += SY_NW) s += __ldcs(colpart + k * n + (long long)peers.rank * m + i);
    }
    /* column sums other ranks computed for my columns */
    if (tid < sp.nsrc) {
        const unsigned long long *f = &peers.mbox[peers.rank]->c_flag[sp.src_rank[tid]];
        const unsigned long long tag = (unsigned long long)cgk_tag(gen, ep);
        unsigned long long t0 = 0ull;
        unsigned int polls = 0;
        while (ld_relaxed_sys(f) != tag)
            if ((++polls & 1023u) == 0u && spin_give_up(&t0, sc)) break;
        (void)ld_acquire_sys(f);
    }
    __syncthreads();
    const double *rx = peers.crecv[peers.rank];
    if (wid < sp.nsrc && Il >= sp.src_col_begin[wid]) s += __ldcg(rx + (long long)sp.src_rank[wid] * m + i);   /* nsrc <= 4 < 8 warps */
    __shared__ double s_part[SY_NW][SY_FR];
    s_part[wid][lane] = s;
    __syncthreads();
    double mine = 0.0;
    if (wid == 0) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < SY_NW; ++q) t += s_part[q][lane];
        y[i] = t;
        mine = warp_sum(t * w[i]);
    }
    double total = 0.0;
    if (cta_partials_finish_sys<SY_NT, false>(mine, partials, dot_out, &sc->ticket_gemv, &total) && tid == 0) {   /* local stores only */
        const unsigned int tag = cgk_tag(gen, ep);
        for (int g = 0; g < peers.ranks; ++g)
            ll_send(&peers.mbox[g]->dot[0][ep & 1u][peers.rank][0], total, tag);
        sc->ep_gemv = ep + 1u;
    }
}

extern "C" void cgk_symv_make_plan(cgk_symplan *sp, int rank, int ranks, int64_t m)
{
    const int G = ranks, mt = (int)(m / SY_T), half = mt / 2;
    sp->nblk = 0; sp->nsrc = 0;
    for (int d = 1; d <= G / 2; ++d) {
        const bool halved = (G % 2 == 0 && d == G / 2);
        /* blocks I read */
        const int h = (rank + d) % G;
        int row_end = mt, col_begin = 0;
        if (halved) { if (rank < h) row_end = half; else col_begin = half; }
        if (row_end > 0 && col_begin < mt) {
            sp->blk_rank[sp->nblk] = h; sp->row_end[sp->nblk] = row_end; sp->col_begin[sp->nblk] = col_begin;
            sp->nblk++;
        }
        /* ranks that read a block of my columns */
        const int g = (rank - d + G) % G;
        int g_row_end = mt, g_col_begin = 0;
        if (halved) { if (g < rank) g_row_end = half; else g_col_begin = half; }
        if (g_row_end > 0 && g_col_begin < mt) {
            sp->src_rank[sp->nsrc] = g; sp->src_col_begin[sp->nsrc] = g_col_begin;
            sp->nsrc++;
        }
    }
}

extern "C" int64_t cgk_symv_tile_list_rank(const cgk_symplan *sp, int rank, int64_t m, int split, int *tiles_xy)
{
    const int64_t mt = m / SY_T;
    int64_t k = 0;
    for (int64_t Il = 0; Il < mt; ++Il)
        for (int64_t J = rank * mt + Il; J < (rank + 1) * mt; ++J)
            for (int part = 0; part < split; ++part) {
                if (tiles_xy) { tiles_xy[2 * k] = (int)Il | (part << 24); tiles_xy[2 * k + 1] = (int)J; }
                ++k;
            }
    for (int b = 0; b < sp->nblk; ++b)
        for (int64_t Il = 0; Il < sp->row_end[b]; ++Il)
            for (int64_t Jl = sp->col_begin[b]; Jl < mt; ++Jl)
                for (int part = 0; part < split; ++part) {
                    if (tiles_xy) { tiles_xy[2 * k] = (int)Il | (part << 24); tiles_xy[2 * k + 1] = (int)(sp->blk_rank[b] * mt + Jl); }
                    ++k;
                }
    return k;
}

extern "C" int cgk_symv_comm(const double *A, const double *v, const double *w, double *y, const void *tiles,
                             int64_t ntiles, double *rowpart, double *colpart, double *partials,
                             double *dot_out, cgk_scalars *sc, int64_t m, int64_t n, int rows_in_flight,
                             int split, const cgk_symplan *sp, const cgk_peers *peers, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (m % SY_T != 0 || n % SY_T != 0 || ntiles < 1 || !sp || !peers || !w) return (int)cudaErrorInvalidValue;
    if (split != 1 && split != 2 && split != 4) return (int)cudaErrorInvalidValue;
    const int2 *tl = (const int2 *)tiles;
    switch (rows_in_flight) {
    case 1: cgk_launch(cgb_halfmat_tile_comm_kernel<1>, dim3((unsigned)ntiles), dim3(SY_NT), 0, st, A, v, tl, rowpart, colpart, sc, (long long)m, (long long)n, split, *peers); break;
    case 2: cgk_launch(cgb_halfmat_tile_comm_kernel<2>, dim3((unsigned)ntiles), dim3(SY_NT), 0, st, A, v, tl, rowpart, colpart, sc, (long long)m, (long long)n, split, *peers); break;
    default: cgk_launch(cgb_halfmat_tile_comm_kernel<4>, dim3((unsigned)ntiles), dim3(SY_NT), 0, st, A, v, tl, rowpart, colpart, sc, (long long)m, (long long)n, split, *peers); break;
    }
    if (sp->nblk > 0)
        cgk_launch(cgb_halfmat_push_kernel, dim3((unsigned)(m / SY_NT), (unsigned)sp->nblk), dim3(SY_NT), 0, st, colpart, sc, (long long)m, (long long)n, split, *sp, *peers);
    cgk_launch(cgb_halfmat_finish_comm_kernel, dim3((unsigned)(m / SY_FR)), dim3(SY_NT), 0, st, rowpart, colpart, w, y, partials, dot_out, sc,
               (long long)m, (long long)n, split, *sp, *peers);
    return (int)cudaGetLastError();
}
