/* cg_kernels.cu -- hand-written sm_100a kernels for the dense FP64 CG iteration.
 *
 * Reference path replaced: the row-block matVec
 * (parallel_cg.c:172-184), the dot products (parallel_cg.c:211-221) and the
 * scale/add/sub vector passes (parallel_cg.c:200-245) as driven by conjugrad()
 * (parallel_cg.c:283-323).  The hot kernel is a bandwidth-bound FP64 GEMV:
 * single-vector, so no tensor cores; everything that matters is bytes in
 * flight, coalescing and not re-reading A.
 *
 * Algorithmic bytes per launch: gemv 8*(m*n + n + m),
 * xr_update 8*6m, p_update 8*3m.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include "cg_kernels.h"
#include "spd_gen.cuh"

/* Launch mode of the calling thread: 1 = kernels of the iteration carry the programmatic-
 * dependent-launch attribute (cg_device.cuh).  Set by the solver around one rank's iterations. */
static thread_local int tl_pdl = 0;
extern "C" void cgk_set_pdl(int on) { tl_pdl = on ? 1 : 0; }
extern "C" int cgk_pdl_enabled(void) { return tl_pdl; }

#define CGK_MAX_VEC_CTAS 592          /* 148 SMs x 4 */

#include "cg_device.cuh"

/* ------------------------------------------------------------------ GEMV (LDG) */
/* One CTA owns R adjacent rows (one contiguous R*n*8-byte panel of A) and walks
 * all n columns; thread t of NT owns VW adjacent columns of every step, so a
 * warp reads 32*VW*8 contiguous bytes per row per load and the CTA NT*VW*8.
 * The direction-vector chunk is loaded once per step into registers and reused
 * for the R rows (v traffic = 1/R of A traffic, served by L1/L2; v <= 1 MiB).
 * U steps are issued back to back: R*U independent 128/256-bit loads in flight
 * per thread.  Row sums: registers -> warp shuffle -> shared -> fixed-order add.
 * Epilogue: y[row] and the CTA's partial of w.y (the fused p.Ap). */
/* Columns [cb, ce) of the CTA's R rows: acc[r] += sum_c a[r][c] * v[c].  cb, ce are multiples
 * of VW.  COH selects ordinary coherent loads of v (peers write other parts of it meanwhile). */
template <int R, int U, int NT, int VW, bool COH, bool HOLD = false>
__device__ __forceinline__ void gemv_accumulate(double (&acc)[R], const double *const (&a)[R],
                                                const double *__restrict__ v, long long cb,
                                                long long ce, int tid)
{
    constexpr long long STEP = (long long)NT * VW;
    constexpr long long TILE = STEP * U;
    const long long main_end = cb + ((ce - cb) / TILE) * TILE;

    for (long long c = cb + (long long)tid * VW; c < main_end; c += TILE) {
        double pv[U][VW];
        double av[R][U][VW];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            if constexpr (COH) Ld<VW>::coherent(v + c + u * STEP, pv[u]);
            else Ld<VW>::keep(v + c + u * STEP, pv[u]);
        }
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if constexpr (HOLD) Ld<VW>::hold(a[r] + c + u * STEP, av[r][u]);
                else Ld<VW>::stream(a[r] + c + u * STEP, av[r][u]);
            }
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int e = 0; e < VW; ++e)
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = fma(av[r][u][e], pv[u][e], acc[r]);
    }
    /* column tail */
    for (long long c = main_end + (long long)tid * VW; c < ce; c += STEP) {
        double pv[VW];
        double av[R][VW];
        if constexpr (COH) Ld<VW>::coherent(v + c, pv);
        else Ld<VW>::keep(v + c, pv);
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if constexpr (HOLD) Ld<VW>::hold(a[r] + c, av[r]);
            else Ld<VW>::stream(a[r] + c, av[r]);
        }
#pragma unroll
        for (int e = 0; e < VW; ++e)
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = fma(av[r][e], pv[e], acc[r]);
    }
}

template <int R, int U, int NT, int VW>
__global__ void __launch_bounds__(NT)
cgb_rowblock_matvec_ldg(const double *__restrict__ A, const double *__restrict__ v,
                const double *__restrict__ w, double *__restrict__ y,
                double *__restrict__ partials, double *__restrict__ dot_out,
                cgk_scalars *__restrict__ sc, long long m, long long n)
{
    const int stop = *(volatile int *)&sc->stop_req;
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->stopped = stop;
        return;
    }
    constexpr int NW = NT / 32;
    __shared__ double s_row[R][NW];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const long long row0 = (long long)blockIdx.x * R;

    const double *a[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        long long row = row0 + r;
        if (row > m - 1) row = m - 1;           /* clamp: reads stay in bounds, result dropped */
        a[r] = A + row * n;
    }
    double acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.0;

    gemv_accumulate<R, U, NT, VW, false>(acc, a, v, 0, n, tid);   /* n % VW == 0 by the plan */

#pragma unroll
    for (int r = 0; r < R; ++r) {
        acc[r] = warp_sum(acc[r]);
        if (lane == 0) s_row[r][wid] = acc[r];
    }
    __syncthreads();
    double contrib = 0.0;
    if (tid < R) {
        double s = 0.0;
#pragma unroll
        for (int q = 0; q < NW; ++q) s += s_row[tid][q];
        const long long row = row0 + tid;
        if (row < m) {
            y[row] = s;
            if (w != nullptr) contrib = s * w[row];
        }
    }
    if (w == nullptr) return;
    if (wid == 0) contrib = warp_sum(contrib);          /* lanes >= R hold 0 */
    cta_partials_finish<NT>(contrib, partials, dot_out, &sc->ticket_gemv);
}

/* Small systems (A within a few L2 sizes; BASELINE configs[1], n = 4096: A = 134 MB against
 * 126 MB of L2): a cyclic stream through a cache slightly too small evicts every line just
 * before its reuse, so the plain kernel pays full HBM traffic every iteration.  Here the
 * first keep_rows rows are loaded evict_last and stay resident from one iteration to the
 * next; the remaining rows stream past them evict_first.  Same arithmetic, same order. */
/* 7 CTAs of 128 threads per SM (<= 72 registers) where the tile is small: at n = 4096 the grid
 * of 1024 CTAs then fits the 148 SMs in ONE wave (6 per SM would leave a second wave of 136) */
template <int R, int U, int NT, int VW>
__global__ void __launch_bounds__(NT, (NT == 128 && R * U * VW <= 16) ? 7 : 0)
cgb_rowblock_matvec_keep(const double *__restrict__ A, const double *__restrict__ v,
                     const double *__restrict__ w, double *__restrict__ y,
                     double *__restrict__ partials, double *__restrict__ dot_out,
                     cgk_scalars *__restrict__ sc, long long m, long long n, long long keep_rows)
{
    /* This is also the GEMV of the programmatic-dependent-launch chain (cg_device.cuh): the
     * first tile of A (never written during a solve) is requested BEFORE the wait, so the
     * kernel's launch and its first memory round trip overlap the tail of the vector kernel;
     * v and w are read after the wait with ordinary (not .nc) loads. */
    constexpr int NW = NT / 32;
    constexpr long long STEP = (long long)NT * VW, TILE = STEP * U;
    __shared__ double s_row[R][NW];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const long long row0 = (long long)blockIdx.x * R;

    const double *a[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        long long row = row0 + r;
        if (row > m - 1) row = m - 1;
        a[r] = A + row * n;
    }
    const bool held = row0 + R <= keep_rows;
    const bool pre = n >= TILE;                       /* a full first tile exists */
    double av0[R][U][VW];
    if (pre) {
        const long long c0 = (long long)tid * VW;
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
            for (int u = 0; u < U; ++u) {
                if (held) Ld<VW>::hold(a[r] + c0 + u * STEP, av0[r][u]);
                else Ld<VW>::stream(a[r] + c0 + u * STEP, av0[r][u]);
            }
    }
    pdl_wait();
    pdl_launch_dependents();
    const int stop = *(volatile int *)&sc->stop_req;
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->stopped = stop;
        return;
    }
    double acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.0;
    if (pre) {                                        /* the loop's first trip, same order of adds */
        const long long c0 = (long long)tid * VW;
        double pv[U][VW];
#pragma unroll
        for (int u = 0; u < U; ++u) Ld<VW>::coherent(v + c0 + u * STEP, pv[u]);
#pragma unroll
        for (int u = 0; u < U; ++u)
#pragma unroll
            for (int e = 0; e < VW; ++e)
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = fma(av0[r][u][e], pv[u][e], acc[r]);
    }
    const long long cb = pre ? TILE : 0;

    if (held) gemv_accumulate<R, U, NT, VW, true, true>(acc, a, v, cb, n, tid);
    else gemv_accumulate<R, U, NT, VW, true, false>(acc, a, v, cb, n, tid);

#pragma unroll
    for (int r = 0; r < R; ++r) {
        acc[r] = warp_sum(acc[r]);
        if (lane == 0) s_row[r][wid] = acc[r];
    }
    __syncthreads();
    double contrib = 0.0;
    if (tid < R) {
        double s = 0.0;
#pragma unroll
        for (int q = 0; q < NW; ++q) s += s_row[tid][q];
        const long long row = row0 + tid;
        if (row < m) {
            y[row] = s;
            if (w != nullptr) contrib = s * __ldcg(w + row);
        }
    }
    if (w == nullptr) return;
    if (wid == 0) contrib = warp_sum(contrib);
    cta_partials_finish<NT>(contrib, partials, dot_out, &sc->ticket_gemv);
}

/* One full tile of the exchange-fused kernel: columns c + u*STEP (u < U) of the CTA's R rows,
 * p first, then A, as in gemv_accumulate; p with ordinary loads (peers write other parts of it). */
template <int R, int U, int NT, int VW>
__device__ __forceinline__ void comm_tile(double (&acc)[R], const double *const (&a)[R], const double *v, long long c)
{
    constexpr long long STEP = (long long)NT * VW;
    double pv[U][VW];
    double av[R][U][VW];
#pragma unroll
    for (int u = 0; u < U; ++u) Ld<VW>::plain(v + c + u * STEP, pv[u]);
#pragma unroll
    for (int r = 0; r < R; ++r)
#pragma unroll
        for (int u = 0; u < U; ++u) Ld<VW>::stream(a[r] + c + u * STEP, av[r][u]);
#pragma unroll
    for (int u = 0; u < U; ++u)
#pragma unroll
        for (int e = 0; e < VW; ++e)
#pragma unroll
            for (int r = 0; r < R; ++r) acc[r] = fma(av[r][u][e], pv[u][e], acc[r]);
}

/* The same product with the p exchange fused in (row-sharded A, ranks > 1; replaces
 * MPI_Allgather + matVec + vecVec + MPI_Allreduce, parallel_cg.c:290-294).  Order of work in
 * a CTA: (1) the rank's own column block, whose slice of p this GPU's p-update kernel wrote
 * before this launch; (2) wait until every peer's slice of event ep has landed in the local
 * p_full (flags in the local mailbox); (3) the remote column blocks.  CTAs of the first wave
 * hide the exchange behind (1); later waves find the flags already set.  The last CTA pushes
 * the rank's p.Ap partial into every rank's mailbox and advances the epoch. */
template <int R, int U, int NT, int VW>
__global__ void __launch_bounds__(NT)
cgb_rowblock_matvec_comm(const double *__restrict__ A, const double *v,
                     const double *w, double *__restrict__ y,
                     double *__restrict__ partials, double *__restrict__ dot_out,
                     cgk_scalars *sc, long long m, long long n, const cgk_peers peers)
{
    pdl_wait();                       /* launch overlap only: everything below depends on the previous kernel */
    pdl_launch_dependents();
    const int stop = *(volatile int *)&sc->stop_req;
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->stopped = stop;
        return;
    }
    const unsigned int ep = *(volatile unsigned int *)&sc->ep_gemv;
    const unsigned int gen = *(volatile unsigned int *)&sc->gen;
    constexpr int NW = NT / 32;
    __shared__ double s_row[R][NW];
    const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const long long row0 = (long long)blockIdx.x * R;

    const double *a[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        long long row = row0 + r;
        if (row > m - 1) row = m - 1;
        a[r] = A + row * n;
    }
    double acc[R];
#pragma unroll
    for (int r = 0; r < R; ++r) acc[r] = 0.0;

    const long long lo = (long long)peers.rank * m, hi = lo + m;
    const unsigned int tag = cgk_tag(gen, ep);
    constexpr long long STEP = (long long)NT * VW;
    constexpr long long TILE = STEP * U;
    /* CGB_PROBE=1 only: thread 0 leaves time stamps (layout in cg_kernels.h); NULL otherwise */
    unsigned long long *pb = (peers.probe && tid == 0 && blockIdx.x < CGK_PROBE_MAX_PRODUCT_CTAS) ? peers.probe + 8ull * blockIdx.x : nullptr;
    if (pb) { pb[0] = global_ns(); pb[5] = sm_id(); pb[6] = ep; pb[7] = 1ull; }
    if (m % TILE == 0) {
        /* The columns in wrapped order lo, lo+TILE, ..., n-TILE, 0, ..., lo-TILE (tiles never
         * straddle a block boundary), as three pieces: the own block -- the plain kernel's loop;
         * ONE tile at which the peers' flags are tested, with its first loads of A already issued
         * (A is local and never waits); the remaining remote tiles -- the plain loop again.
         * The flags were REQUESTED at kernel entry (one relaxed load per lane g < G, nobody waits
         * for it), so for every CTA but the first wave the test is a register compare: round 2's
         * probes (tools/probe_exchange.py) showed the former poll + acquire at this point costing
         * ~1.4 us in every one of the 8192 CTAs, flags long set -- 1 % of the kernel.  The loads
         * of p that follow take their address from the flag words (warp_flags_ready). */
        const unsigned long long early = flags_request(peers, tag);
        const long long tiles = n / TILE;
        const long long local_tiles = m / TILE;
        long long c = lo + (long long)tid * VW;
        for (long long k = 0; k < local_tiles; ++k) {
            comm_tile<R, U, NT, VW>(acc, a, v, c);
            c += TILE;
        }
        if (c >= n) c -= n;
        const double *vq = v;
        if (local_tiles < tiles) {
            double pv0[VW];
            double av0[R][VW];
#pragma unroll
            for (int r = 0; r < R; ++r) Ld<VW>::stream(a[r] + c, av0[r]);
            long long w0 = 0;
            if (pb) { pb[1] = global_ns(); w0 = clock64(); }
            vq = order_after_acquire(vq + warp_flags_ready(peers, tag, early, sc));
            if (pb) { pb[4] = (unsigned long long)(clock64() - w0); pb[2] = global_ns(); }
            Ld<VW>::plain(vq + c, pv0);
#pragma unroll
            for (int e = 0; e < VW; ++e)
#pragma unroll
                for (int r = 0; r < R; ++r) acc[r] = fma(av0[r][e], pv0[e], acc[r]);
            if constexpr (U > 1) {
                double pv[U - 1][VW];
                double av[R][U - 1][VW];
#pragma unroll
                for (int u = 1; u < U; ++u) Ld<VW>::plain(vq + c + u * STEP, pv[u - 1]);
#pragma unroll
                for (int r = 0; r < R; ++r)
#pragma unroll
                    for (int u = 1; u < U; ++u) Ld<VW>::stream(a[r] + c + u * STEP, av[r][u - 1]);
#pragma unroll
                for (int u = 1; u < U; ++u)
#pragma unroll
                    for (int e = 0; e < VW; ++e)
#pragma unroll
                        for (int r = 0; r < R; ++r) acc[r] = fma(av[r][u - 1][e], pv[u - 1][e], acc[r]);
            }
            c += TILE;
            if (c >= n) c -= n;
        }
        for (long long k = local_tiles + 1; k < tiles; ++k) {
            comm_tile<R, U, NT, VW>(acc, a, vq, c);
            c += TILE;
            if (c >= n) c -= n;
        }
    } else {
        /* any shape: own block, wait, the two remote ranges */
        gemv_accumulate<R, U, NT, VW, true>(acc, a, v, lo, hi, tid);
        wait_p_flags(peers, tag, sc);
        gemv_accumulate<R, U, NT, VW, true>(acc, a, v, 0, lo, tid);
        gemv_accumulate<R, U, NT, VW, true>(acc, a, v, hi, n, tid);
    }

    if (pb) pb[3] = global_ns();
#pragma unroll
    for (int r = 0; r < R; ++r) {
        acc[r] = warp_sum(acc[r]);
        if (lane == 0) s_row[r][wid] = acc[r];
    }
    __syncthreads();
    double contrib = 0.0;
    if (tid < R) {
        double s = 0.0;
#pragma unroll
        for (int q = 0; q < NW; ++q) s += s_row[tid][q];
        const long long row = row0 + tid;
        if (row < m) {
            y[row] = s;
            contrib = s * w[row];
        }
    }
    if (wid == 0) contrib = warp_sum(contrib);
    double total = 0.0;
    /* device-scope fences: the CTAs store nothing into peer memory, and the dot the last CTA
     * sends is a self-validating {data, tag} pair (a system-scope fence per CTA costs microseconds) */
    if (cta_partials_finish_sys<NT, false>(contrib, partials, dot_out, &sc->ticket_gemv, &total) && tid == 0) {
        for (int g = 0; g < peers.ranks; ++g)
            ll_send(&peers.mbox[g]->dot[0][ep & 1u][peers.rank][0], total, tag);
        sc->ep_gemv = ep + 1u;
    }
}

/* ------------------------------------------------------------------ GEMV (bulk-async ring) */
/* Persistent, warp-specialised: one producer thread streams [R rows x C cols]
 * tiles of A plus the matching C-column chunk of v into a STAGES-deep shared
 * memory ring with cp.async.bulk (the TMA engine, SASS UBLKCP) completing on
 * mbarriers; NT consumer threads drain the ring with conflict-free 128-bit
 * LDS.  No registers are tied up by loads in flight: bytes in flight per SM =
 * STAGES * (R+1) * C * 8, chosen from the latency-bandwidth product.  A CTA
 * walks row panels blockIdx.x, blockIdx.x+grid, ... so the pipeline never
 * drains between panels (the producer runs ahead into the next panel while the
 * consumers reduce the previous one). */
__device__ __forceinline__ uint32_t smem_u32(const void *p)
{
    return (uint32_t)__cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_LOOP:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra WAIT_DONE;\n"
        "bra WAIT_LOOP;\n"
        "WAIT_DONE:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void *dst, const void *src, uint32_t bytes,
                                         uint64_t *bar, uint64_t policy)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
        ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}

template <int R, int C, int STAGES, int NT>
__global__ void __launch_bounds__(NT + 32)
cgb_rowblock_matvec_bulk(const double *__restrict__ A, const double *__restrict__ v,
                 const double *__restrict__ w, double *__restrict__ y,
                 double *__restrict__ partials, double *__restrict__ dot_out,
                 cgk_scalars *__restrict__ sc, long long m, long long n)
{
    const int stop = *(volatile int *)&sc->stop_req;
    if (stop) {
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->stopped = stop;
        return;
    }
    constexpr int NW = NT / 32;
    constexpr int STAGE_DOUBLES = (R + 1) * C;            /* R row segments + v chunk */
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double *ring = reinterpret_cast<double *>(smem_raw);
    uint64_t *full = reinterpret_cast<uint64_t *>(ring + (size_t)STAGES * STAGE_DOUBLES);
    uint64_t *empty = full + STAGES;
    __shared__ double s_row[R][NW];

    const int tid = threadIdx.x;
    const long long panels = (m + R - 1) / R;
    const long long tiles = n / C;                        /* plan guarantees n % C == 0 */

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], NW); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    if (tid >= NT) {
        /* ---------------- producer warp: one elected lane issues the copies */
        if (tid == NT) {
            uint64_t pol_stream, pol_keep;
            asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol_stream));
            asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol_keep));
            uint32_t it = 0;
            for (long long pnl = blockIdx.x; pnl < panels; pnl += gridDim.x) {
                const long long row0 = pnl * R;
                for (long long t = 0; t < tiles; ++t, ++it) {
                    const int s = it % STAGES;
                    const uint32_t ph = (it / STAGES) & 1u;
                    mbar_wait(&empty[s], ph ^ 1u);
                    double *dst = ring + (size_t)s * STAGE_DOUBLES;
                    mbar_expect_tx(&full[s], (uint32_t)(STAGE_DOUBLES * sizeof(double)));
                    bulk_g2s(dst + (size_t)R * C, v + t * C, C * sizeof(double), &full[s], pol_keep);
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        long long row = row0 + r;
                        if (row > m - 1) row = m - 1;
                        bulk_g2s(dst + (size_t)r * C, A + row * n + t * C, C * sizeof(double),
                                 &full[s], pol_stream);
                    }
                }
            }
        }
        return;
    }

    /* ---------------- consumer warps */
    const int lane = tid & 31, wid = tid >> 5;
    constexpr int CH = C / (NT * 2);                      /* 16-byte chunks per thread per row */
    static_assert(C % (NT * 2) == 0, "tile width must be a multiple of 2*NT");
    uint32_t it = 0;
    bool any_dot = false;
    double cta_dot = 0.0;
    for (long long pnl = blockIdx.x; pnl < panels; pnl += gridDim.x) {
        const long long row0 = pnl * R;
        double acc[R];
#pragma unroll
        for (int r = 0; r < R; ++r) acc[r] = 0.0;
        for (long long t = 0; t < tiles; ++t, ++it) {
            const int s = it % STAGES;
            const uint32_t ph = (it / STAGES) & 1u;
            mbar_wait(&full[s], ph);
            const double2 *tile = reinterpret_cast<const double2 *>(ring + (size_t)s * STAGE_DOUBLES);
            const double2 *vt = tile + (size_t)R * (C / 2);
            double2 pv[CH];
#pragma unroll
            for (int k = 0; k < CH; ++k) pv[k] = vt[tid + k * NT];
#pragma unroll
            for (int r = 0; r < R; ++r) {
#pragma unroll
                for (int k = 0; k < CH; ++k) {
                    const double2 a2 = tile[(size_t)r * (C / 2) + tid + k * NT];
                    acc[r] = fma(a2.x, pv[k].x, acc[r]);
                    acc[r] = fma(a2.y, pv[k].y, acc[r]);
                }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[s]);
        }
        /* panel reduction (consumers only: named barrier 1, NT threads) */
#pragma unroll
        for (int r = 0; r < R; ++r) {
            acc[r] = warp_sum(acc[r]);
            if (lane == 0) s_row[r][wid] = acc[r];
        }
        asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
        if (tid < R) {
            double sum = 0.0;
#pragma unroll
            for (int q = 0; q < NW; ++q) sum += s_row[tid][q];
            const long long row = row0 + tid;
            double contrib = 0.0;
            if (row < m) {
                y[row] = sum;
                if (w != nullptr) contrib = sum * w[row];
            }
            /* rows of one panel in fixed lane order, panels in visiting order */
#pragma unroll
            for (int o = R / 2; o > 0; o >>= 1) contrib += __shfl_xor_sync((1u << R) - 1u, contrib, o);
            cta_dot += contrib;
            any_dot = true;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
    }
    (void)any_dot;
    if (w == nullptr) return;
    /* consumers-only finish: same fixed-order scheme as cta_partials_finish */
    __shared__ int s_last;
    __shared__ double s_red[NW];
    if (tid == 0) {
        partials[blockIdx.x] = cta_dot;
        __threadfence();
        unsigned int t = atomicAdd(&sc->ticket_gemv, 1u);
        s_last = (t == gridDim.x - 1);
    }
    asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
    if (!s_last) return;
    __threadfence();
    double s = 0.0;
    for (unsigned int i = tid; i < gridDim.x; i += NT) s += __ldcg(partials + i);
    s = warp_sum(s);
    if (lane == 0) s_red[wid] = s;
    asm volatile("bar.sync 1, %0;" ::"n"(NT) : "memory");
    if (tid == 0) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < NW; ++q) t += s_red[q];
        *dot_out = t;
        sc->ticket_gemv = 0u;
    }
}

/* ------------------------------------------------------------------ vector phase */
#define VEC_NT 256

/* CTA-level sum of one value per thread -> valid in thread 0 (fixed tree). */
__device__ __forceinline__ double vec_cta_sum(double s)
{
    __shared__ double s_w[VEC_NT / 32];
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) s_w[threadIdx.x >> 5] = s;
    __syncthreads();
    double mine = 0.0;
    if (threadIdx.x == 0)
        for (int q = 0; q < VEC_NT / 32; ++q) mine += s_w[q];
    return mine;
}

/* Thread 0 of the last CTA of a kernel that pushed a p slice: post the slice's tag to every
 * peer (the system fence in cta_partials_finish_sys / below orders it behind the slice). */
__device__ __forceinline__ void post_p_flags(const cgk_peers &peers, unsigned int tag)
{
    for (int g = 0; g < peers.ranks; ++g)
        if (g != peers.rank) st_relaxed_sys(&peers.mbox[g]->p_flag[peers.rank], (unsigned long long)tag);
}

/* r = b - Ax0; p = b - Ax0 (the reference computes it twice, parallel_cg.c:284-285;
 * same value); rr partial.  Not guarded by stop_req: runs before the loop.
 * COMM: the p slice also goes to every peer's p_full (event 0 of the all-gather,
 * parallel_cg.c:290) and the r.r partial to every mailbox (the all-reduce of :287). */
template <bool COMM>
__global__ void __launch_bounds__(VEC_NT)
cgb_init_residual_kernel(const double *__restrict__ b, const double *__restrict__ Ax,
                     double *__restrict__ r, double *p_loc,
                     double *__restrict__ partials, double *__restrict__ rr_out,
                     cgk_scalars *sc, long long m, const cgk_peers peers)
{
    double s = 0.0;
    const long long lo = COMM ? (long long)peers.rank * m : 0;
    for (long long i = (long long)blockIdx.x * VEC_NT + threadIdx.x; i < m;
         i += (long long)gridDim.x * VEC_NT) {
        const double ri = __dsub_rn(b[i], Ax[i]);
        r[i] = ri;
        p_loc[i] = ri;
        if constexpr (COMM) {
            for (int g = 0; g < peers.ranks; ++g)
                if (g != peers.rank) peers.p_full[g][lo + i] = ri;
        }
        s = fma(ri, ri, s);
    }
    const double mine = vec_cta_sum(s);
    if constexpr (COMM) {
        const unsigned int gen = *(volatile unsigned int *)&sc->gen;
        double total = 0.0;
        if (cta_partials_finish_sys<VEC_NT>(mine, partials, rr_out, &sc->ticket_vec, &total) && threadIdx.x == 0) {
            const unsigned int tag = cgk_tag(gen, 0u);
            for (int g = 0; g < peers.ranks; ++g)
                ll_send(&peers.mbox[g]->dot[1][0][peers.rank][0], total, tag);
            post_p_flags(peers, tag);
        }
    } else {
        cta_partials_finish<VEC_NT>(mine, partials, rr_out, &sc->ticket_vec);
    }
}

/* alpha = rho / p.Ap (parallel_cg.c:298); x += alpha*p through a rounded
 * product (scalarVec then vecAdd, :299,303); r -= alpha*Ap likewise (:304,308);
 * partial r.r (:309).  __dmul_rn/__dadd_rn keep the reference's rounding
 * points: no FMA contraction in the element-wise updates.
 * COMM: p.Ap is the rank-ordered sum of the G partials the GEMVs pushed into the local
 * mailbox (MPI_Allreduce :294); the r.r partial is pushed to every mailbox (:313). */
template <bool COMM>
__global__ void __launch_bounds__(VEC_NT)
cgb_xr_update_kernel(double *__restrict__ x, double *__restrict__ r,
                 const double *p_loc, const double *__restrict__ Ap,
                 double *__restrict__ partials, double *__restrict__ rr_out,
                 cgk_scalars *sc, int cur, int guard, long long m, const cgk_peers peers,
                 const double *__restrict__ dot_partials, int n_dot)
{
    /* the first trip's operands are requested before anything is waited for: the scalars (or the
     * peers' partials) and the vectors then cost one memory round trip, not two */
    long long i = (long long)blockIdx.x * VEC_NT + threadIdx.x;
    const bool first = i < m;
    double x0 = 0.0, p0 = 0.0, r0 = 0.0, a0 = 0.0;
    if (first) { x0 = x[i]; p0 = p_loc[i]; r0 = r[i]; }        /* last written two kernels back */
    const double rho = sc->rr[cur];
    pdl_wait();                                                /* Ap, the dot partials, the flags: the product */
    pdl_launch_dependents();
    if (first) a0 = __ldcg(Ap + i);
    if (*(volatile int *)&sc->stop_req) return;
    unsigned int ep = 0u, gen = 0u;
    double pAp;
    unsigned long long *pb = nullptr;
    if constexpr (COMM) {
        ep = *(volatile unsigned int *)&sc->ep_xr;
        gen = *(volatile unsigned int *)&sc->gen;
        if (peers.probe && threadIdx.x == 0 && blockIdx.x < CGK_PROBE_VEC_CTAS) pb = peers.probe + 8ull * (CGK_PROBE_MAX_PRODUCT_CTAS + blockIdx.x);
        long long w0 = 0;
        if (pb) { pb[0] = global_ns(); pb[5] = sm_id(); pb[6] = ep; pb[7] = 1ull; w0 = clock64(); }
        pAp = ll_gather_sum(&peers.mbox[peers.rank]->dot[0][ep & 1u][0][0], peers.ranks, cgk_tag(gen, ep), sc);
        if (pb) { pb[4] = (unsigned long long)(clock64() - w0); pb[1] = global_ns(); }
    } else if (n_dot > 0) {
        /* deferred p.Ap: the GEMV left one partial per CTA (fixed order, same bits in every CTA) */
        pAp = cta_sum_partials<VEC_NT>(dot_partials, n_dot);
        if (blockIdx.x == 0 && threadIdx.x == 0) sc->pAp = pAp;
    } else {
        pAp = sc->pAp;
    }
    const double alpha = rho / pAp;
    if (guard && !isfinite(alpha)) {
        /* every thread of every rank sees the same alpha: all leave x, r alone.  Later kernels
         * read the flags (p_update of this iteration: breakdown; the next GEMV: stop_req). */
        if (blockIdx.x == 0 && threadIdx.x == 0) { sc->breakdown = 3; sc->stop_req = 3; sc->pAp = pAp; }
        return;
    }
    double s = 0.0;
    if (first) {
        x[i] = __dadd_rn(x0, __dmul_rn(p0, alpha));
        const double ri = __dsub_rn(r0, __dmul_rn(a0, alpha));
        r[i] = ri;
        s = fma(ri, ri, s);
        i += (long long)gridDim.x * VEC_NT;
    }
    for (; i < m; i += (long long)gridDim.x * VEC_NT) {
        x[i] = __dadd_rn(x[i], __dmul_rn(p_loc[i], alpha));
        const double ri = __dsub_rn(r[i], __dmul_rn(__ldcg(Ap + i), alpha));
        r[i] = ri;
        s = fma(ri, ri, s);
    }
    const double mine = vec_cta_sum(s);
    if constexpr (COMM) {
        double total = 0.0;
        /* x and r are local: device-scope fences (see cgb_rowblock_matvec_comm) */
        if (cta_partials_finish_sys<VEC_NT, false>(mine, partials, rr_out, &sc->ticket_vec, &total) && threadIdx.x == 0) {
            const unsigned int e = ep + 1u;                       /* r.r events: 0 = init, k+1 = iteration k */
            for (int g = 0; g < peers.ranks; ++g)
                ll_send(&peers.mbox[g]->dot[1][e & 1u][peers.rank][0], total, cgk_tag(gen, e));
            sc->pAp = pAp;
            sc->ep_xr = e;
        }
        if (pb) pb[2] = global_ns();
    } else {
        cta_partials_finish<VEC_NT>(mine, partials, rr_out, &sc->ticket_vec);
    }
}

/* Stop rule and direction update (parallel_cg.c:314-323): if sqrt(beta) < eps the
 * loop breaks BEFORE p is touched (x and r updated, p stale); otherwise
 * p = r + (beta/rsold)*p via a rounded product, and rsold = beta (here: the
 * rr[] ping-pong, parity `cur` baked into the launch).  Thread 0 of CTA 0 does
 * the bookkeeping; every thread derives the same decision from beta.
 * COMM: beta is the rank-ordered sum of the mailbox partials (same bits on every rank, so
 * every rank takes the same branch); the new slice of p is stored into every peer's p_full
 * as it is computed and the last CTA posts the tag the peers' next GEMV waits for. */
template <bool COMM>
__global__ void __launch_bounds__(VEC_NT)
cgb_p_update_kernel(double *p_loc, const double *__restrict__ r,
                cgk_scalars *sc, int cur, double eps, int ignore_stop,
                long long max_iter, long long m, const cgk_peers peers,
                const double *__restrict__ rr_partials, int n_rr)
{
    long long i = (long long)blockIdx.x * VEC_NT + threadIdx.x;
    const bool first = i < m;
    double r0 = 0.0, p0 = 0.0;
    if (first) p0 = p_loc[i];                                  /* last written three kernels back */
    const double rho = sc->rr[cur];
    pdl_wait();                                                /* r, the r.r partials, the flags: xr_update */
    pdl_launch_dependents();
    if (first) r0 = __ldcg(r + i);
    if (*(volatile int *)&sc->stopped || *(volatile int *)&sc->breakdown) return;
    unsigned int ep = 0u, gen = 0u;
    double beta;
    unsigned long long *pb = nullptr;
    if constexpr (COMM) {
        ep = *(volatile unsigned int *)&sc->ep_p;
        gen = *(volatile unsigned int *)&sc->gen;
        const unsigned int e = ep + 1u;
        if (peers.probe && threadIdx.x == 0 && blockIdx.x < CGK_PROBE_VEC_CTAS)
            pb = peers.probe + 8ull * (CGK_PROBE_MAX_PRODUCT_CTAS + CGK_PROBE_VEC_CTAS + blockIdx.x);
        long long w0 = 0;
        if (pb) { pb[0] = global_ns(); pb[5] = sm_id(); pb[6] = ep; pb[7] = 1ull; w0 = clock64(); }
        beta = ll_gather_sum(&peers.mbox[peers.rank]->dot[1][e & 1u][0][0], peers.ranks, cgk_tag(gen, e), sc);
        if (pb) { pb[4] = (unsigned long long)(clock64() - w0); pb[1] = global_ns(); }
    } else if (n_rr > 0) {
        beta = cta_sum_partials<VEC_NT>(rr_partials, n_rr);   /* deferred r.r of xr_update */
    } else {
        beta = sc->rr[cur ^ 1];
    }
    const bool hit = sqrt(beta) < eps;
    const bool brk = hit && !ignore_stop;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const long long k = sc->iters;
        sc->iters = k + 1;
        sc->final_beta = beta;
        if (COMM || n_rr > 0) sc->rr[cur ^ 1] = beta;     /* read as rho by the NEXT iteration only */
        if (hit && sc->first_hit < 0) sc->first_hit = k;
        if (brk) { sc->break_k = k; sc->stop_req = 1; }
        else if (k + 1 >= max_iter) sc->stop_req = 2;
    }
    if (brk) return;
    const double coef = beta / rho;
    const long long lo = COMM ? (long long)peers.rank * m : 0;
    for (bool head = first; i < m; i += (long long)gridDim.x * VEC_NT, head = false) {
        const double pi = head ? __dadd_rn(r0, __dmul_rn(p0, coef)) : __dadd_rn(__ldcg(r + i), __dmul_rn(p_loc[i], coef));
        p_loc[i] = pi;
        if constexpr (COMM) {
            for (int g = 0; g < peers.ranks; ++g)
                if (g != peers.rank) peers.p_full[g][lo + i] = pi;
        }
    }
    if constexpr (COMM) {
        __shared__ int s_last;
        __syncthreads();
        if (threadIdx.x == 0) {
            __threadfence_system();                 /* cumulative over the CTA's peer stores */
            s_last = (atomicAdd(&sc->ticket_vec, 1u) == gridDim.x - 1);
            if (s_last) {
                __threadfence_system();
                post_p_flags(peers, cgk_tag(gen, ep + 1u));
                sc->ticket_vec = 0u;
                sc->ep_p = ep + 1u;
            }
            if (pb) pb[2] = global_ns();
        }
    }
}

/* xr_update and p_update of the peer exchange in one launch (cgk_xrp_update in cg_kernels.h;
 * parallel_cg.c:298-323 with the MPI_Allreduce of :313 in the middle).  Every CTA of the grid is
 * resident at once (the launcher checks), so a CTA may wait for a value another CTA of the same
 * launch produces: the last CTA of the first phase sends the rank's r.r partial to every
 * mailbox, its own included, and all CTAs pick the G partials up from the local mailbox.
 * Flags: stop_req is read at the very start -- no CTA can write it (phase 2) before every CTA
 * has passed phase 1, i.e. has read it.  Epochs: ep_xr is read at the start and both ep_xr and
 * ep_p are advanced by the last CTA of phase 2, again after every CTA has read them. */
__global__ void __launch_bounds__(VEC_NT)
cgb_xrp_update_kernel(double *__restrict__ x, double *__restrict__ r, double *p_loc,
                      const double *__restrict__ Ap, double *__restrict__ partials,
                      cgk_scalars *sc, int cur, int guard, double eps, int ignore_stop,
                      long long max_iter, long long m, const cgk_peers peers)
{
    long long i = (long long)blockIdx.x * VEC_NT + threadIdx.x;
    const bool first = i < m;
    double x0 = 0.0, p0 = 0.0, r0 = 0.0, a0 = 0.0;
    if (first) { x0 = x[i]; p0 = p_loc[i]; r0 = r[i]; }        /* written by the previous launch of this kernel */
    const double rho = sc->rr[cur];
    pdl_wait();                                                /* Ap, the flags: the product */
    pdl_launch_dependents();
    if (first) a0 = __ldcg(Ap + i);
    if (*(volatile int *)&sc->stop_req) return;
    const unsigned int ep = *(volatile unsigned int *)&sc->ep_xr;
    const unsigned int gen = *(volatile unsigned int *)&sc->gen;
    const unsigned int e = ep + 1u;                            /* r.r events: 0 = init, k+1 = iteration k */
    const double pAp = ll_gather_sum(&peers.mbox[peers.rank]->dot[0][ep & 1u][0][0], peers.ranks, cgk_tag(gen, ep), sc);
    const double alpha = rho / pAp;
    if (guard && !isfinite(alpha)) {
        if (blockIdx.x == 0 && threadIdx.x == 0) { sc->breakdown = 3; sc->stop_req = 3; sc->pAp = pAp; }
        return;
    }
    /* ---- phase 1: x += alpha p, r -= alpha Ap, partial of r.r (parallel_cg.c:299-309) */
    double s = 0.0;
    const long long stride = (long long)gridDim.x * VEC_NT;
    if (first) {
        x[i] = __dadd_rn(x0, __dmul_rn(p0, alpha));
        r0 = __dsub_rn(r0, __dmul_rn(a0, alpha));
        r[i] = r0;
        s = fma(r0, r0, s);
    }
    for (long long j = i + stride; j < m; j += stride) {
        x[j] = __dadd_rn(x[j], __dmul_rn(p_loc[j], alpha));
        const double rj = __dsub_rn(r[j], __dmul_rn(__ldcg(Ap + j), alpha));
        r[j] = rj;
        s = fma(rj, rj, s);
    }
    const double mine = vec_cta_sum(s);
    double total = 0.0;
    if (cta_partials_finish_sys<VEC_NT, false>(mine, partials, &sc->rr_local, &sc->ticket_vec, &total) && threadIdx.x == 0) {
        for (int g = 0; g < peers.ranks; ++g)                  /* MPI_Allreduce(beta), :313 */
            ll_send(&peers.mbox[g]->dot[1][e & 1u][peers.rank][0], total, cgk_tag(gen, e));
        sc->pAp = pAp;
    }
    /* ---- phase 2: stop rule, p = r + (beta/rho) p pushed to every rank (:314-323) */
    const double beta = ll_gather_sum(&peers.mbox[peers.rank]->dot[1][e & 1u][0][0], peers.ranks, cgk_tag(gen, e), sc);
    const bool hit = sqrt(beta) < eps;
    const bool brk = hit && !ignore_stop;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        const long long k = sc->iters;
        sc->iters = k + 1;
        sc->final_beta = beta;
        sc->rr[cur ^ 1] = beta;                               /* read as rho by the NEXT iteration only */
        if (hit && sc->first_hit < 0) sc->first_hit = k;
        if (brk) { sc->break_k = k; sc->stop_req = 1; }
        else if (k + 1 >= max_iter) sc->stop_req = 2;
    }
    const bool lastly = !brk;
    if (lastly) {
        const double coef = beta / rho;
        const long long lo = (long long)peers.rank * m;
        if (first) {
            const double pi = __dadd_rn(r0, __dmul_rn(p0, coef));
            p_loc[i] = pi;
            for (int g = 0; g < peers.ranks; ++g)
                if (g != peers.rank) peers.p_full[g][lo + i] = pi;
        }
        for (long long j = i + stride; j < m; j += stride) {
            const double pj = __dadd_rn(r[j], __dmul_rn(p_loc[j], coef));
            p_loc[j] = pj;
            for (int g = 0; g < peers.ranks; ++g)
                if (g != peers.rank) peers.p_full[g][lo + j] = pj;
        }
    }
    /* the epochs advance on a break too (as the two-kernel form's xr_update does): the ticket runs either way */
    __shared__ int s_last2;
    __syncthreads();
    if (threadIdx.x == 0) {
        if (lastly) __threadfence_system();                   /* cumulative over the CTA's peer stores */
        else __threadfence();
        s_last2 = (atomicAdd(&sc->ticket_p, 1u) == gridDim.x - 1);
        if (s_last2) {
            if (lastly) {
                __threadfence_system();
                post_p_flags(peers, cgk_tag(gen, e));
                sc->ep_p = e;
            }
            sc->ticket_p = 0u;
            sc->ep_xr = e;
        }
    }
}

/* See cgk_cg2_step in cg_kernels.h.  Launch k does the bookkeeping of iteration k-1 (its r.r is
 * only complete now) and, unless the solve ends there, the updates of iteration k.  Every CTA
 * derives the same decisions from the same scalars, none from a flag written in this launch. */
template <bool COMM>
__global__ void __launch_bounds__(VEC_NT)
cgb_cg2_step_kernel(double *__restrict__ x, double *r_loc, double *__restrict__ pvec, double *__restrict__ svec,
                const double *__restrict__ w, double *__restrict__ partials, cgk_scalars *sc, double eps,
                int ignore_stop, int guard, long long max_iter, long long m, const cgk_peers peers)
{
    pdl_wait();
    pdl_launch_dependents();
    if (*(volatile int *)&sc->stopped || *(volatile int *)&sc->breakdown) return;
    const unsigned int k = *(volatile unsigned int *)&sc->ep_xr;
    const unsigned int gen = *(volatile unsigned int *)&sc->gen;
    double gamma, delta;
    if constexpr (COMM) {
        delta = ll_gather_sum(&peers.mbox[peers.rank]->dot[0][k & 1u][0][0], peers.ranks, cgk_tag(gen, k), sc);
        gamma = (k == 0u) ? sc->rr[0]
                          : ll_gather_sum(&peers.mbox[peers.rank]->dot[1][k & 1u][0][0], peers.ranks, cgk_tag(gen, k), sc);
    } else {
        delta = sc->pAp;
        gamma = sc->rr[0];
    }
    const bool hit = (k > 0u) && sqrt(gamma) < eps;               /* the reference tests after an update only */
    const bool brk = hit && !ignore_stop;
    const bool spent = (long long)k >= max_iter;
    if (blockIdx.x == 0 && threadIdx.x == 0) {
        if (k > 0u) {
            sc->iters = (long long)k;
            sc->final_beta = gamma;
            if (hit && sc->first_hit < 0) sc->first_hit = (long long)k - 1;
        }
        if (brk) { sc->break_k = (long long)k - 1; sc->stop_req = 1; }
        else if (spent) sc->stop_req = 2;
    }
    if (brk || spent) return;
    const double gprev = sc->cg2_gamma_prev, aprev = sc->cg2_alpha_prev;
    const double beta = (k == 0u) ? 0.0 : gamma / gprev;
    const double alpha = (k == 0u) ? gamma / delta : gamma / (delta - beta * gamma / aprev);
    if (guard && !isfinite(alpha)) {
        if (blockIdx.x == 0 && threadIdx.x == 0) { sc->breakdown = 3; sc->stop_req = 3; }
        return;
    }
    const long long lo = COMM ? (long long)peers.rank * m : 0;
    double acc = 0.0;
    for (long long i = (long long)blockIdx.x * VEC_NT + threadIdx.x; i < m;
         i += (long long)gridDim.x * VEC_NT) {
        const double pi = __dadd_rn(r_loc[i], __dmul_rn(pvec[i], beta));
        const double si = __dadd_rn(__ldcg(w + i), __dmul_rn(svec[i], beta));   /* written by the previous kernel: not .nc */
        pvec[i] = pi;
        svec[i] = si;
        x[i] = __dadd_rn(x[i], __dmul_rn(pi, alpha));
        const double rn = __dsub_rn(r_loc[i], __dmul_rn(si, alpha));
        r_loc[i] = rn;
        if constexpr (COMM) {
            for (int g = 0; g < peers.ranks; ++g)
                if (g != peers.rank) peers.p_full[g][lo + i] = rn;
        }
        acc = fma(rn, rn, acc);
    }
    const double mine = vec_cta_sum(acc);
    double total = 0.0;
    /* one rank: device-scope fences are enough (a system-scope fence costs microseconds) */
    if (cta_partials_finish_sys<VEC_NT, COMM>(mine, partials, COMM ? &sc->rr_local : &sc->rr[0], &sc->ticket_vec, &total) &&
        threadIdx.x == 0) {
        if constexpr (COMM) {
            const unsigned int e = k + 1u;
            for (int g = 0; g < peers.ranks; ++g)
                ll_send(&peers.mbox[g]->dot[1][e & 1u][peers.rank][0], total, cgk_tag(gen, e));
            post_p_flags(peers, cgk_tag(gen, e));
        }
        sc->cg2_gamma_prev = gamma;
        sc->cg2_alpha_prev = alpha;
        sc->pAp = delta;
        sc->ep_xr = k + 1u;
    }
}

/* cg2_step on one rank, deferred reductions: no ticket, no last-CTA phase.  gamma and delta are
 * added up by every CTA from the partials the previous step / the product left; the step's
 * own state moves between the two parity slots of cgk_scalars. */
__global__ void __launch_bounds__(VEC_NT)
cgb_cg2_step_one_kernel(double *__restrict__ x, double *__restrict__ r_loc, double *__restrict__ pvec,
                    double *__restrict__ svec, const double *__restrict__ w, double *partials,
                    cgk_scalars *sc, int cur, double eps, int ignore_stop, int guard, long long max_iter,
                    long long m, const double *__restrict__ dot_partials, int n_dot)
{
    long long i = (long long)blockIdx.x * VEC_NT + threadIdx.x;
    const bool first = i < m;
    double r0 = 0.0, p0 = 0.0, w0 = 0.0, s0 = 0.0, x0 = 0.0;
    if (first) { r0 = r_loc[i]; p0 = pvec[i]; s0 = svec[i]; x0 = x[i]; }     /* the previous step's */
    const unsigned int k = sc->cg2_k[cur];
    const double gprev = sc->cg2_gprev[cur], aprev = sc->cg2_aprev[cur];
    pdl_wait();                                                /* w, delta's partials, the flags: the product */
    pdl_launch_dependents();
    if (first) w0 = __ldcg(w + i);
    if (*(volatile int *)&sc->stopped || *(volatile int *)&sc->breakdown) return;
    double gamma, delta;
    cta_sum_partials2<VEC_NT>(partials + (size_t)cur * CGK_MAX_VEC_CTAS, k == 0u ? 0 : (int)gridDim.x,
                              dot_partials, n_dot, &gamma, &delta);
    if (k == 0u) gamma = sc->rr[0];
    if (n_dot == 0) delta = sc->pAp;
    const bool hit = (k > 0u) && sqrt(gamma) < eps;               /* the reference tests after an update only */
    const bool brk = hit && !ignore_stop;
    const bool spent = (long long)k >= max_iter;
    const bool lead = blockIdx.x == 0 && threadIdx.x == 0;
    if (lead) {
        if (k > 0u) {
            sc->iters = (long long)k;
            sc->final_beta = gamma;
            if (hit && sc->first_hit < 0) sc->first_hit = (long long)k - 1;
        }
        if (brk) { sc->break_k = (long long)k - 1; sc->stop_req = 1; }
        else if (spent) sc->stop_req = 2;
    }
    if (brk || spent) return;
    const double beta = (k == 0u) ? 0.0 : gamma / gprev;
    const double alpha = (k == 0u) ? gamma / delta : gamma / (delta - beta * gamma / aprev);
    if (guard && !isfinite(alpha)) {
        if (lead) { sc->breakdown = 3; sc->stop_req = 3; }
        return;
    }
    if (lead) {
        sc->cg2_k[cur ^ 1] = k + 1u;
        sc->cg2_gprev[cur ^ 1] = gamma;
        sc->cg2_aprev[cur ^ 1] = alpha;
        if (n_dot > 0) sc->pAp = delta;
    }
    double acc = 0.0;
    for (bool head = first; i < m; i += (long long)gridDim.x * VEC_NT, head = false) {
        const double ri = head ? r0 : r_loc[i];
        const double pi = __dadd_rn(ri, __dmul_rn(head ? p0 : pvec[i], beta));
        const double si = __dadd_rn(head ? w0 : __ldcg(w + i), __dmul_rn(head ? s0 : svec[i], beta));
        pvec[i] = pi;
        svec[i] = si;
        x[i] = __dadd_rn(head ? x0 : x[i], __dmul_rn(pi, alpha));
        const double rn = __dsub_rn(ri, __dmul_rn(si, alpha));
        r_loc[i] = rn;
        acc = fma(rn, rn, acc);
    }
    const double mine = vec_cta_sum(acc);
    if (threadIdx.x == 0) partials[(size_t)(cur ^ 1) * CGK_MAX_VEC_CTAS + blockIdx.x] = mine;
}

__global__ void cgb_reset_kernel(cgk_scalars *sc, unsigned int gen)
{
    sc->rr[0] = sc->rr[1] = 0.0;
    sc->pAp = sc->pAp_local = sc->rr_local = 0.0;
    sc->final_beta = 0.0; sc->rsold0 = 0.0;
    sc->iters = 0; sc->first_hit = -1; sc->break_k = -1;
    sc->stop_req = 0; sc->stopped = 0;
    sc->ticket_gemv = 0u; sc->ticket_vec = 0u;
    sc->gen = gen; sc->ep_gemv = sc->ep_xr = sc->ep_p = 0u;
    sc->comm_error = 0;
    sc->breakdown = 0;
    sc->cg2_gamma_prev = 0.0; sc->cg2_alpha_prev = 0.0;
    sc->cg2_k[0] = sc->cg2_k[1] = 0u;
    sc->cg2_gprev[0] = sc->cg2_gprev[1] = 0.0;
    sc->cg2_aprev[0] = sc->cg2_aprev[1] = 0.0;
    sc->ticket_p = 0u;
}

template <bool COMM>
__global__ void cgb_after_init_kernel(cgk_scalars *sc, int guard, const cgk_peers peers)
{
    if constexpr (COMM) {
        const unsigned int gen = *(volatile unsigned int *)&sc->gen;
        const double rr0 = ll_gather_sum(&peers.mbox[peers.rank]->dot[1][0][0][0], peers.ranks, cgk_tag(gen, 0u), sc);
        if (threadIdx.x == 0) sc->rr[0] = rr0;
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        sc->rsold0 = sc->rr[0];
        sc->final_beta = sc->rr[0];
        /* guard: x0 already solves the system exactly -> nothing to iterate (the reference
         * would divide 0/0 in its first iteration, parallel_cg.c:298) */
        if (guard && sc->rr[0] == 0.0) { sc->breakdown = 4; sc->stop_req = 4; }
    }
}

/* ------------------------------------------------------------------ generator */
__device__ __forceinline__ double gen_elem(uint64_t key, uint64_t n, uint64_t i, uint64_t j,
                                           unsigned flags)
{
    double val = 0.5 * (cgb_gen_u(key, i * n + j) + cgb_gen_u(key, j * n + i));
    if (i == j) {
        /* n*I (generateSPDmatrix.m:17), or with CGB_GEN_KAPPA_LOG2(k) the spread diagonal
         * n*(1 + (2^k - 1) t^2), t = i/(n-1): every operation individually rounded, so the
         * CPU twin reproduces the bits; k = 0 gives exactly n */
        const unsigned k = (flags >> 8) & 0x3fu;
        double d = 1.0;
        if (k != 0u && n > 1) {
            const double t = __ddiv_rn((double)i, (double)(n - 1));
            d = __dadd_rn(1.0, __dmul_rn(__dsub_rn(ldexp(1.0, (int)k), 1.0), __dmul_rn(t, t)));
        }
        val = __dadd_rn(val, __dmul_rn((double)n, d));
    }
    if (flags & CGB_GEN_ROUND4) val = __ddiv_rn(rint(__dmul_rn(val, 1.0e4)), 1.0e4);
    return val;
}

__global__ void __launch_bounds__(256)
cgb_generate_A_kernel(double *__restrict__ A, long long n, long long row0, long long m,
                  uint64_t key, unsigned flags)
{
    const long long total = m * n;
    for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total;
         e += (long long)gridDim.x * 256) {
        const long long i = e / n, j = e - i * n;
        A[e] = gen_elem(key, (uint64_t)n, (uint64_t)(row0 + i), (uint64_t)j, flags);
    }
}

/* CGB_GEN_SPECTRUM (spd_gen.cuh): A = H diag(lambda) H element by element */
__global__ void __launch_bounds__(256)
cgb_generate_spectrum_kernel(double *__restrict__ A, long long n, long long row0, long long m,
                             uint64_t key, unsigned flags, double inv, double s4)
{
    const long long total = m * n;
    for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < total;
         e += (long long)gridDim.x * 256) {
        const long long i = e / n, j = e - i * n;
        double val = cgb_gen_spectrum_elem(key, (uint64_t)n, (uint64_t)(row0 + i), (uint64_t)j, flags, inv, s4);
        if (flags & CGB_GEN_ROUND4) val = __ddiv_rn(rint(__dmul_rn(val, 1.0e4)), 1.0e4);
        A[e] = val;
    }
}

__global__ void __launch_bounds__(256)
cgb_generate_b_kernel(double *__restrict__ b, long long n, long long row0, long long m,
                  uint64_t key, unsigned flags)
{
    for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < m;
         i += (long long)gridDim.x * 256) {
        double val = cgb_gen_u(key, (uint64_t)n * (uint64_t)n + (uint64_t)(row0 + i));
        if (flags & CGB_GEN_ROUND4) val = __ddiv_rn(rint(__dmul_rn(val, 1.0e4)), 1.0e4);
        b[i] = val;
    }
}

__global__ void __launch_bounds__(256) cgb_l2_flush_kernel(double2 *buf, long long n2)
{
    for (long long e = (long long)blockIdx.x * 256 + threadIdx.x; e < n2;
         e += (long long)gridDim.x * 256) {
        double2 t = buf[e];
        t.x += 1.0; t.y -= 1.0;
        buf[e] = t;
    }
}

/* ------------------------------------------------------------------ launchers */
static inline int64_t ceil_div(int64_t a, int64_t b) { return (a + b - 1) / b; }

extern "C" int64_t cgk_gemv_max_ctas(int64_t m) { return m < 1 ? 1 : m; }
extern "C" int64_t cgk_vec_max_ctas(int64_t m) { (void)m; return CGK_MAX_VEC_CTAS; }

static int vec_grid(int64_t m)
{
    int64_t g = ceil_div(m, VEC_NT);
    if (g < 1) g = 1;
    if (g > CGK_MAX_VEC_CTAS) g = CGK_MAX_VEC_CTAS;
    return (int)g;
}

typedef void (*gemv_fn)(const double *, const double *, const double *, double *, double *,
                        double *, cgk_scalars *, long long, long long);

template <int R, int U, int NT>
static gemv_fn pick_ldg_vec(int vec)
{
    if (R * U * vec > 64) return nullptr;                    /* would spill: not instantiated */
    switch (vec) {
    case 4: if constexpr (R * U * 4 <= 64) return cgb_rowblock_matvec_ldg<R, U, NT, 4>; else return nullptr;
    case 2: if constexpr (R * U * 2 <= 64) return cgb_rowblock_matvec_ldg<R, U, NT, 2>; else return nullptr;
    default: return nullptr;
    }
}
template <int R, int U>
static gemv_fn pick_ldg_nt(int nt, int vec)
{
    switch (nt) {
    case 128: return pick_ldg_vec<R, U, 128>(vec);
    case 256: return pick_ldg_vec<R, U, 256>(vec);
    case 512: return pick_ldg_vec<R, U, 512>(vec);
    default: return nullptr;
    }
}
template <int R>
static gemv_fn pick_ldg_u(int u, int nt, int vec)
{
    switch (u) {
    case 1: return pick_ldg_nt<R, 1>(nt, vec);
    case 2: return pick_ldg_nt<R, 2>(nt, vec);
    case 4: return pick_ldg_nt<R, 4>(nt, vec);
    default: return nullptr;
    }
}
static gemv_fn pick_ldg(int r, int u, int nt, int vec)
{
    if (vec == 1) return cgb_rowblock_matvec_ldg<4, 1, 128, 1>;      /* odd n / unaligned: scalar loads */
    switch (r) {
    case 1: return pick_ldg_u<1>(u, nt, vec);
    case 2: return pick_ldg_u<2>(u, nt, vec);
    case 4: return pick_ldg_u<4>(u, nt, vec);
    case 8: return pick_ldg_u<8>(u, nt, vec);
    default: return nullptr;
    }
}

/* L2-resident variants (small n) */
typedef void (*gemv_keep_fn)(const double *, const double *, const double *, double *, double *,
                             double *, cgk_scalars *, long long, long long, long long);
struct keep_cfg { int r, u, nt, vec; gemv_keep_fn fn; };
static const keep_cfg k_keep[] = {
    {4, 1, 128, 4, cgb_rowblock_matvec_keep<4, 1, 128, 4>},
    /* the same 16 doubles in flight per thread at a finer row granularity (2x / 4x the CTAs): selectable, never the
     * default -- measured on one B200 for n = 1024 .. 7680 (profiles/r02_rows_per_cta_sweep.txt) (2,2) is within
     * +-3 % of (4,1) except at n = 4096 (-16 %: two waves instead of one) and (1,4) is 10-30 % slower everywhere,
     * i.e. the sizes between one and two waves of (4,1) CTAs are NOT limited by wave quantisation */
    {2, 2, 128, 4, cgb_rowblock_matvec_keep<2, 2, 128, 4>},
    {1, 4, 128, 4, cgb_rowblock_matvec_keep<1, 4, 128, 4>},
    {4, 2, 128, 4, cgb_rowblock_matvec_keep<4, 2, 128, 4>},
    {8, 1, 128, 4, cgb_rowblock_matvec_keep<8, 1, 128, 4>},
    {2, 1, 256, 4, cgb_rowblock_matvec_keep<2, 1, 256, 4>},
    {4, 2, 256, 4, cgb_rowblock_matvec_keep<4, 2, 256, 4>},
};
static const keep_cfg *find_keep(int r, int u, int nt, int vec)
{
    for (const keep_cfg &c : k_keep)
        if (c.r == r && c.u == u && c.nt == nt && c.vec == vec) return &c;
    return nullptr;
}

/* bulk variants: (R, C, STAGES, NT) */
struct bulk_cfg { int r, c, stages, nt; gemv_fn fn; };
static const bulk_cfg k_bulk[] = {
    {4, 512, 4, 128, cgb_rowblock_matvec_bulk<4, 512, 4, 128>},
    {4, 512, 8, 128, cgb_rowblock_matvec_bulk<4, 512, 8, 128>},
    {8, 512, 4, 128, cgb_rowblock_matvec_bulk<8, 512, 4, 128>},
    {8, 512, 5, 128, cgb_rowblock_matvec_bulk<8, 512, 5, 128>},
    {8, 256, 8, 128, cgb_rowblock_matvec_bulk<8, 256, 8, 128>},
    {4, 1024, 4, 256, cgb_rowblock_matvec_bulk<4, 1024, 4, 256>},
    {8, 1024, 2, 256, cgb_rowblock_matvec_bulk<8, 1024, 2, 256>},
    {8, 1024, 3, 256, cgb_rowblock_matvec_bulk<8, 1024, 3, 256>},
    {16, 256, 5, 128, cgb_rowblock_matvec_bulk<16, 256, 5, 128>},
    {16, 512, 3, 128, cgb_rowblock_matvec_bulk<16, 512, 3, 128>},
};
static const bulk_cfg *find_bulk(int r, int c, int stages, int nt)
{
    for (const bulk_cfg &b : k_bulk)
        if (b.r == r && b.c == c && b.stages == stages && b.nt == nt) return &b;
    return nullptr;
}
static int bulk_smem(const bulk_cfg *b)
{
    return b->stages * (b->r + 1) * b->c * 8 + 2 * b->stages * 8;
}

extern "C" int cgk_gemv_make_plan(cgk_gemv_plan *plan, int64_t m, int64_t n, const void *A,
                                  const void *v, int sm_count)
{
    if (!plan || m < 1 || n < 1) return (int)cudaErrorInvalidValue;
    if (sm_count <= 0) sm_count = 148;
    const uintptr_t pa = (uintptr_t)A, pv = (uintptr_t)v;
    int maxvec = 1;
    if (n % 4 == 0 && pa % 32 == 0 && pv % 32 == 0) maxvec = 4;
    else if (n % 2 == 0 && pa % 16 == 0 && pv % 16 == 0) maxvec = 2;
    int variant = plan->variant, rows = plan->rows_per_cta, unroll = plan->unroll;
    int threads = plan->threads, vec = plan->vec;
    if (vec == 0 || vec > maxvec) vec = maxvec;
    if (vec == 3) vec = 2;
    if (variant == CGK_GEMV_AUTO) variant = CGK_GEMV_LDG;

    if (variant == CGK_GEMV_BULK) {
        const int r = rows ? rows : 8, st = plan->stages ? plan->stages : 4;
        const int nt = threads ? threads : 128, c = plan->tile_cols ? plan->tile_cols : 512;
        const bulk_cfg *b = find_bulk(r, c, st, nt);
        if (b && n % c == 0 && pa % 16 == 0 && pv % 16 == 0) {
            plan->variant = CGK_GEMV_BULK;
            plan->rows_per_cta = r; plan->unroll = 0; plan->threads = nt; plan->vec = 2;
            plan->stages = st; plan->tile_cols = c; plan->smem_bytes = bulk_smem(b);
            int occ = 1;
            cudaFuncSetAttribute(b->fn, cudaFuncAttributeMaxDynamicSharedMemorySize, plan->smem_bytes);
            if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, b->fn, nt + 32, plan->smem_bytes) != cudaSuccess || occ < 1) {
                (void)cudaGetLastError();
                occ = 1;
            }
            const int64_t panels = ceil_div(m, r);
            const int64_t g = (int64_t)sm_count * occ;
            plan->grid = (int)(panels < g ? panels : g);
            return 0;
        }
        rows = unroll = threads = 0;       /* shape not tileable: streaming-load kernel */
    }

    plan->variant = CGK_GEMV_LDG;
    plan->stages = 0; plan->tile_cols = 0; plan->smem_bytes = 0;
    if (vec == 1) {
        plan->rows_per_cta = 4; plan->unroll = 1; plan->threads = 128; plan->vec = 1;
    } else {
        /* defaults from the B200 sweep (tools/tune_gemv.py, profiles/r01_tune_gemv.md): every
         * reasonable shape sits within ~1 % of the same plateau (7.4 TB/s at n = 65536); the
         * picks below are the fastest measured per size class */
        if (rows == 0 && unroll == 0 && threads == 0) {
            if (n >= 32768 && m >= 8192 && vec == 4) { rows = 8; unroll = 2; threads = 512; }
            else if (n >= 8192) { rows = 2; unroll = 1; threads = 256; }
            else { rows = 4; unroll = 1; threads = 128; }
        }
        if (rows == 0) rows = 4;
        if (threads == 0) threads = (n >= 8192) ? 256 : 128;
        if (unroll == 0) unroll = 2;
        while (rows > 1 && rows > m) rows >>= 1;
        plan->rows_per_cta = rows; plan->unroll = unroll; plan->threads = threads; plan->vec = vec;
    }
    if (!pick_ldg(plan->rows_per_cta, plan->unroll, plan->threads, plan->vec))
        return (int)cudaErrorInvalidValue;
    plan->grid = (int)ceil_div(m, plan->rows_per_cta);
    /* L2 residency: only when A is too large to stay in L2 by itself but small enough for the
     * kept part to matter (<= 3 L2 sizes; measured: no gain at 4.3), and a keep variant of the chosen shape exists */
    plan->keep_rows = 0;
    if (plan->l2_bytes > 0 && plan->keep_pct >= 0 &&
        find_keep(plan->rows_per_cta, plan->unroll, plan->threads, plan->vec)) {
        const double a_bytes = 8.0 * (double)m * (double)n;
        const int pct = plan->keep_pct > 0 ? plan->keep_pct : 60;
        if (a_bytes > 0.8 * (double)plan->l2_bytes && a_bytes <= 3.0 * (double)plan->l2_bytes) {
            long long rows = (long long)((double)plan->l2_bytes * pct / 100.0 / (8.0 * (double)n));
            rows -= rows % plan->rows_per_cta;
            if (rows > m) rows = m;
            plan->keep_rows = rows > 0 ? rows : 0;
        }
    }
    return 0;
}

extern "C" int cgk_gemv(const cgk_gemv_plan *plan, const double *A, const double *v,
                        const double *w, double *y, double *partials, double *dot_out,
                        cgk_scalars *sc, int64_t m, int64_t n, void *stream)
{
    cudaStream_t st = (cudaStream_t)stream;
    if (plan->variant == CGK_GEMV_BULK) {
        const bulk_cfg *b = find_bulk(plan->rows_per_cta, plan->tile_cols, plan->stages, plan->threads);
        if (!b) return (int)cudaErrorInvalidValue;
        b->fn<<<plan->grid, plan->threads + 32, plan->smem_bytes, st>>>(A, v, w, y, partials, dot_out,
                                                                        sc, (long long)m, (long long)n);
        return (int)cudaGetLastError();
    }
    /* the L2-keep kernel doubles as the GEMV of the dependent-launch chain (small systems) */
    if (plan->keep_rows > 0 ||
        (tl_pdl && plan->l2_bytes > 0 && 8.0 * (double)m * (double)n <= 3.0 * (double)plan->l2_bytes &&
         find_keep(plan->rows_per_cta, plan->unroll, plan->threads, plan->vec))) {
        const keep_cfg *k = find_keep(plan->rows_per_cta, plan->unroll, plan->threads, plan->vec);
        if (!k) return (int)cudaErrorInvalidValue;
        cgk_launch(k->fn, dim3(plan->grid), dim3(plan->threads), 0, st, A, v, w, y, partials, dot_out, sc, (long long)m,
                   (long long)n, (long long)plan->keep_rows);
        return (int)cudaGetLastError();
    }
    gemv_fn fn = pick_ldg(plan->rows_per_cta, plan->unroll, plan->threads, plan->vec);
    if (!fn) return (int)cudaErrorInvalidValue;
    fn<<<plan->grid, plan->threads, 0, st>>>(A, v, w, y, partials, dot_out, sc, (long long)m, (long long)n);
    return (int)cudaGetLastError();
}

static const cgk_peers k_no_peers = {};

/* comm variants: the size-class defaults of the plain kernel plus narrower-load fallbacks */
typedef void (*gemv_comm_fn)(const double *, const double *, const double *, double *, double *,
                             double *, cgk_scalars *, long long, long long, const cgk_peers);
struct comm_cfg { int r, u, nt, vec; gemv_comm_fn fn; };
static const comm_cfg k_comm[] = {
    {4, 2, 256, 4, cgb_rowblock_matvec_comm<4, 2, 256, 4>},
    {2, 1, 256, 4, cgb_rowblock_matvec_comm<2, 1, 256, 4>},
    {2, 2, 256, 4, cgb_rowblock_matvec_comm<2, 2, 256, 4>},
    {4, 1, 128, 4, cgb_rowblock_matvec_comm<4, 1, 128, 4>},
    {4, 2, 128, 2, cgb_rowblock_matvec_comm<4, 2, 128, 2>},
    {4, 1, 128, 1, cgb_rowblock_matvec_comm<4, 1, 128, 1>},
};
static const comm_cfg *find_comm(int r, int u, int nt, int vec)
{
    for (const comm_cfg &c : k_comm)
        if (c.r == r && c.u == u && c.nt == nt && c.vec == vec) return &c;
    return nullptr;
}

extern "C" int cgk_gemv_make_plan_comm(cgk_gemv_plan *plan, int64_t m, int64_t n, const void *A,
                                       const void *v, int sm_count)
{
    (void)sm_count;
    if (!plan || m < 1 || n < 1) return (int)cudaErrorInvalidValue;
    const uintptr_t pa = (uintptr_t)A, pv = (uintptr_t)v;
    int vec = 1;                       /* block boundaries are multiples of m */
    if (n % 4 == 0 && m % 4 == 0 && pa % 32 == 0 && pv % 32 == 0) vec = 4;
    else if (n % 2 == 0 && m % 2 == 0 && pa % 16 == 0 && pv % 16 == 0) vec = 2;
    if (plan->vec != 0 && plan->vec < vec) vec = plan->vec == 3 ? 2 : plan->vec;
    const comm_cfg *c = nullptr;
    if (plan->rows_per_cta && plan->unroll && plan->threads)
        c = find_comm(plan->rows_per_cta, plan->unroll, plan->threads, vec);
    if (!c) {
        if (vec == 4) {
            /* (4,2,256): 3 CTAs per SM.  The one-rank kernel's (8,2,512) shape was measured here too
             * (round 2, 2 GPUs, with the flag test moved out of the unrolled loops so that it fits
             * 128 registers): 1 % slower at n = 65536 (7.31 against 7.38-7.44 TB/s) and 2 % slower at
             * n = 16384, so it is not instantiated */
            if (n >= 8192) c = find_comm(4, 2, 256, 4);     /* measured on 2 GPUs: beats (2,1,256) at n = 16384 */
            else c = find_comm(4, 1, 128, 4);
        } else if (vec == 2) c = find_comm(4, 2, 128, 2);
        else c = find_comm(4, 1, 128, 1);
    }
    if (!c) return (int)cudaErrorInvalidValue;
    plan->variant = CGK_GEMV_LDG;
    plan->rows_per_cta = c->r; plan->unroll = c->u; plan->threads = c->nt; plan->vec = c->vec;
    plan->stages = 0; plan->tile_cols = 0; plan->smem_bytes = 0;
    plan->grid = (int)ceil_div(m, c->r);
    return 0;
}

extern "C" int cgk_gemv_comm(const cgk_gemv_plan *plan, const double *A, const double *v,
                             const double *w, double *y, double *partials, double *dot_out,
                             cgk_scalars *sc, int64_t m, int64_t n, const cgk_peers *peers,
                             void *stream)
{
    const comm_cfg *c = find_comm(plan->rows_per_cta, plan->unroll, plan->threads, plan->vec);
    if (!c || !peers || !w) return (int)cudaErrorInvalidValue;
    cgk_launch(c->fn, dim3(plan->grid), dim3(plan->threads), 0, (cudaStream_t)stream, A, v, w, y, partials, dot_out, sc,
                                                                  (long long)m, (long long)n, *peers);
    return (int)cudaGetLastError();
}

extern "C" int cgk_init_residual(const double *b, const double *Ax, double *r, double *p_loc,
                                 double *partials, double *rr_out, cgk_scalars *sc, int64_t m,
                                 const cgk_peers *peers, void *stream)
{
    if (peers)
        cgb_init_residual_kernel<true><<<vec_grid(m), VEC_NT, 0, (cudaStream_t)stream>>>(b, Ax, r, p_loc, partials,
                                                                                    rr_out, sc, (long long)m, *peers);
    else
        cgb_init_residual_kernel<false><<<vec_grid(m), VEC_NT, 0, (cudaStream_t)stream>>>(b, Ax, r, p_loc, partials,
                                                                                     rr_out, sc, (long long)m, k_no_peers);
    return (int)cudaGetLastError();
}

extern "C" int cgk_xr_update(double *x, double *r, const double *p_loc, const double *Ap,
                             double *partials, double *rr_out, cgk_scalars *sc, int cur, int guard,
                             int64_t m, const cgk_peers *peers, const double *dot_partials, int n_dot,
                             void *stream)
{
    if (peers)
        cgk_launch(cgb_xr_update_kernel<true>, dim3(vec_grid(m)), dim3(VEC_NT), 0, (cudaStream_t)stream, x, r, p_loc, Ap, partials,
                   rr_out, sc, cur, guard, (long long)m, *peers, (const double *)nullptr, 0);
    else
        cgk_launch(cgb_xr_update_kernel<false>, dim3(vec_grid(m)), dim3(VEC_NT), 0, (cudaStream_t)stream, x, r, p_loc, Ap, partials,
                   rr_out, sc, cur, guard, (long long)m, k_no_peers, dot_partials, n_dot);
    return (int)cudaGetLastError();
}

extern "C" int cgk_xrp_supported(int64_t m, int sm_count)
{
    static int per_sm = -1;
    if (per_sm < 0) {
        int occ = 0;
        if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, cgb_xrp_update_kernel, VEC_NT, 0) != cudaSuccess) { (void)cudaGetLastError(); occ = 0; }
        per_sm = occ;
    }
    if (sm_count <= 0) sm_count = 148;
    /* half of what fits: the tail of the product may still hold part of the SMs when this kernel starts */
    return (int64_t)vec_grid(m) <= (int64_t)sm_count * per_sm / 2;
}

extern "C" int cgk_xrp_update(double *x, double *r, double *p_loc, const double *Ap, double *partials,
                              cgk_scalars *sc, int cur, int guard, double eps, int ignore_stop,
                              int64_t max_iter, int64_t m, const cgk_peers *peers, void *stream)
{
    if (!peers) return (int)cudaErrorInvalidValue;
    cgk_launch(cgb_xrp_update_kernel, dim3(vec_grid(m)), dim3(VEC_NT), 0, (cudaStream_t)stream, x, r, p_loc, Ap, partials, sc, cur,
               guard, eps, ignore_stop, (long long)max_iter, (long long)m, *peers);
    return (int)cudaGetLastError();
}

extern "C" int cgk_vec_grid(int64_t m) { return vec_grid(m); }

extern "C" int cgk_p_update(double *p_loc, const double *r, cgk_scalars *sc, int cur, double eps,
                            int ignore_stop, int64_t max_iter, int64_t m, const cgk_peers *peers,
                            const double *rr_partials, int n_rr, void *stream)
{
    if (peers)
        cgk_launch(cgb_p_update_kernel<true>, dim3(vec_grid(m)), dim3(VEC_NT), 0, (cudaStream_t)stream, p_loc, r, sc, cur, eps,
                   ignore_stop, (long long)max_iter, (long long)m, *peers, (const double *)nullptr, 0);
    else
        cgk_launch(cgb_p_update_kernel<false>, dim3(vec_grid(m)), dim3(VEC_NT), 0, (cudaStream_t)stream, p_loc, r, sc, cur, eps,
                   ignore_stop, (long long)max_iter, (long long)m, k_no_peers, rr_partials, n_rr);
    return (int)cudaGetLastError();
}

extern "C" int cgk_cg2_step(double *x, double *r_loc, double *pvec, double *svec, const double *w,
                            double *partials, cgk_scalars *sc, double eps, int ignore_stop, int guard,
                            int64_t max_iter, int64_t m, const cgk_peers *peers, void *stream)
{
    if (peers)
        cgk_launch(cgb_cg2_step_kernel<true>, dim3(vec_grid(m)), dim3(VEC_NT), 0, (cudaStream_t)stream, x, r_loc, pvec, svec, w,
                   partials, sc, eps, ignore_stop, guard, (long long)max_iter, (long long)m, *peers);
    else
        cgb_cg2_step_kernel<false><<<vec_grid(m), VEC_NT, 0, (cudaStream_t)stream>>>(x, r_loc, pvec, svec, w, partials, sc, eps,
                                                                                ignore_stop, guard, (long long)max_iter, (long long)m, k_no_peers);
    return (int)cudaGetLastError();
}

extern "C" int cgk_cg2_step_one(double *x, double *r_loc, double *pvec, double *svec, const double *w, double *partials,
                                cgk_scalars *sc, int cur, double eps, int ignore_stop, int guard, int64_t max_iter,
                                int64_t m, const double *dot_partials, int n_dot, void *stream)
{
    cgk_launch(cgb_cg2_step_one_kernel, dim3(vec_grid(m)), dim3(VEC_NT), 0, (cudaStream_t)stream, x, r_loc, pvec, svec, w, partials,
               sc, cur, eps, ignore_stop, guard, (long long)max_iter, (long long)m, dot_partials, n_dot);
    return (int)cudaGetLastError();
}

extern "C" int cgk_reset(cgk_scalars *sc, unsigned int gen, void *stream)
{
    cgb_reset_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(sc, gen);
    return (int)cudaGetLastError();
}

extern "C" int cgk_after_init(cgk_scalars *sc, int guard, const cgk_peers *peers, void *stream)
{
    if (peers) cgb_after_init_kernel<true><<<1, 32, 0, (cudaStream_t)stream>>>(sc, guard, *peers);
    else cgb_after_init_kernel<false><<<1, 32, 0, (cudaStream_t)stream>>>(sc, guard, k_no_peers);
    return (int)cudaGetLastError();
}

extern "C" int cgk_generate_A(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                              unsigned flags, void *stream)
{
    int64_t g = ceil_div(m * n, 256 * 8);
    if (g < 1) g = 1;
    if (g > 148 * 32) g = 148 * 32;
    cgb_generate_A_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(A, (long long)n, (long long)row0,
                                                               (long long)m, cgb_gen_key(seed), flags);
    return (int)cudaGetLastError();
}

extern "C" int cgk_generate_spectrum_A(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                                       unsigned flags, double inv, double s4, void *stream)
{
    int64_t g = ceil_div(m * n, 256 * 8);
    if (g < 1) g = 1;
    if (g > 148 * 32) g = 148 * 32;
    cgb_generate_spectrum_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(A, (long long)n, (long long)row0, (long long)m,
                                                                          cgb_gen_key(seed), flags, inv, s4);
    return (int)cudaGetLastError();
}

extern "C" int cgk_generate_b(double *b, int64_t n, int64_t row0, int64_t m, uint64_t seed,
                              unsigned flags, void *stream)
{
    int64_t g = ceil_div(m, 256);
    if (g < 1) g = 1;
    if (g > 592) g = 592;
    cgb_generate_b_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(b, (long long)n, (long long)row0,
                                                               (long long)m, cgb_gen_key(seed), flags);
    return (int)cudaGetLastError();
}

extern "C" int cgk_l2_flush(void *buf, int64_t bytes, void *stream)
{
    cgb_l2_flush_kernel<<<148 * 8, 256, 0, (cudaStream_t)stream>>>((double2 *)buf, (long long)(bytes / 16));
    return (int)cudaGetLastError();
}

extern "C" const char *cgk_error_string(int err) { return cudaGetErrorString((cudaError_t)err); }
