/* cg_device.cuh -- device helpers shared by the kernel files (cg_kernels.cu, cg_symv.cu):
 * the streaming / cached / coherent global loads, the warp sum and the fixed-order CTA
 * partials reduction. */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "cg_kernels.h"

/* ------------------------------------------------------------------ loads */
/* A is streamed exactly once per GEMV: read-only path, do not allocate in L1,
 * evict-first in L2 (256-bit form) so the re-used direction vector stays hot. */
template <int VW> struct Ld;
template <> struct Ld<1> {
    static __device__ __forceinline__ void stream(const double *p, double (&o)[1]) {
        asm("ld.global.nc.L1::no_allocate.f64 %0, [%1];" : "=d"(o[0]) : "l"(p));
    }
    static __device__ __forceinline__ void hold(const double *p, double (&o)[1]) {
        asm("ld.global.nc.L1::no_allocate.L2::evict_last.f64 %0, [%1];" : "=d"(o[0]) : "l"(p));
    }
    static __device__ __forceinline__ void keep(const double *p, double (&o)[1]) {
        o[0] = __ldg(p);
    }
    static __device__ __forceinline__ void coherent(const double *p, double (&o)[1]) {
        asm volatile("ld.global.f64 %0, [%1];" : "=d"(o[0]) : "l"(p) : "memory");
    }
    static __device__ __forceinline__ void plain(const double *p, double (&o)[1]) {
        asm("ld.global.f64 %0, [%1];" : "=d"(o[0]) : "l"(p));
    }
};
template <> struct Ld<2> {
    static __device__ __forceinline__ void stream(const double *p, double (&o)[2]) {
        asm("ld.global.nc.L1::no_allocate.v2.f64 {%0,%1}, [%2];" : "=d"(o[0]), "=d"(o[1]) : "l"(p));
    }
    static __device__ __forceinline__ void hold(const double *p, double (&o)[2]) {
        asm("ld.global.nc.L1::no_allocate.L2::evict_last.v2.f64 {%0,%1}, [%2];" : "=d"(o[0]), "=d"(o[1]) : "l"(p));
    }
    static __device__ __forceinline__ void keep(const double *p, double (&o)[2]) {
        asm("ld.global.nc.v2.f64 {%0,%1}, [%2];" : "=d"(o[0]), "=d"(o[1]) : "l"(p));
    }
    static __device__ __forceinline__ void coherent(const double *p, double (&o)[2]) {
        asm volatile("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(o[0]), "=d"(o[1]) : "l"(p) : "memory");
    }
    static __device__ __forceinline__ void plain(const double *p, double (&o)[2]) {
        asm("ld.global.v2.f64 {%0,%1}, [%2];" : "=d"(o[0]), "=d"(o[1]) : "l"(p));
    }
};
template <> struct Ld<4> {      /* 256-bit LDG: new on sm_100 */
    static __device__ __forceinline__ void stream(const double *p, double (&o)[4]) {
        unsigned long long a, b, c, d;
        asm("ld.global.nc.L1::no_allocate.L2::evict_first.v4.b64 {%0,%1,%2,%3}, [%4];"
            : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
        o[0] = __longlong_as_double(a); o[1] = __longlong_as_double(b);
        o[2] = __longlong_as_double(c); o[3] = __longlong_as_double(d);
    }
    /* rows of A that are meant to survive in L2 until the next iteration (small n) */
    static __device__ __forceinline__ void hold(const double *p, double (&o)[4]) {
        unsigned long long a, b, c, d;
        asm("ld.global.nc.L1::no_allocate.L2::evict_last.v4.b64 {%0,%1,%2,%3}, [%4];"
            : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
        o[0] = __longlong_as_double(a); o[1] = __longlong_as_double(b);
        o[2] = __longlong_as_double(c); o[3] = __longlong_as_double(d);
    }
    static __device__ __forceinline__ void keep(const double *p, double (&o)[4]) {
        unsigned long long a, b, c, d;
        asm("ld.global.nc.v4.b64 {%0,%1,%2,%3}, [%4];"
            : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
        o[0] = __longlong_as_double(a); o[1] = __longlong_as_double(b);
        o[2] = __longlong_as_double(c); o[3] = __longlong_as_double(d);
    }
    /* the direction vector while peers are still writing other parts of it: an ordinary
     * (coherent) load, ordered after the acquire that observed the peer's flag */
    static __device__ __forceinline__ void coherent(const double *p, double (&o)[4]) {
        unsigned long long a, b, c, d;
        asm volatile("ld.global.v4.b64 {%0,%1,%2,%3}, [%4];"
                     : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p) : "memory");
        o[0] = __longlong_as_double(a); o[1] = __longlong_as_double(b);
        o[2] = __longlong_as_double(c); o[3] = __longlong_as_double(d);
    }
    /* the same ordinary (not .nc) load without the compiler barrier: free to be scheduled among
     * the loads of A.  Ordering against a preceding flag acquire is kept by address dependency:
     * the caller launders the base pointer through an `asm volatile` placed after the acquire
     * (order_after_acquire below), so no such load can be hoisted above it. */
    static __device__ __forceinline__ void plain(const double *p, double (&o)[4]) {
        unsigned long long a, b, c, d;
        asm("ld.global.v4.b64 {%0,%1,%2,%3}, [%4];"
            : "=l"(a), "=l"(b), "=l"(c), "=l"(d) : "l"(p));
        o[0] = __longlong_as_double(a); o[1] = __longlong_as_double(b);
        o[2] = __longlong_as_double(c); o[3] = __longlong_as_double(d);
    }
};
__device__ __forceinline__ const double *order_after_acquire(const double *p)
{
    asm volatile("" : "+l"(p) : : "memory");
    return p;
}

/* Programmatic dependent launch (one rank; cgk_set_pdl): a kernel launched with the attribute
 * may start while its predecessor in the stream is still running.  Rule kept by every kernel
 * of the iteration: before pdl_wait() only data that the IMMEDIATE predecessor does not write
 * (the matrix, vectors last written two kernels back); pdl_wait() returns once the
 * predecessor has completed and flushed; pdl_launch_dependents() comes AFTER the wait, so the
 * successor cannot start before this kernel's own predecessor is complete (completion is then
 * transitive along the chain).  Without the launch attribute both are no-ops. */
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

extern "C" int cgk_pdl_enabled(void);
template <typename... P, typename... Args>
static inline void cgk_launch(void (*kern)(P...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args)
{
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = cgk_pdl_enabled() ? 1 : 0;
    (void)cudaLaunchKernelEx(&cfg, kern, P(args)...);
}

__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

/* Fixed-order two-stage reduction of one partial per CTA: the last CTA to
 * arrive (ticket) sums partials[0..grid) with a fixed thread->index mapping and
 * a fixed tree, so the result is bitwise reproducible run to run (no FP64
 * atomics).  `mine` is the CTA's partial, valid in thread 0. */
template <int NT>
__device__ __forceinline__ void cta_partials_finish(double mine, double *partials,
                                                    double *out, unsigned int *ticket)
{
    __shared__ int s_last;
    __shared__ double s_red[NT / 32];
    const int tid = threadIdx.x;
    if (out == nullptr) {
        /* deferred: the consumer kernel adds the partials itself (cta_sum_partials), which
         * takes the store -> fence -> ticket -> reload chain off the end of this kernel */
        if (tid == 0) partials[blockIdx.x] = mine;
        return;
    }
    if (tid == 0) {
        partials[blockIdx.x] = mine;
        __threadfence();
        unsigned int t = atomicAdd(ticket, 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    double s = 0.0;
    for (unsigned int i = tid; i < gridDim.x; i += NT) s += __ldcg(partials + i);
    s = warp_sum(s);
    if ((tid & 31) == 0) s_red[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < NT / 32; ++q) t += s_red[q];
        *out = t;
        *ticket = 0u;
    }
}


/* The consumer side of a deferred reduction: the whole CTA adds partials[0..count) of the
 * PREVIOUS kernel with a fixed thread -> index mapping and a fixed tree and every thread gets
 * the total.  Every CTA (of any kernel with the same NT) computes the same bits, so all take
 * the same branches on it. */
template <int NT>
__device__ __forceinline__ double cta_sum_partials(const double *partials, int count)
{
    __shared__ double s_cp[NT / 32];
    __shared__ double s_total;
    const int tid = threadIdx.x;
    double s = 0.0;
#pragma unroll 4
    for (int i = tid; i < count; i += NT) s += __ldcg(partials + i);
    s = warp_sum(s);
    if ((tid & 31) == 0) s_cp[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < NT / 32; ++q) t += s_cp[q];
        s_total = t;
    }
    __syncthreads();
    return s_total;
}

/* Two such sums in one pass (one pair of barriers). */
template <int NT>
__device__ __forceinline__ void cta_sum_partials2(const double *pa, int na, const double *pb, int nb,
                                                  double *sum_a, double *sum_b)
{
    __shared__ double s_cp2[2][NT / 32];
    __shared__ double s_tot2[2];
    const int tid = threadIdx.x;
    double a = 0.0, b = 0.0;
#pragma unroll 4
    for (int i = tid; i < na; i += NT) a += __ldcg(pa + i);
#pragma unroll 4
    for (int i = tid; i < nb; i += NT) b += __ldcg(pb + i);
    a = warp_sum(a);
    b = warp_sum(b);
    if ((tid & 31) == 0) { s_cp2[0][tid >> 5] = a; s_cp2[1][tid >> 5] = b; }
    __syncthreads();
    if (tid < 2) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < NT / 32; ++q) t += s_cp2[tid][q];
        s_tot2[tid] = t;
    }
    __syncthreads();
    *sum_a = s_tot2[0];
    *sum_b = s_tot2[1];
}

/* ------------------------------------------------------------------ peer exchange */
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long ld_acquire_sys(const unsigned long long *p)
{
    unsigned long long v;
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_relaxed_sys(unsigned long long *p, unsigned long long v)
{
    asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long global_ns(void)
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ unsigned int sm_id(void)
{
    unsigned int v;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(v));
    return v;
}
/* tag of exchange event e of solve generation gen (never 0 for gen & 0x3ff != 0) */
__device__ __forceinline__ unsigned int cgk_tag(unsigned int gen, unsigned int e)
{
    return ((gen & 0x3ffu) << 22) | (e & 0x3fffffu);
}
/* Called every 1024 failed polls: gives up once any waiter has given up or the peer has been
 * silent for CGK_SPIN_TIMEOUT_NS, so a dead rank becomes an error code instead of a hung GPU. */
static __device__ __noinline__ bool spin_give_up(unsigned long long *t0, cgk_scalars *sc)
{
    if (*(volatile int *)&sc->comm_error) return true;
    const unsigned long long now = global_ns();
    if (*t0 == 0ull) { *t0 = now; return false; }
    if (now - *t0 > CGK_SPIN_TIMEOUT_NS) {
        *(volatile int *)&sc->comm_error = 1;
        __threadfence();
        return true;
    }
    return false;
}
/* LL-style scalar: two 8-byte words, each {32 bits of the double | tag << 32}; 8-byte stores
 * are single NVLink writes, so a reader that sees the tag in both words has the value. */
__device__ __forceinline__ void ll_send(unsigned long long *slot, double v, unsigned int tag)
{
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    const unsigned long long t = (unsigned long long)tag << 32;
    st_relaxed_sys(slot, (b & 0xffffffffull) | t);
    st_relaxed_sys(slot + 1, (b >> 32) | t);
}
/* All threads call it.  Threads g < G poll the LOCAL slots[g][2] until both words carry `tag`;
 * thread 0 adds the G values in rank order -- every rank computes the same bits. */
__device__ __forceinline__ double ll_gather_sum(const unsigned long long *slots, int G,
                                                unsigned int tag, cgk_scalars *sc)
{
    __shared__ double s_val[CGK_MAX_RANKS];
    __shared__ double s_sum;
    if ((int)threadIdx.x < G) {
        const unsigned long long *s = slots + 2 * threadIdx.x;
        unsigned long long w0, w1, t0 = 0ull;
        unsigned int polls = 0;
        for (;;) {
            w0 = ld_relaxed_sys(s);
            w1 = ld_relaxed_sys(s + 1);
            if ((unsigned int)(w0 >> 32) == tag && (unsigned int)(w1 >> 32) == tag) break;
            if ((++polls & 1023u) == 0u && spin_give_up(&t0, sc)) break;
        }
        s_val[threadIdx.x] = __longlong_as_double((long long)((w0 & 0xffffffffull) | (w1 << 32)));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int g = 0; g < G; ++g) t += s_val[g];
        s_sum = t;
    }
    __syncthreads();
    return s_sum;
}
/* Threads g < G (g != rank) wait until peer g's p slice of event `tag` has landed here. */
__device__ __forceinline__ void wait_p_flags(const cgk_peers &peers, unsigned int tag, cgk_scalars *sc)
{
    const int g = threadIdx.x;
    if (g < peers.ranks && g != peers.rank) {
        const unsigned long long *f = &peers.mbox[peers.rank]->p_flag[g];
        unsigned long long t0 = 0ull;
        unsigned int polls = 0;
        while (ld_relaxed_sys(f) != (unsigned long long)tag)
            if ((++polls & 1023u) == 0u && spin_give_up(&t0, sc)) break;
        (void)ld_acquire_sys(f);
    }
    __syncthreads();
}
/* Warp-level form for use inside a streaming loop: no CTA barrier, the warps stay decoupled. */
__device__ __forceinline__ void warp_wait_p_flags(const cgk_peers &peers, unsigned int tag, cgk_scalars *sc)
{
    const int g = threadIdx.x & 31;
    if (g < peers.ranks && g != peers.rank) {
        const unsigned long long *f = &peers.mbox[peers.rank]->p_flag[g];
        unsigned long long t0 = 0ull;
        unsigned int polls = 0;
        /* poll relaxed (no L1 invalidation per poll), then acquire once */
        while (ld_relaxed_sys(f) != (unsigned long long)tag)
            if ((++polls & 1023u) == 0u && spin_give_up(&t0, sc)) break;
        (void)ld_acquire_sys(f);
    }
    __syncwarp();
}
/* The same wait for a CTA that asked for the flags long ago.  `early` is what lane g's relaxed load of
 * p_flag[g] returned when it was issued at kernel entry (lanes >= G and the own rank: the tag itself);
 * by now -- the own column block is done -- the value is in a register, so the test costs no memory
 * round trip.  Only a warp that still sees an old tag polls (the first wave).  Returns 0 in every lane,
 * computed FROM the observed flag words: the caller adds it to the base address of its loads of p, so
 * those loads cannot issue before the flag values have arrived (a true address dependency -- the
 * hardware's ordering, with no fence and no second load), and the peer's fence.sys put its slice in
 * this GPU's memory before the flag. */
__device__ __forceinline__ unsigned int warp_flags_ready(const cgk_peers &peers, unsigned int tag, unsigned long long early,
                                                         cgk_scalars *sc)
{
    const int g = threadIdx.x & 31;
    const bool mine = g < peers.ranks && g != peers.rank;
    unsigned long long seen = mine ? early : (unsigned long long)tag;
    if (__any_sync(0xffffffffu, seen != (unsigned long long)tag)) {
        if (mine && seen != (unsigned long long)tag) {
            const unsigned long long *f = &peers.mbox[peers.rank]->p_flag[g];
            unsigned long long t0 = 0ull;
            unsigned int polls = 0;
            while ((seen = ld_relaxed_sys(f)) != (unsigned long long)tag)
                if ((++polls & 1023u) == 0u && spin_give_up(&t0, sc)) { seen = (unsigned long long)tag; break; }
            (void)ld_acquire_sys(f);
        }
        __syncwarp();
    }
    const unsigned long long d = seen ^ (unsigned long long)tag;              /* 0 */
    return __reduce_or_sync(0xffffffffu, (unsigned int)d | (unsigned int)(d >> 32));
}
/* lane g's early request for p_flag[g] (see warp_flags_ready) */
__device__ __forceinline__ unsigned long long flags_request(const cgk_peers &peers, unsigned int tag)
{
    const int g = threadIdx.x & 31;
    if (g < peers.ranks && g != peers.rank) return ld_relaxed_sys(&peers.mbox[peers.rank]->p_flag[g]);
    return (unsigned long long)tag;
}
/* cta_partials_finish for kernels that stored into peer memory: thread 0's fence is
 * system-scope and cumulative over the CTA's stores (all before the preceding barrier), so
 * whatever the last CTA posts afterwards is ordered behind every CTA's peer stores.  Returns
 * true in all threads of the last CTA; *total is valid in its thread 0. */
template <int NT, bool SYS = true>
__device__ __forceinline__ bool cta_partials_finish_sys(double mine, double *partials, double *out,
                                                        unsigned int *ticket, double *total)
{
    __shared__ int s_last;
    __shared__ double s_red[NT / 32];
    const int tid = threadIdx.x;
    if (tid == 0) {
        partials[blockIdx.x] = mine;
        if constexpr (SYS) __threadfence_system(); else __threadfence();
        unsigned int t = atomicAdd(ticket, 1u);
        s_last = (t == gridDim.x - 1);
    }
    __syncthreads();
    if (!s_last) return false;
    __threadfence();
    double s = 0.0;
    for (unsigned int i = tid; i < gridDim.x; i += NT) s += __ldcg(partials + i);
    s = warp_sum(s);
    if ((tid & 31) == 0) s_red[tid >> 5] = s;
    __syncthreads();
    if (tid == 0) {
        double t = 0.0;
#pragma unroll
        for (int q = 0; q < NT / 32; ++q) t += s_red[q];
        *out = t;
        *total = t;
        *ticket = 0u;
        if constexpr (SYS) __threadfence_system();
    }
    return true;
}

