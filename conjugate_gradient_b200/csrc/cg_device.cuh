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
};

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

