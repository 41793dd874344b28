/* cg_stage.c -- pageable host memory to HBM at PCIe rate.
 *
 * A caller shaped like the reference's main() hands conjugrad() arrays that came from malloc
 * (memAllocate(), serialConjugate.c:62-64,76-84; parallel_cg.c:97-99,138-146).  The driver's own
 * path for such a source is a single-threaded bounce through a small internal buffer (~11 GB/s
 * measured, against ~55 GB/s from pinned memory), which made the upload of A the whole cost of a
 * call.  Here the copy is done by a team of host threads, each with its own pair of pinned
 * staging slots and its own copy stream: a thread claims the next piece of the source, waits
 * until the DMA that last used its slot has drained (one event per slot), packs the piece into
 * the slot with memcpy and issues ONE cudaMemcpyAsync / cudaMemcpy2DAsync for it.  The threads
 * never wait for each other; the pieces are disjoint, so their order does not matter.
 *
 * The source is described as a list of 2-D runs (rows of `width` bytes, `spitch` apart on the
 * host, `dpitch` apart on the device): a contiguous block is a run of chunk-wide rows, the tile
 * runs of the asserted-symmetric upload (cgb_load_host_symmetric) are passed as they are.
 *
 * Environment (read at every call; for measurements, tools/h2d_ceiling.py):
 *   CGB_STAGE=0            never stage (the driver's pageable path, as in round 1)
 *   CGB_STAGE=stage        always the ring
 *   CGB_STAGE=register     pin the caller's pages in place, piece by piece, ahead of the DMA
 *                          (unset / auto: the ring)
 *   CGB_STAGE_THREADS=T    copy threads (default: min(8, online CPUs / ranks), at least 2)
 *   CGB_STAGE_CHUNK_MB=C   bytes per piece (default for the ring 8 MiB on one rank, 1 MiB on several; 32 MiB for in-place pinning);
 *   CGB_STAGE_CHUNK_KB=K   the same in KiB (>= 64)
 */
#define _GNU_SOURCE
#include <cuda_runtime_api.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "../../include/cgb200.h"
#include "cg_kernels.h"
#include "cg_internal.h"

#define STAGE_SLOTS 2

int cgb_host_is_pageable(const void *p)
{
    struct cudaPointerAttributes at;
    memset(&at, 0, sizeof at);
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { (void)cudaGetLastError(); return 1; }
    return at.type == cudaMemoryTypeUnregistered;
}

static int env_int(const char *name, int dflt, int lo, int hi)
{
    const char *e = getenv(name);
    if (!e || !*e) return dflt;
    char *end = NULL;
    long v = strtol(e, &end, 10);
    if (end == e || v < lo || v > hi) return dflt;
    return (int)v;
}

/* How a pageable source travels: the staging ring, unless the environment says otherwise.  Pinning the
 * caller's pages in place (CGB_STAGE=register) moves no extra bytes through host memory, but
 * cudaHostRegister serialises between processes: measured with two ranks uploading at once, 4-6 GB/s
 * per GPU against 28 GB/s through the ring (one rank alone: 43 against 53) -- so it stays an option. */
int cgb_stage_mode_for(int ranks)
{
    const char *e = getenv("CGB_STAGE");
    (void)ranks;
    if (!e || !*e || !strcmp(e, "auto")) return CGB_H2D_STAGED;
    if (!strcmp(e, "0") || !strcmp(e, "off")) return CGB_H2D_DRIVER;
    if (!strcmp(e, "register")) return CGB_H2D_REGISTERED;
    return CGB_H2D_STAGED;
}
int cgb_stage_mode(void) { return cgb_stage_mode_for(1); }

static int stage_threads(const cgb_ctx *c)
{
    long cpus = sysconf(_SC_NPROCESSORS_ONLN);
    cpu_set_t set;
    if (sched_getaffinity(0, sizeof set, &set) == 0) { int k = CPU_COUNT(&set); if (k > 0 && k < cpus) cpus = k; }
    if (cpus < 1) cpus = 1;
    int share = (int)(cpus / (c->ranks > 0 ? c->ranks : 1));
    int t = share < 8 ? share : 8;         /* B200 box sweep (tools/h2d_ceiling.py): 8 threads x 8 MiB = 52.8 GB/s of 55.5 pinned */
    if (t < 2) t = 2;
    t = env_int("CGB_STAGE_THREADS", t, 1, CGB_STAGE_MAX_THREADS);
    return t;
}

void cgb_stage_release(cgb_ctx *c)
{
    if (c->stage_pin) cudaFreeHost(c->stage_pin);
    c->stage_pin = NULL; c->stage_pin_bytes = 0;
    for (int t = 0; t < CGB_STAGE_MAX_THREADS; ++t) {
        for (int s = 0; s < STAGE_SLOTS; ++s)
            if (c->stage_ev[t][s]) { cudaEventDestroy(c->stage_ev[t][s]); c->stage_ev[t][s] = NULL; }
        if (c->stage_stream[t]) { cudaStreamDestroy(c->stage_stream[t]); c->stage_stream[t] = NULL; }
    }
    (void)cudaGetLastError();
}

/* pinned slots, streams and events are kept in the context from one call to the next:
 * cudaHostAlloc of the ring costs more than the upload of a small system */
static cudaError_t stage_reserve(cgb_ctx *c, int T, size_t chunk)
{
    const size_t need = (size_t)T * STAGE_SLOTS * chunk;
    if (c->stage_pin_bytes < need) {
        if (c->stage_pin) cudaFreeHost(c->stage_pin);
        c->stage_pin = NULL; c->stage_pin_bytes = 0;
        cudaError_t e = cudaHostAlloc(&c->stage_pin, need, cudaHostAllocDefault);
        if (e != cudaSuccess) return e;
        c->stage_pin_bytes = need;
    }
    for (int t = 0; t < T; ++t) {
        if (!c->stage_stream[t]) {
            cudaError_t e = cudaStreamCreateWithFlags(&c->stage_stream[t], cudaStreamNonBlocking);
            if (e != cudaSuccess) return e;
        }
        for (int s = 0; s < STAGE_SLOTS; ++s)
            if (!c->stage_ev[t][s]) {
                cudaError_t e = cudaEventCreateWithFlags(&c->stage_ev[t][s], cudaEventDisableTiming);
                if (e != cudaSuccess) return e;
            }
    }
    return cudaSuccess;
}

typedef struct {
    cgb_ctx *c;
    const cgb_stage_run *runs;
    int64_t nruns;
    size_t chunk;
    int mode;
    pthread_mutex_t mu;
    int64_t next_run;          /* cursor of the piece generator (under mu) */
    int64_t next_row;
    size_t next_col;
    cudaError_t err;           /* first failure of any thread (under mu)   */
} stage_job;

typedef struct { stage_job *job; int t; } stage_arg;

typedef struct { char *dst; const char *src; size_t width, rows, dpitch, spitch; } piece;

static int next_piece(stage_job *j, piece *p)
{
    int have = 0;
    pthread_mutex_lock(&j->mu);
    while (j->err == cudaSuccess && j->next_run < j->nruns) {
        const cgb_stage_run *r = &j->runs[j->next_run];
        if (j->next_row >= (int64_t)r->rows || r->width == 0) { j->next_run++; j->next_row = 0; j->next_col = 0; continue; }
        if (r->width > j->chunk) {
            /* a row wider than a slot: pieces of one row, cut along its width */
            size_t w = r->width - j->next_col;
            if (w > j->chunk) w = j->chunk;
            p->dst = (char *)r->dst + (size_t)j->next_row * r->dpitch + j->next_col;
            p->src = (const char *)r->src + (size_t)j->next_row * r->spitch + j->next_col;
            p->width = w; p->rows = 1; p->dpitch = p->spitch = w;
            j->next_col += w;
            if (j->next_col >= r->width) { j->next_col = 0; j->next_row++; }
            have = 1;
            break;
        }
        size_t per = j->chunk / r->width;
        size_t rows = r->rows - (size_t)j->next_row;
        if (rows > per) rows = per;
        p->dst = (char *)r->dst + (size_t)j->next_row * r->dpitch;
        p->src = (const char *)r->src + (size_t)j->next_row * r->spitch;
        p->width = r->width; p->rows = rows; p->dpitch = r->dpitch; p->spitch = r->spitch;
        j->next_row += (int64_t)rows;
        have = 1;
        break;
    }
    pthread_mutex_unlock(&j->mu);
    return have;
}

static void job_fail(stage_job *j, cudaError_t e)
{
    pthread_mutex_lock(&j->mu);
    if (j->err == cudaSuccess) j->err = e;
    pthread_mutex_unlock(&j->mu);
}

static cudaError_t dma_piece(const piece *p, const void *from, size_t from_pitch, cudaStream_t st)
{
    if (p->rows == 1 || (p->dpitch == p->width && from_pitch == p->width))
        return cudaMemcpyAsync(p->dst, from, p->width * p->rows, cudaMemcpyHostToDevice, st);
    return cudaMemcpy2DAsync(p->dst, p->dpitch, from, from_pitch, p->width, p->rows, cudaMemcpyHostToDevice, st);
}

static void *stage_main(void *arg)
{
    stage_arg *a = arg;
    stage_job *j = a->job;
    cgb_ctx *c = j->c;
    const int t = a->t;
    cudaError_t e = cudaSetDevice(c->device);
    if (e != cudaSuccess) { job_fail(j, e); return NULL; }
    cudaStream_t st = c->stage_stream[t];
    char *slot[STAGE_SLOTS];
    int used[STAGE_SLOTS];
    for (int s = 0; s < STAGE_SLOTS; ++s) { slot[s] = (char *)c->stage_pin + ((size_t)t * STAGE_SLOTS + (size_t)s) * j->chunk; used[s] = 0; }
    /* register mode: the pieces of the caller's own pages this thread has pinned and not yet released */
    struct { void *p; int on; } reg[STAGE_SLOTS];
    memset(reg, 0, sizeof reg);
    piece p;
    int s = 0;
    while (next_piece(j, &p)) {
        if (used[s]) {
            e = cudaEventSynchronize(c->stage_ev[t][s]);      /* the DMA out of this slot has finished */
            if (e != cudaSuccess) { job_fail(j, e); break; }
            if (reg[s].on) { cudaHostUnregister(reg[s].p); reg[s].on = 0; }
        }
        if (j->mode == CGB_H2D_REGISTERED && p.rows == 1) {
            /* page-aligned interior of the piece pinned in place; the unaligned ends are few bytes
             * and go through the slot */
            const size_t pg = 4096;
            uintptr_t lo = ((uintptr_t)p.src + pg - 1) & ~(uintptr_t)(pg - 1);
            uintptr_t hi = ((uintptr_t)p.src + p.width) & ~(uintptr_t)(pg - 1);
            if (hi > lo && cudaHostRegister((void *)lo, hi - lo, cudaHostRegisterReadOnly) == cudaSuccess) {
                reg[s].p = (void *)lo; reg[s].on = 1;
                const size_t head = lo - (uintptr_t)p.src, tail = (uintptr_t)p.src + p.width - hi;
                e = cudaMemcpyAsync(p.dst + head, (void *)lo, hi - lo, cudaMemcpyHostToDevice, st);
                if (e == cudaSuccess && head) { memcpy(slot[s], p.src, head); e = cudaMemcpyAsync(p.dst, slot[s], head, cudaMemcpyHostToDevice, st); }
                if (e == cudaSuccess && tail) { memcpy(slot[s] + pg, (void *)hi, tail); e = cudaMemcpyAsync(p.dst + head + (hi - lo), slot[s] + pg, tail, cudaMemcpyHostToDevice, st); }
                if (e == cudaSuccess) e = cudaEventRecord(c->stage_ev[t][s], st);
                if (e != cudaSuccess) { job_fail(j, e); break; }
                used[s] = 1; s = (s + 1) % STAGE_SLOTS;
                continue;
            }
            (void)cudaGetLastError();            /* registration refused: stage this piece */
        }
        for (size_t r = 0; r < p.rows; ++r) memcpy(slot[s] + r * p.width, p.src + r * p.spitch, p.width);
        e = dma_piece(&p, slot[s], p.width, st);
        if (e == cudaSuccess) e = cudaEventRecord(c->stage_ev[t][s], st);
        if (e != cudaSuccess) { job_fail(j, e); break; }
        used[s] = 1; s = (s + 1) % STAGE_SLOTS;
    }
    for (s = 0; s < STAGE_SLOTS; ++s) {
        if (used[s]) { e = cudaEventSynchronize(c->stage_ev[t][s]); if (e != cudaSuccess) job_fail(j, e); }
        if (reg[s].on) cudaHostUnregister(reg[s].p);
    }
    return NULL;
}

/* The ring's shape for a staged upload of `total` bytes when nothing in the environment fixes it: pure arithmetic
 * (tests/test_plans.py pins the decisions measured on 1, 2, 4 and 8 GPUs).  In: the team size the host allows, the
 * best piece size once the ring exists (8 MiB alone, 1 MiB with other ranks), the ring the context already holds.
 * Out: the path (the driver's below 64 MiB), threads, bytes per slot. */
void cgb_stage_plan(size_t total, int ranks, int team, size_t max_chunk, size_t have_ring, int *mode_out, int *threads_out, size_t *chunk_out)
{
    int T = team;
    *mode_out = CGB_H2D_STAGED; *threads_out = T; *chunk_out = max_chunk;
    if (total < ((size_t)64 << 20)) { *mode_out = CGB_H2D_DRIVER; return; }
    const size_t ring_max = (size_t)T * STAGE_SLOTS * max_chunk;
    const int solo = ranks <= 1;          /* several ranks: 1 MiB slots, ring_max is small already (measured on 2, 4, 8 GPUs) */
    size_t ring;
    if (solo && total >= ((size_t)16 << 30)) ring = ring_max;
    else {
        ring = total / 40;
        if (ring < ((size_t)2 << 20)) ring = (size_t)2 << 20;
        if (solo && ring > ((size_t)16 << 20)) ring = (size_t)16 << 20;
    }
    if (ring > ring_max) ring = ring_max;
    if (have_ring > ring) ring = have_ring < ring_max ? have_ring : ring_max;
    /* A small ring is cut into FEWER, LARGER slots: pieces under ~2 MiB move slowly whatever the thread count
     * (profiles/r02_h2d_1gpu_sweep.json, 16 MiB of ring either way: 8 threads x 1 MiB 33.7 GB/s, 4 threads x 2 MiB
     * 46.6; 2 threads never pass 24.5), so a thread is added per 4 MiB of ring, from two up to the team's size. */
    if (solo) {
        int Tr = (int)((ring + ((size_t)4 << 20) - 1) / ((size_t)4 << 20));
        if (Tr < 2) Tr = 2;
        if (Tr < T) T = Tr;
    }
    size_t chunk = ring / ((size_t)T * STAGE_SLOTS) & ~(size_t)0xffff;              /* multiple of 64 KiB */
    if (chunk < ((size_t)256 << 10)) {                                              /* few bytes: fewer threads, not smaller pieces */
        chunk = (size_t)256 << 10;
        T = (int)(ring / (STAGE_SLOTS * chunk));
        if (T < 1) T = 1;
    }
    *threads_out = T; *chunk_out = chunk;
}

/* All runs to the device; returns when every byte has arrived.  *path_out says how. */
int cgb_stage_upload(cgb_ctx *c, const cgb_stage_run *runs, int64_t nruns, int pageable, int *path_out)
{
    size_t total = 0;
    for (int64_t i = 0; i < nruns; ++i) total += runs[i].width * runs[i].rows;
    int mode = pageable ? cgb_stage_mode_for(c->ranks) : CGB_H2D_DIRECT;
    /* piece size of the ring: alone on the host 8 MiB pieces measured best (52.3 GB/s of 55.5 pinned; 2 MiB: 43-47);
     * with other ranks uploading at the same time the host's memory bandwidth is shared and small pieces -- rings
     * small enough to be read back by the DMA engine out of the last-level cache -- do better: two ranks 37 GB/s
     * per GPU with 1-2 MiB against 28 with 8 MiB; eight ranks (4 threads each) 13.6-14.4 with 1 MiB against 10.4-11
     * with 2 MiB and 8.8 with 8 MiB (profiles/r02_h2d_2gpu_modes.json, r02_h2d_g8.json) */
    const size_t max_chunk = (size_t)(mode == CGB_H2D_REGISTERED ? 32 : (c->ranks > 1 ? 1 : 8)) << 20;
    size_t chunk = (size_t)env_int("CGB_STAGE_CHUNK_MB", (int)(max_chunk >> 20), 1, 256) << 20;
    int chunk_given = getenv("CGB_STAGE_CHUNK_MB") != NULL;
    {
        const int kb = env_int("CGB_STAGE_CHUNK_KB", 0, 64, 256 << 10);       /* finer than a MiB, for the sweep */
        if (kb > 0) { chunk = (size_t)kb << 10; chunk_given = 1; }
    }
    int T = stage_threads(c);
    if (mode == CGB_H2D_STAGED && !chunk_given) {
        /* The ring is pinned memory, and pinning is not free: cudaHostAlloc measured 0.37-0.45 ms per MiB on the bench
         * boxes, so the full 8 x 2 x 8 MiB ring costs 45-58 ms -- more than the whole upload of an n = 8192 system
         * (537 MB: 10 ms at the pinned rate, 50 ms through the driver).  A caller shaped like the reference's main()
         * calls conjugrad() ONCE, so the first call is the one that counts.  Measured (tools/first_call.py,
         * profiles/r02_first_call_slots.txt): a ring of 13-16 MiB worked by 4 threads with ~2 MiB slots already moves
         * 44-47 GB/s (the full ring: 50-52), and a larger ring's pinning is not paid back below ~20 GB of upload
         * (n = 16384, 2.1 GB: first call 53 ms with 13 MiB of ring against 83 ms with total / 40 = 54 MiB).  So:
         *   - total / 40, between 2 and 16 MiB, below 16 GiB of upload;
         *   - the full ring (threads x 2 x the best piece size) for uploads of 16 GiB and more;
         *   - a ring only ever grows; a later call uses what is there.
         * Growing to the full ring on a context's third upload was tried and dropped: the growth cost 60-160 ms
         * (free, pin 128 MiB, more streams) and the uploads after it were no faster (n = 16384: 44.4 against 45.0 ms).
         * Below 64 MiB the driver's own path (11 GB/s, no set-up) wins. */
        cgb_stage_plan(total, c->ranks, T, max_chunk, c->stage_pin_bytes, &mode, &T, &chunk);
    }
    if (mode != CGB_H2D_DIRECT && mode != CGB_H2D_DRIVER && total < 2 * chunk) mode = CGB_H2D_DRIVER;   /* small: not worth a thread team */
    if (path_out) *path_out = mode;
    if (mode == CGB_H2D_DIRECT || mode == CGB_H2D_DRIVER) {
        /* pinned (or registered, or managed) source: the DMA engine reads it in place; pieces of at
         * most 256 MiB so that a pageable source still pipelines through the driver's bounce buffer */
        const size_t slab = (size_t)256 << 20;
        for (int64_t i = 0; i < nruns; ++i) {
            const cgb_stage_run *r = &runs[i];
            if (r->rows == 0 || r->width == 0) continue;
            cudaError_t e = cudaSuccess;
            if (r->rows == 1 || (r->dpitch == r->width && r->spitch == r->width)) {
                const size_t bytes = r->width * r->rows;
                for (size_t off = 0; off < bytes && e == cudaSuccess; off += slab)
                    e = cudaMemcpyAsync((char *)r->dst + off, (const char *)r->src + off, bytes - off < slab ? bytes - off : slab,
                                        cudaMemcpyHostToDevice, c->stream);
            } else
                e = cudaMemcpy2DAsync(r->dst, r->dpitch, r->src, r->spitch, r->width, r->rows, cudaMemcpyHostToDevice, c->stream);
            if (e != cudaSuccess) {
                cgb_set_error("CUDA error: %s in `%s` (cg_stage.c:%d)", cudaGetErrorString(e), "host to device copy", __LINE__);
                (void)cudaGetLastError();
                return CGB_ERR_CUDA;
            }
        }
        cudaError_t e = cudaStreamSynchronize(c->stream);
        if (e != cudaSuccess) {
            cgb_set_error("CUDA error: %s in `%s` (cg_stage.c:%d)", cudaGetErrorString(e), "host to device copy", __LINE__);
            return CGB_ERR_CUDA;
        }
        return CGB_OK;
    }

    {
        size_t pieces = (total + chunk - 1) / chunk;
        if ((size_t)T > pieces) T = (int)pieces;
        if (T < 1) T = 1;
    }
    cudaError_t e = stage_reserve(c, T, chunk);
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        cgb_set_error("CUDA error: %s in `%s` (cg_stage.c:%d)", cudaGetErrorString(e), "pinned staging ring", __LINE__);
        return e == cudaErrorMemoryAllocation ? CGB_ERR_NOMEM : CGB_ERR_CUDA;
    }
    /* the copy streams start after whatever the solver stream still has in flight on these arrays */
    e = cudaStreamSynchronize(c->stream);
    if (e != cudaSuccess) { cgb_set_error("CUDA error: %s in `%s` (cg_stage.c:%d)", cudaGetErrorString(e), "stream sync", __LINE__); return CGB_ERR_CUDA; }

    stage_job job;
    memset(&job, 0, sizeof job);
    job.c = c; job.runs = runs; job.nruns = nruns; job.chunk = chunk; job.mode = mode; job.err = cudaSuccess;
    pthread_mutex_init(&job.mu, NULL);
    pthread_t th[CGB_STAGE_MAX_THREADS];
    stage_arg arg[CGB_STAGE_MAX_THREADS];
    char started[CGB_STAGE_MAX_THREADS];
    for (int t = 0; t < T; ++t) { arg[t].job = &job; arg[t].t = t; }
    for (int t = 1; t < T; ++t) started[t] = pthread_create(&th[t], NULL, stage_main, &arg[t]) == 0;
    stage_main(&arg[0]);                       /* the caller is a member of the team */
    for (int t = 1; t < T; ++t) if (started[t]) pthread_join(th[t], NULL);
    /* a thread that could not be started: thread 0's loop has already drained its share */
    pthread_mutex_destroy(&job.mu);
    c->stage_threads_used = T;
    if (job.err != cudaSuccess) {
        (void)cudaGetLastError();
        cgb_set_error("CUDA error: %s in `%s` (cg_stage.c:%d)", cudaGetErrorString(job.err), "staged host to device copy", __LINE__);
        return CGB_ERR_CUDA;
    }
    return CGB_OK;
}
