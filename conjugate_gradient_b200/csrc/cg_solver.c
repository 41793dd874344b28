/* cg_solver.c -- host side of libcgb200.so, in C like the reference.
 *
 * Replaces conjugrad() of the reference (serialConjugate.c:180-259,
 * parallel_cg.c:248-345): the per-rank solver state lives on one B200, the
 * iteration is a short chain of hand-written kernels (cg_kernels.cu) plus, for
 * ranks > 1, the three collectives of the reference re-expressed in NCCL over
 * NVLink:  MPI_Allgather(p) parallel_cg.c:290-291 -> ncclAllGather,
 *          MPI_Allreduce(p.Ap) :294 and MPI_Allreduce(r.r) :287,313 -> ncclAllReduce.
 * All solver scalars stay in device memory; the host only replays a captured
 * CUDA graph of `chunk` iterations and polls one flag between replays.
 */
#define _GNU_SOURCE
#include <cuda_runtime_api.h>
#include <nccl.h>
#include <math.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include "../../include/cgb200.h"
#include "cg_kernels.h"
#include "cg_internal.h"
#include "spd_gen.cuh"

static __thread char tl_err[512];

const char *cgb_last_error(void) { return tl_err; }
int cgb_version(void) { return CGB_VERSION; }

void cgb_set_error(const char *fmt, const char *a, const char *b, int line)
{
    snprintf(tl_err, sizeof tl_err, fmt, a, b, line);
}

#define CU(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            cgb_set_error("CUDA error: %s in `%s` (cg_solver.c:%d)", cudaGetErrorString(e_), #call, __LINE__); \
            return CGB_ERR_CUDA;                                                         \
        }                                                                                \
    } while (0)
#define KL(call)                                                                         \
    do {                                                                                 \
        int e_ = (call);                                                                 \
        if (e_ != 0) {                                                                   \
            cgb_set_error("kernel launch error: %s in `%s` (cg_solver.c:%d)", cgk_error_string(e_), #call, __LINE__); \
            return CGB_ERR_CUDA;                                                         \
        }                                                                                \
    } while (0)
#define NC(call)                                                                         \
    do {                                                                                 \
        ncclResult_t e_ = (call);                                                        \
        if (e_ != ncclSuccess) {                                                         \
            cgb_set_error("NCCL error: %s in `%s` (cg_solver.c:%d)", ncclGetErrorString(e_), #call, __LINE__); \
            return CGB_ERR_NCCL;                                                         \
        }                                                                                \
    } while (0)
#define FAIL(code, msg)                                                                  \
    do {                                                                                 \
        cgb_set_error("%s%s (cg_solver.c:%d)", msg, "", __LINE__);                       \
        return (code);                                                                   \
    } while (0)

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

int cgb_device_count(void)
{
    int c = 0;
    if (cudaGetDeviceCount(&c) != cudaSuccess) { (void)cudaGetLastError(); return -1; }
    return c;
}

void cgb_default_options(cgb_options *o)
{
    memset(o, 0, sizeof *o);
    o->eps = CGB_EPSILON;
    o->use_graph = 1;
}

/* ------------------------------------------------------------------ guarded device arrays
 * With CGB_DEBUG_GUARDS=1 every solver array sits between two 256-byte bands of a fixed byte,
 * which cgb_debug_check_guards() reads back: a kernel writing one element outside its array is
 * caught by the tests (compute-sanitizer's job where it is available).  Off by default: plain
 * cudaMalloc / cudaFree. */
static cudaError_t galloc(cgb_ctx *c, void **out, size_t bytes, const char *name)
{
    *out = NULL;
    if (!c->guards_on) return cudaMalloc(out, bytes);
    if (c->n_guards >= CGB_MAX_GUARDS) return cudaErrorMemoryAllocation;
    char *base = NULL;
    cudaError_t e = cudaMalloc((void **)&base, bytes + 2 * CGB_GUARD_BYTES);
    if (e != cudaSuccess) return e;
    e = cudaMemset(base, CGB_GUARD_BYTE, CGB_GUARD_BYTES);
    if (e == cudaSuccess) e = cudaMemset(base + CGB_GUARD_BYTES + bytes, CGB_GUARD_BYTE, CGB_GUARD_BYTES);
    if (e != cudaSuccess) { cudaFree(base); return e; }
    c->guards[c->n_guards].user = base + CGB_GUARD_BYTES;
    c->guards[c->n_guards].bytes = bytes;
    c->guards[c->n_guards].name = name;
    c->n_guards++;
    *out = base + CGB_GUARD_BYTES;
    return cudaSuccess;
}

static void gfree(cgb_ctx *c, void *user)
{
    if (!user) return;
    for (int i = 0; i < c->n_guards; ++i)
        if (c->guards[i].user == user) {
            cudaFree((char *)user - CGB_GUARD_BYTES);
            c->guards[i] = c->guards[--c->n_guards];
            return;
        }
    cudaFree(user);
}

int cgb_debug_check_guards(cgb_ctx *c)
{
    if (!c) FAIL(CGB_ERR_ARG, "cgb_debug_check_guards: null context");
    if (!c->guards_on) FAIL(CGB_ERR_STATE, "guard bands are off: create the context with CGB_DEBUG_GUARDS=1 in the environment");
    CU(cudaSetDevice(c->device));
    CU(cudaStreamSynchronize(c->stream));
    unsigned char h[2][CGB_GUARD_BYTES];
    for (int i = 0; i < c->n_guards; ++i) {
        char *u = c->guards[i].user;
        CU(cudaMemcpy(h[0], u - CGB_GUARD_BYTES, CGB_GUARD_BYTES, cudaMemcpyDeviceToHost));
        CU(cudaMemcpy(h[1], u + c->guards[i].bytes, CGB_GUARD_BYTES, cudaMemcpyDeviceToHost));
        for (int side = 0; side < 2; ++side)
            for (int k = 0; k < CGB_GUARD_BYTES; ++k)
                if (h[side][k] != CGB_GUARD_BYTE) {
                    snprintf(tl_err, sizeof tl_err, "guard band %s array `%s` (%zu bytes) was overwritten at byte %d",
                             side ? "after" : "before", c->guards[i].name, c->guards[i].bytes, side ? k : k - CGB_GUARD_BYTES);
                    return CGB_ERR_STATE;
                }
    }
    if (c->win) {                                     /* the slack after [mailbox | p_full | crecv] */
        const size_t used = CGK_WINDOW_HEADER_BYTES + 2 * (size_t)c->cap_n * sizeof(double);
        CU(cudaMemcpy(h[0], (char *)c->win + used, CGB_GUARD_BYTES, cudaMemcpyDeviceToHost));
        for (int k = 0; k < CGB_GUARD_BYTES; ++k)
            if (h[0][k] != CGB_GUARD_BYTE) {
                snprintf(tl_err, sizeof tl_err, "guard band after the peer window was overwritten at byte %d", k);
                return CGB_ERR_STATE;
            }
    }
    return CGB_OK;
}

/* ------------------------------------------------------------------ contexts */
static int ctx_init_common(cgb_ctx *c, int device)
{
    {
        const char *e = getenv("CGB_DEBUG_GUARDS");
        c->guards_on = e && *e && *e != '0';
    }
    struct cudaDeviceProp prop;
    CU(cudaSetDevice(device));
    CU(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) {
        snprintf(tl_err, sizeof tl_err,
                 "device %d (%s, sm_%d%d) is not a Blackwell sm_100 GPU: libcgb200 has no other code path",
                 device, prop.name, prop.major, prop.minor);
        return CGB_ERR_CUDA;
    }
    c->device = device;
    c->sm_count = prop.multiProcessorCount;
    c->l2_bytes = (int64_t)prop.l2CacheSize;
    {
        /* experiments: CGB_L2_SETASIDE=1 reserves the maximum persisting L2 carve-out */
        const char *e = getenv("CGB_L2_SETASIDE");
        if (e && *e == '1' && prop.persistingL2CacheMaxSize > 0)
            (void)cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, (size_t)prop.persistingL2CacheMaxSize);
        (void)cudaGetLastError();
    }
    snprintf(c->name, sizeof c->name, "%.63s", prop.name);
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&c->ev[0]));
    CU(cudaEventCreate(&c->ev[1]));
    CU(cudaEventCreate(&c->ev[2]));
    CU(galloc(c, (void **)&c->sc, sizeof(cgk_scalars), "scalars"));
    CU(cudaMemset(c->sc, 0, sizeof(cgk_scalars)));
    CU(cudaHostAlloc((void **)&c->h_sc, sizeof(cgk_scalars), cudaHostAllocDefault));
    CU(galloc(c, (void **)&c->partials_vec, 2 * (size_t)cgk_vec_max_ctas(1) * sizeof(double), "partials_vec"));
    return CGB_OK;
}

int cgb_create(cgb_ctx **out, int device)
{
    if (!out) FAIL(CGB_ERR_ARG, "cgb_create: null output pointer");
    cgb_ctx *c = calloc(1, sizeof *c);
    if (!c) FAIL(CGB_ERR_NOMEM, "cgb_create: out of host memory");
    c->rank = 0; c->ranks = 1;
    int rc = ctx_init_common(c, device);
    if (rc != CGB_OK) { cgb_destroy(c); return rc; }
    *out = c;
    return CGB_OK;
}

int cgb_unique_id(void *id128)
{
    if (!id128) FAIL(CGB_ERR_ARG, "cgb_unique_id: null buffer");
    ncclUniqueId id;
    NC(ncclGetUniqueId(&id));
    memcpy(id128, &id, sizeof id);
    return CGB_OK;
}

int cgb_create_rank(cgb_ctx **out, int device, int rank, int ranks, const void *id128)
{
    if (!out || ranks < 1 || rank < 0 || rank >= ranks) FAIL(CGB_ERR_ARG, "cgb_create_rank: bad rank/ranks");
    if (ranks > 1 && !id128) FAIL(CGB_ERR_ARG, "cgb_create_rank: ranks > 1 needs a unique id");
    cgb_ctx *c = calloc(1, sizeof *c);
    if (!c) FAIL(CGB_ERR_NOMEM, "cgb_create_rank: out of host memory");
    c->rank = rank; c->ranks = ranks;
    int rc = ctx_init_common(c, device);
    if (rc != CGB_OK) { cgb_destroy(c); return rc; }
    if (ranks > 1) {
        ncclUniqueId id;
        memcpy(&id, id128, sizeof id);
        ncclResult_t e = ncclCommInitRank(&c->comm, ranks, id, rank);
        if (e != ncclSuccess) {
            snprintf(tl_err, sizeof tl_err, "NCCL error: %s in ncclCommInitRank", ncclGetErrorString(e));
            cgb_destroy(c);
            return CGB_ERR_NCCL;
        }
        c->has_comm = 1;
    }
    *out = c;
    return CGB_OK;
}

static void drop_graph(cgb_ctx *c);
static void close_peers(cgb_ctx *c);
void cgb_unmap_peers(cgb_ctx *c)
{
    if (!c) return;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    drop_graph(c);                    /* the captured kernels point into the peers' windows */
    close_peers(c);
    (void)cudaGetLastError();
}

static void drop_graph(cgb_ctx *c)
{
    if (c->graph_exec) { cudaGraphExecDestroy(c->graph_exec); c->graph_exec = NULL; }
    if (c->graph) { cudaGraphDestroy(c->graph); c->graph = NULL; }
    memset(&c->graph_key, 0, sizeof c->graph_key);
}

/* Fault injection for the tests of the agreement below: CGB_DEBUG_FAIL_<step>=<rank> makes that rank
 * (of a team of several) fail that local step with CGB_ERR_NOMEM, as if it had run out of memory. */
static int injected_failure(const cgb_ctx *c, const char *step)
{
    char name[64];
    snprintf(name, sizeof name, "CGB_DEBUG_FAIL_%s", step);
    const char *e = getenv(name);
    if (!e || !*e || c->ranks < 2 || atoi(e) != c->rank) return 0;
    snprintf(tl_err, sizeof tl_err, "rank %d: injected failure of step %s (%s is set)", c->rank, step, name);
    return 1;
}

/* Every rank contributes its status (CGB_OK or a negative code) and all ranks get the smallest one:
 * the ranks agree BEFORE a collective step whether to enter it, so that a rank which failed on its
 * own (out of device memory, a bad input file, a failed copy) never leaves the others inside an NCCL
 * call or a spin on a mailbox -- NCCL has no time-out.  One 4-byte all-reduce; ranks == 1: nothing. */
int cgb_ranks_agree(cgb_ctx *c, int rc_local)
{
    if (c->ranks < 2 || !c->has_comm || c->broken) return rc_local;
    if (!c->d_agree) {
        if (cudaMalloc((void **)&c->d_agree, 2 * sizeof(int)) != cudaSuccess) { (void)cudaGetLastError(); c->d_agree = NULL; }
    }
    int agreed = rc_local;
    if (!c->d_agree ||
        cudaMemcpyAsync(c->d_agree, &rc_local, sizeof(int), cudaMemcpyHostToDevice, c->stream) != cudaSuccess ||
        ncclAllReduce(c->d_agree, c->d_agree + 1, 1, ncclInt, ncclMin, c->comm, c->stream) != ncclSuccess ||
        cudaMemcpyAsync(&agreed, c->d_agree + 1, sizeof(int), cudaMemcpyDeviceToHost, c->stream) != cudaSuccess ||
        cudaStreamSynchronize(c->stream) != cudaSuccess) {
        (void)cudaGetLastError();
        c->broken = 1;
        if (rc_local == CGB_OK) cgb_set_error("%s%s (cg_solver.c:%d)", "the ranks could not agree on a status (NCCL or CUDA failure)", "", __LINE__);
        return rc_local != CGB_OK ? rc_local : CGB_ERR_NCCL;
    }
    if (agreed != CGB_OK && rc_local == CGB_OK)
        snprintf(tl_err, sizeof tl_err, "rank %d: another rank failed (status %d) before a collective step; all ranks give up together", c->rank, agreed);
    return agreed;
}

/* Unmaps the peers' windows.  No collective here: a solve ends with the all-gather of x, so
 * once cgb_solve() has returned on a rank no peer touches its window until the next
 * (collective) solve or re-allocation. */
static void close_peers(cgb_ctx *c)
{
    for (int g = 0; g < CGK_MAX_RANKS; ++g) {
        if (c->peer_ipc[g] && c->peer_win[g]) cudaIpcCloseMemHandle(c->peer_win[g]);
        c->peer_win[g] = NULL; c->peer_ipc[g] = 0;
    }
    memset(&c->peers, 0, sizeof c->peers);
    c->peer_ok = 0;
    (void)cudaGetLastError();
}

static void free_problem(cgb_ctx *c)
{
    drop_graph(c);
    int ipc = 0;
    for (int g = 0; g < CGK_MAX_RANKS; ++g) ipc |= c->peer_ipc[g];
    close_peers(c);
    /* The window was exported with cudaIpcGetMemHandle and other PROCESSES hold it open: freeing it
     * before they have closed their mapping is undefined (CUDA runtime, cudaIpcCloseMemHandle).  Every
     * rank unmaps first (above), then all meet, then each frees its own.  Ranks that are threads of
     * one process map each other by pointer and need no rendezvous (and may be torn down one after the
     * other by a single thread, cg_dropin.c: team_release).  A communicator already known to be broken
     * is not waited on. */
    if (ipc && c->win) (void)cgb_ranks_agree(c, CGB_OK);
    gfree(c, c->A); gfree(c, c->b); gfree(c, c->r); gfree(c, c->Ap); gfree(c, c->svec);
    c->svec = NULL;
    if (c->win) cudaFree(c->win); else gfree(c, c->p_full);
    gfree(c, c->x_full); gfree(c, c->partials_gemv);
    gfree(c, c->sy_rowpart); gfree(c, c->sy_colpart); gfree(c, c->sy_tiles);
    c->sy_rowpart = c->sy_colpart = NULL; c->sy_tiles = NULL; c->sy_ntiles = 0; c->sym_state = 0;
    c->A = c->b = c->r = c->Ap = c->p_full = c->x_full = c->partials_gemv = NULL;
    c->win = NULL; c->win_bytes = 0;
    c->n = c->m = 0; c->cap_n = c->cap_m = 0; c->loaded = 0;
}

/* Collective over the ranks: every rank publishes its window (CUDA IPC handle, or the raw
 * pointer for ranks that are threads of one process) through one small ncclAllGather, maps
 * the others', and the ranks agree (all-reduce MIN) on whether the peer path is usable.  This
 * is the only set-up the fused exchange needs; it replaces nothing in the reference, where
 * MPI owns the transport. */
typedef struct {
    cudaIpcMemHandle_t handle;        /* 64 B */
    uint64_t ptr;
    int64_t pid;
    int32_t device;
    int32_t exported;
    char pad[128 - 64 - 8 - 8 - 4 - 4];
} window_card;

static int exchange_windows(cgb_ctx *c, int64_t n)
{
    const int G = c->ranks;
    cudaStream_t st = c->stream;
    close_peers(c);
    snprintf(c->peer_why, sizeof c->peer_why, "%s", "");
    int usable = G <= CGK_MAX_RANKS;
    const char *off = getenv("CGB_DISABLE_PEER");
    if (off && *off && *off != '0') { usable = 0; snprintf(c->peer_why, sizeof c->peer_why, "CGB_DISABLE_PEER is set"); }
    if (G > CGK_MAX_RANKS) snprintf(c->peer_why, sizeof c->peer_why, "more than %d ranks", CGK_MAX_RANKS);

    window_card mine;
    memset(&mine, 0, sizeof mine);
    mine.ptr = (uint64_t)(uintptr_t)c->win;
    mine.pid = (int64_t)getpid();
    mine.device = c->device;
    if (usable) {
        if (cudaIpcGetMemHandle(&mine.handle, c->win) == cudaSuccess) mine.exported = 1;
        else (void)cudaGetLastError();       /* still usable by ranks of this process */
    }
    window_card *d_cards = NULL, *h_cards = NULL;
    int *d_ok = NULL;
    CU(cudaMalloc((void **)&d_cards, (size_t)G * sizeof(window_card) + 64));
    h_cards = malloc((size_t)G * sizeof(window_card));
    if (!h_cards) { cudaFree(d_cards); FAIL(CGB_ERR_NOMEM, "out of host memory"); }
    d_ok = (int *)((char *)d_cards + (size_t)G * sizeof(window_card));
    int rc = CGB_OK;
#define XW(call) do { if ((call) != cudaSuccess) { cgb_set_error("CUDA error: %s in `%s` (cg_solver.c:%d)", cudaGetErrorString(cudaGetLastError()), #call, __LINE__); rc = CGB_ERR_CUDA; goto out; } } while (0)
#define XN(call) do { ncclResult_t e_ = (call); if (e_ != ncclSuccess) { cgb_set_error("NCCL error: %s in `%s` (cg_solver.c:%d)", ncclGetErrorString(e_), #call, __LINE__); rc = CGB_ERR_NCCL; goto out; } } while (0)
    XW(cudaMemcpyAsync(d_cards + c->rank, &mine, sizeof mine, cudaMemcpyHostToDevice, st));
    XN(ncclAllGather(d_cards + c->rank, d_cards, sizeof(window_card), ncclChar, c->comm, st));
    XW(cudaMemcpyAsync(h_cards, d_cards, (size_t)G * sizeof(window_card), cudaMemcpyDeviceToHost, st));
    XW(cudaStreamSynchronize(st));

    int ok = usable;
    for (int g = 0; g < G && ok; ++g) {
        if (g == c->rank) { c->peer_win[g] = c->win; continue; }
        const window_card *k = &h_cards[g];
        if (k->pid == mine.pid) {                       /* a thread of this process */
            if (k->device == c->device) { ok = 0; snprintf(c->peer_why, sizeof c->peer_why, "two ranks share GPU %d", c->device); break; }
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, c->device, k->device) != cudaSuccess || !can) {
                (void)cudaGetLastError();
                ok = 0; snprintf(c->peer_why, sizeof c->peer_why, "GPU %d cannot access GPU %d", c->device, k->device);
                break;
            }
            cudaError_t e = cudaDeviceEnablePeerAccess(k->device, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) {
                ok = 0; snprintf(c->peer_why, sizeof c->peer_why, "cudaDeviceEnablePeerAccess: %s", cudaGetErrorString(e));
            }
            (void)cudaGetLastError();
            c->peer_win[g] = (void *)(uintptr_t)k->ptr;
        } else {
            void *q = NULL;
            cudaError_t e = k->exported ? cudaIpcOpenMemHandle(&q, k->handle, cudaIpcMemLazyEnablePeerAccess)
                                        : cudaErrorNotSupported;
            if (e != cudaSuccess) {
                (void)cudaGetLastError();
                ok = 0; snprintf(c->peer_why, sizeof c->peer_why, "cudaIpcOpenMemHandle(rank %d): %s", g, cudaGetErrorString(e));
                break;
            }
            c->peer_win[g] = q; c->peer_ipc[g] = 1;
        }
    }
    /* agree: MIN over the ranks */
    XW(cudaMemcpyAsync(d_ok, &ok, sizeof ok, cudaMemcpyHostToDevice, st));
    XN(ncclAllReduce(d_ok, d_ok, 1, ncclInt, ncclMin, c->comm, st));
    int all_ok = 0;
    XW(cudaMemcpyAsync(&all_ok, d_ok, sizeof ok, cudaMemcpyDeviceToHost, st));
    XW(cudaStreamSynchronize(st));
    if (all_ok) {
        c->peers.rank = c->rank; c->peers.ranks = G;
        for (int g = 0; g < G; ++g) {
            c->peers.mbox[g] = (cgk_mailbox *)c->peer_win[g];
            c->peers.p_full[g] = (double *)((char *)c->peer_win[g] + CGK_WINDOW_HEADER_BYTES);
            c->peers.crecv[g] = c->peers.p_full[g] + n;             /* [ranks][m] = n doubles; c->n is not set yet */
        }
        c->peer_ok = 1;
    } else {
        if (ok) snprintf(c->peer_why, sizeof c->peer_why, "another rank could not map its peers");
        int keep_why = 1;
        char why[sizeof c->peer_why];
        snprintf(why, sizeof why, "%s", c->peer_why);
        close_peers(c);
        if (keep_why) snprintf(c->peer_why, sizeof c->peer_why, "%s", why);
    }
out:
#undef XW
#undef XN
    free(h_cards);
    cudaFree(d_cards);
    return rc;
}

int cgb_destroy(cgb_ctx *c)
{
    if (!c) return CGB_OK;
    cudaSetDevice(c->device);
    if (c->stream) cudaStreamSynchronize(c->stream);
    free_problem(c);
    cgb_stage_release(c);
    cudaFree(c->d_agree);
    if (c->has_comm) { if (c->broken) ncclCommAbort(c->comm); else ncclCommDestroy(c->comm); }
    for (int i = 0; i < 3; ++i) if (c->ev[i]) cudaEventDestroy(c->ev[i]);
    for (int i = 0; i < c->n_gemv_ev; ++i) cudaEventDestroy(c->gemv_ev[i]);
    free(c->gemv_ev);
    gfree(c, c->sc); gfree(c, c->partials_vec); cudaFree(c->flush_buf); cudaFree(c->probe);
    if (c->h_sc) cudaFreeHost(c->h_sc);
    if (c->stream) cudaStreamDestroy(c->stream);
    (void)cudaGetLastError();
    free(c);
    return CGB_OK;
}

int cgb_device_info(cgb_ctx *c, int *device, int *sm_count, int64_t *free_bytes,
                    int64_t *total_bytes, char *name64)
{
    if (!c) FAIL(CGB_ERR_ARG, "cgb_device_info: null context");
    size_t fr = 0, tot = 0;
    CU(cudaSetDevice(c->device));
    CU(cudaMemGetInfo(&fr, &tot));
    if (device) *device = c->device;
    if (sm_count) *sm_count = c->sm_count;
    if (free_bytes) *free_bytes = (int64_t)fr;
    if (total_bytes) *total_bytes = (int64_t)tot;
    if (name64) snprintf(name64, 64, "%s", c->name);
    return CGB_OK;
}

/* ------------------------------------------------------------------ problem */
/* The device arrays of an n x n system on this rank; no collective inside. */
static int alloc_problem(cgb_ctx *c, int64_t n, int64_t m)
{
    if (injected_failure(c, "ALLOC")) return CGB_ERR_NOMEM;
    const size_t abytes = (size_t)m * (size_t)n * sizeof(double);
    cudaError_t e = galloc(c, (void **)&c->A, abytes, "A");
    if (e != cudaSuccess) {
        (void)cudaGetLastError();
        snprintf(tl_err, sizeof tl_err, "can't allocate memory for matrix block: %zu bytes on GPU %d (%s)",
                 abytes, c->device, cudaGetErrorString(e));
        return CGB_ERR_NOMEM;
    }
#define AL(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { (void)cudaGetLastError(); \
        cgb_set_error("can't allocate memory for vector: %s in `%s` (cg_solver.c:%d)", cudaGetErrorString(e_), #call, __LINE__); \
        return e_ == cudaErrorMemoryAllocation ? CGB_ERR_NOMEM : CGB_ERR_CUDA; } } while (0)
    AL(galloc(c, (void **)&c->b, (size_t)m * sizeof(double), "b"));
    AL(galloc(c, (void **)&c->r, (size_t)m * sizeof(double), "r"));
    AL(galloc(c, (void **)&c->Ap, (size_t)m * sizeof(double), "Ap"));
    AL(galloc(c, (void **)&c->svec, (size_t)m * sizeof(double), "s"));
    if (c->ranks > 1) {
        /* [mailbox | p_full (n) | column-sum receive area (ranks x m = n)] in one allocation of whole
         * 2 MiB pages: one IPC handle */
        const size_t gran = (size_t)2 << 20;
        const size_t used = CGK_WINDOW_HEADER_BYTES + 2 * (size_t)n * sizeof(double);
        c->win_bytes = (used + CGB_GUARD_BYTES + gran - 1) / gran * gran;
        AL(cudaMalloc(&c->win, c->win_bytes));
        AL(cudaMemsetAsync(c->win, 0, c->win_bytes, c->stream));
        AL(cudaMemsetAsync((char *)c->win + used, CGB_GUARD_BYTE, CGB_GUARD_BYTES, c->stream));
        AL(cudaStreamSynchronize(c->stream));
        c->p_full = (double *)((char *)c->win + CGK_WINDOW_HEADER_BYTES);
    } else {
        AL(galloc(c, (void **)&c->p_full, (size_t)n * sizeof(double), "p"));
    }
    AL(galloc(c, (void **)&c->x_full, (size_t)n * sizeof(double), "x"));
    AL(galloc(c, (void **)&c->partials_gemv, (size_t)cgk_gemv_max_ctas(m) * sizeof(double), "partials_gemv"));
#undef AL
    return CGB_OK;
}

int cgb_ensure_problem(cgb_ctx *c, int64_t n)
{
    if (n < 1) FAIL(CGB_ERR_ARG, "matrix order must be >= 1");
    if (n > CGB_MAX_ORDER) FAIL(CGB_ERR_ARG, "matrix order out of range (n * n * 8 bytes must be addressable)");
    if (n % c->ranks != 0) {
        /* parallel_cg.c:86-90: "%d is not divisible by %d" */
        snprintf(tl_err, sizeof tl_err, "%lld is not divisible by %d", (long long)n, c->ranks);
        return CGB_ERR_NOT_DIVISIBLE;
    }
    const int64_t m = n / c->ranks;
    CU(cudaSetDevice(c->device));
    if (n != c->cap_n || m != c->cap_m) {
        free_problem(c);
        int rc = alloc_problem(c, n, m);                   /* local */
        rc = cgb_ranks_agree(c, rc);                       /* all ranks allocated, or none goes on */
        if (rc == CGB_OK && c->ranks > 1) {
            rc = exchange_windows(c, n);                   /* collective */
            rc = cgb_ranks_agree(c, rc);
        }
        if (rc != CGB_OK) {
            char keep[sizeof tl_err];
            snprintf(keep, sizeof keep, "%s", tl_err);
            free_problem(c);
            snprintf(tl_err, sizeof tl_err, "%s", keep);
            return rc;
        }
        c->cap_n = n; c->cap_m = m;
    }
    c->n = n; c->m = m;
    return CGB_OK;
}

static int load_host_body(cgb_ctx *c, int64_t n, const double *A_rows, const double *b_rows, const double *x0_full);
int cgb_load_host(cgb_ctx *c, int64_t n, const double *A_rows, const double *b_rows,
                  const double *x0_full)
{
    if (!c || !A_rows || !b_rows || !x0_full) FAIL(CGB_ERR_ARG, "cgb_load_host: null argument");
    int rc = cgb_ensure_problem(c, n);
    if (rc != CGB_OK) return rc;
    /* every rank learns whether every rank's copy arrived: nobody enters the (collective) solve alone */
    return cgb_ranks_agree(c, load_host_body(c, n, A_rows, b_rows, x0_full));
}

static int load_host_body(cgb_ctx *c, int64_t n, const double *A_rows, const double *b_rows, const double *x0_full)
{
    int rc = CGB_OK;
    if (injected_failure(c, "LOAD")) return CGB_ERR_NOMEM;
    const int64_t m = c->m;
    double t0 = now_s();
    CU(cudaMemcpyAsync(c->x_full, x0_full, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CU(cudaMemcpyAsync(c->b, b_rows, (size_t)m * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    /* A: a pinned source is read in place by the DMA engine; a malloc'ed one (the reference's
     * memAllocate(), serialConjugate.c:62-64) goes through the staging ring of cg_stage.c */
    const size_t total = (size_t)m * (size_t)n * sizeof(double);
    {
        cgb_stage_run runs[1];                               /* the contiguous block: one row of `total` bytes */
        const int nr = 1;
        runs[0].dst = c->A; runs[0].src = A_rows; runs[0].width = total; runs[0].rows = 1; runs[0].dpitch = runs[0].spitch = total;
        rc = cgb_stage_upload(c, runs, nr, cgb_host_is_pageable(A_rows), &c->h2d_path);
        if (rc != CGB_OK) return rc;
    }
    CU(cudaStreamSynchronize(c->stream));
    c->h2d_seconds = now_s() - t0;
    c->h2d_bytes = (int64_t)(total + (size_t)(n + m) * sizeof(double));
    c->loaded = 1;
    c->sym_state = 0;
    c->tiles_only = 0;
    return CGB_OK;
}

/* Can this context run the symmetric product on an n x n system (prepare_symv's shape rules)? */
static int symv_shape_ok(const cgb_ctx *c, int64_t n)
{
    const int T = cgk_symv_tile();
    if (n < 1 || n % c->ranks != 0) return 0;
    if (n % T != 0 || (n / c->ranks) % T != 0) return 0;
    return c->ranks == 1 || c->peer_ok;
}

/* A caller that asserts a_ij == a_ji (cgb_options.symmetric = 2) need not send what the
 * symmetric product never reads: only the rank's tiles of the tile list travel (for one rank
 * the 512 x 512 tiles on and above the diagonal, n (n + 512) / 2 values instead of n^2).
 * Strided copies, one per run of adjacent tiles in a tile row. */
static int load_host_symmetric_body(cgb_ctx *c, int64_t n, const double *A_rows, const double *b_rows, const double *x0_full);
int cgb_load_host_symmetric(cgb_ctx *c, int64_t n, const double *A_rows, const double *b_rows,
                            const double *x0_full)
{
    if (!c || !A_rows || !b_rows || !x0_full) FAIL(CGB_ERR_ARG, "cgb_load_host_symmetric: null argument");
    int rc = cgb_ensure_problem(c, n);
    if (rc != CGB_OK) return rc;
    if (!symv_shape_ok(c, n)) {          /* the same answer on every rank */
        snprintf(tl_err, sizeof tl_err, "cgb_load_host_symmetric needs n / ranks to be a multiple of %d and, on several ranks, "
                 "the peer-memory exchange (n = %lld, ranks = %d)", cgk_symv_tile(), (long long)n, c->ranks);
        return CGB_ERR_ARG;
    }
    return cgb_ranks_agree(c, load_host_symmetric_body(c, n, A_rows, b_rows, x0_full));
}

static int load_host_symmetric_body(cgb_ctx *c, int64_t n, const double *A_rows, const double *b_rows, const double *x0_full)
{
    int rc = CGB_OK;
    const int64_t m = c->m, T = cgk_symv_tile();
    const int multi = c->ranks > 1;
    cgk_symplan sp;
    if (multi) cgk_symv_make_plan(&sp, c->rank, c->ranks, m);
    const int64_t cnt = multi ? cgk_symv_tile_list_rank(&sp, c->rank, m, 1, NULL) : cgk_symv_tile_list(n, 1, NULL);
    int *tl = malloc((size_t)cnt * 2 * sizeof(int));
    if (!tl) FAIL(CGB_ERR_NOMEM, "out of host memory for the tile list");
    if (multi) cgk_symv_tile_list_rank(&sp, c->rank, m, 1, tl); else cgk_symv_tile_list(n, 1, tl);
    double t0 = now_s();
    cudaError_t ce = cudaMemcpyAsync(c->x_full, x0_full, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, c->stream);
    if (ce == cudaSuccess) ce = cudaMemcpyAsync(c->b, b_rows, (size_t)m * sizeof(double), cudaMemcpyHostToDevice, c->stream);
    /* one 2-D run per stretch of adjacent tiles in a tile row (cut at 4 MiB of row width) */
    cgb_stage_run *runs = malloc((size_t)(cnt > 0 ? cnt : 1) * sizeof *runs);
    if (!runs) { free(tl); FAIL(CGB_ERR_NOMEM, "out of host memory for the tile runs"); }
    int64_t sent = 0, nr = 0;
    const int64_t max_tiles = ((int64_t)4 << 20) / (T * (int64_t)sizeof(double)) > 0 ? ((int64_t)4 << 20) / (T * (int64_t)sizeof(double)) : 1;
    for (int64_t k = 0; k < cnt; ) {
        const int64_t I = tl[2 * k] & 0xffffff, J0 = tl[2 * k + 1];
        int64_t q = k + 1;
        while (q < cnt && q - k < max_tiles && (tl[2 * q] & 0xffffff) == I && tl[2 * q + 1] == J0 + (q - k)) ++q;
        const size_t off = (size_t)(I * T) * (size_t)n + (size_t)(J0 * T);
        runs[nr].dst = c->A + off; runs[nr].src = A_rows + off;
        runs[nr].width = (size_t)((q - k) * T) * sizeof(double); runs[nr].rows = (size_t)T;
        runs[nr].dpitch = runs[nr].spitch = (size_t)n * sizeof(double);
        ++nr;
        sent += (q - k) * T * T;
        k = q;
    }
    free(tl);
    if (ce != cudaSuccess) {
        (void)cudaGetLastError();
        free(runs);
        snprintf(tl_err, sizeof tl_err, "cgb_load_host_symmetric: %s", cudaGetErrorString(ce));
        return CGB_ERR_CUDA;
    }
    rc = cgb_stage_upload(c, runs, nr, cgb_host_is_pageable(A_rows), &c->h2d_path);
    free(runs);
    if (rc != CGB_OK) return rc;
    CU(cudaStreamSynchronize(c->stream));
    c->h2d_seconds = now_s() - t0;
    c->h2d_bytes = (sent + n + m) * (int64_t)sizeof(double);
    c->loaded = 1;
    c->sym_state = 1;          /* asserted by the caller, as cgb_options.symmetric = 2 */
    c->tiles_only = 1;
    return CGB_OK;
}

/* rows [row0, row0 + m) of the generated matrix in the family the flags name (spd_gen.cuh) */
static int generate_rows(double *A, int64_t n, int64_t row0, int64_t m, uint64_t seed, unsigned flags, cudaStream_t st)
{
    if (flags & CGB_GEN_GRAM) return cgk_generate_gram_A(A, n, row0, m, seed, flags, st);
    if (flags & CGB_GEN_SPECTRUM) {
        double inv = 0.0, s4 = 0.0;
        cgb_gen_spectrum_scalars(cgb_gen_key(seed), (uint64_t)n, flags, &inv, &s4);      /* two O(n) sums, host, index order */
        return cgk_generate_spectrum_A(A, n, row0, m, seed, flags, inv, s4, st);
    }
    return cgk_generate_A(A, n, row0, m, seed, flags, st);
}
static int generate_body(cgb_ctx *c, int64_t n, uint64_t seed, unsigned flags);
int cgb_generate(cgb_ctx *c, int64_t n, uint64_t seed, unsigned flags)
{
    if (!c) FAIL(CGB_ERR_ARG, "cgb_generate: null context");
    if ((flags & CGB_GEN_GRAM) && (flags & CGB_GEN_SPECTRUM)) FAIL(CGB_ERR_ARG, "cgb_generate: CGB_GEN_GRAM and CGB_GEN_SPECTRUM are different families, pick one");
    int rc = cgb_ensure_problem(c, n);
    if (rc != CGB_OK) return rc;
    return cgb_ranks_agree(c, generate_body(c, n, seed, flags));
}

static int generate_body(cgb_ctx *c, int64_t n, uint64_t seed, unsigned flags)
{
    const int64_t row0 = (int64_t)c->rank * c->m;
    KL(generate_rows(c->A, n, row0, c->m, seed, flags, c->stream));
    KL(cgk_generate_b(c->b, n, row0, c->m, seed, flags, c->stream));
    CU(cudaMemsetAsync(c->x_full, 0, (size_t)n * sizeof(double), c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->h2d_seconds = 0.0;
    c->h2d_bytes = 0;
    c->h2d_path = CGB_H2D_NONE;
    c->loaded = 1;
    c->tiles_only = 0;
    c->sym_state = 1;          /* 0.5 (u_ij + u_ji): IEEE addition commutes, exactly symmetric */
    return CGB_OK;
}

int cgb_set_x0(cgb_ctx *c, const double *x0_full)
{
    if (!c || !x0_full) FAIL(CGB_ERR_ARG, "cgb_set_x0: null argument");
    if (!c->loaded) FAIL(CGB_ERR_STATE, "cgb_set_x0: no problem loaded");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(c->x_full, x0_full, (size_t)c->n * sizeof(double), cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return CGB_OK;
}

int cgb_read_x(cgb_ctx *c, double *x_full)
{
    if (!c || !x_full) FAIL(CGB_ERR_ARG, "cgb_read_x: null argument");
    if (!c->loaded) FAIL(CGB_ERR_STATE, "cgb_read_x: no problem loaded");
    CU(cudaSetDevice(c->device));
    double t0 = now_s();
    CU(cudaMemcpyAsync(x_full, c->x_full, (size_t)c->n * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    c->d2h_seconds = now_s() - t0;
    return CGB_OK;
}

int cgb_debug_read(cgb_ctx *c, int what, int64_t offset, int64_t count, double *dst)
{
    if (!c || !dst || offset < 0 || count < 0) FAIL(CGB_ERR_ARG, "cgb_debug_read: bad argument");
    if (!c->loaded) FAIL(CGB_ERR_STATE, "cgb_debug_read: no problem loaded");
    const double *src = NULL; int64_t lim = 0;
    switch (what) {
    case 0: src = c->A; lim = c->m * c->n; break;
    case 1: src = c->b; lim = c->m; break;
    case 2: src = c->x_full; lim = c->n; break;
    case 3: src = c->r; lim = c->m; break;
    case 4: src = c->p_full; lim = c->n; break;
    case 5: src = c->Ap; lim = c->m; break;
    default: FAIL(CGB_ERR_ARG, "cgb_debug_read: unknown array");
    }
    if (offset + count > lim) FAIL(CGB_ERR_ARG, "cgb_debug_read: range out of bounds");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(dst, src + offset, (size_t)count * sizeof(double), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return CGB_OK;
}

int64_t cgb_debug_probe_read(cgb_ctx *c, unsigned long long *dst, int64_t max_words)
{
    if (!c || !dst || max_words < 0) { cgb_set_error("%s%s (cg_solver.c:%d)", "cgb_debug_probe_read: bad argument", "", __LINE__); return CGB_ERR_ARG; }
    if (!c->probe) { cgb_set_error("%s%s (cg_solver.c:%d)", "no probe data: solve on several ranks with CGB_PROBE=1 in the environment", "", __LINE__); return CGB_ERR_STATE; }
    const int64_t w = max_words < CGK_PROBE_WORDS ? max_words : CGK_PROBE_WORDS;
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(dst, c->probe, (size_t)w * sizeof(unsigned long long), cudaMemcpyDeviceToHost, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return w;
}

/* Guard-band self test: one double stored at element `offset` of an array, where offset may lie up
 * to 32 elements outside it.  Refused unless the context was created with CGB_DEBUG_GUARDS=1. */
int cgb_debug_poke(cgb_ctx *c, int what, int64_t offset, double value)
{
    if (!c) FAIL(CGB_ERR_ARG, "cgb_debug_poke: null context");
    if (!c->guards_on) FAIL(CGB_ERR_STATE, "cgb_debug_poke: only with CGB_DEBUG_GUARDS=1");
    if (!c->loaded) FAIL(CGB_ERR_STATE, "cgb_debug_poke: no problem loaded");
    double *dst = NULL; int64_t lim = 0;
    switch (what) {
    case 1: dst = c->b; lim = c->m; break;
    case 2: dst = c->x_full; lim = c->n; break;
    case 3: dst = c->r; lim = c->m; break;
    case 5: dst = c->Ap; lim = c->m; break;
    default: FAIL(CGB_ERR_ARG, "cgb_debug_poke: unknown array");
    }
    if (offset < -32 || offset >= lim + 32) FAIL(CGB_ERR_ARG, "cgb_debug_poke: offset beyond the guard band");
    CU(cudaSetDevice(c->device));
    CU(cudaMemcpyAsync(dst + offset, &value, sizeof value, cudaMemcpyHostToDevice, c->stream));
    CU(cudaStreamSynchronize(c->stream));
    return CGB_OK;
}

/* ------------------------------------------------------------------ solver */
static void plan_request_from_options(cgk_gemv_plan *p, const cgb_options *o)
{
    memset(p, 0, sizeof *p);
    p->keep_pct = o->gemv_l2_keep_pct;
    p->variant = o->gemv_variant; p->rows_per_cta = o->gemv_rows; p->unroll = o->gemv_unroll;
    p->threads = o->gemv_threads; p->vec = o->gemv_vec; p->stages = o->gemv_stages;
    p->tile_cols = o->gemv_tile_cols;
}

/* cgb_options.symmetric: verify a_ij == a_ji once per loaded matrix, allocate the tile list and
 * the two partial-sum workspaces, and switch the loop's matVec to cg_symv.cu. */
static int prepare_symv(cgb_ctx *c, const cgb_options *o)
{
    const int T = cgk_symv_tile();
    const int multi = c->ranks > 1;
    if (multi && c->comm_mode != CGB_COMM_PEER)
        FAIL(CGB_ERR_STATE, "cgb_options.symmetric on several ranks needs the peer-memory exchange (CGB_COMM_PEER)");
    if (c->n % T != 0 || c->m % T != 0) {
        snprintf(tl_err, sizeof tl_err, "cgb_options.symmetric needs n (and n / ranks) to be a multiple of %d (n = %lld, ranks = %d)",
                 T, (long long)c->n, c->ranks);
        return CGB_ERR_ARG;
    }
    cudaStream_t st = c->stream;
    if (multi && c->sym_state < 0)
        FAIL(CGB_ERR_ARG, "cgb_options.symmetric: the matrix is not exactly symmetric (a_ij != a_ji somewhere)");
    /* several ranks, loaded matrix, symmetric = 1: a row block cannot verify a_ij == a_ji by itself -- the
     * ranks do it together (verify_symmetry_ranks, collective) once cgb_solve has let them agree to go on */
    c->sym_pending = multi && c->sym_state == 0 && o->symmetric != 2;
    if (multi && o->symmetric == 2 && c->sym_state == 0) c->sym_state = 1;
    if (c->sym_state == 0 && o->symmetric == 2) c->sym_state = 1;
    if (c->sym_state == 0 && !multi) {
        int *d_flag = NULL, h_flag = 1;
        CU(cudaMalloc((void **)&d_flag, sizeof(int)));
        int e = cgk_symmetry_check(c->A, c->n, d_flag, st);
        cudaError_t ce = cudaMemcpyAsync(&h_flag, d_flag, sizeof(int), cudaMemcpyDeviceToHost, st);
        if (ce == cudaSuccess) ce = cudaStreamSynchronize(st);
        cudaFree(d_flag);
        if (e != 0 || ce != cudaSuccess) { (void)cudaGetLastError(); FAIL(CGB_ERR_CUDA, "symmetry check failed to run"); }
        c->sym_state = h_flag ? -1 : 1;
    }
    if (c->sym_state < 0)
        FAIL(CGB_ERR_ARG, "cgb_options.symmetric: the matrix is not exactly symmetric (a_ij != a_ji somewhere)");
    if (!c->sy_tiles) {
        if (multi) cgk_symv_make_plan(&c->sy_plan, c->rank, c->ranks, c->m);
        const int64_t ntl = multi ? c->m / T : c->n / T;      /* tile rows of colpart; rowpart has n / T x m */
        const int64_t whole = multi ? cgk_symv_tile_list_rank(&c->sy_plan, c->rank, c->m, 1, NULL) : cgk_symv_tile_list(c->n, 1, NULL);
        int split = cgk_symv_pick_split(whole, c->sm_count);
        const char *es = getenv("CGB_SYMV_SPLIT");
        if (es && (atoi(es) == 1 || atoi(es) == 2 || atoi(es) == 4)) split = atoi(es);
        c->symv_split = split;
        const int64_t cnt = whole * split;
        int *h = malloc((size_t)cnt * 2 * sizeof(int));
        if (!h) FAIL(CGB_ERR_NOMEM, "out of host memory for the tile list");
        if (multi) cgk_symv_tile_list_rank(&c->sy_plan, c->rank, c->m, split, h); else cgk_symv_tile_list(c->n, split, h);
        cudaError_t ce = galloc(c, &c->sy_tiles, (size_t)cnt * 2 * sizeof(int), "symv tile list");
        if (ce == cudaSuccess) ce = cudaMemcpy(c->sy_tiles, h, (size_t)cnt * 2 * sizeof(int), cudaMemcpyHostToDevice);
        free(h);
        if (ce == cudaSuccess) ce = galloc(c, (void **)&c->sy_rowpart, (size_t)(c->n / T) * (size_t)c->m * sizeof(double), "symv row partials");
        if (ce == cudaSuccess) ce = galloc(c, (void **)&c->sy_colpart, (size_t)split * (size_t)ntl * (size_t)c->n * sizeof(double), "symv column partials");
        if (ce != cudaSuccess) {
            (void)cudaGetLastError();
            gfree(c, c->sy_tiles); gfree(c, c->sy_rowpart); gfree(c, c->sy_colpart);
            c->sy_tiles = NULL; c->sy_rowpart = c->sy_colpart = NULL;
            FAIL(CGB_ERR_NOMEM, "can't allocate memory for the symmetric-product workspaces");
        }
        c->sy_ntiles = cnt;
    }
    c->symv_ru = (o->gemv_unroll == 1 || o->gemv_unroll == 2) ? o->gemv_unroll : 4;   /* 4 rows in flight measured best */
    c->use_symv = 1;
    return CGB_OK;
}

/* The matVec of one iteration with whatever exchange it carries: y = A_blk v into c->Ap and the
 * (all-reduced, or pushed) dot w.y, where v is the gathered vector and w the rank's slice of it.
 * Classic CG: v = p (parallel_cg.c:290-294).  Fused-reductions CG: v = r. */
static int enqueue_product(cgb_ctx *c, cudaEvent_t ev_a, cudaEvent_t ev_b)
{
    const int multi = c->ranks > 1;
    double *v_loc = c->p_full + (size_t)c->rank * (size_t)c->m;
    cudaStream_t st = c->stream;
    if (c->comm_mode == CGB_COMM_PEER) {
        /* the collectives live inside the kernels: slices and dot partials are pushed into the
         * peers' windows over NVLink, consumers spin on their own mailbox */
        if (ev_a) CU(cudaEventRecord(ev_a, st));
        if (c->use_symv) {
            KL(cgk_symv_comm(c->A, c->p_full, v_loc, c->Ap, c->sy_tiles, c->sy_ntiles, c->sy_rowpart, c->sy_colpart,
                             c->partials_gemv, &c->sc->pAp_local, c->sc, c->m, c->n, c->symv_ru, c->symv_split,
                             &c->sy_plan, &c->peers, st));
            c->launches += 2;
        } else
        KL(cgk_gemv_comm(&c->plan, c->A, c->p_full, v_loc, c->Ap, c->partials_gemv, &c->sc->pAp_local,
                         c->sc, c->m, c->n, &c->peers, st));                     /* :290-294 */
        if (ev_b) CU(cudaEventRecord(ev_b, st));
        c->launches += 1;
        return CGB_OK;
    }
    if (multi)                                                   /* MPI_Allgather(p), :290-291 */
        NC(ncclAllGather(v_loc, c->p_full, (size_t)c->m, ncclDouble, c->comm, st));
    if (ev_a) CU(cudaEventRecord(ev_a, st));
    if (c->use_symv) {                                           /* upper tiles only, two kernels */
        KL(cgk_symv(c->A, c->p_full, v_loc, c->Ap, c->sy_tiles, c->sy_ntiles, c->sy_rowpart, c->sy_colpart,
                    c->partials_gemv, c->n_dot ? NULL : &c->sc->pAp, c->sc, c->n, c->symv_ru, c->symv_split, st));
        c->launches += 1;
    } else
    KL(cgk_gemv(&c->plan, c->A, c->p_full, v_loc, c->Ap, c->partials_gemv,
                multi ? &c->sc->pAp_local : (c->n_dot ? NULL : &c->sc->pAp), c->sc, c->m, c->n, st));   /* :292-293 */
    if (ev_b) CU(cudaEventRecord(ev_b, st));
    if (multi)                                                   /* MPI_Allreduce(alpha), :294 */
        NC(ncclAllReduce(&c->sc->pAp_local, &c->sc->pAp, 1, ncclDouble, ncclSum, c->comm, st));
    c->launches += 1;
    return CGB_OK;
}

/* One CG iteration on the solver stream; `cur` is the parity of the loop index
 * (selects which rr[] slot is rho).  Mirrors parallel_cg.c:290-323. */
static int enqueue_iteration_body(cgb_ctx *c, const cgb_options *o, int64_t max_iter, int cur,
                                  cudaEvent_t ev_a, cudaEvent_t ev_b);
static int enqueue_iteration(cgb_ctx *c, const cgb_options *o, int64_t max_iter, int cur,
                             cudaEvent_t ev_a, cudaEvent_t ev_b)
{
    cgk_set_pdl(c->pdl);              /* one rank: the iteration's kernels overlap their launches */
    const int rc = enqueue_iteration_body(c, o, max_iter, cur, ev_a, ev_b);
    cgk_set_pdl(0);
    return rc;
}

static int enqueue_iteration_body(cgb_ctx *c, const cgb_options *o, int64_t max_iter, int cur,
                                  cudaEvent_t ev_a, cudaEvent_t ev_b)
{
    const int multi = c->ranks > 1;
    double *p_loc = c->p_full + (size_t)c->rank * (size_t)c->m;
    double *x_loc = c->x_full + (size_t)c->rank * (size_t)c->m;
    cudaStream_t st = c->stream;
    const cgk_peers *peers = c->comm_mode == CGB_COMM_PEER ? &c->peers : NULL;
    if (o->cg_variant == CGB_CG_FUSED_REDUCTIONS) {
        /* [step k: bookkeeping of k-1, updates of k, new r pushed] [w = A r, delta = r.w] */
        if (c->defer_rr)      /* one rank: no ticket phase, parity-slot state (cgk_cg2_step_one) */
            KL(cgk_cg2_step_one(x_loc, p_loc, c->r, c->svec, c->Ap, c->partials_vec, c->sc, cur, o->eps, o->ignore_stop,
                                o->guard, max_iter, c->m, c->partials_gemv, c->n_dot, st));
        else
        KL(cgk_cg2_step(x_loc, p_loc, c->r, c->svec, c->Ap, c->partials_vec, c->sc, o->eps, o->ignore_stop,
                        o->guard, max_iter, c->m, peers, st));
        c->launches += 1;
        return enqueue_product(c, ev_a, ev_b);
    }
    int rc = enqueue_product(c, ev_a, ev_b);
    if (rc != CGB_OK) return rc;
    if (peers && c->use_xrp) {
        /* both vector kernels and the all-reduce between them in one launch (cgk_xrp_update) */
        KL(cgk_xrp_update(x_loc, c->r, p_loc, c->Ap, c->partials_vec, c->sc, cur, o->guard, o->eps, o->ignore_stop,
                          max_iter, c->m, peers, st));                                     /* :298-323 */
        c->launches += 1;
        return CGB_OK;
    }
    if (peers) {
        KL(cgk_xr_update(x_loc, c->r, p_loc, c->Ap, c->partials_vec, &c->sc->rr_local, c->sc, cur,
                         o->guard, c->m, peers, NULL, 0, st));                             /* :298-313 */
        KL(cgk_p_update(p_loc, c->r, c->sc, cur, o->eps, o->ignore_stop, max_iter, c->m, peers, NULL, 0, st)); /* :314-323 */
        c->launches += 2;
        return CGB_OK;
    }
    KL(cgk_xr_update(x_loc, c->r, p_loc, c->Ap, c->partials_vec,
                     multi ? &c->sc->rr_local : (c->defer_rr ? NULL : &c->sc->rr[cur ^ 1]), c->sc, cur, o->guard, c->m, NULL,
                     c->partials_gemv, c->n_dot, st));                                     /* :298-309 */
    if (multi)                                                   /* MPI_Allreduce(beta), :313 */
        NC(ncclAllReduce(&c->sc->rr_local, &c->sc->rr[cur ^ 1], 1, ncclDouble, ncclSum, c->comm, st));
    KL(cgk_p_update(p_loc, c->r, c->sc, cur, o->eps, o->ignore_stop, max_iter, c->m, NULL,
                    c->defer_rr ? c->partials_vec : NULL, c->defer_rr ? cgk_vec_grid(c->m) : 0, st)); /* :314-323 */
    c->launches += 2;
    return CGB_OK;
}

static int build_graph(cgb_ctx *c, const cgb_options *o, int64_t max_iter, int chunk)
{
    struct cgb_graph_key key;
    memset(&key, 0, sizeof key);
    key.n = c->n; key.m = c->m; key.max_iter = max_iter; key.eps = o->eps;
    key.ignore_stop = o->ignore_stop; key.chunk = chunk; key.plan = c->plan; key.guard = o->guard;
    key.A = c->A; key.win = c->win; key.comm = c->comm_mode;
    key.symv = c->use_symv; key.symv_ru = c->symv_ru; key.symv_split = c->symv_split;
    key.cg_variant = o->cg_variant; key.n_dot = c->n_dot; key.defer_rr = c->defer_rr; key.pdl = c->pdl;
    key.xrp = c->use_xrp;
    if (c->graph_exec && memcmp(&key, &c->graph_key, sizeof key) == 0) return CGB_OK;
    drop_graph(c);
    const int64_t before = c->launches;
    CU(cudaStreamBeginCapture(c->stream, cudaStreamCaptureModeThreadLocal));
    int rc = CGB_OK;
    for (int k = 0; k < chunk && rc == CGB_OK; ++k)
        rc = enqueue_iteration(c, o, max_iter, k & 1, NULL, NULL);
    cudaError_t e = cudaStreamEndCapture(c->stream, &c->graph);
    c->graph_launches = c->launches - before;          /* kernels of this library in one replay */
    c->launches = before;
    if (rc != CGB_OK) { if (c->graph) { cudaGraphDestroy(c->graph); c->graph = NULL; } return rc; }
    CU(e);
    CU(cudaGraphInstantiate(&c->graph_exec, c->graph, 0));
    c->graph_key = key;
    return CGB_OK;
}

static int pick_chunk(const cgb_ctx *c, const cgb_options *o)
{
    int chunk = o->check_every;
    if (chunk <= 0) {
        /* keep the host poll (~20 us) under ~10 % of a replay: estimate the
         * iteration from the GEMV's bytes at ~6 TB/s plus ~6 us of launch gaps */
        double t_iter = 8.0 * (double)c->m * (double)c->n / 6.0e12 + 6.0e-6;
        chunk = (int)ceil(200.0e-6 / t_iter);
        if (chunk > 64) chunk = 64;
    }
    if (chunk < 2) chunk = 2;
    if (chunk & 1) ++chunk;
    return chunk;
}

static int ensure_gemv_events(cgb_ctx *c, int pairs)
{
    if (pairs * 2 <= c->n_gemv_ev) return CGB_OK;
    cudaEvent_t *ne = realloc(c->gemv_ev, (size_t)pairs * 2 * sizeof *ne);
    if (!ne) FAIL(CGB_ERR_NOMEM, "out of host memory for events");
    c->gemv_ev = ne;
    while (c->n_gemv_ev < pairs * 2) {
        CU(cudaEventCreate(&c->gemv_ev[c->n_gemv_ev]));
        c->n_gemv_ev++;
    }
    return CGB_OK;
}

/* cgb_options.symmetric = 1 on a matrix loaded on several ranks: is a_ij == a_ji bit for bit?  Collective.
 * Each rank compares its diagonal block exactly and fingerprints its G column blocks (cgk_symmetry_hash);
 * one all-gather of G + 1 words per rank; every rank compares S[g][h] with S[h][g] and reaches the same
 * verdict.  Sets c->sym_state to 1 or -1. */
static int verify_symmetry_ranks(cgb_ctx *c)
{
    const int G = c->ranks;
    cudaStream_t st = c->stream;
    unsigned long long *d = NULL, h[CGK_MAX_RANKS * (CGK_MAX_RANKS + 2)];
    int *d_flag = NULL, h_flag = 1;
    if (G > CGK_MAX_RANKS) FAIL(CGB_ERR_ARG, "cgb_options.symmetric = 1 on more ranks than the peer exchange supports");
    CU(cudaMalloc((void **)&d, (size_t)(G + 1) * (size_t)(G + 2) * sizeof *d));
    unsigned long long *mine = d + (size_t)G * (size_t)(G + 2);           /* send buffer: G sums, NaN flag, diagonal verdict */
    d_flag = (int *)(mine + G + 1);
    int rc = CGB_OK;
    if (cgk_symmetry_hash(c->A, c->n, (int64_t)c->rank * c->m, c->m, G, mine, st) != 0 ||
        cudaMemsetAsync(d_flag, 0, sizeof(unsigned long long), st) != cudaSuccess ||
        cgk_symmetry_check_ld(c->A + (size_t)c->rank * (size_t)c->m, c->m, c->n, d_flag, st) != 0) rc = CGB_ERR_CUDA;
    if (rc == CGB_OK && ncclAllGather(mine, d, (size_t)(G + 2), ncclUint64, c->comm, st) != ncclSuccess) { rc = CGB_ERR_NCCL; c->broken = 1; }
    if (rc == CGB_OK && (cudaMemcpyAsync(h, d, (size_t)G * (size_t)(G + 2) * sizeof *h, cudaMemcpyDeviceToHost, st) != cudaSuccess ||
                         cudaStreamSynchronize(st) != cudaSuccess)) rc = CGB_ERR_CUDA;
    cudaFree(d);
    (void)h_flag;
    if (rc != CGB_OK) { (void)cudaGetLastError(); FAIL(rc, "the cross-rank symmetry check failed to run"); }
    int sym = 1;
    for (int g = 0; g < G; ++g) {
        const unsigned long long *sg = h + (size_t)g * (size_t)(G + 2);
        if (sg[G] != 0ull || (sg[G + 1] & 0xffffffffull) != 0ull) sym = 0;            /* a NaN, or its diagonal block differs */
        for (int q = 0; q < G; ++q)
            if (sg[q] != h[(size_t)q * (size_t)(G + 2) + (size_t)g]) sym = 0;
    }
    c->sym_state = sym ? 1 : -1;
    return CGB_OK;
}

/* What cgb_solve settles on this rank alone before the first collective step: options, transport,
 * launch plan, workspaces of the symmetric product, the captured graph.  Anything here may fail on
 * one rank only (device memory), which is why cgb_solve lets the ranks agree on the outcome. */
typedef struct {
    cgb_options o;
    int64_t max_iter;
    const cgk_peers *peers;
    cgk_gemv_plan req;
    int use_graph, chunk, timed_pairs;
} solve_prep;

static int solve_prepare(cgb_ctx *c, const cgb_options *opt, solve_prep *sp)
{
    if (injected_failure(c, "PREPARE")) return CGB_ERR_NOMEM;
    cgb_options o;
    if (opt) o = *opt; else cgb_default_options(&o);
    if (!(o.eps > 0.0)) o.eps = CGB_EPSILON;
    const int64_t max_iter = o.max_iter > 0 ? o.max_iter : c->n;      /* parallel_cg.c:288 */
    const int multi = c->ranks > 1;
    cudaStream_t st = c->stream;
    CU(cudaSetDevice(c->device));

    /* transport of the per-iteration exchange (every rank passes the same option) */
    if (o.cg_variant != CGB_CG_CLASSIC && o.cg_variant != CGB_CG_FUSED_REDUCTIONS) FAIL(CGB_ERR_ARG, "unknown cg_variant");
    c->comm_mode = 0;
    if (multi) {
        if (o.comm == CGB_COMM_PEER && !c->peer_ok) {
            snprintf(tl_err, sizeof tl_err, "CGB_COMM_PEER requested but the peer windows are not mapped: %s", c->peer_why);
            return CGB_ERR_STATE;
        }
        c->comm_mode = (o.comm == CGB_COMM_NCCL || !c->peer_ok) ? CGB_COMM_NCCL : CGB_COMM_PEER;
    }
    if (c->comm_mode == CGB_COMM_PEER) {
        /* CGB_PROBE=1: the exchange kernels leave per-CTA time stamps (cg_kernels.h, cgb_debug_probe_read) */
        const char *e = getenv("CGB_PROBE");
        const int want = e && *e && *e != '0';
        if (want && !c->probe) {
            CU(cudaMalloc((void **)&c->probe, (size_t)CGK_PROBE_WORDS * sizeof(unsigned long long)));
            CU(cudaMemsetAsync(c->probe, 0, (size_t)CGK_PROBE_WORDS * sizeof(unsigned long long), st));
        }
        if ((c->peers.probe != NULL) != (want != 0)) drop_graph(c);     /* the captured kernels carry the pointer */
        c->peers.probe = want ? c->probe : NULL;
    }
    const cgk_peers *peers = c->comm_mode == CGB_COMM_PEER ? &c->peers : NULL;
    if (multi && !peers && o.cg_variant == CGB_CG_FUSED_REDUCTIONS)
        FAIL(CGB_ERR_STATE, "cg_variant = CGB_CG_FUSED_REDUCTIONS on several ranks needs the peer-memory exchange (CGB_COMM_PEER)");

    cgk_gemv_plan req;
    plan_request_from_options(&req, &o);
    req.l2_bytes = c->l2_bytes;
    c->plan = req;
    if (peers) KL(cgk_gemv_make_plan_comm(&c->plan, c->m, c->n, c->A, c->p_full, c->sm_count));
    else KL(cgk_gemv_make_plan(&c->plan, c->m, c->n, c->A, c->p_full, c->sm_count));
    if ((int64_t)c->plan.grid > cgk_gemv_max_ctas(c->m)) FAIL(CGB_ERR_STATE, "gemv grid exceeds workspace");
    c->use_symv = 0;
    if (c->tiles_only && !o.symmetric)
        FAIL(CGB_ERR_STATE, "the matrix came through cgb_load_host_symmetric: only the symmetric product's tiles are on the "
                            "device, solve with cgb_options.symmetric set");
    if (o.symmetric) {
        int rc = prepare_symv(c, &o);
        if (rc != CGB_OK) return rc;
    }

    /* one rank, classic recurrence: the two dot products are finished by their consumers
     * (cg_kernels.h, "Deferred reductions"); CGB_DEFER=0 keeps the producers' last-CTA sums */
    c->n_dot = 0; c->defer_rr = 0; c->pdl = 0;
    {
        const char *e = getenv("CGB_PDL");
        /* one rank; or the peer exchange (all kernels ours) where an iteration is short (measured
         * on 2 GPUs: +5 % at n = 4096, nothing to -2 % at n = 16384); not between NCCL's kernels */
        const int short_iter = 8.0 * (double)c->m * (double)c->n <= 3.0 * (double)c->l2_bytes;
        if ((!multi || (c->comm_mode == CGB_COMM_PEER && short_iter)) && !(e && *e == '0')) c->pdl = 1;
        if (multi && c->comm_mode == CGB_COMM_PEER && e && *e == '2') c->pdl = 1;      /* measurement: at any size */
    }
    {
        const char *e = getenv("CGB_DEFER");
        if (!multi && !(e && *e == '0')) {
            c->defer_rr = 1;
            if (c->use_symv) c->n_dot = cgk_symv_finish_grid(c->n);
            else if (c->plan.variant == CGK_GEMV_LDG && c->plan.grid <= 4096) c->n_dot = c->plan.grid;
        }
    }

    {
        /* peer exchange, classic recurrence: the two vector kernels as ONE launch when CGB_XRP=1.  Off by
         * default on the measurements of round 2 (8 GPUs, graph replay): n = 65536: 1618.1 -> 1619.3 it/s,
         * n = 16384: 15995 -> 15570 it/s -- with programmatic dependent launch the second kernel's launch
         * already hides behind the first, and the merged form makes every CTA wait twice */
        const char *e = getenv("CGB_XRP");
        c->use_xrp = peers && o.cg_variant == CGB_CG_CLASSIC && (e && *e == '1') && cgk_xrp_supported(c->m, c->sm_count);
    }
    const int use_graph = o.use_graph && !o.time_gemv;
    int chunk = 0;
    if (use_graph) {
        chunk = pick_chunk(c, &o);
        int rc = build_graph(c, &o, max_iter, chunk);
        if (rc != CGB_OK && c->pdl) {          /* a driver that cannot capture programmatic edges */
            (void)cudaGetLastError();
            c->pdl = 0;
            rc = build_graph(c, &o, max_iter, chunk);
        }
        if (rc != CGB_OK) return rc;
    }
    int timed_pairs = 0;
    if (o.time_gemv) {
        timed_pairs = max_iter < 1024 ? (int)max_iter : 1024;
        int rc = ensure_gemv_events(c, timed_pairs);
        if (rc != CGB_OK) return rc;
    }
    sp->o = o; sp->max_iter = max_iter; sp->peers = peers; sp->req = req;
    sp->use_graph = use_graph; sp->chunk = chunk; sp->timed_pairs = timed_pairs;
    return CGB_OK;
}

int cgb_solve(cgb_ctx *c, const cgb_options *opt, cgb_stats *stats)
{
    if (!c) FAIL(CGB_ERR_ARG, "cgb_solve: null context");
    if (!c->loaded) FAIL(CGB_ERR_STATE, "cgb_solve: no problem loaded (cgb_load_host / cgb_generate first)");
    if (c->broken) FAIL(CGB_ERR_COMM, "cgb_solve: an earlier collective step of this context failed; destroy it and create a new one");
    solve_prep sp;
    memset(&sp, 0, sizeof sp);
    int prc = cgb_ranks_agree(c, solve_prepare(c, opt, &sp));     /* all ranks enter the solve, or none */
    if (prc != CGB_OK) return prc;
    if (c->sym_pending) {
        c->sym_pending = 0;
        int vrc = verify_symmetry_ranks(c);               /* collective; every rank reaches the same verdict */
        if (vrc != CGB_OK) return vrc;
        if (c->sym_state < 0)
            FAIL(CGB_ERR_ARG, "cgb_options.symmetric: the matrix is not exactly symmetric (a_ij != a_ji somewhere)");
    }
    const cgb_options o = sp.o;
    const int64_t max_iter = sp.max_iter;
    const int multi = c->ranks > 1;
    cudaStream_t st = c->stream;
    const cgk_peers *peers = sp.peers;
    const cgk_gemv_plan req = sp.req;
    const int use_graph = sp.use_graph, chunk = sp.chunk, timed_pairs = sp.timed_pairs;
    double *p_loc = c->p_full + (size_t)c->rank * (size_t)c->m;
    if ((++c->gen & 0x3ffu) == 0u) ++c->gen;              /* tags never collide with zeroed memory */
    c->launches = 0;
    CU(cudaStreamSynchronize(st));
    const double t0 = now_s();
    CU(cudaEventRecord(c->ev[0], st));
    /* ---- init: Ax0, r = p = b - Ax0, rho = r.r   (parallel_cg.c:283-287) */
    if (peers)      /* the MPI_Barrier of parallel_cg.c:278: no rank pushes into a window whose
                     * owner has not entered this solve (late ranks never meet a spin time-out) */
        NC(ncclAllReduce(&c->sc->rr_local, &c->sc->rr_local, 1, ncclDouble, ncclSum, c->comm, st));
    KL(cgk_reset(c->sc, c->gen, st));
    {
        cgk_gemv_plan ip = req;                     /* same request, v = x0 */
        KL(cgk_gemv_make_plan(&ip, c->m, c->n, c->A, c->x_full, c->sm_count));
        KL(cgk_gemv(&ip, c->A, c->x_full, NULL, c->Ap, c->partials_gemv, NULL, c->sc, c->m, c->n, st));
    }
    KL(cgk_init_residual(c->b, c->Ap, c->r, p_loc, c->partials_vec,
                         multi ? &c->sc->rr_local : &c->sc->rr[0], c->sc, c->m, peers, st));
    if (multi && !peers)
        NC(ncclAllReduce(&c->sc->rr_local, &c->sc->rr[0], 1, ncclDouble, ncclSum, c->comm, st));
    KL(cgk_after_init(c->sc, o.guard, peers, st));
    c->launches += 4;
    if (o.cg_variant == CGB_CG_FUSED_REDUCTIONS) {
        /* w0 = A r0 and delta0 = r0.w0 belong to the set-up; s starts at 0 (beta0 = 0) */
        CU(cudaMemsetAsync(c->svec, 0, (size_t)c->m * sizeof(double), st));
        int rc = enqueue_product(c, NULL, NULL);
        if (rc != CGB_OK) return rc;
    }
    CU(cudaEventRecord(c->ev[1], st));

    /* ---- loop */
    int64_t issued = 0;
    int gemv_pairs_used = 0;
    for (;;) {
        if (use_graph) {
            CU(cudaGraphLaunch(c->graph_exec, st));
            issued += chunk;
            c->launches += c->graph_launches;
        } else {
            /* stream mode: two iterations per poll keeps the rr[] parity aligned */
            for (int k = 0; k < 2; ++k) {
                cudaEvent_t ea = NULL, eb = NULL;
                if (o.time_gemv && gemv_pairs_used < timed_pairs) {
                    ea = c->gemv_ev[2 * gemv_pairs_used]; eb = c->gemv_ev[2 * gemv_pairs_used + 1];
                }
                int rc = enqueue_iteration(c, &o, max_iter, (int)((issued + k) & 1), ea, eb);
                if (rc != CGB_OK) return rc;
                if (ea && issued + k < max_iter) gemv_pairs_used++;
            }
            issued += 2;
        }
        /* fixed-iteration mode: no polling (the fused-reductions step k closes iteration k-1) */
        if (o.ignore_stop && issued < max_iter + (o.cg_variant == CGB_CG_FUSED_REDUCTIONS ? 1 : 0)) continue;
        CU(cudaMemcpyAsync(c->h_sc, c->sc, sizeof(cgk_scalars), cudaMemcpyDeviceToHost, st));
        CU(cudaStreamSynchronize(st));
        if (c->h_sc->stop_req != 0 || c->h_sc->comm_error != 0) break;
        if (issued > max_iter + 2 * (int64_t)(chunk > 2 ? chunk : 2) + 4) FAIL(CGB_ERR_STATE, "solver loop did not terminate");
    }
    CU(cudaEventRecord(c->ev[2], st));
    /* every rank returns the full x, as the reference's replicated x (parallel_cg.c:299-303) */
    if (multi)
        NC(ncclAllGather(c->x_full + (size_t)c->rank * (size_t)c->m, c->x_full, (size_t)c->m, ncclDouble, c->comm, st));
    CU(cudaStreamSynchronize(st));
    const double t1 = now_s();
    if (c->h_sc->comm_error != 0) {
        c->broken = 1;                 /* the ranks are no longer in lockstep: no further collectives on this context */
        snprintf(tl_err, sizeof tl_err, "rank %d: a peer rank did not answer within %.0f s during the solve "
                 "(iteration %lld); results are invalid", c->rank, CGK_SPIN_TIMEOUT_NS * 1e-9, (long long)c->h_sc->iters);
        return CGB_ERR_COMM;
    }

    if (stats) {
        float ms_init = 0.f, ms_loop = 0.f;
        CU(cudaEventElapsedTime(&ms_init, c->ev[0], c->ev[1]));
        CU(cudaEventElapsedTime(&ms_loop, c->ev[1], c->ev[2]));
        memset(stats, 0, sizeof *stats);
        stats->n = c->n; stats->m = c->m; stats->rank = c->rank; stats->ranks = c->ranks;
        stats->iterations = c->h_sc->iters;
        stats->break_k = c->h_sc->break_k;
        stats->first_hit_k = c->h_sc->first_hit;
        stats->final_beta = c->h_sc->final_beta;
        stats->final_residual = sqrt(c->h_sc->final_beta);
        stats->rsold0 = c->h_sc->rsold0;
        stats->solve_seconds = t1 - t0;
        stats->loop_ms = ms_loop; stats->init_ms = ms_init;
        stats->h2d_seconds = c->h2d_seconds; stats->d2h_seconds = c->d2h_seconds; stats->h2d_bytes = c->h2d_bytes;
        stats->gemv_bytes = 8.0 * ((double)c->m * (double)c->n + (double)c->n + (double)c->m);
        if (c->use_symv)      /* declared change of algorithmic bytes: the rank's half of its row block
                               * (the upper triangle on one rank), p once, Ap once */
            stats->gemv_bytes = 8.0 * (0.5 * (double)c->m * ((double)c->n + 1.0) + (double)c->n + (double)c->m);
        stats->kernel_launches = c->launches;
        stats->gemv_variant = c->plan.variant; stats->gemv_rows = c->plan.rows_per_cta;
        stats->gemv_unroll = c->plan.unroll; stats->gemv_threads = c->plan.threads;
        stats->gemv_vec = c->plan.vec; stats->gemv_stages = c->plan.stages;
        stats->gemv_tile_cols = c->plan.tile_cols; stats->gemv_grid = c->plan.grid;
        stats->graph_chunk = chunk;
        stats->comm = c->comm_mode;
        if (c->use_symv) { stats->gemv_variant = CGB_GEMV_SYMV; stats->gemv_unroll = c->symv_ru; stats->gemv_grid = (int)c->sy_ntiles; }
        stats->gemv_keep_rows = c->plan.keep_rows;
        stats->exit_reason = c->h_sc->stop_req;
        stats->cg_variant = o.cg_variant;
        stats->pdl = c->pdl;
        stats->h2d_path = c->h2d_path;
        if (gemv_pairs_used > 0) {
            double sum = 0.0, mn = 1e300;
            for (int i = 0; i < gemv_pairs_used; ++i) {
                float ms = 0.f;
                CU(cudaEventElapsedTime(&ms, c->gemv_ev[2 * i], c->gemv_ev[2 * i + 1]));
                sum += ms; if (ms < mn) mn = ms;
            }
            stats->gemv_ms_avg = sum / gemv_pairs_used;
            stats->gemv_ms_min = mn;
            stats->gemv_timed = gemv_pairs_used;
        }
    }
    if (c->guards_on) return cgb_debug_check_guards(c);      /* CGB_DEBUG_GUARDS=1: every solve is checked */
    return CGB_OK;
}

int cgb_conjugrad_rank(cgb_ctx *c, const double *local_A, const double *local_b,
                       double *local_x, int64_t local_row, const cgb_options *opt,
                       cgb_stats *stats)
{
    if (!c || !local_A || !local_b || !local_x || local_row < 1 || local_row > CGB_MAX_ORDER)
        FAIL(CGB_ERR_ARG, "cgb_conjugrad_rank: bad argument");
    const int64_t n = local_row * c->ranks;
    int rc = CGB_OK;
    if (opt && opt->symmetric == 2) {            /* asserted symmetric: the unread half stays on the host */
        rc = cgb_ensure_problem(c, n);
        if (rc != CGB_OK) return rc;
        const int peer = c->ranks == 1 || (opt->comm != CGB_COMM_NCCL && c->peer_ok);
        rc = (peer && symv_shape_ok(c, n)) ? cgb_load_host_symmetric(c, n, local_A, local_b, local_x)
                                           : cgb_load_host(c, n, local_A, local_b, local_x);
    } else rc = cgb_load_host(c, n, local_A, local_b, local_x);
    if (rc != CGB_OK) return rc;
    rc = cgb_solve(c, opt, stats);
    if (rc != CGB_OK) return rc;
    rc = cgb_read_x(c, local_x);
    if (rc == CGB_OK && stats) { stats->h2d_seconds = c->h2d_seconds; stats->d2h_seconds = c->d2h_seconds; stats->h2d_bytes = c->h2d_bytes; }
    return rc;
}

/* ------------------------------------------------------------------ measurement */
int cgb_time_gemv(cgb_ctx *c, const cgb_options *opt, int warmup, int reps, int flush_l2,
                  double *ms_avg, double *ms_min)
{
    if (!c || reps < 1) FAIL(CGB_ERR_ARG, "cgb_time_gemv: bad argument");
    if (!c->loaded) FAIL(CGB_ERR_STATE, "cgb_time_gemv: no problem loaded");
    cgb_options o;
    if (opt) o = *opt; else cgb_default_options(&o);
    CU(cudaSetDevice(c->device));
    cgk_gemv_plan plan;
    plan_request_from_options(&plan, &o);
    KL(cgk_gemv_make_plan(&plan, c->m, c->n, c->A, c->p_full, c->sm_count));
    c->use_symv = 0;
    if (o.symmetric && c->ranks == 1) {
        int rc = prepare_symv(c, &o);
        if (rc != CGB_OK) return rc;
    }
    cudaStream_t st = c->stream;
    const size_t flush_bytes = (size_t)512 << 20;
    if (flush_l2 && !c->flush_buf) {
        CU(cudaMalloc(&c->flush_buf, flush_bytes));
        CU(cudaMemsetAsync(c->flush_buf, 0, flush_bytes, st));
    }
    KL(cgk_reset(c->sc, c->gen, st));
    double *p_loc = c->p_full + (size_t)c->rank * (size_t)c->m;
    for (int i = 0; i < warmup; ++i)
        if (c->use_symv)
            KL(cgk_symv(c->A, c->p_full, p_loc, c->Ap, c->sy_tiles, c->sy_ntiles, c->sy_rowpart, c->sy_colpart,
                        c->partials_gemv, &c->sc->pAp_local, c->sc, c->n, c->symv_ru, c->symv_split, st));
        else
        KL(cgk_gemv(&plan, c->A, c->p_full, p_loc, c->Ap, c->partials_gemv, &c->sc->pAp_local, c->sc, c->m, c->n, st));
    double sum = 0.0, mn = 1e300;
    for (int i = 0; i < reps; ++i) {
        if (flush_l2) KL(cgk_l2_flush(c->flush_buf, (int64_t)flush_bytes, st));
        CU(cudaEventRecord(c->ev[0], st));
        if (c->use_symv)
            KL(cgk_symv(c->A, c->p_full, p_loc, c->Ap, c->sy_tiles, c->sy_ntiles, c->sy_rowpart, c->sy_colpart,
                        c->partials_gemv, &c->sc->pAp_local, c->sc, c->n, c->symv_ru, c->symv_split, st));
        else
        KL(cgk_gemv(&plan, c->A, c->p_full, p_loc, c->Ap, c->partials_gemv, &c->sc->pAp_local, c->sc, c->m, c->n, st));
        CU(cudaEventRecord(c->ev[1], st));
        CU(cudaEventSynchronize(c->ev[1]));
        float ms = 0.f;
        CU(cudaEventElapsedTime(&ms, c->ev[0], c->ev[1]));
        sum += ms; if (ms < mn) mn = ms;
    }
    if (ms_avg) *ms_avg = sum / reps;
    if (ms_min) *ms_min = mn;
    c->plan = plan;
    return CGB_OK;
}

/* ------------------------------------------------------------------ single-kernel operators */
static int dev_dup(cgb_ctx *c, double **d, const double *h, int64_t count, cudaStream_t st)
{
    CU(galloc(c, (void **)d, (size_t)(count > 0 ? count : 1) * sizeof(double), "operator scratch"));
    if (h) CU(cudaMemcpyAsync(*d, h, (size_t)count * sizeof(double), cudaMemcpyHostToDevice, st));
    return CGB_OK;
}

int cgb_op_matvec(cgb_ctx *c, const double *A, const double *v, const double *w, double *y,
                  double *dot, int64_t m, int64_t n, const cgb_options *opt)
{
    if (!c || !A || !v || !y || m < 1 || n < 1 || m > CGB_MAX_ORDER || n > CGB_MAX_ORDER) FAIL(CGB_ERR_ARG, "cgb_op_matvec: bad argument");
    cgb_options o;
    if (opt) o = *opt; else cgb_default_options(&o);
    CU(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    double *dA = NULL, *dv = NULL, *dw = NULL, *dy = NULL, *dpart = NULL;
    int rc;
    if ((rc = dev_dup(c, &dA, A, m * n, st)) || (rc = dev_dup(c, &dv, v, n, st)) ||
        (rc = dev_dup(c, &dy, NULL, m, st)) || (rc = dev_dup(c, &dpart, NULL, cgk_gemv_max_ctas(m), st)))
        goto out;
    if (w && (rc = dev_dup(c, &dw, w, m, st))) goto out;
    cgk_gemv_plan plan;
    plan_request_from_options(&plan, &o);
    rc = CGB_ERR_CUDA;
    if (cgk_gemv_make_plan(&plan, m, n, dA, dv, c->sm_count) != 0) { cgb_set_error("%s%s (cg_solver.c:%d)", "no gemv plan for this shape", "", __LINE__); goto out; }
    if (cgk_reset(c->sc, c->gen, st) != 0) goto out;
    if (cgk_gemv(&plan, dA, dv, dw, dy, dpart, &c->sc->pAp, c->sc, m, n, st) != 0) {
        cgb_set_error("%s%s (cg_solver.c:%d)", "gemv launch failed: ", cudaGetErrorString(cudaGetLastError()), __LINE__);
        goto out;
    }
    if (cudaMemcpyAsync(y, dy, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaMemcpyAsync(c->h_sc, c->sc, sizeof(cgk_scalars), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    {
        cudaError_t e = cudaStreamSynchronize(st);
        if (e != cudaSuccess) { cgb_set_error("CUDA error: %s in `%s` (cg_solver.c:%d)", cudaGetErrorString(e), "gemv", __LINE__); goto out; }
    }
    if (dot) *dot = w ? c->h_sc->pAp : 0.0;
    c->plan = plan;
    rc = CGB_OK;
out:
    if (rc == CGB_OK && c->guards_on) rc = cgb_debug_check_guards(c);
    gfree(c, dA); gfree(c, dv); gfree(c, dw); gfree(c, dy); gfree(c, dpart);
    return rc;
}

int cgb_op_init_residual(cgb_ctx *c, const double *b, const double *Ax, double *r, double *p,
                         double *rr, int64_t m)
{
    if (!c || !b || !Ax || !r || !p || m < 1) FAIL(CGB_ERR_ARG, "cgb_op_init_residual: bad argument");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    double *db = NULL, *dAx = NULL, *dr = NULL, *dp = NULL;
    int rc;
    if ((rc = dev_dup(c, &db, b, m, st)) || (rc = dev_dup(c, &dAx, Ax, m, st)) ||
        (rc = dev_dup(c, &dr, NULL, m, st)) || (rc = dev_dup(c, &dp, NULL, m, st)))
        goto out;
    rc = CGB_ERR_CUDA;
    if (cgk_reset(c->sc, c->gen, st) != 0) goto out;
    if (cgk_init_residual(db, dAx, dr, dp, c->partials_vec, &c->sc->rr[0], c->sc, m, NULL, st) != 0) goto out;
    if (cudaMemcpyAsync(r, dr, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaMemcpyAsync(p, dp, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaMemcpyAsync(c->h_sc, c->sc, sizeof(cgk_scalars), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaStreamSynchronize(st) != cudaSuccess) goto out;
    if (rr) *rr = c->h_sc->rr[0];
    rc = CGB_OK;
out:
    if (rc == CGB_ERR_CUDA) cgb_set_error("%s%s (cg_solver.c:%d)", "CUDA error in init_residual op: ", cudaGetErrorString(cudaGetLastError()), __LINE__);
    if (rc == CGB_OK && c->guards_on) rc = cgb_debug_check_guards(c);
    gfree(c, db); gfree(c, dAx); gfree(c, dr); gfree(c, dp);
    return rc;
}

int cgb_op_xr_update(cgb_ctx *c, double rho, double pAp, double *x, double *r, const double *p,
                     const double *Ap, double *rr, int64_t m)
{
    if (!c || !x || !r || !p || !Ap || m < 1) FAIL(CGB_ERR_ARG, "cgb_op_xr_update: bad argument");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    double *dx = NULL, *dr = NULL, *dp = NULL, *dAp = NULL;
    int rc;
    if ((rc = dev_dup(c, &dx, x, m, st)) || (rc = dev_dup(c, &dr, r, m, st)) ||
        (rc = dev_dup(c, &dp, p, m, st)) || (rc = dev_dup(c, &dAp, Ap, m, st)))
        goto out;
    rc = CGB_ERR_CUDA;
    if (cgk_reset(c->sc, c->gen, st) != 0) goto out;
    memset(c->h_sc, 0, sizeof *c->h_sc);
    c->h_sc->rr[0] = rho; c->h_sc->pAp = pAp; c->h_sc->first_hit = -1; c->h_sc->break_k = -1;
    if (cudaMemcpyAsync(c->sc, c->h_sc, sizeof(cgk_scalars), cudaMemcpyHostToDevice, st) != cudaSuccess) goto out;
    if (cgk_xr_update(dx, dr, dp, dAp, c->partials_vec, &c->sc->rr[1], c->sc, 0, 0, m, NULL, NULL, 0, st) != 0) goto out;
    if (cudaMemcpyAsync(x, dx, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaMemcpyAsync(r, dr, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaMemcpyAsync(c->h_sc, c->sc, sizeof(cgk_scalars), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaStreamSynchronize(st) != cudaSuccess) goto out;
    if (rr) *rr = c->h_sc->rr[1];
    rc = CGB_OK;
out:
    if (rc == CGB_ERR_CUDA) cgb_set_error("%s%s (cg_solver.c:%d)", "CUDA error in xr_update op: ", cudaGetErrorString(cudaGetLastError()), __LINE__);
    if (rc == CGB_OK && c->guards_on) rc = cgb_debug_check_guards(c);
    gfree(c, dx); gfree(c, dr); gfree(c, dp); gfree(c, dAp);
    return rc;
}

int cgb_op_p_update(cgb_ctx *c, double rho, double beta, double eps, double *p, const double *r,
                    int *stopped, int64_t m)
{
    if (!c || !p || !r || m < 1) FAIL(CGB_ERR_ARG, "cgb_op_p_update: bad argument");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    double *dp = NULL, *dr = NULL;
    int rc;
    if ((rc = dev_dup(c, &dp, p, m, st)) || (rc = dev_dup(c, &dr, r, m, st))) goto out;
    rc = CGB_ERR_CUDA;
    memset(c->h_sc, 0, sizeof *c->h_sc);
    c->h_sc->rr[0] = rho; c->h_sc->rr[1] = beta; c->h_sc->first_hit = -1; c->h_sc->break_k = -1;
    if (cudaMemcpyAsync(c->sc, c->h_sc, sizeof(cgk_scalars), cudaMemcpyHostToDevice, st) != cudaSuccess) goto out;
    if (cgk_p_update(dp, dr, c->sc, 0, eps, 0, (int64_t)1 << 40, m, NULL, NULL, 0, st) != 0) goto out;
    if (cudaMemcpyAsync(p, dp, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaMemcpyAsync(c->h_sc, c->sc, sizeof(cgk_scalars), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    if (cudaStreamSynchronize(st) != cudaSuccess) goto out;
    if (stopped) *stopped = c->h_sc->stop_req == 1;
    rc = CGB_OK;
out:
    if (rc == CGB_ERR_CUDA) cgb_set_error("%s%s (cg_solver.c:%d)", "CUDA error in p_update op: ", cudaGetErrorString(cudaGetLastError()), __LINE__);
    if (rc == CGB_OK && c->guards_on) rc = cgb_debug_check_guards(c);
    gfree(c, dp); gfree(c, dr);
    return rc;
}

int cgb_op_generate(cgb_ctx *c, int64_t n, int64_t row0, int64_t m, uint64_t seed, unsigned flags,
                    double *A_rows, double *b_rows)
{
    if (!c || n < 1 || n > CGB_MAX_ORDER || m < 1 || row0 < 0 || row0 + m > n) FAIL(CGB_ERR_ARG, "cgb_op_generate: bad argument");
    CU(cudaSetDevice(c->device));
    cudaStream_t st = c->stream;
    double *dA = NULL, *db = NULL;
    int rc = CGB_OK;
    if (A_rows) {
        if ((rc = dev_dup(c, &dA, NULL, m * n, st))) goto out;
        rc = CGB_ERR_CUDA;
        if (generate_rows(dA, n, row0, m, seed, flags, st) != 0) goto out;
        if (cudaMemcpyAsync(A_rows, dA, (size_t)m * (size_t)n * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    }
    if (b_rows) {
        if ((rc = dev_dup(c, &db, NULL, m, st))) goto out;
        rc = CGB_ERR_CUDA;
        if (cgk_generate_b(db, n, row0, m, seed, flags, st) != 0) goto out;
        if (cudaMemcpyAsync(b_rows, db, (size_t)m * sizeof(double), cudaMemcpyDeviceToHost, st) != cudaSuccess) goto out;
    }
    rc = cudaStreamSynchronize(st) == cudaSuccess ? CGB_OK : CGB_ERR_CUDA;
out:
    if (rc == CGB_ERR_CUDA) cgb_set_error("%s%s (cg_solver.c:%d)", "CUDA error in generate op: ", cudaGetErrorString(cudaGetLastError()), __LINE__);
    if (rc == CGB_OK && c->guards_on) rc = cgb_debug_check_guards(c);
    gfree(c, dA); gfree(c, db);
    return rc;
}
