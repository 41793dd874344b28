/* cg_internal.h -- private to libcgb200.so (not installed). */
#ifndef CG_INTERNAL_H
#define CG_INTERNAL_H
#include <cuda_runtime_api.h>
#include <nccl.h>
#include <stdint.h>
#include "cg_kernels.h"

struct cgb_graph_key {
    int64_t n, m, max_iter;
    double eps;
    int ignore_stop, chunk;
    cgk_gemv_plan plan;
    const void *A;
    const void *win;        /* peer window the captured kernels point into */
    int comm, guard;
    int symv, symv_ru, symv_split, cg_variant;
    int n_dot, defer_rr, pdl, xrp;
};

/* Largest matrix order the entry points accept: n * n * 8 must stay far inside size_t / int64_t
 * (2^24 squared is 2.2 PB of FP64 -- beyond any GPU set this library runs on). */
#define CGB_MAX_ORDER ((int64_t)1 << 24)

#define CGB_MAX_GUARDS 24

/* pageable host memory -> HBM (cg_stage.c) */
#define CGB_STAGE_MAX_THREADS 32
#define CGB_STAGE_REGISTER_FROM_RANKS 4
typedef struct cgb_stage_run {      /* rows of `width` bytes: host rows `spitch` apart, device rows `dpitch` apart */
    void *dst; const void *src;
    size_t width, rows, dpitch, spitch;
} cgb_stage_run;
#define CGB_GUARD_BYTES 256
#define CGB_GUARD_BYTE 0xA5

struct cgb_ctx {
    int device, rank, ranks, sm_count, has_comm, loaded;
    int64_t l2_bytes;
    char name[64];
    cudaStream_t stream;
    cudaEvent_t ev[3];
    cudaEvent_t *gemv_ev;
    int n_gemv_ev;
    ncclComm_t comm;
    /* problem (device) */
    int64_t n, m, cap_n, cap_m;
    double *A;              /* m x n row block, row-major                       */
    double *b, *r, *Ap;     /* m (fused-reductions CG: r holds the direction p, Ap holds w = A r) */
    double *svec;           /* m: s = A p of the fused-reductions recurrence     */
    double *p_full;         /* n: gathered direction; own slice at rank*m; with ranks > 1 it
                             * lives inside `win`, the window the peers map       */
    double *x_full;         /* n: x0 on entry; own slice updated; gathered at end */
    double *partials_gemv;  /* one per gemv CTA                                 */
    double *partials_vec;   /* one per vector-kernel CTA                        */
    cgk_scalars *sc;        /* device scalars                                   */
    cgk_scalars *h_sc;      /* pinned mirror                                    */
    void *flush_buf;
    cgk_gemv_plan plan;
    cudaGraph_t graph;
    cudaGraphExec_t graph_exec;
    struct cgb_graph_key graph_key;
    int64_t launches;
    double h2d_seconds, d2h_seconds;
    int64_t h2d_bytes;
    int64_t graph_launches;           /* kernels per replay of the captured graph */
    int n_dot, defer_rr;              /* deferred reductions of the current solve (cg_kernels.h) */
    int pdl;                          /* iteration kernels launched with programmatic dependent launch */
    int use_xrp;                      /* peer exchange: xr_update and p_update as one launch (cgk_xrp_update) */
    /* peer-memory exchange (ranks > 1): this rank's window [cgk_mailbox | p_full], the peers'
     * windows as mapped here (same-process peer access or CUDA IPC), and the table the kernels get */
    void *win;
    size_t win_bytes;
    void *peer_win[CGK_MAX_RANKS];
    int peer_ipc[CGK_MAX_RANKS];      /* 1: opened with cudaIpcOpenMemHandle (close it)  */
    int peer_ok;                      /* every rank mapped every window                  */
    cgk_peers peers;
    unsigned long long *probe;        /* CGB_PROBE=1: per-CTA time stamps of the exchange kernels (CGK_PROBE_WORDS) */
    int *d_agree;                     /* 2 ints: the status all-reduce of cgb_ranks_agree */
    int broken;                       /* a collective failed or a peer stopped answering: no further collectives on this context */
    unsigned int gen;                 /* solve generation, in lockstep on all ranks      */
    int comm_mode;                    /* CGB_COMM_* in use by the current / last solve   */
    char peer_why[160];               /* why peer_ok == 0                                */
    /* symmetric product (cg_symv.cu) */
    int sym_state;                    /* 0 not checked, 1 exactly symmetric, -1 not      */
    int sym_pending;                  /* several ranks: the cross-rank check is due at the next solve */
    int tiles_only;                   /* cgb_load_host_symmetric: only the tiles the symmetric product reads are on the device */
    int use_symv, symv_ru, symv_split;
    double *sy_rowpart, *sy_colpart;  /* (n / T) * n each                                */
    void *sy_tiles;
    int64_t sy_ntiles;
    cgk_symplan sy_plan;              /* ranks > 1: which half of the row block this rank reads */
    /* staging ring of cg_stage.c: T threads x 2 pinned slots, one copy stream per thread */
    void *stage_pin;
    size_t stage_pin_bytes;
    cudaStream_t stage_stream[CGB_STAGE_MAX_THREADS];
    cudaEvent_t stage_ev[CGB_STAGE_MAX_THREADS][2];
    int stage_threads_used;
    int h2d_path;                     /* CGB_H2D_* of the last upload */
    /* guard bands (CGB_DEBUG_GUARDS=1 at cgb_create): 256 B of a fixed byte before and after every
     * solver array, verified by cgb_debug_check_guards -- the out-of-bounds-write check of the tests */
    int guards_on, n_guards;
    struct { void *user; size_t bytes; const char *name; } guards[CGB_MAX_GUARDS];
};

void cgb_set_error(const char *fmt, const char *a, const char *b, int line);
/* (Re)allocates the device arrays (and, with ranks > 1, the peer window: collective) for order n. */
int cgb_ensure_problem(cgb_ctx *c, int64_t n);
/* MIN over the ranks of a status: everybody learns whether everybody is fine (cg_solver.c). */
int cgb_ranks_agree(cgb_ctx *c, int rc_local);
/* Unmaps the peers' windows of a context (team tear-down: all unmap before any frees). */
void cgb_unmap_peers(cgb_ctx *c);
/* cg_stage.c */
int  cgb_host_is_pageable(const void *p);
int  cgb_stage_mode(void);
int  cgb_stage_mode_for(int ranks);
int  cgb_stage_upload(cgb_ctx *c, const cgb_stage_run *runs, int64_t nruns, int pageable, int *path_out);
void cgb_stage_release(cgb_ctx *c);
void cgb_stage_plan(size_t total, int ranks, int team, size_t max_chunk, size_t have_ring, int *mode_out, int *threads_out, size_t *chunk_out);

#endif
