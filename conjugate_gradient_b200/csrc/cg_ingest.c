/* cg_ingest.c -- the reference's text inputs straight onto the GPUs.
 *
 * The reference reads matrixA.txt with one fscanf("%f%*c") per value on rank 0
 * (initialize(), parallel_cg.c:147-168) and then scatters the rows
 * (MPI_Scatter, parallel_cg.c:112-113); point-to-point_cg.c streams each rank's strip as it
 * reads (scatterRow, :183-235).  Both are >95 % of the published wall time.
 * Here every rank finds the bytes of ITS rows in the (memory-mapped) file, moves them through
 * two pinned staging buffers into HBM and lets the GPU tokenise and convert them
 * (cg_ingest.cu); the full matrix never exists on the host.  b and x0 are n values each and
 * use the host parser (cg_io.c).
 *
 * Environment (read at every call; for measurements, tools/ingest_rate.py):
 *   CGB_INGEST_TIMING=1      one line on stderr per load: where the time went
 *   CGB_INGEST_CHUNK_MIB=C   bytes per staging chunk (default 8 MiB, 1..256)
 *   CGB_INGEST_TEAM=T        members of the copy team (default min(8, CPUs / ranks), chunk / 512 KiB at most)
 *   CGB_INGEST_NO_POOL=1     no team: threads started and joined per chunk (cgb_parallel_copy), as in round 1
 */
#define _GNU_SOURCE
#include <cuda_runtime_api.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include "../../include/cgb200.h"
#include "cg_kernels.h"
#include "cg_internal.h"
#include "cg_io.h"

/* Two pinned buffers of one chunk each.  Pinning costs ~0.42 ms per MiB and a
 * one-shot caller -- the CLI -- pays it on every run: 2 x 32 MiB were 27 ms of a 0.19 s ingest of
 * an n = 8192 matrix; 8 MiB pieces still move at the PCIe rate (profiles/r02_h2d_1gpu_sweep.json).
 * CGB_INGEST_CHUNK_MIB (1..256) overrides it for measurements. */
#define CHUNK_BYTES_DEFAULT ((size_t)8 << 20)
#define CHUNK_SLACK 4096
#define EXC_CAP 65536u

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

#define CU(call)                                                                         \
    do {                                                                                 \
        cudaError_t e_ = (call);                                                         \
        if (e_ != cudaSuccess) {                                                         \
            cgb_set_error("CUDA error: %s in `%s` (cg_ingest.c:%d)", cudaGetErrorString(e_), #call, __LINE__); \
            rc = CGB_ERR_CUDA; goto out;                                                 \
        }                                                                                \
    } while (0)

static int io_error(const char *what, const char *path)
{
    cgb_set_error("%s%s (cg_ingest.c:%d)", what, path, __LINE__);
    return CGB_ERR_IO;
}

/* Irregular text (a BOM, "1-2", inf/nan the fast paths refuse): the sequential strtod loop of
 * cg_io.c decides, with the reference's memory footprint (the whole matrix on the host). */
static int load_block_on_host(cgb_ctx *c, const char *fileA, int64_t first, int64_t count)
{
    const int64_t nn = c->n * c->n;
    double *hA = malloc((size_t)nn * sizeof(double));
    if (!hA) { cgb_set_error("%s%s (cg_ingest.c:%d)", "can't allocate memory for vector", "", __LINE__); return CGB_ERR_NOMEM; }
    int64_t got = 0;
    int rc = cgb_read_values(fileA, hA, nn, &got);
    if (rc != CGB_OK) {
        char msg[160];
        snprintf(msg, sizeof msg, "expected %lld values, found %lld in ", (long long)nn, (long long)got);
        free(hA);
        return io_error(msg, fileA);
    }
    cudaError_t e = cudaMemcpyAsync(c->A, hA + first, (size_t)count * sizeof(double), cudaMemcpyHostToDevice, c->stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(c->stream);
    free(hA);
    if (e != cudaSuccess) { cgb_set_error("CUDA error: %s in `%s` (cg_ingest.c:%d)", cudaGetErrorString(e), "upload", __LINE__); return CGB_ERR_CUDA; }
    return CGB_OK;
}

/* Staging memory of one text load: ONE pinned allocation (two chunk buffers) and ONE device allocation (two
 * chunk buffers, the per-CTA token counts and offsets, the exception list, the running state), carved up at
 * 256-byte borders -- nine allocator calls each way were 11 ms of a 50 ms load (CGB_INGEST_TIMING). */
typedef struct {
    unsigned char *pin_base, *dev_base;
    unsigned char *pin[2], *dtxt[2];
    cudaEvent_t ev[2];
    unsigned int *d_counts;
    long long *d_offsets, *d_exc_idx;
    unsigned long long *d_exc_pos;
    cgk_ingest_state *d_state;
} ingest_buffers;

static size_t up256(size_t x) { return (x + 255) & ~(size_t)255; }

static int ingest_buffers_get(ingest_buffers *b, size_t buf_bytes, int64_t max_blocks, cudaStream_t st)
{
    int rc = CGB_OK;
    const size_t chunk = up256(buf_bytes);
    size_t off = 2 * chunk;
    const size_t o_counts = off;  off += up256((size_t)max_blocks * sizeof(unsigned int));
    const size_t o_offsets = off; off += up256((size_t)max_blocks * sizeof(long long));
    const size_t o_exc_pos = off; off += up256((size_t)EXC_CAP * sizeof(unsigned long long));
    const size_t o_exc_idx = off; off += up256((size_t)EXC_CAP * sizeof(long long));
    const size_t o_state = off;   off += up256(sizeof(cgk_ingest_state));
    CU(cudaHostAlloc((void **)&b->pin_base, 2 * chunk, cudaHostAllocDefault));
    CU(cudaMalloc((void **)&b->dev_base, off));
    for (int s = 0; s < 2; ++s) {
        b->pin[s] = b->pin_base + (size_t)s * chunk;
        b->dtxt[s] = b->dev_base + (size_t)s * chunk;
        CU(cudaEventCreateWithFlags(&b->ev[s], cudaEventDisableTiming));
    }
    b->d_counts = (unsigned int *)(b->dev_base + o_counts);
    b->d_offsets = (long long *)(b->dev_base + o_offsets);
    b->d_exc_pos = (unsigned long long *)(b->dev_base + o_exc_pos);
    b->d_exc_idx = (long long *)(b->dev_base + o_exc_idx);
    b->d_state = (cgk_ingest_state *)(b->dev_base + o_state);
    CU(cudaMemsetAsync(b->d_state, 0, sizeof *b->d_state, st));
out:
    return rc;
}

static void ingest_buffers_put(ingest_buffers *b)
{
    if (b->pin_base) cudaFreeHost(b->pin_base);
    cudaFree(b->dev_base);
    for (int s = 0; s < 2; ++s)
        if (b->ev[s]) cudaEventDestroy(b->ev[s]);
    memset(b, 0, sizeof *b);
}

/* The host threads count tokens (cgb_text_locate) while the calling thread pins and allocates. */
typedef struct { const cgb_text *tx; int64_t first, count, total; size_t beg, end; } locate_job;
static void *locate_main(void *arg)
{
    locate_job *j = arg;
    j->total = cgb_text_locate(j->tx, j->first, j->count, &j->beg, &j->end);
    return NULL;
}

static int load_text_body(cgb_ctx *c, int64_t n, const char *fileA, const char *fileB, const char *fileX);
int cgb_load_text(cgb_ctx *c, int64_t n, const char *fileA, const char *fileB, const char *fileX)
{
    if (!c || !fileA || !fileB || !fileX) { cgb_set_error("%s%s (cg_ingest.c:%d)", "cgb_load_text: null argument", "", __LINE__); return CGB_ERR_ARG; }
    int rc = cgb_ensure_problem(c, n);
    if (rc != CGB_OK) return rc;
    /* a rank whose part of the input is bad must not leave the others in the solve: all learn of it here */
    return cgb_ranks_agree(c, load_text_body(c, n, fileA, fileB, fileX));
}

static int load_text_body(cgb_ctx *c, int64_t n, const char *fileA, const char *fileB, const char *fileX)
{
    int rc = CGB_OK;
    const double t0 = now_s();
    const int64_t m = c->m, first = (int64_t)c->rank * m * n, count = m * n;
    cudaStream_t st = c->stream;

    double *hv = NULL;
    ingest_buffers B;
    cgb_copy_pool *pool = NULL;
    cgk_ingest_state h_state;
    cgb_text tx;
    int have_text = 0;
    memset(&tx, 0, sizeof tx);
    memset(&B, 0, sizeof B);

    /* ---- b (the rank's m entries) and x0 (all n): n values each, host parser */
    hv = malloc((size_t)n * sizeof(double));
    if (!hv) { cgb_set_error("%s%s (cg_ingest.c:%d)", "can't allocate memory for vector", "", __LINE__); rc = CGB_ERR_NOMEM; goto out; }
    {
        int64_t got = 0;
        if (cgb_read_values(fileB, hv, n, &got) != CGB_OK) {
            rc = io_error(got == 0 ? "Could not open file or no values in " : "too few values in ", fileB); goto out;
        }
        CU(cudaMemcpyAsync(c->b, hv + (size_t)c->rank * (size_t)m, (size_t)m * sizeof(double), cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
        if (cgb_read_values(fileX, hv, n, &got) != CGB_OK) {
            rc = io_error(got == 0 ? "Could not open file or no values in " : "too few values in ", fileX); goto out;
        }
        CU(cudaMemcpyAsync(c->x_full, hv, (size_t)n * sizeof(double), cudaMemcpyHostToDevice, st));
        CU(cudaStreamSynchronize(st));
    }

    /* ---- A: locate the rank's bytes, stream them, parse on the GPU */
    const double t_vec = now_s();
    double t_loc = t_vec, t_alloc = t_vec;
    if (cgb_text_open(fileA, &tx) != 0) { rc = io_error("Could not open file ", fileA); goto out; }
    have_text = 1;
    size_t CHUNK_BYTES = CHUNK_BYTES_DEFAULT;
    {
        const char *e = getenv("CGB_INGEST_CHUNK_MIB");
        const long v = (e && *e) ? atol(e) : 0;
        if (v >= 1 && v <= 256) CHUNK_BYTES = (size_t)v << 20;
    }
    const size_t buf_bytes = CHUNK_BYTES + CHUNK_SLACK;
    locate_job lj = {&tx, first, count, 0, 0, 0};
    pthread_t lth;
    const int threaded = pthread_create(&lth, NULL, locate_main, &lj) == 0;
    const int rc_buf = ingest_buffers_get(&B, buf_bytes, cgk_ingest_blocks((int64_t)buf_bytes), st);
    t_alloc = now_s();
    if (threaded) pthread_join(lth, NULL); else locate_main(&lj);
    t_loc = now_s();
    const size_t beg = lj.beg, end = lj.end;
    const int64_t total = lj.total;
    if (total < 0) { rc = load_block_on_host(c, fileA, first, count); goto done; }
    /* fewer tokens than values: either a short file or tokens like "1-2" that the sequential
     * strtod loop reads as two values -- that loop is the authority (and words the error) */
    if (total < n * n) { rc = load_block_on_host(c, fileA, first, count); goto done; }
    if (rc_buf != CGB_OK) { rc = rc_buf; goto out; }
    unsigned char **pin = B.pin, **dtxt = B.dtxt;
    cudaEvent_t *ev = B.ev;
    unsigned int *d_counts = B.d_counts;
    long long *d_offsets = B.d_offsets, *d_exc_idx = B.d_exc_idx;
    unsigned long long *d_exc_pos = B.d_exc_pos;
    cgk_ingest_state *d_state = B.d_state;

    {
        /* the copy team spins between chunks: never more members than this rank's share of the CPUs */
        int team = cgb_copy_pool_threads(CHUNK_BYTES);
        const long share = sysconf(_SC_NPROCESSORS_ONLN) / (c->ranks > 0 ? c->ranks : 1);
        if (team > share) team = (int)share;
        { const char *e = getenv("CGB_INGEST_TEAM"); const long v = (e && *e) ? atol(e) : 0; if (v >= 1 && v <= 64) team = (int)v; }
        pool = getenv("CGB_INGEST_NO_POOL") ? NULL : cgb_copy_pool_start(team);
    }
    {
        size_t pos = beg;
        int s = 0, used[2] = {0, 0};
        while (pos < end) {
            size_t cend = end - pos > CHUNK_BYTES ? pos + CHUNK_BYTES : end;
            /* cut at a separator so that no token straddles two chunks (a token is far shorter than the slack) */
            while (cend < end && !cgb_is_sep_byte((unsigned char)tx.p[cend]) && cend - pos < CHUNK_BYTES + CHUNK_SLACK / 2) ++cend;
            if (cend < end && !cgb_is_sep_byte((unsigned char)tx.p[cend])) { rc = load_block_on_host(c, fileA, first, count); goto done; }
            const size_t len = cend - pos;
            const size_t padded = (len + 15) / 16 * 16;
            if (used[s]) CU(cudaEventSynchronize(ev[s]));         /* its previous upload has left the pinned buffer */
            cgb_copy_pool_run(pool, pin[s], tx.p + pos, len);    /* mapped file -> pinned staging buffer */
            memset(pin[s] + len, '\n', padded - len + 64);
            CU(cudaMemcpyAsync(dtxt[s], pin[s], padded + 64, cudaMemcpyHostToDevice, st));
            CU(cudaEventRecord(ev[s], st));
            used[s] = 1;
            int e = cgk_ingest_chunk(dtxt[s], (int64_t)padded, (int64_t)pos, c->A, count, d_counts, d_offsets,
                                     d_state, d_exc_pos, d_exc_idx, EXC_CAP, st);
            if (e != 0) { cgb_set_error("kernel launch error: %s in `%s` (cg_ingest.c:%d)", cgk_error_string(e), "cgk_ingest_chunk", __LINE__); rc = CGB_ERR_CUDA; goto out; }
            pos = cend;
            s ^= 1;
        }
    }
    CU(cudaMemcpyAsync(&h_state, d_state, sizeof h_state, cudaMemcpyDeviceToHost, st));
    CU(cudaStreamSynchronize(st));
    if (h_state.base != count) {
        /* the tokeniser on the device and the one that located the range disagree: do not guess */
        rc = load_block_on_host(c, fileA, first, count); goto done;
    }
    if (h_state.exc_count > EXC_CAP) { rc = load_block_on_host(c, fileA, first, count); goto done; }
    if (h_state.exc_count > 0) {
        /* tokens outside the exact fast path: strtod on the host, scattered back */
        const unsigned int ne = h_state.exc_count;
        unsigned long long *hpos = malloc((size_t)ne * sizeof *hpos);
        long long *hidx = malloc((size_t)ne * sizeof *hidx);
        double *hval = malloc((size_t)ne * sizeof *hval);
        double *dval = NULL;
        int bad = !(hpos && hidx && hval);
        if (!bad) {
            if (cudaMemcpy(hpos, d_exc_pos, (size_t)ne * sizeof *hpos, cudaMemcpyDeviceToHost) != cudaSuccess ||
                cudaMemcpy(hidx, d_exc_idx, (size_t)ne * sizeof *hidx, cudaMemcpyDeviceToHost) != cudaSuccess) bad = 2;
        }
        for (unsigned int i = 0; i < ne && !bad; ++i) {
            const char *p = tx.p + hpos[i], *e = p;
            while (e < tx.p + tx.len && !cgb_is_sep_byte((unsigned char)*e)) ++e;
            if (!cgb_token_to_double(p, e, &hval[i])) bad = 3;   /* not a number as one token: sequential semantics decide */
        }
        if (!bad) {
            if (cudaMalloc((void **)&dval, (size_t)ne * sizeof(double)) != cudaSuccess ||
                cudaMemcpy(dval, hval, (size_t)ne * sizeof(double), cudaMemcpyHostToDevice) != cudaSuccess ||
                cudaMemcpy(d_exc_idx, hidx, (size_t)ne * sizeof *hidx, cudaMemcpyHostToDevice) != cudaSuccess ||
                cgk_scatter(c->A, d_exc_idx, dval, ne, st) != 0 || cudaStreamSynchronize(st) != cudaSuccess) bad = 2;
        }
        cudaFree(dval);
        free(hpos); free(hidx); free(hval);
        if (bad == 3) { rc = load_block_on_host(c, fileA, first, count); goto done; }
        if (bad) { (void)cudaGetLastError(); cgb_set_error("%s%s (cg_ingest.c:%d)", "resolving parser exceptions failed", "", __LINE__); rc = CGB_ERR_CUDA; goto out; }
    }
done:
    if (rc == CGB_OK) {
        c->h2d_seconds = now_s() - t0;
        if (getenv("CGB_INGEST_TIMING"))                        /* where a text load spends its time (tools/ingest_rate.py) */
            fprintf(stderr, "cgb_load_text rank %d: b+x0 %.4f s, buffers %.4f s while locating, locate joined after %.4f s, stream+parse %.4f s, total %.4f s\n",
                    c->rank, t_vec - t0, t_alloc - t_vec, t_loc - t_vec, now_s() - t_loc, c->h2d_seconds);
        c->loaded = 1;
        c->sym_state = 0;
        c->tiles_only = 0;
        c->h2d_bytes = 0;
        c->h2d_path = CGB_H2D_NONE;
    }
out:
    if (have_text) cgb_text_close(&tx);
    free(hv);
    cgb_copy_pool_stop(pool);
    ingest_buffers_put(&B);
    return rc;
}
