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
 */
#define _GNU_SOURCE
#include <cuda_runtime_api.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "../../include/cgb200.h"
#include "cg_kernels.h"
#include "cg_internal.h"
#include "cg_io.h"

#define CHUNK_BYTES ((size_t)32 << 20)
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
    unsigned char *pin[2] = {NULL, NULL}, *dtxt[2] = {NULL, NULL};
    cudaEvent_t ev[2] = {NULL, NULL};
    unsigned int *d_counts = NULL;
    long long *d_offsets = NULL, *d_exc_idx = NULL;
    unsigned long long *d_exc_pos = NULL;
    cgk_ingest_state *d_state = NULL, h_state;
    cgb_text tx;
    int have_text = 0;
    memset(&tx, 0, sizeof tx);

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
    if (cgb_text_open(fileA, &tx) != 0) { rc = io_error("Could not open file ", fileA); goto out; }
    have_text = 1;
    size_t beg = 0, end = 0;
    const int64_t total = cgb_text_locate(&tx, first, count, &beg, &end);
    if (total < 0) { rc = load_block_on_host(c, fileA, first, count); goto done; }
    /* fewer tokens than values: either a short file or tokens like "1-2" that the sequential
     * strtod loop reads as two values -- that loop is the authority (and words the error) */
    if (total < n * n) { rc = load_block_on_host(c, fileA, first, count); goto done; }

    const size_t buf_bytes = CHUNK_BYTES + CHUNK_SLACK;
    const int64_t max_blocks = cgk_ingest_blocks((int64_t)buf_bytes);
    for (int s = 0; s < 2; ++s) {
        CU(cudaHostAlloc((void **)&pin[s], buf_bytes, cudaHostAllocDefault));
        CU(cudaMalloc((void **)&dtxt[s], buf_bytes));
        CU(cudaEventCreateWithFlags(&ev[s], cudaEventDisableTiming));
    }
    CU(cudaMalloc((void **)&d_counts, (size_t)max_blocks * sizeof(unsigned int)));
    CU(cudaMalloc((void **)&d_offsets, (size_t)max_blocks * sizeof(long long)));
    CU(cudaMalloc((void **)&d_exc_pos, (size_t)EXC_CAP * sizeof(unsigned long long)));
    CU(cudaMalloc((void **)&d_exc_idx, (size_t)EXC_CAP * sizeof(long long)));
    CU(cudaMalloc((void **)&d_state, sizeof *d_state));
    CU(cudaMemsetAsync(d_state, 0, sizeof *d_state, st));

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
            cgb_parallel_copy(pin[s], tx.p + pos, len);          /* mapped file -> pinned staging buffer */
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
        c->loaded = 1;
        c->sym_state = 0;
        c->tiles_only = 0;
        c->h2d_bytes = 0;
        c->h2d_path = CGB_H2D_NONE;
    }
out:
    if (have_text) cgb_text_close(&tx);
    free(hv);
    for (int s = 0; s < 2; ++s) {
        if (pin[s]) cudaFreeHost(pin[s]);
        cudaFree(dtxt[s]);
        if (ev[s]) cudaEventDestroy(ev[s]);
    }
    cudaFree(d_counts); cudaFree(d_offsets); cudaFree(d_exc_pos); cudaFree(d_exc_idx); cudaFree(d_state);
    return rc;
}
