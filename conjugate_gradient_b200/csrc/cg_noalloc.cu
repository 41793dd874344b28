/* cg_noalloc.cu -- cgb_matvec_noalloc(): the row-block matVec with the fused dot for callers that
 * may not allocate device memory (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * It is a second door to an existing kernel (cgk_gemv's streaming-load product, the
 * one cgb_op_matvec and the solver use); no existing code path changes.
 *
 * The product needs scratch memory: one partial of the fused dot w.y per CTA, a ticket counter
 * and the stop flag the kernel reads (cgk_scalars).  cgb_op_matvec cudaMallocs them per call; here
 * they live in ONE file-scope device array of a stated, modest size that every GPU which loads the
 * library holds until the process exits:
 *
 *     [ cgk_scalars, padded to 512 B | CGB_NOALLOC_MAX_CTAS doubles of dot partials ]  = 1 MiB + 512 B
 *
 * In host code the array's NAME is a host-side stand-in, not a device address (a kernel handed
 * it would fault or, where the GPU can reach host memory, quietly use the wrong memory), so the
 * device address is asked for once per device, with that device current, through
 * cudaGetSymbolAddress, and that pointer is what the kernel gets.
 *
 * Every call shares the array and finds the previous call's contents in it: the scalar block
 * (ticket, stop flag) is cleared on the caller's stream before each launch; the partials need no
 * clearing (every CTA of the launch writes its own slot before the last CTA adds them up).  Calls
 * that overlap on one GPU -- two streams, two threads -- race on it: serialise them.
 */
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <mutex>

#include "../../include/cgb200.h"
#include "cg_kernels.h"

#define NOALLOC_HEADER_BYTES 512
static_assert(sizeof(cgk_scalars) <= NOALLOC_HEADER_BYTES, "scalar block outgrew its slot in the scratch array");
static_assert(NOALLOC_HEADER_BYTES % 256 == 0, "partials must stay 256-byte aligned");

static __device__ __align__(256) unsigned char cgb_noalloc_scratch[NOALLOC_HEADER_BYTES + (size_t)CGB_NOALLOC_MAX_CTAS * sizeof(double)];

extern "C" void cgb_set_error(const char *fmt, const char *a, const char *b, int line);

#define NOALLOC_MAX_DEVICES 64
static std::mutex g_mu;
static void *g_scratch_addr[NOALLOC_MAX_DEVICES];      /* device address of the array on each device, 0 = not asked yet */
static int g_sm_count[NOALLOC_MAX_DEVICES];

static int fail(int code, const char *msg, const char *detail, int line)
{
    cgb_set_error("cgb_matvec_noalloc: %s%s (cg_noalloc.cu:%d)", msg, detail, line);
    return code;
}

/* set-up, once per device, with that device current: the scratch array's device address */
static int scratch_for_current_device(void **scratch, int *sm_count)
{
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(CGB_ERR_CUDA, "no CUDA device: ", cudaGetErrorString(e), __LINE__); }
    if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return fail(CGB_ERR_ARG, "device ordinal out of range", "", __LINE__);
    std::lock_guard<std::mutex> lock(g_mu);
    if (!g_scratch_addr[dev]) {
        cudaDeviceProp prop;
        e = cudaGetDeviceProperties(&prop, dev);
        if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(CGB_ERR_CUDA, "cudaGetDeviceProperties: ", cudaGetErrorString(e), __LINE__); }
        if (prop.major < 10) return fail(CGB_ERR_CUDA, "not a Blackwell sm_100 GPU: libcgb200 has no other code path", "", __LINE__);
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, cgb_noalloc_scratch);
        if (e != cudaSuccess || !p) { (void)cudaGetLastError(); return fail(CGB_ERR_CUDA, "cudaGetSymbolAddress(scratch): ", cudaGetErrorString(e), __LINE__); }
        g_scratch_addr[dev] = p;
        g_sm_count[dev] = prop.multiProcessorCount;
    }
    *scratch = g_scratch_addr[dev];
    *sm_count = g_sm_count[dev];
    return CGB_OK;
}

static bool on_device(const void *p)
{
    cudaPointerAttributes at;
    memset(&at, 0, sizeof at);
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { (void)cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeDevice || at.type == cudaMemoryTypeManaged;
}

extern "C" int cgb_matvec_noalloc(const double *d_A, const double *d_v, const double *d_w, double *d_y,
                                  double *d_dot, int64_t m, int64_t n, void *stream)
{
    if (!d_A || !d_v || !d_y || m < 1 || n < 1) return fail(CGB_ERR_ARG, "null pointer or empty shape", "", __LINE__);
    if ((d_w == nullptr) != (d_dot == nullptr)) return fail(CGB_ERR_ARG, "d_w and d_dot go together (both, or neither)", "", __LINE__);
    if (m > CGB_NOALLOC_MAX_ROWS) {
        char msg[320];
        snprintf(msg, sizeof msg, "m = %lld rows exceeds CGB_NOALLOC_MAX_ROWS = %lld (the static scratch array holds %d dot partials); "
                 "use cgb_op_matvec / the solver for larger blocks", (long long)m, (long long)CGB_NOALLOC_MAX_ROWS, CGB_NOALLOC_MAX_CTAS);
        return fail(CGB_ERR_ARG, msg, "", __LINE__);
    }
    if (n > ((int64_t)1 << 24)) return fail(CGB_ERR_ARG, "n out of range", "", __LINE__);
    if (!on_device(d_A) || !on_device(d_v) || !on_device(d_y) || (d_w && (!on_device(d_w) || !on_device(d_dot))))
        return fail(CGB_ERR_ARG, "every pointer must be device memory of the current device (this call copies and allocates nothing)", "", __LINE__);

    void *scratch = nullptr;
    int sms = 0;
    int rc = scratch_for_current_device(&scratch, &sms);
    if (rc != CGB_OK) return rc;
    cgk_scalars *sc = (cgk_scalars *)scratch;
    double *partials = (double *)((unsigned char *)scratch + NOALLOC_HEADER_BYTES);

    cgk_gemv_plan plan;
    memset(&plan, 0, sizeof plan);                       /* library defaults; streaming-load kernel; no L2 policy */
    if (cgk_gemv_make_plan(&plan, m, n, d_A, d_v, sms) != 0) return fail(CGB_ERR_ARG, "no launch plan for this shape", "", __LINE__);
    if (plan.grid > CGB_NOALLOC_MAX_CTAS) return fail(CGB_ERR_ARG, "launch needs more dot partials than the static scratch array holds", "", __LINE__);

    cudaStream_t st = (cudaStream_t)stream;
    /* what must start at zero: the ticket counter and the stop flag (the whole scalar block) */
    cudaError_t e = cudaMemsetAsync(sc, 0, sizeof(cgk_scalars), st);
    if (e != cudaSuccess) { (void)cudaGetLastError(); return fail(CGB_ERR_CUDA, "cudaMemsetAsync(scratch header): ", cudaGetErrorString(e), __LINE__); }
    const int le = cgk_gemv(&plan, d_A, d_v, d_w, d_y, partials, d_dot, sc, m, n, st);
    if (le != 0) return fail(CGB_ERR_CUDA, "kernel launch: ", cgk_error_string(le), __LINE__);
    return CGB_OK;                                       /* asynchronous: the results are ready when the stream is */
}
