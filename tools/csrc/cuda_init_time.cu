// Bare cost of becoming a CUDA process on this box: cudaFree(0) (driver + context), then one 537 MB cudaMalloc.
#include <cuda_runtime.h>
#include <stdio.h>
#include <time.h>
static double now() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec + 1e-9 * ts.tv_nsec; }
int main() {
    double t0 = now();
    cudaError_t e = cudaFree(0);
    double t1 = now();
    void *p = nullptr;
    cudaMalloc(&p, (size_t)8192 * 8192 * 8);
    double t2 = now();
    printf("cudaFree(0) %.3f s (%s), cudaMalloc(537 MB) %.3f s\n", t1 - t0, cudaGetErrorString(e), t2 - t1);
    return 0;
}
