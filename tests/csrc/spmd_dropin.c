/* spmd_dropin.c -- TEST CODE: SPMD callers shaped like the reference's MPI program,
 * with threads standing in for the MPI ranks (no MPI in this image).  "Rank 0" reads the files and hands
 * every rank its row block of A, its block of b and the FULL x, exactly what MPI_Scatter / MPI_Bcast deliver at
 * parallel_cg.c:111-115; every rank then makes the reference's own call
 *     conjugrad(local_matrixA, local_vectorB, local_vectorX, local_row, myrank, procsnum)    (parallel_cg.c:118)
 * through `#define conjugrad conjugrad_mpi`.  Every rank must end with the same full x (the reference's
 * replicated x); rank 0's copy is printed in hexadecimal floating point.
 * usage: spmd_dropin np n matrixA.txt vectorb.txt initialguess.txt */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "cgb200.h"
#define conjugrad conjugrad_mpi

typedef struct { double *local_matrixA, *local_vectorB, *local_vectorX; int local_row, myrank, procsnum, ok; } rank_t;

static int slurp(const char *path, double *dst, long count)
{
    FILE *f = fopen(path, "r");
    if (!f) { printf("Could not open file\n"); return -1; }
    long k = 0;
    while (k < count && fscanf(f, "%lf%*c", &dst[k]) == 1) ++k;
    fclose(f);
    return k == count ? 0 : -1;
}

static void *rank_main(void *arg)
{
    rank_t *r = arg;
    conjugrad(r->local_matrixA, r->local_vectorB, r->local_vectorX, r->local_row, r->myrank, r->procsnum);
    cgb_stats st;
    r->ok = cgb_last_stats(&st) == CGB_OK;
    cgb_release_rank();
    return NULL;
}

int main(int argc, char **argv)
{
    if (argc != 6) { fprintf(stderr, "usage: %s np n A.txt b.txt x0.txt\n", argv[0]); return 2; }
    const int np = atoi(argv[1]);
    const long n = atol(argv[2]);
    if (np < 1 || np > 16 || n < 1) return 2;
    if (n % np != 0) { printf("%ld is not divisible by %d\n", n, np); return 1; }      /* parallel_cg.c:86-90 */
    const long m = n / np;
    double *A = malloc((size_t)n * (size_t)n * sizeof(double)), *b = malloc((size_t)n * sizeof(double)), *x0 = malloc((size_t)n * sizeof(double));
    if (!A || !b || !x0) return 3;
    if (slurp(argv[3], A, n * n) || slurp(argv[4], b, n) || slurp(argv[5], x0, n)) return 4;

    rank_t rk[16];
    pthread_t th[16];
    for (int g = 0; g < np; ++g) {
        rk[g].local_matrixA = A + (size_t)g * (size_t)m * (size_t)n;      /* the scattered row block */
        rk[g].local_vectorB = b + (size_t)g * (size_t)m;
        rk[g].local_vectorX = malloc((size_t)n * sizeof(double));          /* every rank: the full vector */
        memcpy(rk[g].local_vectorX, x0, (size_t)n * sizeof(double));
        rk[g].local_row = (int)m; rk[g].myrank = g; rk[g].procsnum = np; rk[g].ok = 0;
        pthread_create(&th[g], NULL, rank_main, &rk[g]);
    }
    int ok = 1;
    for (int g = 0; g < np; ++g) { pthread_join(th[g], NULL); ok &= rk[g].ok; }
    if (!ok) { fprintf(stderr, "a rank did not solve\n"); return 5; }
    for (int g = 1; g < np; ++g)
        if (memcmp(rk[g].local_vectorX, rk[0].local_vectorX, (size_t)n * sizeof(double)) != 0) { fprintf(stderr, "rank %d holds a different x\n", g); return 6; }
    for (long i = 0; i < n; ++i) printf("x %a\n", rk[0].local_vectorX[i]);
    return 0;
}
