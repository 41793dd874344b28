/* serial_dropin.c -- TEST CODE: a caller shaped like the reference's serial program.
 * It owns A, b, x on the host, reads them from the three text files with "%lf%*c" (the float->double form
 * of the reference's token rule), and makes the reference's own call: conjugrad(matrixA, vectorB, vectorX)
 * (serialConjugate.c:40,68), with the one added line that gives the library its run-time ROWS.
 * Prints the solution one value per line in hexadecimal floating point (exact) after the library's own
 * "average clock execution time" line.
 * usage: serial_dropin n matrixA.txt vectorb.txt initialguess.txt */
#include <stdio.h>
#include <stdlib.h>
#include "cgb200.h"

static int slurp(const char *path, double *dst, long count)
{
    FILE *f = fopen(path, "r");
    if (!f) { printf("Could not open file\n"); return -1; }
    long k = 0;
    while (k < count && fscanf(f, "%lf%*c", &dst[k]) == 1) ++k;
    fclose(f);
    return k == count ? 0 : -1;
}

int main(int argc, char **argv)
{
    if (argc != 5) { fprintf(stderr, "usage: %s n A.txt b.txt x0.txt\n", argv[0]); return 2; }
    const long n = atol(argv[1]);
    if (n < 1) return 2;
    double *matrixA = malloc((size_t)n * (size_t)n * sizeof(double));
    double *vectorB = malloc((size_t)n * sizeof(double));
    double *vectorX = malloc((size_t)n * sizeof(double));
    if (!matrixA || !vectorB || !vectorX) return 3;
    if (slurp(argv[2], matrixA, n * n) || slurp(argv[3], vectorB, n) || slurp(argv[4], vectorX, n)) return 4;

    cgb_set_dimension(n);
    conjugrad(matrixA, vectorB, vectorX);

    cgb_stats st;
    if (cgb_last_stats(&st) != CGB_OK) { fprintf(stderr, "no solve happened: %s\n", cgb_last_error()); return 5; }
    printf("iterations %lld\n", (long long)st.iterations);
    for (long i = 0; i < n; ++i) printf("x %a\n", vectorX[i]);
    free(matrixA); free(vectorB); free(vectorX);
    cgb_shutdown();
    return 0;
}
