/* cg_io.h -- text inputs of the reference CLI (plain C, host only).
 * Replaces initialize() (serialConjugate.c:85-105, parallel_cg.c:147-168) and
 * printer() (serialConjugate.c:260-276), and gives dimensions.txt -- shipped by
 * the reference but read by none of its programs -- its evident meaning:
 * four integers, one per line: rows(A) cols(A) rows(b) cols(b). */
#ifndef CG_IO_H
#define CG_IO_H
#include <stdint.h>
#include <stdio.h>
#include "../../include/cgb200.h"
#ifdef __cplusplus
extern "C" {
#endif
/* printer() (serialConjugate.c:260-276): "%f  " per element, newline after the last column. */
void cgb_print_vector(FILE *out, const double *v, int64_t rows, int64_t cols);
/* cgb_write_values() with the choice to append (row blocks of a matrix written in turn). */
int cgb_write_values_mode(const char *path, const double *v, int64_t count, int fmt_decimals, int append);
#ifdef __cplusplus
}
#endif
#endif
