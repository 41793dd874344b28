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
/* ---- pieces of the host parser the GPU ingest path (cg_ingest.c) shares */
typedef struct cgb_text { char *p; size_t len; int mapped; } cgb_text;
int  cgb_text_open(const char *path, cgb_text *t);          /* mmap (read-only) or slurp; 0 = ok */
void cgb_text_close(cgb_text *t);
int  cgb_is_sep_byte(unsigned char ch);                     /* whitespace , ;                    */
/* One token [p, e) -> double exactly as strtod would; 0 when it is not a plain number. */
int  cgb_token_to_double(const char *p, const char *e, double *out);
/* Byte range [*beg, *end) that holds values first .. first+count-1 (it starts at the first
 * byte of value `first`).  Returns the number of values in the file, or -1 when the text has
 * bytes the fast tokeniser does not handle (a BOM): use cgb_read_values() then. */
int64_t cgb_text_locate(const cgb_text *t, int64_t first, int64_t count, size_t *beg, size_t *end);
/* counting passes over a file cgb_text_locate has made in this process (the ranks of one process share a pass) */
unsigned long cgb_text_locate_passes(void);

/* memcpy split between up to 8 threads (CGB_IO_THREADS caps it; below 2 MiB a plain memcpy). */
void cgb_parallel_copy(void *dst, const void *src, size_t len);

/* The same copy by a team of threads that is started once and reused for every chunk of a text load
 * (cgb_copy_pool_threads(chunk) members, the caller included).  start returns NULL when no team could be
 * formed: run(NULL, ...) is cgb_parallel_copy. */
typedef struct cgb_copy_pool cgb_copy_pool;
int  cgb_copy_pool_threads(size_t chunk);
cgb_copy_pool *cgb_copy_pool_start(int threads);
void cgb_copy_pool_run(cgb_copy_pool *p, void *dst, const void *src, size_t len);
void cgb_copy_pool_stop(cgb_copy_pool *p);

/* One value as printf("%.{decimals}f\n") would write it, by integer arithmetic (the writer's fast path);
 * returns the bytes written (at most 48), or 0 when the value is left to printf (inf, nan, >= 2^52, decimals > 17). */
size_t cgb_format_fixed(char *out, double v, int decimals);

/* cgb_write_values() with the choice to append (row blocks of a matrix written in turn). */
int cgb_write_values_mode(const char *path, const double *v, int64_t count, int fmt_decimals, int append);
#ifdef __cplusplus
}
#endif
#endif
