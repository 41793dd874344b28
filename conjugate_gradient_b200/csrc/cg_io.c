/* cg_io.c -- see cg_io.h. */
#define _GNU_SOURCE
#include "cg_io.h"
#include "../../include/cgb200.h"
#include <errno.h>
#include <stdlib.h>
#include <string.h>

static char *slurp(const char *path, size_t *len)
{
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    size_t cap = (size_t)1 << 20, n = 0;
    char *buf = malloc(cap + 1);
    if (!buf) { fclose(f); return NULL; }
    for (;;) {
        size_t got = fread(buf + n, 1, cap - n, f);
        n += got;
        if (got == 0) break;
        if (n == cap) {
            cap *= 2;
            char *nb = realloc(buf, cap + 1);
            if (!nb) { free(buf); fclose(f); return NULL; }
            buf = nb;
        }
    }
    fclose(f);
    buf[n] = '\0';
    *len = n;
    return buf;
}

/* separators the reference's "%f%*c" tolerates: one arbitrary character after
 * each number plus fscanf's leading-whitespace skip.  We accept whitespace,
 * ',' and ';' and skip a UTF-8 BOM (initialguess1.txt carries one). */
static const char *skip_sep(const char *p, const char *end)
{
    while (p < end) {
        unsigned char ch = (unsigned char)*p;
        if (ch == ' ' || ch == '\t' || ch == '\n' || ch == '\r' || ch == ',' || ch == ';' || ch == '\f' || ch == '\v') { ++p; continue; }
        if (ch == 0xEF && p + 2 < end && (unsigned char)p[1] == 0xBB && (unsigned char)p[2] == 0xBF) { p += 3; continue; }
        break;
    }
    return p;
}

static int64_t parse_all(const char *buf, size_t len, double *dst, int64_t cap, int *malformed)
{
    const char *p = buf, *end = buf + len;
    int64_t k = 0;
    *malformed = 0;
    for (;;) {
        p = skip_sep(p, end);
        if (p >= end) break;
        char *q;
        errno = 0;
        double v = strtod(p, &q);
        if (q == p) { *malformed = 1; break; }
        if (dst && k < cap) dst[k] = v;
        ++k;
        p = q;
        if (dst && k >= cap) break;
    }
    return k;
}

int cgb_read_values(const char *path, double *dst, int64_t count, int64_t *got)
{
    size_t len = 0;
    char *buf = slurp(path, &len);
    if (got) *got = 0;
    if (!buf) return CGB_ERR_IO;
    int bad = 0;
    int64_t k = parse_all(buf, len, dst, count, &bad);
    free(buf);
    if (got) *got = k;
    return (k == count && !bad) ? CGB_OK : CGB_ERR_IO;
}

int64_t cgb_count_values(const char *path)
{
    size_t len = 0;
    char *buf = slurp(path, &len);
    if (!buf) return -1;
    int bad = 0;
    int64_t k = parse_all(buf, len, NULL, 0, &bad);
    free(buf);
    return bad ? -1 : k;
}

int cgb_read_dimensions(const char *path, int64_t dims[4])
{
    double v[4];
    int64_t got = 0;
    if (cgb_read_values(path, v, 4, &got) != CGB_OK) return CGB_ERR_IO;
    for (int i = 0; i < 4; ++i) {
        if (v[i] < 1 || v[i] != (double)(int64_t)v[i]) return CGB_ERR_IO;
        dims[i] = (int64_t)v[i];
    }
    return CGB_OK;
}

void cgb_print_vector(FILE *out, const double *v, int64_t rows, int64_t cols)
{
    for (int64_t i = 0; i < rows; ++i)
        for (int64_t j = 0; j < cols; ++j) {
            fprintf(out, "%f  ", v[i * cols + j]);
            if (j == cols - 1) fprintf(out, "\n");
        }
}

int cgb_write_values(const char *path, const double *v, int64_t count, int fmt_decimals)
{
    FILE *f = fopen(path, "w");
    if (!f) return CGB_ERR_IO;
    for (int64_t i = 0; i < count; ++i) {
        if (fmt_decimals >= 0) fprintf(f, "%.*f\n", fmt_decimals, v[i]);
        else fprintf(f, "%.17g\n", v[i]);
    }
    return fclose(f) == 0 ? CGB_OK : CGB_ERR_IO;
}
