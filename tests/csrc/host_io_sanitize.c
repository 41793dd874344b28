/* host_io_sanitize.c -- TEST CODE.  Drives the host text path of the product (cg_io.c: the
 * multi-threaded exact parser, the writer, the token locator the GPU ingest shares) and is built
 * together with it under -fsanitize=address,undefined and under -fsanitize=thread
 * (tests/test_host_sanitizers.py).  Every value is checked against strtod on the same token, so a
 * run is also a parity run; the sanitizers add what the GPU box cannot give (compute-sanitizer is
 * closed there) for the part of the path that is plain C.
 * usage: host_io_sanitize <scratch dir> */
#define _GNU_SOURCE
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "../../include/cgb200.h"
#include "../../conjugate_gradient_b200/csrc/cg_io.h"

#ifdef CGB_TEST_THREADS_FAIL
/* built with -Dpthread_create=cgb_test_failing_create: every thread creation in cg_io.c is refused, the
 * jobs must then run on the calling thread with the same results */
#include <errno.h>
#include <pthread.h>
int cgb_test_failing_create(pthread_t *t, const pthread_attr_t *a, void *(*fn)(void *), void *arg)
{
    (void)t; (void)a; (void)fn; (void)arg;
    return EAGAIN;
}
#endif

static uint64_t rng_state = 0x9E3779B97F4A7C15ull;
static uint64_t rnd(void)
{
    rng_state ^= rng_state << 13; rng_state ^= rng_state >> 7; rng_state ^= rng_state << 17;
    return rng_state;
}

static int fails = 0;
#define CHECK(cond, ...) do { if (!(cond)) { fprintf(stderr, "FAIL %s:%d: ", __FILE__, __LINE__); fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); ++fails; } } while (0)

/* count values in mixed notations and separators; returns the text and the expected doubles */
static char *make_text(int64_t count, double **want, size_t *len_out)
{
    static const char *seps[] = {"\n", " ", "\r\n", ",", "\t", " \n", ";", "  "};
    char *buf = malloc((size_t)count * 32 + 16);
    double *w = malloc((size_t)count * sizeof(double));
    size_t len = 0;
    for (int64_t i = 0; i < count; ++i) {
        const uint64_t r = rnd();
        const double u = (double)(r >> 11) / 9007199254740992.0;
        char tok[64];
        switch (r & 7) {
        case 0: snprintf(tok, sizeof tok, "%.4f", u); break;                       /* generateSPDmatrix.m:29 */
        case 1: snprintf(tok, sizeof tok, "%.1f", floor(u * 10)); break;           /* the x0 format, :41 */
        case 2: snprintf(tok, sizeof tok, "%.17g", (u - 0.5) * 1e6); break;
        case 3: snprintf(tok, sizeof tok, "%.6e", u * 1e-30); break;
        case 4: snprintf(tok, sizeof tok, "%d", (int)(u * 2000) - 1000); break;
        case 5: snprintf(tok, sizeof tok, "+%.3f", u * 65536.0); break;
        case 6: snprintf(tok, sizeof tok, "%.15g", 65536.0 + u); break;            /* a diagonal entry */
        default: snprintf(tok, sizeof tok, "%.12E", (u - 0.5) * 1e200); break;
        }
        w[i] = strtod(tok, NULL);
        const size_t tl = strlen(tok);
        memcpy(buf + len, tok, tl); len += tl;
        const char *s = seps[(r >> 8) & 7];
        memcpy(buf + len, s, strlen(s)); len += strlen(s);
    }
    *want = w; *len_out = len;
    return buf;
}

static void write_file(const char *path, const char *buf, size_t len)
{
    FILE *f = fopen(path, "wb");
    if (!f || fwrite(buf, 1, len, f) != len) { perror(path); exit(2); }
    fclose(f);
}

static void round_trip(const char *dir, int64_t count, const char *threads)
{
    char path[4096], path2[4096];
    snprintf(path, sizeof path, "%s/vals_%lld.txt", dir, (long long)count);
    snprintf(path2, sizeof path2, "%s/vals_%lld_out.txt", dir, (long long)count);
    setenv("CGB_IO_THREADS", threads, 1);
    double *want = NULL; size_t len = 0;
    char *txt = make_text(count, &want, &len);
    write_file(path, txt, len);

    /* exact-size destination: an overrun of one element is an ASan report */
    double *got = malloc((size_t)count * sizeof(double));
    int64_t n = -1;
    int rc = cgb_read_values(path, got, count, &n);
    CHECK(rc == CGB_OK && n == count, "read rc=%d n=%lld (count %lld, threads %s)", rc, (long long)n, (long long)count, threads);
    for (int64_t i = 0; i < count && rc == CGB_OK; ++i)
        if (memcmp(&got[i], &want[i], sizeof(double)) != 0) { CHECK(0, "value %lld: %a != %a", (long long)i, got[i], want[i]); break; }
    CHECK(cgb_count_values(path) == count, "count_values");

    /* asking for more than the file holds is an error, never a read past the text */
    double *big = malloc((size_t)(count + 3) * sizeof(double));
    rc = cgb_read_values(path, big, count + 3, &n);
    CHECK(rc == CGB_ERR_IO && n == count, "short file: rc=%d n=%lld", rc, (long long)n);
    free(big);
    /* a prefix */
    if (count > 5) {
        double *pre = malloc((size_t)(count / 2) * sizeof(double));
        rc = cgb_read_values(path, pre, count / 2, &n);
        CHECK(rc == CGB_OK, "prefix rc=%d", rc);
        CHECK(memcmp(pre, want, (size_t)(count / 2) * sizeof(double)) == 0, "prefix differs");
        free(pre);
    }

    /* the locator the GPU ingest uses: any [first, first+cnt) range converts to the same values */
    cgb_text t;
    CHECK(cgb_text_open(path, &t) == 0, "text_open");
    for (int k = 0; k < 8 && count > 0; ++k) {
        const int64_t first = (int64_t)(rnd() % (uint64_t)count);
        const int64_t cnt = 1 + (int64_t)(rnd() % (uint64_t)(count - first));
        size_t b = 0, e = 0;
        const int64_t total = cgb_text_locate(&t, first, cnt, &b, &e);
        CHECK(total == count && b <= e && e <= t.len, "locate total=%lld b=%zu e=%zu len=%zu", (long long)total, b, e, t.len);
        int64_t j = first;
        for (size_t p = b; p < e && total == count; ) {
            while (p < e && cgb_is_sep_byte((unsigned char)t.p[p])) ++p;
            size_t q = p;
            while (q < e && !cgb_is_sep_byte((unsigned char)t.p[q])) ++q;
            if (q == p) break;
            double v = 0.0;
            CHECK(cgb_token_to_double(t.p + p, t.p + q, &v) && memcmp(&v, &want[j], 8) == 0, "token %lld in range", (long long)j);
            ++j; p = q;
        }
        CHECK(j == first + cnt, "range [%lld,+%lld) yielded %lld values", (long long)first, (long long)cnt, (long long)(j - first));
    }
    cgb_text_close(&t);

    /* writer -> parser round trip at full precision, and the appending form */
    rc = cgb_write_values(path2, want, count, 17);
    CHECK(rc == CGB_OK, "write rc=%d", rc);
    rc = cgb_read_values(path2, got, count, &n);
    CHECK(rc == CGB_OK && n == count, "reread rc=%d", rc);
    for (int64_t i = 0; i < count && rc == CGB_OK; ++i)
        if (!(got[i] == want[i] || fabs(got[i] - want[i]) <= 1e-17 * fabs(want[i]) + 1e-17)) { CHECK(0, "%%.17f round trip at %lld: %a vs %a", (long long)i, got[i], want[i]); break; }
    const int64_t h = count / 2;
    rc = cgb_write_values_mode(path2, want, h, 4, 0);
    if (rc == CGB_OK) rc = cgb_write_values_mode(path2, want + h, count - h, 4, 1);
    CHECK(rc == CGB_OK && cgb_count_values(path2) == count, "append form");
    free(got); free(want); free(txt);
}

/* ranks of one process locating their rows in one file at the same time: one counting pass, the same answers */
typedef struct { const char *path; int64_t first, count, total; size_t beg, end; } rank_job;
static void *rank_locate(void *arg)
{
    rank_job *j = arg;
    cgb_text t;
    if (cgb_text_open(j->path, &t) != 0) { j->total = -99; return NULL; }
    j->total = cgb_text_locate(&t, j->first, j->count, &j->beg, &j->end);
    cgb_text_close(&t);
    return NULL;
}

int main(int argc, char **argv)
{
    const char *dir = argc > 1 ? argv[1] : "/tmp";
    char path[4096];
    /* sizes around the thread-split thresholds (1 MiB of text per thread) and tiny ones */
    const int64_t counts[] = {1, 2, 7, 1000, 70000, 400000};
    const char *threads[] = {"1", "3", "8", "64"};
    for (unsigned c = 0; c < sizeof counts / sizeof *counts; ++c)
        for (unsigned t = 0; t < sizeof threads / sizeof *threads; ++t)
            round_trip(dir, counts[c], threads[t]);

    /* malformed inputs: status codes, no out-of-bounds access */
    double v[8]; int64_t n = 0;
    snprintf(path, sizeof path, "%s/missing.txt", dir);
    CHECK(cgb_read_values(path, v, 4, &n) == CGB_ERR_IO, "missing file");
    snprintf(path, sizeof path, "%s/empty.txt", dir);
    write_file(path, "", 0);
    CHECK(cgb_read_values(path, v, 1, &n) == CGB_ERR_IO && n == 0, "empty file");
    CHECK(cgb_count_values(path) == 0, "empty count");
    snprintf(path, sizeof path, "%s/bad.txt", dir);
    write_file(path, "1.0\n2.0\nabc\n4.0\n", 16);
    CHECK(cgb_read_values(path, v, 4, &n) == CGB_ERR_IO, "bad token");
    snprintf(path, sizeof path, "%s/bom.txt", dir);
    write_file(path, "\xEF\xBB\xBF" "0.5\n0.25", 11);                                   /* initialguess1.txt has a BOM */
    CHECK(cgb_read_values(path, v, 2, &n) == CGB_OK && v[0] == 0.5 && v[1] == 0.25, "BOM");
    snprintf(path, sizeof path, "%s/nosep.txt", dir);
    write_file(path, "3.5", 3);                                                         /* no trailing separator */
    CHECK(cgb_read_values(path, v, 1, &n) == CGB_OK && v[0] == 3.5, "no trailing separator");
    int64_t dims[4];
    snprintf(path, sizeof path, "%s/dimensions.txt", dir);
    write_file(path, "2\n2\n2\n1\n", 8);
    CHECK(cgb_read_dimensions(path, dims) == CGB_OK && dims[0] == 2 && dims[3] == 1, "dimensions");
    write_file(path, "2\n2\n", 4);
    CHECK(cgb_read_dimensions(path, dims) != CGB_OK, "short dimensions");

    /* the ingest's staging copy: every byte, every split, exact-size buffers */
    {
        /* 8 MiB + 3 and 32 MiB + 5: a quotient len / threads that is a whole number of pages with a remainder left over */
        const size_t sizes[] = {0, 1, 4095, 4096, 4097, (size_t)1 << 20, ((size_t)3 << 20) + 17, ((size_t)32 << 20) + 2053,
                                ((size_t)8 << 20) + 3, ((size_t)32 << 20) + 5, ((size_t)2 << 20) + 1, (size_t)8 * 4096 + 7};
        for (unsigned t = 0; t < sizeof threads / sizeof *threads; ++t) {
            setenv("CGB_IO_THREADS", threads[t], 1);
            for (unsigned k = 0; k < sizeof sizes / sizeof *sizes; ++k) {
                const size_t len = sizes[k];
                unsigned char *src = malloc(len + 1), *dst = malloc(len + 1);
                for (size_t i = 0; i < len; ++i) src[i] = (unsigned char)(i * 2654435761u >> 13);
                memset(dst, 0xEE, len + 1);
                cgb_parallel_copy(dst, src, len);
                CHECK(memcmp(dst, src, len) == 0 && dst[len] == 0xEE, "parallel copy of %zu bytes (threads %s)", len, threads[t]);
                free(src); free(dst);
            }
        }
        /* the team that lives for a whole text load: the same sizes in a row through ONE team, every team size */
        const int teams[] = {2, 3, 8, 16};
        for (unsigned t = 0; t < sizeof teams / sizeof *teams; ++t) {
            cgb_copy_pool *pool = cgb_copy_pool_start(teams[t]);
#ifdef CGB_TEST_THREADS_FAIL
            CHECK(pool == NULL, "no team when no thread can be created (%d)", teams[t]);   /* run(NULL, ..) still copies */
#else
            CHECK(pool != NULL, "copy team of %d", teams[t]);
#endif
            for (int round = 0; round < 2; ++round)
                for (unsigned k = 0; k < sizeof sizes / sizeof *sizes; ++k) {
                    const size_t len = sizes[k];
                    unsigned char *src = malloc(len + 1), *dst = malloc(len + 1);
                    for (size_t i = 0; i < len; ++i) src[i] = (unsigned char)((i + k) * 2654435761u >> 11);
                    memset(dst, 0xEE, len + 1);
                    cgb_copy_pool_run(pool, dst, src, len);
                    CHECK(memcmp(dst, src, len) == 0 && dst[len] == 0xEE, "team copy of %zu bytes (team %d)", len, teams[t]);
                    free(src); free(dst);
                }
            cgb_copy_pool_stop(pool);
        }
        CHECK(cgb_copy_pool_start(1) == NULL, "a team of one is no team");
        cgb_copy_pool_stop(NULL);
        CHECK(cgb_copy_pool_threads((size_t)8 << 20) >= 1 && cgb_copy_pool_threads((size_t)8 << 20) <= 8, "team size");
        CHECK(cgb_copy_pool_threads(1000) == 1, "team size for tiny chunks");
    }

#ifndef CGB_TEST_THREADS_FAIL
    {
        snprintf(path, sizeof path, "%s/shared.txt", dir);
        const int64_t cnt = 400000;
        double *v = malloc((size_t)cnt * sizeof *v);
        for (int64_t i = 0; i < cnt; ++i) v[i] = (double)(i % 9973) / 16.0;
        CHECK(cgb_write_values(path, v, cnt, 4) == CGB_OK, "write shared");
        free(v);
        unsetenv("CGB_LOCATE_NO_CACHE");
        enum { R = 6 };
        for (int round = 0; round < 2; ++round) {
            rank_job job[R], solo[R];
            pthread_t th[R];
            const unsigned long before = cgb_text_locate_passes();
            for (int r = 0; r < R; ++r) { job[r].path = path; job[r].first = cnt / R * r; job[r].count = cnt / R; job[r].total = 0; job[r].beg = job[r].end = 0; }
            for (int r = 0; r < R; ++r) pthread_create(&th[r], NULL, rank_locate, &job[r]);
            for (int r = 0; r < R; ++r) pthread_join(th[r], NULL);
            CHECK(cgb_text_locate_passes() - before == (round == 0 ? 1u : 0u), "one counting pass for %d ranks (round %d: %lu)", R, round, cgb_text_locate_passes() - before);
            setenv("CGB_LOCATE_NO_CACHE", "1", 1);
            for (int r = 0; r < R; ++r) { solo[r] = job[r]; rank_locate(&solo[r]); }
            unsetenv("CGB_LOCATE_NO_CACHE");
            for (int r = 0; r < R; ++r)
                CHECK(job[r].total == cnt && solo[r].total == cnt && job[r].beg == solo[r].beg && job[r].end == solo[r].end && job[r].beg < job[r].end,
                      "rank %d: shared [%zu,%zu) alone [%zu,%zu)", r, job[r].beg, job[r].end, solo[r].beg, solo[r].end);
        }
    }
#endif

    /* the writer's integer-arithmetic "%.{d}f": byte for byte printf's, wherever it answers at all */
    {
        uint64_t st = 0x9E3779B97F4A7C15ull;
        long covered = 0, total = 0;
        const double fixed[] = {0.0, -0.0, 0.03125, 0.09375, 0.5, 1.5, 2.5, -2.5, 0.00005, 0.00015, 0.99995, 0.999949999999,
                                8192.5118, 4096.00005, 1e-5, 4.9e-324, 2.2250738585072014e-308, 1e15, 9.2e14, 9.3e14,
                                4503599627370495.5, 4503599627370496.0, 1e22, 1.7976931348623157e308, 0.125, 0.375, 1e-17,
                                5e-18, 123456789.987654321, 0.1, 0.2, 0.3, 0.7, 1.0 / 3.0, 2.0 / 3.0, 1e-300};
        for (int rep = 0; rep < 400000; ++rep) {
            double v;
            if (rep < (int)(sizeof fixed / sizeof *fixed)) v = fixed[rep];
            else {
                st ^= st << 13; st ^= st >> 7; st ^= st << 17;
                const int kind = (int)(st % 5);
                if (kind == 0) { uint64_t b = st * 0x2545F4914F6CDD1Dull; memcpy(&v, &b, 8); }              /* any bit pattern */
                else if (kind == 1) v = (double)(st >> 11) / 9007199254740992.0;                              /* U(0,1), 53 bits */
                else if (kind == 2) v = (double)(int64_t)(st >> 40) / (double)(1 << (int)((st >> 8) % 20));   /* exact ties at k / 2^j */
                else if (kind == 3) v = ((double)(st >> 11) / 9007199254740992.0 - 0.5) * 65536.0;
                else v = (double)((st >> 20) % 100000000ull) / 10000.0 + ((st & 1) ? 0.00005 : 0.0);
            }
            for (int d = 0; d <= 18; d += (rep < 64 ? 1 : 4)) {
                char a[512], b[64];
                memset(b, 0x7f, sizeof b);
                const size_t k = cgb_format_fixed(b, v, d);
                ++total;
                if (!k) continue;
                ++covered;
                const int w = snprintf(a, sizeof a, "%.*f\n", d, v);
                CHECK(k <= 48 && (size_t)w == k && memcmp(a, b, k) == 0 && b[k] == 0x7f, "format_fixed(%a, %d) = %.*s, printf %s", v, d, (int)k, b, a);
            }
        }
        CHECK(covered * 10 > total * 6, "the fast path covers most values (%ld of %ld)", covered, total);
        char b[64];
        CHECK(cgb_format_fixed(b, 1.0 / 0.0, 4) == 0 && cgb_format_fixed(b, 0.0 / 0.0, 4) == 0 && cgb_format_fixed(b, 1e300, 4) == 0
              && cgb_format_fixed(b, 1.0, 18) == 0 && cgb_format_fixed(b, 1.0, -1) == 0, "values left to printf");
    }

    FILE *nul = fopen("/dev/null", "w");
    const double m[6] = {1, 2, 3, 4, 5, 6};
    cgb_print_vector(nul, m, 2, 3);
    fclose(nul);
    if (fails) { fprintf(stderr, "%d check(s) failed\n", fails); return 1; }
    printf("host_io_sanitize: ok\n");
    return 0;
}
