/* cg_io.c -- see cg_io.h. */
#define _GNU_SOURCE
#include "cg_io.h"
#include "../../include/cgb200.h"
#include <errno.h>
#include <fcntl.h>
#include <pthread.h>
#include <sched.h>
#include <time.h>
#include <stdlib.h>
#include <string.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <unistd.h>

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

/* ---- parallel ingest (text parsing is >95 % of the reference's wall time,
 * fscanf("%f%*c") at ~4 M values/s, serialConjugate.c:92-98).  The file is cut at separator
 * characters into one span per thread; every thread converts its tokens with Clinger's exact
 * fast path -- an integer mantissa below 2^53 times or divided by a power of ten up to 1e22
 * is ONE correctly rounded IEEE operation, hence the same double strtod returns -- and hands
 * anything else (more than 15 digits, large exponents, inf/nan, hex) to strtod itself.  Files
 * whose tokens are not plain numbers fall back to the sequential strtod loop above, so the
 * accepted grammar and the error behaviour are exactly those of parse_all(). */
static const double k_pow10[23] = {1e0, 1e1, 1e2, 1e3, 1e4, 1e5, 1e6, 1e7, 1e8, 1e9, 1e10, 1e11,
                                   1e12, 1e13, 1e14, 1e15, 1e16, 1e17, 1e18, 1e19, 1e20, 1e21, 1e22};

/* The byte loops below run once per byte of a multi-GB text.  The exported names are reached
 * through the PLT from inside a -fPIC library and are never inlined (~5 ns per byte, 0.2 GB/s per
 * thread, measured: locating the rows of an n = 8192 matrix took 0.37 s on 8 cores), so the file
 * uses these file-local forms and the exported ones are wrappers for the other translation units. */
static inline int is_sep(unsigned char ch)
{
    return (unsigned)(ch - '\t') <= 4u || ch == ' ' || ch == ',' || ch == ';';   /* \t \n \v \f \r are 9..13 */
}

int cgb_is_sep_byte(unsigned char ch) { return is_sep(ch); }

#if defined(__SSE2__)
#include <emmintrin.h>
/* 16 bytes at p: bit i of the result = byte i is a separator; *high gets bit i = byte i >= 0x80 */
static inline unsigned sep_mask16(const char *p, unsigned *high)
{
    const __m128i v = _mm_loadu_si128((const __m128i *)p);
    const __m128i w = _mm_sub_epi8(v, _mm_set1_epi8('\t'));
    __m128i m = _mm_cmpeq_epi8(_mm_min_epu8(w, _mm_set1_epi8(4)), w);            /* (ch - 9) <= 4, unsigned */
    m = _mm_or_si128(m, _mm_cmpeq_epi8(v, _mm_set1_epi8(' ')));
    m = _mm_or_si128(m, _mm_cmpeq_epi8(v, _mm_set1_epi8(',')));
    m = _mm_or_si128(m, _mm_cmpeq_epi8(v, _mm_set1_epi8(';')));
    *high = (unsigned)_mm_movemask_epi8(v);
    return (unsigned)_mm_movemask_epi8(m);
}
#endif

/* One token [p, e): returns 1 and *out if it is a plain decimal number, else 0. */
static inline int token_to_double(const char *p, const char *e, double *out);
int cgb_token_to_double(const char *p, const char *e, double *out) { return token_to_double(p, e, out); }

static inline int token_to_double(const char *p, const char *e, double *out)
{
    const char *s = p;
    int neg = 0;
    if (s < e && (*s == '+' || *s == '-')) { neg = (*s == '-'); ++s; }
    uint64_t w = 0;
    int digits = 0, sig = 0, exp10 = 0, seen_point = 0;
    for (; s < e; ++s) {
        const unsigned d = (unsigned)(*s - '0');
        if (d <= 9u) {
            ++digits;
            if (w != 0 || d != 0) {
                if (++sig > 19) goto slow;
                w = w * 10u + d;
            }
            if (seen_point) --exp10;
        } else if (*s == '.' && !seen_point) seen_point = 1;
        else break;
    }
    if (digits == 0) goto slow;
    if (s < e && (*s == 'e' || *s == 'E')) {
        const char *t = s + 1;
        int eneg = 0, ev = 0, ed = 0;
        if (t < e && (*t == '+' || *t == '-')) { eneg = (*t == '-'); ++t; }
        for (; t < e && (unsigned)(*t - '0') <= 9u; ++t) { if (ev < 10000) ev = ev * 10 + (*t - '0'); ++ed; }
        if (ed == 0) goto slow;
        exp10 += eneg ? -ev : ev;
        s = t;
    }
    if (s != e) goto slow;                                 /* trailing characters: not a plain number */
    if (w == 0) { *out = neg ? -0.0 : 0.0; return 1; }
    if (sig <= 15 && exp10 >= -22 && exp10 <= 22) {
        double v = (double)w;                              /* exact: w < 10^15 < 2^53 */
        v = exp10 < 0 ? v / k_pow10[-exp10] : v * k_pow10[exp10];
        *out = neg ? -v : v;
        return 1;
    }
slow: {
        char tmp[80];
        const size_t len = (size_t)(e - p);
        if (len == 0 || len >= sizeof tmp) return 0;
        memcpy(tmp, p, len);
        tmp[len] = '\0';
        char *q;
        const double v = strtod(tmp, &q);
        if (q != tmp + len) return 0;                      /* strtod would stop inside the token */
        *out = v;
        return 1;
    }
}

typedef struct {
    const char *beg, *end;
    double *dst;              /* NULL: count only */
    int64_t count, offset, cap;
    int irregular;
} span_job;

/* Token starts in [beg, end): a non-separator byte whose predecessor is a separator (or beg, which is a cut:
 * "after a separator").  *junk is set when a byte >= 0x80 was seen (a BOM: the sequential path decides). */
static inline int64_t count_range(const char *beg, const char *end, unsigned *junk_out)
{
    int64_t k = 0;
    unsigned prev_sep = 1, junk = 0;
    const char *p = beg;
#if defined(__SSE2__)
    for (; end - p >= 16; p += 16) {
        unsigned high;
        const unsigned sep = sep_mask16(p, &high);
        const unsigned starts = ~sep & ((sep << 1) | prev_sep) & 0xFFFFu;
        k += __builtin_popcount(starts);
        prev_sep = (sep >> 15) & 1u;
        junk |= high;
    }
#endif
    for (; p < end; ++p) {
        const unsigned char ch = (unsigned char)*p;
        const unsigned sep = (unsigned)is_sep(ch);
        junk |= ch & 0x80u;
        k += (int64_t)(~sep & prev_sep & 1u);
        prev_sep = sep;
    }
    *junk_out |= junk;
    return k;
}

static void *span_count(void *arg)
{
    span_job *j = arg;
    unsigned junk = 0;
    j->count = count_range(j->beg, j->end, &junk);
    if (junk) j->irregular = 1;
    return NULL;
}

static void *span_parse(void *arg)
{
    span_job *j = arg;
    const char *p = j->beg;
    int64_t k = j->offset;
    while (p < j->end) {
        while (p < j->end && is_sep((unsigned char)*p)) ++p;
        if (p >= j->end) break;
        const char *e = p;
        while (e < j->end && !is_sep((unsigned char)*e)) ++e;
        if (k >= j->cap) break;                            /* caller wanted only the first `cap` values */
        if (!token_to_double(p, e, &j->dst[k])) { j->irregular = 1; return NULL; }
        ++k;
        p = e;
    }
    return NULL;
}

#define IO_MAX_THREADS 64
static int io_threads(size_t len)
{
    long nc = sysconf(_SC_NPROCESSORS_ONLN);
    {
        /* a process pinned to fewer CPUs than are online (taskset, a container's cpuset) gets threads for those:
         * the copy team's workers spin between chunks and must not outnumber the CPUs they may run on */
        cpu_set_t set;
        if (sched_getaffinity(0, sizeof set, &set) == 0) { const int k = CPU_COUNT(&set); if (k > 0 && k < nc) nc = k; }
    }
    const char *e = getenv("CGB_IO_THREADS");
    if (e && *e) nc = atol(e);
    if (nc < 1) nc = 1;
    if (nc > IO_MAX_THREADS) nc = IO_MAX_THREADS;
    size_t by_size = len / ((size_t)1 << 20) + 1;          /* >= 1 MiB of text per thread */
    return (int)((size_t)nc < by_size ? (size_t)nc : by_size);
}

/* fn over jobs[0..T): job 0 on the calling thread, the others on their own threads; a thread that
 * cannot be created (process limits) has its job run here instead, so the result never depends on it. */
static void run_jobs(void *(*fn)(void *), void *jobs, size_t stride, int T)
{
    pthread_t th[IO_MAX_THREADS];
    char started[IO_MAX_THREADS];
    for (int t = 1; t < T; ++t) started[t] = pthread_create(&th[t], NULL, fn, (char *)jobs + (size_t)t * stride) == 0;
    fn(jobs);
    for (int t = 1; t < T; ++t) {
        if (started[t]) pthread_join(th[t], NULL);
        else fn((char *)jobs + (size_t)t * stride);
    }
}

/* memcpy by several threads: one thread moves ~2.5 GB/s out of a freshly mapped file (page faults
 * included), which was the limit of the GPU text ingest's staging copy (cg_ingest.c). */
typedef struct { char *dst; const char *src; size_t len; } copy_job;
static void *copy_main(void *arg)
{
    copy_job *j = arg;
    memcpy(j->dst, j->src, j->len);
    return NULL;
}
void cgb_parallel_copy(void *dst, const void *src, size_t len)
{
    int T = io_threads(len);
    if (T > 8) T = 8;
    if (T <= 1) { memcpy(dst, src, len); return; }
    copy_job job[8];
    /* page-sized pieces that COVER len: the quotient is rounded up first (len / T rounded down left up to
     * T - 1 tail bytes uncopied whenever len / T was a multiple of the page size: 8 MiB + 3 bytes on 8 threads) */
    const size_t per = ((len + (size_t)T - 1) / (size_t)T + 4095) & ~(size_t)4095;
    int used = 0;
    for (size_t off = 0; off < len && used < 8; off += per, ++used) {
        job[used].dst = (char *)dst + off; job[used].src = (const char *)src + off;
        job[used].len = len - off < per ? len - off : per;
    }
    run_jobs(copy_main, job, sizeof job[0], used);
}

/* A team of copy threads that lives for one text load: cgb_parallel_copy starts and joins its threads for
 * every chunk, which is a third of the copy time at 8 MiB per chunk (20 against 33 GB/s on the build
 * container).  Between chunks the workers SPIN on the chunk counter for up to POOL_SPIN_NS (the caller is
 * back within tens of microseconds: it only issues one copy and three launches) and only then sleep on a
 * condition variable -- waking 15 sleepers through futexes cost more than the 0.25 ms a chunk takes
 * (measured on the 16-vCPU bench box: 8 MiB chunks on 16 threads 53-61 ms against 20 ms with 8 threads). */
#define POOL_SPIN_NS 200000L
struct cgb_copy_pool;
typedef struct { struct cgb_copy_pool *pool; int share; } pool_arg;
struct cgb_copy_pool {
    int workers;                               /* threads besides the caller */
    pthread_t th[IO_MAX_THREADS];
    pthread_mutex_t mu;
    pthread_cond_t go;
    unsigned long epoch;                       /* one per chunk; atomic, stored under mu */
    int pending;                               /* workers still copying; atomic */
    int quit, sleepers;                        /* quit: atomic, stored under mu; sleepers: under mu */
    char *dst; const char *src; size_t len, per;   /* published before the epoch (release / acquire) */
    pool_arg arg[IO_MAX_THREADS];
};

static inline void cpu_relax(void)
{
#if defined(__x86_64__) || defined(__i386__)
    __builtin_ia32_pause();
#endif
}

static long mono_ns(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (long)ts.tv_sec * 1000000000L + ts.tv_nsec;
}

static void pool_copy_share(int share, char *dst, const char *src, size_t len, size_t per)
{
    const size_t off = (size_t)share * per;
    if (off < len) memcpy(dst + off, src + off, len - off < per ? len - off : per);
}

static void *pool_main(void *a)
{
    struct cgb_copy_pool *p = ((pool_arg *)a)->pool;
    const int share = ((pool_arg *)a)->share;
    unsigned long seen = 0;
    for (;;) {
        unsigned long e;
        long t_begin = 0;
        unsigned spins = 0;
        while ((e = __atomic_load_n(&p->epoch, __ATOMIC_ACQUIRE)) == seen && !__atomic_load_n(&p->quit, __ATOMIC_ACQUIRE)) {
            cpu_relax();
            if ((++spins & 63u) != 0) continue;
            const long now = mono_ns();
            if (!t_begin) t_begin = now;
            if (now - t_begin < POOL_SPIN_NS) continue;
            pthread_mutex_lock(&p->mu);                    /* nothing for a long while: sleep until told */
            ++p->sleepers;
            while (__atomic_load_n(&p->epoch, __ATOMIC_ACQUIRE) == seen && !__atomic_load_n(&p->quit, __ATOMIC_ACQUIRE))
                pthread_cond_wait(&p->go, &p->mu);
            --p->sleepers;
            pthread_mutex_unlock(&p->mu);
        }
        if (e == seen) return NULL;                        /* quit */
        seen = e;
        pool_copy_share(share, p->dst, p->src, p->len, p->per);
        __atomic_fetch_sub(&p->pending, 1, __ATOMIC_ACQ_REL);
    }
}

cgb_copy_pool *cgb_copy_pool_start(int threads)
{
    if (threads > IO_MAX_THREADS) threads = IO_MAX_THREADS;
    if (threads < 2) return NULL;
    cgb_copy_pool *p = calloc(1, sizeof *p);
    if (!p) return NULL;
    pthread_mutex_init(&p->mu, NULL);
    pthread_cond_init(&p->go, NULL);
    for (int t = 0; t < threads - 1; ++t) {
        p->arg[p->workers].pool = p; p->arg[p->workers].share = p->workers + 1;     /* share 0 is the caller's */
        if (pthread_create(&p->th[p->workers], NULL, pool_main, &p->arg[p->workers]) != 0) break;
        ++p->workers;                           /* a thread that cannot be started just makes the team smaller */
    }
    if (p->workers == 0) { cgb_copy_pool_stop(p); return NULL; }
    return p;
}

void cgb_copy_pool_run(cgb_copy_pool *p, void *dst, const void *src, size_t len)
{
    if (!p) { cgb_parallel_copy(dst, src, len); return; }
    if (len < ((size_t)1 << 20)) { memcpy(dst, src, len); return; }
    const size_t shares = (size_t)p->workers + 1;
    const size_t per = ((len + shares - 1) / shares + 4095) & ~(size_t)4095;        /* pages, covering len */
    p->dst = dst; p->src = src; p->len = len; p->per = per;
    __atomic_store_n(&p->pending, p->workers, __ATOMIC_RELAXED);
    pthread_mutex_lock(&p->mu);
    __atomic_store_n(&p->epoch, p->epoch + 1, __ATOMIC_RELEASE);
    if (p->sleepers > 0) pthread_cond_broadcast(&p->go);
    pthread_mutex_unlock(&p->mu);
    pool_copy_share(0, (char *)dst, (const char *)src, len, per);
    unsigned spins = 0;
    while (__atomic_load_n(&p->pending, __ATOMIC_ACQUIRE) > 0) {
        cpu_relax();
        if ((++spins & 4095u) == 0) sched_yield();          /* a worker without a CPU of its own gets ours */
    }
}

void cgb_copy_pool_stop(cgb_copy_pool *p)
{
    if (!p) return;
    pthread_mutex_lock(&p->mu);
    __atomic_store_n(&p->quit, 1, __ATOMIC_RELEASE);
    pthread_cond_broadcast(&p->go);
    pthread_mutex_unlock(&p->mu);
    for (int t = 0; t < p->workers; ++t) pthread_join(p->th[t], NULL);
    pthread_cond_destroy(&p->go);
    pthread_mutex_destroy(&p->mu);
    free(p);
}

/* threads a copy team should have for chunks of `chunk` bytes: the online CPUs (CGB_IO_THREADS caps it), at
 * most 8 -- on the 16-vCPU bench box teams of 4 to 8 stream an 8 MiB chunk in the same time (stream + parse of
 * the n = 8192 file 12.8-12.9 ms) and a team of 16 is slower (16.5 ms): the caller's CUDA calls and the driver's
 * own threads want CPUs too -- and no thread with less than 512 KiB to move */
int cgb_copy_pool_threads(size_t chunk)
{
    int T = io_threads((size_t)IO_MAX_THREADS << 20);
    if (T > 8) T = 8;
    const size_t by_size = chunk / ((size_t)512 << 10);
    if ((size_t)T > by_size) T = (int)by_size;
    return T < 1 ? 1 : T;
}

/* Returns the number of values, or -1 when the file needs the sequential path. */
static int64_t parse_parallel(const char *buf, size_t len, double *dst, int64_t cap)
{
    const int T = io_threads(len);
    span_job job[IO_MAX_THREADS];
    const char *end = buf + len, *cut = buf;
    for (int t = 0; t < T; ++t) {
        const char *nxt = (t == T - 1) ? end : buf + (len / (size_t)T) * (size_t)(t + 1);
        if (nxt < cut) nxt = cut;
        while (nxt < end && !is_sep((unsigned char)*nxt)) ++nxt;              /* cut only at a separator */
        memset(&job[t], 0, sizeof job[t]);
        job[t].beg = cut; job[t].end = nxt;
        cut = nxt;
    }
    run_jobs(span_count, job, sizeof job[0], T);
    int64_t total = 0;
    for (int t = 0; t < T; ++t) {
        if (job[t].irregular) return -1;
        job[t].offset = total;
        total += job[t].count;
    }
    if (!dst) return total;
    for (int t = 0; t < T; ++t) { job[t].dst = dst; job[t].cap = cap; }
    run_jobs(span_parse, job, sizeof job[0], T);
    for (int t = 0; t < T; ++t)
        if (job[t].irregular) return -1;
    return total;
}

/* Whole file, read-only: mmap when possible (no copy of a multi-GB text), else slurp. */
typedef cgb_text text_t;

/* Which file a mapping shows: (device, inode, size, mtime, ctime), kept beside the mappings this file hands out
 * (cgb_text's layout is shared with callers and stays three fields).  The token locator uses it to recognise the
 * same unchanged file when several ranks of one process ask about it. */
typedef struct { dev_t dev; ino_t ino; off_t size; struct timespec mtim, ctim; } file_id;
#define MAP_SLOTS 32
static pthread_mutex_t g_map_mu = PTHREAD_MUTEX_INITIALIZER;
static struct { const char *p; file_id id; } g_maps[MAP_SLOTS];

static void map_register(const char *p, const struct stat *sb)
{
    pthread_mutex_lock(&g_map_mu);
    for (int i = 0; i < MAP_SLOTS; ++i)
        if (!g_maps[i].p) {
            g_maps[i].p = p;
            g_maps[i].id.dev = sb->st_dev; g_maps[i].id.ino = sb->st_ino; g_maps[i].id.size = sb->st_size;
            g_maps[i].id.mtim = sb->st_mtim; g_maps[i].id.ctim = sb->st_ctim;
            break;
        }                                                  /* table full: this mapping is simply not shared */
    pthread_mutex_unlock(&g_map_mu);
}

static int map_lookup(const char *p, file_id *id)
{
    int found = 0;
    pthread_mutex_lock(&g_map_mu);
    for (int i = 0; i < MAP_SLOTS && !found; ++i)
        if (g_maps[i].p == p) { *id = g_maps[i].id; found = 1; }
    pthread_mutex_unlock(&g_map_mu);
    return found;
}

static void map_forget(const char *p)
{
    pthread_mutex_lock(&g_map_mu);
    for (int i = 0; i < MAP_SLOTS; ++i)
        if (g_maps[i].p == p) { g_maps[i].p = NULL; break; }
    pthread_mutex_unlock(&g_map_mu);
}

static int same_file(const file_id *a, const file_id *b)
{
    return a->dev == b->dev && a->ino == b->ino && a->size == b->size &&
           a->mtim.tv_sec == b->mtim.tv_sec && a->mtim.tv_nsec == b->mtim.tv_nsec &&
           a->ctim.tv_sec == b->ctim.tv_sec && a->ctim.tv_nsec == b->ctim.tv_nsec;
}

int cgb_text_open(const char *path, cgb_text *t)
{
    memset(t, 0, sizeof *t);
    int fd = open(path, O_RDONLY);
    if (fd < 0) return -1;
    struct stat sb;
    if (fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode) && sb.st_size > 0) {
        void *m = mmap(NULL, (size_t)sb.st_size, PROT_READ, MAP_PRIVATE, fd, 0);
        if (m != MAP_FAILED) {
            (void)madvise(m, (size_t)sb.st_size, MADV_SEQUENTIAL);
            t->p = m; t->len = (size_t)sb.st_size; t->mapped = 1;
            map_register(m, &sb);
            close(fd);
            return 0;
        }
    }
    close(fd);
    t->p = slurp(path, &t->len);
    return t->p ? 0 : -1;
}

void cgb_text_close(cgb_text *t)
{
    if (t->mapped) { map_forget(t->p); munmap(t->p, t->len); } else free(t->p);
    t->p = NULL;
}

/* k-th (0-based) token start inside [beg, end), or end. */
static const char *span_seek(const char *beg, const char *end, int64_t k)
{
    const char *p = beg;
#if defined(__SSE2__)
    /* whole 16-byte groups whose token starts all lie before the k-th are skipped by count */
    unsigned prev_sep = 1;
    while (end - p >= 16) {
        unsigned high;
        const unsigned sep = sep_mask16(p, &high);
        const int64_t c = __builtin_popcount(~sep & ((sep << 1) | prev_sep) & 0xFFFFu);
        if (c > k) break;                                  /* the k-th start is inside this group */
        k -= c;
        prev_sep = (sep >> 15) & 1u;
        p += 16;
    }
    if (!prev_sep)                                         /* inside a token counted already: step over its tail */
        while (p < end && !is_sep((unsigned char)*p)) ++p;
#endif
    for (;;) {
        while (p < end && is_sep((unsigned char)*p)) ++p;
        if (p >= end || k == 0) return p;
        while (p < end && !is_sep((unsigned char)*p)) ++p;
        --k;
    }
}

/* ---- locating a rank's rows.  The file is cut into blocks of LOC_BLOCK bytes (each cut moved forward to the next
 * separator, so no token straddles two blocks); the host threads count the token starts of every block; a rank's
 * first and one-past-last token are then found by a prefix sum over the blocks and a seek inside ONE block.
 *
 * The counts of the last file located are shared between the ranks of one process: `cg_b200 --gpus G` runs G
 * ranks as threads and every one of them needs the counts of the WHOLE file -- G passes over the same bytes at the
 * same time (8 x 470 MB for the n = 8192 file where one rank's own share is 59 MB).  The first caller counts while
 * the others wait on the mutex, then each seeks in its own mapping.  The entry is used only for the same
 * unchanged file (file_id) and never for slurped (unmapped) text; CGB_LOCATE_NO_CACHE=1 turns it off
 * (measurements of a single load: tools/ingest_rate.py). */
#define LOC_BLOCK ((size_t)1 << 20)
static pthread_mutex_t g_loc_mu = PTHREAD_MUTEX_INITIALIZER;
static struct {
    int valid, irregular;
    file_id id;
    size_t nblocks;
    int64_t *count;                                        /* token starts per block */
} g_loc;
static unsigned long g_loc_passes;                         /* counting passes made so far (under g_loc_mu) */

unsigned long cgb_text_locate_passes(void)
{
    pthread_mutex_lock(&g_loc_mu);
    const unsigned long k = g_loc_passes;
    pthread_mutex_unlock(&g_loc_mu);
    return k;
}

/* where block k begins: its nominal position moved forward to the next separator (a pure function of the text) */
static inline const char *block_cut(const char *buf, size_t len, size_t k)
{
    if (k == 0) return buf;
    const char *end = buf + len;
    if (k > len / LOC_BLOCK) return end;                  /* also keeps k * LOC_BLOCK from overflowing */
    const char *p = buf + k * LOC_BLOCK;
    while (p < end && !is_sep((unsigned char)*p)) ++p;
    return p;
}

typedef struct { const char *buf; size_t len, k0, k1; int64_t *count; int irregular; } block_job;
static void *block_count(void *arg)
{
    block_job *j = arg;
    unsigned junk = 0;
    const char *beg = block_cut(j->buf, j->len, j->k0);
    for (size_t k = j->k0; k < j->k1; ++k) {
        const char *nxt = block_cut(j->buf, j->len, k + 1);
        j->count[k] = count_range(beg, nxt, &junk);
        beg = nxt;
    }
    if (junk) j->irregular = 1;
    return NULL;
}

int64_t cgb_text_locate(const cgb_text *tx, int64_t first, int64_t count, size_t *beg_out, size_t *end_out)
{
    const char *buf = tx->p;
    const size_t len = tx->len;
    const size_t nblocks = len / LOC_BLOCK + 1;
    file_id id;
    const char *off = getenv("CGB_LOCATE_NO_CACHE");
    const int share = tx->mapped && !(off && *off && *off != '0') && map_lookup(buf, &id);
    int64_t *cnt = NULL;                                  /* the counts this call works with */
    int irregular = 0;
    *beg_out = *end_out = len;
    pthread_mutex_lock(&g_loc_mu);
    if (share && g_loc.valid && same_file(&g_loc.id, &id) && g_loc.nblocks == nblocks) {
        cnt = malloc(nblocks * sizeof *cnt);              /* a copy: the entry may be replaced once the mutex is gone */
        if (cnt) { memcpy(cnt, g_loc.count, nblocks * sizeof *cnt); irregular = g_loc.irregular; }
    }
    if (!cnt) {
        if (!share) pthread_mutex_unlock(&g_loc_mu);       /* nothing to share: count without holding the others up */
        cnt = malloc(nblocks * sizeof *cnt);
        if (!cnt) { if (share) pthread_mutex_unlock(&g_loc_mu); return -1; }   /* -1: the sequential path, which reports */
        int T = io_threads(len);
        if ((size_t)T > nblocks) T = (int)nblocks;
        block_job job[IO_MAX_THREADS];
        for (int t = 0; t < T; ++t) {
            job[t].buf = buf; job[t].len = len; job[t].count = cnt; job[t].irregular = 0;
            job[t].k0 = nblocks * (size_t)t / (size_t)T; job[t].k1 = nblocks * (size_t)(t + 1) / (size_t)T;
        }
        run_jobs(block_count, job, sizeof job[0], T);
        for (int t = 0; t < T; ++t) irregular |= job[t].irregular;
        if (!share) pthread_mutex_lock(&g_loc_mu);
        ++g_loc_passes;
        if (share) {
            int64_t *keep = malloc(nblocks * sizeof *keep);
            if (keep) {
                memcpy(keep, cnt, nblocks * sizeof *keep);
                free(g_loc.count);
                g_loc.count = keep; g_loc.nblocks = nblocks; g_loc.irregular = irregular; g_loc.id = id; g_loc.valid = 1;
            }
        }
    }
    pthread_mutex_unlock(&g_loc_mu);
    if (irregular) { free(cnt); return -1; }
    int64_t total = 0;
    for (size_t k = 0; k < nblocks; ++k) total += cnt[k];
    if (first < 0 || count < 0 || first + count > total) { free(cnt); return total; }
    const char *end = buf + len, *pb = end, *pe = end;
    const int64_t last = first + count;                    /* first token AFTER the range */
    int64_t before = 0;
    for (size_t k = 0; k < nblocks; ++k) {
        if (first >= before && first < before + cnt[k])
            pb = span_seek(block_cut(buf, len, k), block_cut(buf, len, k + 1), first - before);
        if (last >= before && last < before + cnt[k])
            pe = span_seek(block_cut(buf, len, k), block_cut(buf, len, k + 1), last - before);
        before += cnt[k];
    }
    if (count == 0) pe = pb;
    free(cnt);
    *beg_out = (size_t)(pb - buf);
    *end_out = (size_t)(pe - buf);
    return total;
}

int cgb_read_values(const char *path, double *dst, int64_t count, int64_t *got)
{
    text_t tx;
    if (got) *got = 0;
    if (!path || count < 0 || (count > 0 && !dst)) return CGB_ERR_ARG;
    if (cgb_text_open(path, &tx) != 0) return CGB_ERR_IO;
    int64_t k = parse_parallel(tx.p, tx.len, dst, count);
    int bad = 0;
    if (k < 0) {                                           /* irregular tokens: exact sequential semantics */
        cgb_text_close(&tx);
        size_t len = 0;
        char *buf = slurp(path, &len);
        if (!buf) return CGB_ERR_IO;
        k = parse_all(buf, len, dst, count, &bad);
        free(buf);
    } else {
        cgb_text_close(&tx);
        if (k > count) k = count;                          /* parse_all stops after `count` values too */
    }
    if (got) *got = k;
    return (k == count && !bad) ? CGB_OK : CGB_ERR_IO;
}

int64_t cgb_count_values(const char *path)
{
    text_t tx;
    if (cgb_text_open(path, &tx) != 0) return -1;
    int64_t k = parse_parallel(tx.p, tx.len, NULL, 0);
    cgb_text_close(&tx);
    if (k >= 0) return k;
    size_t len = 0;
    char *buf = slurp(path, &len);
    if (!buf) return -1;
    int bad = 0;
    k = parse_all(buf, len, NULL, 0, &bad);
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

/* generateSPDmatrix.m:28-43 writes one "%.4f\n" per value with fprintf; the same bytes are
 * produced here by one formatter thread per slice and written in order. */
/* "%.{d}f\n" without printf: glibc's conversion goes through multi-precision arithmetic, ~430 ns per value
 * (18 M values/s on 8 cores: the 470 MB matrix file of n = 8192 took 3.6 s to format).  For a finite double
 * below 2^63 / 10^d the digits follow from integer arithmetic alone: v = m 2^e with a 53-bit m, so
 * v 10^d = (m 10^d) >> -e, exact in 128 bits for d <= 17, and the bits shifted out decide the rounding
 * exactly as printf does in the default rounding mode (to nearest, ties to the even digit: 0.03125 -> "0.0312").
 * Returns the number of bytes written, or 0 when the value is not covered (inf, nan, huge): snprintf then. */
static size_t format_fixed(char *out, double v, int d)
{
    static const uint64_t p10[18] = {1ull, 10ull, 100ull, 1000ull, 10000ull, 100000ull, 1000000ull, 10000000ull, 100000000ull,
                                     1000000000ull, 10000000000ull, 100000000000ull, 1000000000000ull, 10000000000000ull,
                                     100000000000000ull, 1000000000000000ull, 10000000000000000ull, 100000000000000000ull};
    if (d < 0 || d > 17) return 0;
    uint64_t bits;
    memcpy(&bits, &v, sizeof bits);
    const int neg = (int)(bits >> 63);
    const int bexp = (int)((bits >> 52) & 0x7ff);
    uint64_t m = bits & 0xfffffffffffffull;
    if (bexp == 0x7ff) return 0;                           /* inf, nan */
    int e;                                                 /* v = m * 2^e */
    if (bexp == 0) e = -1074; else { m |= 1ull << 52; e = bexp - 1075; }
    uint64_t q;
    if (m == 0) q = 0;
    else if (e >= 0) return 0;                             /* >= 2^52: printf's business */
    else {
        const int sh = -e;                                 /* 1 .. 1074 */
        const unsigned __int128 w = (unsigned __int128)m * p10[d];          /* < 2^110 */
        if (sh >= 120) q = 0;                              /* w < 2^110 <= half of 2^sh: rounds to zero, never a tie */
        else {
            const unsigned __int128 hi = w >> sh;
            if (hi >> 63) return 0;                        /* integer part does not fit: printf's business */
            q = (uint64_t)hi;
            const unsigned __int128 rem = w - (hi << sh), half = (unsigned __int128)1 << (sh - 1);
            if (rem > half || (rem == half && (q & 1ull))) ++q;
        }
    }
    const uint64_t ip = q / p10[d];
    uint64_t fp = q % p10[d];
    char tmp[24];
    int k = 0;
    uint64_t t = ip;
    do { tmp[k++] = (char)('0' + t % 10); t /= 10; } while (t);
    char *o = out;
    if (neg) *o++ = '-';
    while (k) *o++ = tmp[--k];
    if (d > 0) {
        *o++ = '.';
        for (int i = d - 1; i >= 0; --i) { o[i] = (char)('0' + fp % 10); fp /= 10; }
        o += d;
    }
    *o++ = '\n';
    return (size_t)(o - out);
}

/* for the tests: one value as "%.{d}f\n" by the fast path (0 = not covered) */
size_t cgb_format_fixed(char *out, double v, int decimals) { return format_fixed(out, v, decimals); }

typedef struct { const double *v; int64_t count; int decimals; char *buf; size_t len; int fail; } fmt_job;

static void *fmt_main(void *arg)
{
    fmt_job *j = arg;
    size_t cap = (size_t)j->count * 28 + 64, n = 0;
    j->buf = malloc(cap);
    if (!j->buf) { j->fail = 1; return NULL; }
    for (int64_t i = 0; i < j->count; ++i) {
        if (cap - n < 400) {                               /* "%.4f" of 1e308 is 314 characters */
            cap = cap * 2 + 512;
            char *nb = realloc(j->buf, cap);
            if (!nb) { j->fail = 1; return NULL; }
            j->buf = nb;
        }
        if (j->decimals >= 0) {
            const size_t k = format_fixed(j->buf + n, j->v[i], j->decimals);
            n += k ? k : (size_t)snprintf(j->buf + n, cap - n, "%.*f\n", j->decimals, j->v[i]);
        } else
            n += (size_t)snprintf(j->buf + n, cap - n, "%.17g\n", j->v[i]);
    }
    j->len = n;
    return NULL;
}

int cgb_write_values(const char *path, const double *v, int64_t count, int fmt_decimals)
{
    return cgb_write_values_mode(path, v, count, fmt_decimals, 0);
}

int cgb_write_values_mode(const char *path, const double *v, int64_t count, int fmt_decimals, int append)
{
    FILE *f = fopen(path, append ? "a" : "w");
    if (!f) return CGB_ERR_IO;
    int rc = CGB_OK;
    const int64_t slab = (int64_t)1 << 22;                 /* values per round: bounds the buffers */
    for (int64_t base = 0; base < count && rc == CGB_OK; base += slab) {
        const int64_t cnt = count - base < slab ? count - base : slab;
        int T = io_threads((size_t)cnt * 16);
        fmt_job job[IO_MAX_THREADS];
        for (int t = 0; t < T; ++t) {
            const int64_t lo = cnt * t / T, hi = cnt * (t + 1) / T;
            memset(&job[t], 0, sizeof job[t]);
            job[t].v = v + base + lo; job[t].count = hi - lo; job[t].decimals = fmt_decimals;
        }
        run_jobs(fmt_main, job, sizeof job[0], T);
        for (int t = 0; t < T; ++t) {
            if (job[t].fail || (job[t].len && fwrite(job[t].buf, 1, job[t].len, f) != job[t].len)) rc = CGB_ERR_IO;
            free(job[t].buf);
        }
    }
    if (fclose(f) != 0) rc = CGB_ERR_IO;
    return rc;
}
