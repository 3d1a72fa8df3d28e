/*
 * lda_io.c -- the engine's file contract, host side (plain C).
 *
 *   read_settings   oni-lda-c/lda-estimate.c:331-350
 *   read_data       oni-lda-c/lda-data.c:22-59
 *   save_lda_model  oni-lda-c/lda-model.c:193-217
 *   save_gamma      oni-lda-c/lda-estimate.c:88-104
 *
 * Same grammar and byte-for-byte the same output text as the reference; unlike
 * the reference every fopen is checked and failures are reported.
 */
#define _GNU_SOURCE
#include <ctype.h>
#include <errno.h>
#include <math.h>
#include <pthread.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <fcntl.h>
#include <sys/stat.h>
#include <sys/types.h>
#include <unistd.h>

#include "../../include/oni_lda.h"

static __thread char g_err[512];

int oni_set_error(int code, const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof g_err, fmt, ap);
    va_end(ap);
    return code;
}

const char *oni_last_error(void) { return g_err; }
const char *oni_version(void) { return "oni-lda-b200 0.1"; }

/* ---------------------------------------------------------------- settings */
int oni_lda_read_settings(const char *path, oni_lda_settings *out)
{
    if (!path || !out) return oni_set_error(ONI_ERR_ARG, "read_settings: null argument");
    FILE *f = fopen(path, "r");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot open settings file %s: %s", path, strerror(errno));
    char action[100] = "";
    int v1 = 0, v3 = 0, ok = 0;
    float v2 = 0, v4 = 0;
    /* positional, exactly the five fscanf patterns of lda-estimate.c:336-340 */
    ok += fscanf(f, "var max iter %d\n", &v1);
    ok += fscanf(f, "var convergence %f\n", &v2);
    ok += fscanf(f, "em max iter %d\n", &v3);
    ok += fscanf(f, "em convergence %f\n", &v4);
    ok += fscanf(f, "alpha %99s", action);
    fclose(f);
    if (ok != 5) return oni_set_error(ONI_ERR_IO, "settings file %s: expected 5 fields, parsed %d", path, ok);
    out->var_max_iter = v1;
    out->var_converged = v2;
    out->em_max_iter = v3;
    out->em_converged = v4;
    out->estimate_alpha = (strcmp(action, "fixed") == 0) ? 0 : 1; /* :341-348 */
    return ONI_OK;
}

/* ------------------------------------------------------------------ corpus */
/* Whole-file read + hand-rolled integer scan: the grammar is "N" followed by N
 * "w:c" tokens, whitespace (blank or newline) separated, lda-data.c:35-51. */
typedef struct { int fd; char *dst; size_t off, n; int ok; } slurp_job;

static void *slurp_slice(void *arg)
{
    slurp_job *j = arg;
    size_t done = 0;
    j->ok = 1;
    while (done < j->n) {
        ssize_t r = pread(j->fd, j->dst + done, j->n - done, (off_t)(j->off + done));
        if (r <= 0) { j->ok = 0; break; }
        done += (size_t)r;
    }
    return NULL;
}

/* The whole file in one NUL-terminated buffer; a large file is read by a few threads, one slice
 * each (a 432 MB model.dat out of the page cache: 0.25 s with one fread, ~0.05 s this way). */
static char *slurp(const char *path, size_t *len)
{
    int fd = open(path, O_RDONLY);
    if (fd < 0) return NULL;
    struct stat st;
    if (fstat(fd, &st) != 0 || st.st_size < 0) { close(fd); return NULL; }
    const size_t n = (size_t)st.st_size;
    char *buf = malloc(n + 1);
    if (!buf) { close(fd); errno = ENOMEM; return NULL; }
    enum { MAXT = 16 };
    long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
    int nt = (int)(ncpu < 1 ? 1 : (ncpu > MAXT ? MAXT : ncpu));
    if (n < ((size_t)8 << 20)) nt = 1;
    slurp_job job[MAXT];
    pthread_t th[MAXT];
    int ok = 1;
    for (int t = 0; t < nt; t++) {
        const size_t a = n / (size_t)nt * (size_t)t, b = (t == nt - 1) ? n : n / (size_t)nt * (size_t)(t + 1);
        job[t].fd = fd; job[t].dst = buf + a; job[t].off = a; job[t].n = b - a;
        if (nt == 1 || pthread_create(&th[t], NULL, slurp_slice, &job[t]) != 0) { slurp_slice(&job[t]); th[t] = 0; }
    }
    for (int t = 0; t < nt; t++) {
        if (nt > 1 && th[t]) pthread_join(th[t], NULL);
        ok &= job[t].ok;
    }
    close(fd);
    if (!ok) { free(buf); errno = EIO; return NULL; }
    buf[n] = 0;
    *len = n;
    return buf;
}

static inline const char *skip_ws(const char *p)
{
    while (*p == ' ' || *p == '\n' || *p == '\t' || *p == '\r') p++;
    return p;
}

/* parses an optionally signed decimal; returns NULL if none */
static inline const char *scan_int(const char *p, long long *out)
{
    int neg = 0;
    if (*p == '-') { neg = 1; p++; } else if (*p == '+') p++;
    if (*p < '0' || *p > '9') return NULL;
    long long v = 0;
    while (*p >= '0' && *p <= '9') { v = v * 10 + (*p - '0'); p++; }
    *out = neg ? -v : v;
    return p;
}

/* Parallel path for the shape lda_pre.py writes (one document per line, lda_pre.py:94):
 * the buffer is cut at newlines into one piece per thread.  Pass 1 counts each piece's documents
 * (non-blank lines) and pairs (':' characters), a prefix sum gives every piece its place in the
 * final arrays, pass 2 parses straight into them -- no per-thread arrays, no merge copy.  A line
 * that does not hold exactly "N" and N "w:c" tokens, or a piece whose parse disagrees with its
 * counts, makes the caller fall back to the sequential parser below, which accepts any whitespace
 * like fscanf does. */
typedef struct {
    const char *beg, *end;
    size_t ndocs, npairs;        /* pass 1 */
    int64_t *len;                /* pass 2: pairs per document, at this piece's place */
    int32_t *word, *count;
    long long nterms;
    int ok;
} parse_piece;

static void *count_piece(void *arg)
{
    parse_piece *pc = arg;
    size_t docs = 0, pairs = 0;
    int seen = 0;
    for (const char *p = pc->beg; p < pc->end; p++) {
        const char ch = *p;
        if (ch == '\n') { docs += (size_t)seen; seen = 0; }
        else if (ch == ':') { pairs++; seen = 1; }
        else if (ch != ' ' && ch != '\t' && ch != '\r') seen = 1;
    }
    pc->ndocs = docs + (size_t)seen;
    pc->npairs = pairs;
    return NULL;
}

static void *parse_lines(void *arg)
{
    parse_piece *pc = arg;
    const char *p = pc->beg;
    size_t nd = 0, np = 0;
    long long nterms = 0;
    pc->ok = 0;
    while (p < pc->end) {
        while (p < pc->end && (*p == ' ' || *p == '\t' || *p == '\r')) p++;
        if (p >= pc->end) break;
        if (*p == '\n') { p++; continue; }  /* blank line */
        long long n;
        const char *q = scan_int(p, &n);
        if (!q || n < 0) return NULL;
        p = q;
        if (nd >= pc->ndocs || np + (size_t)n > pc->npairs) return NULL;  /* disagrees with pass 1 */
        for (long long i = 0; i < n; i++) {
            long long w, c;
            while (*p == ' ' || *p == '\t') p++;
            q = scan_int(p, &w);
            if (!q || *q != ':' || w < 0 || w > 0x7ffffffe) return NULL;
            q = scan_int(q + 1, &c);
            if (!q) return NULL;
            p = q;
            pc->word[np] = (int32_t)w;
            pc->count[np] = (int32_t)c;
            np++;
            if (w >= nterms) nterms = w + 1;
        }
        while (*p == ' ' || *p == '\t' || *p == '\r') p++;
        if (*p == '\n') p++;
        else if (*p != 0) return NULL;  /* more tokens on the line: not line-shaped */
        pc->len[nd++] = n;
    }
    pc->nterms = nterms;
    pc->ok = (nd == pc->ndocs && np == pc->npairs);
    return NULL;
}

static void run_pieces(parse_piece *pc, int nt, void *(*fn)(void *))
{
    pthread_t th[32];
    for (int t = 0; t < nt; t++)
        if (nt == 1 || pthread_create(&th[t], NULL, fn, &pc[t]) != 0) { fn(&pc[t]); th[t] = 0; }
    for (int t = 0; t < nt; t++)
        if (nt > 1 && th[t]) pthread_join(th[t], NULL);
}

static int read_corpus_parallel(char *buf, size_t len, oni_lda_corpus *out)
{
    enum { MAXT = 32 };
    long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
    int nt = (int)(ncpu < 1 ? 1 : (ncpu > MAXT ? MAXT : ncpu));
    if (len < ((size_t)1 << 20)) nt = 1;
    parse_piece pc[MAXT];
    memset(pc, 0, sizeof pc);
    const char *cut = buf;
    for (int t = 0; t < nt; t++) {
        pc[t].beg = cut;
        const char *want = (t == nt - 1) ? buf + len : buf + (len / (size_t)nt) * (size_t)(t + 1);
        if (want < cut) want = cut;
        while (want < buf + len && *want != '\n') want++;
        if (want < buf + len) want++;  /* the piece ends after its newline */
        pc[t].end = cut = want;
    }
    run_pieces(pc, nt, count_piece);
    size_t ndocs = 0, npairs = 0;
    for (int t = 0; t < nt; t++) { ndocs += pc[t].ndocs; npairs += pc[t].npairs; }
    if (ndocs > 0x7fffffffu) return 1;
    int64_t *ptr = malloc(sizeof(int64_t) * (ndocs + 1));
    int32_t *word = malloc(sizeof(int32_t) * (npairs ? npairs : 1));
    int32_t *count = malloc(sizeof(int32_t) * (npairs ? npairs : 1));
    if (!ptr || !word || !count) { free(ptr); free(word); free(count); return 1; }
    size_t d = 0, k = 0;
    for (int t = 0; t < nt; t++) {
        pc[t].len = ptr + 1 + d; pc[t].word = word + k; pc[t].count = count + k;
        d += pc[t].ndocs; k += pc[t].npairs;
    }
    run_pieces(pc, nt, parse_lines);
    int ok = 1;
    long long nterms = 0;
    for (int t = 0; t < nt; t++) {
        ok &= pc[t].ok;
        if (pc[t].nterms > nterms) nterms = pc[t].nterms;
    }
    if (!ok) { free(ptr); free(word); free(count); return 1; }  /* not line-shaped: the sequential parser decides */
    ptr[0] = 0;
    for (size_t i = 1; i <= ndocs; i++) ptr[i] += ptr[i - 1];  /* lengths -> offsets */
    out->num_docs = (int32_t)ndocs; out->num_terms = (int32_t)nterms; out->nnz = (int64_t)npairs;
    out->doc_ptr = ptr; out->word = word; out->count = count;
    return 0;
}

int oni_lda_read_corpus(const char *path, oni_lda_corpus *out)
{
    if (!path || !out) return oni_set_error(ONI_ERR_ARG, "read_corpus: null argument");
    memset(out, 0, sizeof *out);
    size_t len = 0;
    char *buf = slurp(path, &len);
    if (!buf) return oni_set_error(ONI_ERR_IO, "cannot read corpus %s: %s", path, strerror(errno));
    if (read_corpus_parallel(buf, len, out) == 0) { free(buf); return ONI_OK; }
    memset(out, 0, sizeof *out);

    /* sequential, whitespace-tolerant: "N" then N "w:c" tokens, lda-data.c:35-51 */
    /* pass 1: count documents and pairs so each array is allocated once */
    size_t ndocs = 0, npairs = 0;
    {
        const char *p = skip_ws(buf);
        while (*p) {
            long long n;
            const char *q = scan_int(p, &n);
            if (!q || n < 0) { const size_t at = (size_t)(p - buf); free(buf); return oni_set_error(ONI_ERR_IO, "%s: bad document length at byte %zu", path, at); }
            p = q;
            for (long long i = 0; i < n; i++) {
                long long w, c;
                p = skip_ws(p);
                q = scan_int(p, &w);
                if (!q || *q != ':') { const size_t at = (size_t)(p - buf); free(buf); return oni_set_error(ONI_ERR_IO, "%s: bad word:count at byte %zu", path, at); }
                q = scan_int(q + 1, &c);
                if (!q) { const size_t at = (size_t)(p - buf); free(buf); return oni_set_error(ONI_ERR_IO, "%s: bad count at byte %zu", path, at); }
                p = q;
            }
            ndocs++;
            npairs += (size_t)n;
            p = skip_ws(p);
        }
    }
    if (ndocs > 0x7fffffffu) { free(buf); return oni_set_error(ONI_ERR_ARG, "%s: too many documents", path); }
    int64_t *ptr = malloc(sizeof(int64_t) * (ndocs + 1));
    int32_t *word = malloc(sizeof(int32_t) * (npairs ? npairs : 1));
    int32_t *count = malloc(sizeof(int32_t) * (npairs ? npairs : 1));
    if (!ptr || !word || !count) {
        free(buf); free(ptr); free(word); free(count);
        return oni_set_error(ONI_ERR_NOMEM, "read_corpus: out of memory");
    }
    /* pass 2: fill */
    size_t d = 0, k = 0;
    long long nterms = 0;
    const char *p = skip_ws(buf);
    ptr[0] = 0;
    while (*p) {
        long long n = 0, w = 0, c = 0;
        p = scan_int(p, &n);
        for (long long i = 0; i < n; i++) {
            p = skip_ws(p);
            p = scan_int(p, &w);
            p = scan_int(p + 1, &c);
            if (w < 0 || w > 0x7ffffffe) {
                free(buf); free(ptr); free(word); free(count);
                return oni_set_error(ONI_ERR_IO, "%s: word id %lld out of range", path, w);
            }
            word[k] = (int32_t)w;
            count[k] = (int32_t)c;
            if (w >= nterms) nterms = w + 1; /* lda-data.c:49 */
            k++;
        }
        ptr[++d] = (int64_t)k;
        p = skip_ws(p);
    }
    free(buf);
    out->num_docs = (int32_t)ndocs;
    out->num_terms = (int32_t)nterms;
    out->nnz = (int64_t)npairs;
    out->doc_ptr = ptr;
    out->word = word;
    out->count = count;
    return ONI_OK;
}

void oni_lda_free_corpus(oni_lda_corpus *c)
{
    if (!c) return;
    free((void *)c->doc_ptr);
    free((void *)c->word);
    free((void *)c->count);
    memset(c, 0, sizeof *c);
}

/* ----------------------------------------------------------------- writers */
/* "%5.10f" without printf.  The reference prints every beta and gamma value with
 * this format (lda-model.c:205, lda-estimate.c:97-100); at flow-day size that is
 * 21 M values and printf's generic path dominates the whole run.  The conversion
 * below is exact: v = m * 2^e with a 53-bit m, so m * 10^10 fits 128 bits and the
 * 10-decimal value is that product shifted by e, rounded half to even on the true
 * remainder -- the same digits glibc prints.  Anything outside the fast range
 * (non-finite, |v| >= 2^62 / 1e10 ... or so small that the shift leaves 128 bits)
 * goes through snprintf. */
typedef unsigned __int128 u128;

static inline int fmt_u64(char *out, unsigned long long v)
{
    char tmp[24];
    int n = 0;
    do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    for (int i = 0; i < n; i++) out[i] = tmp[n - 1 - i];
    return n;
}

/* writes the text of printf("%5.10f", v) at out (no terminator); returns its length */
static int format_fixed10(char *out, double v)
{
    union { double d; unsigned long long u; } b;
    b.d = v;
    const int neg = (int)(b.u >> 63);
    const int ex = (int)((b.u >> 52) & 0x7ff);
    unsigned long long m = b.u & 0xfffffffffffffULL;
    if (ex == 0x7ff) return snprintf(out, 64, "%5.10f", v);  /* inf / nan */
    int e;                                                   /* v = m * 2^e */
    if (ex == 0) e = -1074; else { m |= 1ULL << 52; e = ex - 1075; }
    u128 scaled;  /* round(|v| * 1e10) */
    if (e >= 0) {
        if (e > 10) return snprintf(out, 64, "%5.10f", v);   /* |v| >= 2^63: rare, let libc do it */
        scaled = ((u128)m * 10000000000ULL) << e;
    } else {
        const int sh = -e;
        const u128 prod = (u128)m * 10000000000ULL;          /* < 2^87 */
        if (sh >= 128) scaled = 0;                           /* |v| < 2^-75: rounds to zero */
        else {
            scaled = prod >> sh;
            const u128 rem = prod & ((((u128)1) << sh) - 1);
            const u128 half = ((u128)1) << (sh - 1);
            if (rem > half || (rem == half && (scaled & 1))) scaled++;  /* half to even */
        }
    }
    const unsigned long long ip = (unsigned long long)(scaled / 10000000000ULL);
    unsigned long long fp = (unsigned long long)(scaled % 10000000000ULL);
    int n = 0;
    if (neg) out[n++] = '-';
    n += fmt_u64(out + n, ip);
    out[n++] = '.';
    for (int i = 9; i >= 0; i--) { out[n + i] = (char)('0' + fp % 10); fp /= 10; }
    n += 10;
    return n;  /* always >= 12 characters, so the minimum width 5 never pads */
}

/* A matrix as text: rows [r0, r1), every value with `lead` before it (beta: a blank
 * before every value; gamma: a blank between values), '\n' after each row. */
typedef struct {
    const double *a;
    int cols, r0, r1, lead_first;
    char *buf;
    size_t len;
} fmt_job;

static void *fmt_rows(void *arg)
{
    fmt_job *j = arg;
    char *p = j->buf;
    for (int r = j->r0; r < j->r1; r++) {
        const double *row = j->a + (size_t)r * j->cols;
        for (int c = 0; c < j->cols; c++) {
            if (c > 0 || j->lead_first) *p++ = ' ';
            p += format_fixed10(p, row[c]);
        }
        *p++ = '\n';
    }
    j->len = (size_t)(p - j->buf);
    return NULL;
}

static int write_matrix(const char *path, const double *a, int rows, int cols, int lead_first)
{
    FILE *f = fopen(path, "w");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", path, strerror(errno));
    /* values the fast formatter hands to libc (non-finite, |v| >= 2^63) can be hundreds of
     * characters long; a matrix holding any is written the plain way */
    int plain = 0;
    for (size_t i = 0, n = (size_t)rows * cols; i < n; i++)
        if (!(fabs(a[i]) < 9.0e18)) { plain = 1; break; }
    if (plain) {
        setvbuf(f, NULL, _IOFBF, 1 << 20);
        for (int r = 0; r < rows; r++) {
            for (int c = 0; c < cols; c++)
                fprintf(f, (c > 0 || lead_first) ? " %5.10f" : "%5.10f", a[(size_t)r * cols + c]);
            fputc('\n', f);
        }
        if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %s", path);
        return ONI_OK;
    }
    /* blocks of rows formatted by a few threads, written in order */
    enum { MAXT = 8 };
    const size_t per_value = 32;  /* sign + up to 19 digits + '.' + 10 + blank, generous */
    const int block = (int)(((size_t)1 << 22) / ((size_t)cols * per_value + 1)) + 1;  /* ~4 MB of text */
    long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
    int nt = (int)(ncpu < 1 ? 1 : (ncpu > MAXT ? MAXT : ncpu));
    if (rows <= block) nt = 1;
    fmt_job job[MAXT];
    pthread_t th[MAXT];
    int rc = ONI_OK;
    for (int t = 0; t < nt; t++) {
        job[t].buf = malloc((size_t)block * ((size_t)cols * per_value + 1));
        if (!job[t].buf) { rc = oni_set_error(ONI_ERR_NOMEM, "save: out of memory"); nt = t; break; }
    }
    for (int r = 0; rc == ONI_OK && r < rows; ) {
        int used = 0;
        for (int t = 0; t < nt && r < rows; t++, used++) {
            job[t].a = a; job[t].cols = cols; job[t].lead_first = lead_first;
            job[t].r0 = r; job[t].r1 = r + block < rows ? r + block : rows;
            r = job[t].r1;
            if (nt == 1) fmt_rows(&job[t]);
            else if (pthread_create(&th[t], NULL, fmt_rows, &job[t]) != 0) { fmt_rows(&job[t]); th[t] = 0; }
        }
        for (int t = 0; t < used; t++) {
            if (nt > 1 && th[t]) pthread_join(th[t], NULL);
            if (fwrite(job[t].buf, 1, job[t].len, f) != job[t].len) rc = oni_set_error(ONI_ERR_IO, "write failed on %s", path);
        }
    }
    for (int t = 0; t < nt; t++) free(job[t].buf);
    if (fclose(f) != 0 && rc == ONI_OK) rc = oni_set_error(ONI_ERR_IO, "write failed on %s", path);
    return rc;
}

int oni_lda_save_model(const char *root, int32_t k, int32_t v, const double *logbeta, double alpha)
{
    char name[4096];
    if (!root) return oni_set_error(ONI_ERR_ARG, "save_model: null root");
    if (logbeta) {
        snprintf(name, sizeof name, "%s.beta", root);
        int rc = write_matrix(name, logbeta, k, v, 1); /* " %5.10f" per value, lda-model.c:205 */
        if (rc != ONI_OK) return rc;
    }
    snprintf(name, sizeof name, "%s.other", root);
    FILE *f = fopen(name, "w");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", name, strerror(errno));
    fprintf(f, "num_topics %d\n", k);   /* lda-model.c:213-215 */
    fprintf(f, "num_terms %d\n", v);
    fprintf(f, "alpha %5.10f\n", alpha);
    if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %s", name);
    return ONI_OK;
}

int oni_lda_save_gamma(const char *path, int32_t d, int32_t k, const double *gamma)
{
    if (!path || !gamma) return oni_set_error(ONI_ERR_ARG, "save_gamma: null argument");
    /* "%5.10f" then " %5.10f" (lda-estimate.c:97-100) */
    return write_matrix(path, gamma, d, k, 0);
}

/* ----------------------------------------------------------- model.dat writer */
/* The inverse of oni_lda_read_corpus: "N w:c w:c ...\n" per document, the text lda_pre.py:94
 * writes and read_data (lda-data.c:35-51) parses.  Documents are cut into a few ranges, each
 * formatted by its own thread into its own buffer; the buffers are written in order. */
typedef struct { const oni_lda_corpus *c; int32_t d0, d1; char *buf; size_t len; int failed; } wc_job;

static char *put_u32(char *p, uint32_t v)
{
    char tmp[10];
    int n = 0;
    do { tmp[n++] = (char)('0' + v % 10); v /= 10; } while (v);
    while (n) *p++ = tmp[--n];
    return p;
}

static void *write_corpus_range(void *arg)
{
    wc_job *j = (wc_job *)arg;
    const oni_lda_corpus *c = j->c;
    const int64_t a = c->doc_ptr[j->d0], b = c->doc_ptr[j->d1];
    /* per pair: blank + <=10 digits + ':' + <=10 digits; per document: <=10 digits + newline */
    const size_t cap = (size_t)(b - a) * 22 + (size_t)(j->d1 - j->d0) * 11 + 16;
    j->buf = (char *)malloc(cap);
    if (!j->buf) { j->failed = 1; return NULL; }
    char *p = j->buf;
    for (int32_t d = j->d0; d < j->d1; d++) {
        const int64_t lo = c->doc_ptr[d], hi = c->doc_ptr[d + 1];
        p = put_u32(p, (uint32_t)(hi - lo));
        for (int64_t i = lo; i < hi; i++) {
            *p++ = ' ';
            p = put_u32(p, (uint32_t)c->word[i]);
            *p++ = ':';
            p = put_u32(p, (uint32_t)c->count[i]);
        }
        *p++ = '\n';
    }
    j->len = (size_t)(p - j->buf);
    return NULL;
}

int oni_lda_write_corpus(const char *path, const oni_lda_corpus *c)
{
    if (!path || !c || !c->doc_ptr || c->num_docs < 0) return oni_set_error(ONI_ERR_ARG, "write_corpus: bad argument");
    for (int64_t i = 0; i < c->nnz; i++)
        if (c->word[i] < 0 || c->count[i] < 0) return oni_set_error(ONI_ERR_ARG, "write_corpus: negative word id or count");
    FILE *f = fopen(path, "w");
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", path, strerror(errno));
    enum { MAXT = 16 };
    long ncpu = sysconf(_SC_NPROCESSORS_ONLN);
    int nt = (int)(ncpu < 1 ? 1 : (ncpu > MAXT ? MAXT : ncpu));
    if (c->nnz < (1 << 16)) nt = 1;
    wc_job job[MAXT];
    pthread_t th[MAXT];
    int rc = ONI_OK;
    for (int t = 0; t < nt; t++) {
        /* ranges balanced by pairs */
        const int64_t target0 = c->nnz * t / nt, target1 = c->nnz * (t + 1) / nt;
        int32_t d0 = 0, d1 = c->num_docs;
        if (t > 0) { int32_t lo = 0, hi = c->num_docs; while (lo < hi) { int32_t m = lo + (hi - lo) / 2; if (c->doc_ptr[m] < target0) lo = m + 1; else hi = m; } d0 = lo; }
        if (t < nt - 1) { int32_t lo = 0, hi = c->num_docs; while (lo < hi) { int32_t m = lo + (hi - lo) / 2; if (c->doc_ptr[m] < target1) lo = m + 1; else hi = m; } d1 = lo; }
        job[t].c = c; job[t].d0 = d0; job[t].d1 = d1 < d0 ? d0 : d1; job[t].buf = NULL; job[t].len = 0; job[t].failed = 0;
        if (nt == 1 || pthread_create(&th[t], NULL, write_corpus_range, &job[t]) != 0) { write_corpus_range(&job[t]); th[t] = 0; }
    }
    for (int t = 0; t < nt; t++) {
        if (nt > 1 && th[t]) pthread_join(th[t], NULL);
        if (job[t].failed) rc = oni_set_error(ONI_ERR_NOMEM, "write_corpus: out of memory");
        else if (rc == ONI_OK && fwrite(job[t].buf, 1, job[t].len, f) != job[t].len)
            rc = oni_set_error(ONI_ERR_IO, "write failed on %s", path);
        free(job[t].buf);
    }
    if (fclose(f) != 0 && rc == ONI_OK) rc = oni_set_error(ONI_ERR_IO, "write failed on %s", path);
    return rc;
}

/* Test hook: the fast formatter against the C library on one value. */
int oni_internal_format_fixed10(double v, char *out64)
{
    if (!(fabs(v) < 9.0e18)) return -1;  /* the writer sends these through fprintf */
    int n = format_fixed10(out64, v);
    out64[n] = 0;
    return n;
}

/* ------------------------------------------------------------- lda_pre.py */
/* The corpus indexing stage of the pipeline (lda_pre.py:30-94), byte-identical outputs:
 *   <dir>doc_wc.dat  rows "ip,word,count"                       (input, ml_ops.sh:59-62)
 *   <dir>words.dat   "<word id>,<word>"   ids from 0, first appearance   (lda_pre.py:33-44)
 *   <dir>doc.dat     "<doc id>,<ip>"      ids from 1, first appearance   (lda_pre.py:57-73)
 *   <dir>model.dat   "<pairs> w:c w:c ..." one line per document in doc.dat order; a word
 *                    repeated inside a document stays a separate pair     (lda_pre.py:66-94)
 * `dir` is used as a prefix exactly like the script's sys.argv[1] (ml_ops.sh passes "path/").
 * Fields are split at commas; a field wrapped in double quotes may contain commas (the csv
 * module's default dialect); the count is copied through as text. */
static FILE *open_out(const char *path)
{
    FILE *f = fopen(path, "w");
    if (f) setvbuf(f, NULL, _IOFBF, 1 << 20);
    return f;
}

typedef struct { const char *s; uint32_t len; int32_t id; } pre_key;
typedef struct { pre_key *slot; size_t cap, used; } pre_map;

static uint64_t pre_hash(const char *s, uint32_t n)
{
    uint64_t h = 1469598103934665603ULL;
    for (uint32_t i = 0; i < n; i++) { h ^= (unsigned char)s[i]; h *= 1099511628211ULL; }
    return h;
}

static int pre_map_grow(pre_map *m)
{
    size_t ncap = m->cap ? m->cap * 2 : (1u << 16);
    pre_key *ns = calloc(ncap, sizeof(pre_key));
    if (!ns) return 0;
    for (size_t i = 0; i < m->cap; i++)
        if (m->slot[i].s) {
            size_t j = pre_hash(m->slot[i].s, m->slot[i].len) & (ncap - 1);
            while (ns[j].s) j = (j + 1) & (ncap - 1);
            ns[j] = m->slot[i];
        }
    free(m->slot);
    m->slot = ns; m->cap = ncap;
    return 1;
}

/* id of the key, inserting it with id = next_id if new (*is_new says which) */
static int32_t pre_map_get(pre_map *m, const char *s, uint32_t n, int32_t next_id, int *is_new)
{
    if ((m->used + 1) * 10 >= m->cap * 7 && !pre_map_grow(m)) return -1;
    size_t j = pre_hash(s, n) & (m->cap - 1);
    while (m->slot[j].s) {
        if (m->slot[j].len == n && memcmp(m->slot[j].s, s, n) == 0) { *is_new = 0; return m->slot[j].id; }
        j = (j + 1) & (m->cap - 1);
    }
    m->slot[j].s = s; m->slot[j].len = n; m->slot[j].id = next_id;
    m->used++;
    *is_new = 1;
    return next_id;
}

/* one csv field starting at p (stops at ',' or end of line); returns the char after it */
static const char *pre_field(const char *p, const char *eol, const char **fs, uint32_t *fl)
{
    if (p < eol && *p == '"') {
        const char *q = p + 1;
        while (q < eol && *q != '"') q++;
        *fs = p + 1; *fl = (uint32_t)(q - p - 1);
        p = (q < eol) ? q + 1 : q;
        while (p < eol && *p != ',') p++;
    } else {
        const char *q = p;
        while (q < eol && *q != ',') q++;
        *fs = p; *fl = (uint32_t)(q - p);
        p = q;
    }
    return (p < eol) ? p + 1 : p;
}

int oni_lda_pre(const char *dir)
{
    if (!dir) return oni_set_error(ONI_ERR_ARG, "lda_pre: null directory prefix");
    char name[4096];
    snprintf(name, sizeof name, "%sdoc_wc.dat", dir);
    size_t len = 0;
    char *buf = slurp(name, &len);
    if (!buf) return oni_set_error(ONI_ERR_IO, "cannot read %s: %s", name, strerror(errno));
    int rc = ONI_OK;
    pre_map words = {0}, docs = {0};
    size_t nrows = 0, cap_rows = 1 << 20;
    int32_t *row_doc = malloc(sizeof(int32_t) * cap_rows), *row_word = malloc(sizeof(int32_t) * cap_rows);
    const char **row_cnt = malloc(sizeof(char *) * cap_rows);
    uint32_t *row_cnt_len = malloc(sizeof(uint32_t) * cap_rows);
    FILE *fw = NULL, *fd = NULL, *fm = NULL;
    int64_t *doc_ptr = NULL, *cursor = NULL;
    uint32_t *order = NULL;
    if (!row_doc || !row_word || !row_cnt || !row_cnt_len) { rc = oni_set_error(ONI_ERR_NOMEM, "lda_pre: out of memory"); goto done; }
    snprintf(name, sizeof name, "%swords.dat", dir);
    fw = open_out(name);
    snprintf(name, sizeof name, "%sdoc.dat", dir);
    fd = open_out(name);
    if (!fw || !fd) { rc = oni_set_error(ONI_ERR_IO, "lda_pre: cannot write under %s: %s", dir, strerror(errno)); goto done; }
    int32_t n_words = 0, n_docs = 0;
    for (const char *p = buf, *end = buf + len; p < end;) {
        const char *eol = memchr(p, '\n', (size_t)(end - p));
        if (!eol) eol = end;
        const char *line_end = (eol > p && eol[-1] == '\r') ? eol - 1 : eol;
        if (line_end > p) {
            const char *f[3]; uint32_t fl[3];
            const char *q = p;
            int nf = 0;
            while (nf < 3) {  /* the first three fields; anything after them is ignored */
                const int had_comma = (q > p && q[-1] == ',');
                if (q >= line_end && !(nf > 0 && had_comma)) break;
                q = pre_field(q, line_end, &f[nf], &fl[nf]);
                nf++;
            }
            if (nf < 3) { rc = oni_set_error(ONI_ERR_IO, "doc_wc.dat row %zu: expected ip,word,count", nrows + 1); goto done; }
            if (nrows == cap_rows) {
                cap_rows *= 2;
                row_doc = realloc(row_doc, sizeof(int32_t) * cap_rows); row_word = realloc(row_word, sizeof(int32_t) * cap_rows);
                row_cnt = realloc(row_cnt, sizeof(char *) * cap_rows); row_cnt_len = realloc(row_cnt_len, sizeof(uint32_t) * cap_rows);
                if (!row_doc || !row_word || !row_cnt || !row_cnt_len) { rc = oni_set_error(ONI_ERR_NOMEM, "lda_pre: out of memory"); goto done; }
            }
            int is_new;
            int32_t w = pre_map_get(&words, f[1], fl[1], n_words, &is_new);
            if (w < 0) { rc = oni_set_error(ONI_ERR_NOMEM, "lda_pre: out of memory"); goto done; }
            if (is_new) { fprintf(fw, "%d,%.*s\n", w, (int)fl[1], f[1]); n_words++; }           /* lda_pre.py:40 */
            int32_t d = pre_map_get(&docs, f[0], fl[0], n_docs + 1, &is_new);                       /* ids from 1, :61 */
            if (d < 0) { rc = oni_set_error(ONI_ERR_NOMEM, "lda_pre: out of memory"); goto done; }
            if (is_new) { fprintf(fd, "%d,%.*s\n", d, (int)fl[0], f[0]); n_docs++; }              /* :71 */
            row_doc[nrows] = d - 1; row_word[nrows] = w; row_cnt[nrows] = f[2]; row_cnt_len[nrows] = fl[2];
            nrows++;
        }
        p = eol + 1;
    }
    /* rows grouped by document, order of appearance kept (stable counting sort) */
    doc_ptr = calloc((size_t)n_docs + 1, sizeof(int64_t));
    cursor = malloc(sizeof(int64_t) * ((size_t)n_docs + 1));
    order = malloc(sizeof(uint32_t) * (nrows ? nrows : 1));
    if (!doc_ptr || !cursor || !order) { rc = oni_set_error(ONI_ERR_NOMEM, "lda_pre: out of memory"); goto done; }
    for (size_t i = 0; i < nrows; i++) doc_ptr[row_doc[i] + 1]++;
    for (int32_t d = 0; d < n_docs; d++) doc_ptr[d + 1] += doc_ptr[d];
    memcpy(cursor, doc_ptr, sizeof(int64_t) * ((size_t)n_docs + 1));
    for (size_t i = 0; i < nrows; i++) order[cursor[row_doc[i]]++] = (uint32_t)i;
    snprintf(name, sizeof name, "%smodel.dat", dir);
    fm = open_out(name);
    if (!fm) { rc = oni_set_error(ONI_ERR_IO, "cannot write %s: %s", name, strerror(errno)); goto done; }
    for (int32_t d = 0; d < n_docs; d++) {
        fprintf(fm, "%lld", (long long)(doc_ptr[d + 1] - doc_ptr[d]));                             /* lda_pre.py:94 */
        for (int64_t k = doc_ptr[d]; k < doc_ptr[d + 1]; k++) {
            const uint32_t i = order[k];
            fprintf(fm, " %d:%.*s", row_word[i], (int)row_cnt_len[i], row_cnt[i]);               /* :66 */
        }
        fputc('\n', fm);
    }
done:
    if (fw && fclose(fw) != 0 && rc == ONI_OK) rc = oni_set_error(ONI_ERR_IO, "lda_pre: write failed (words.dat)");
    if (fd && fclose(fd) != 0 && rc == ONI_OK) rc = oni_set_error(ONI_ERR_IO, "lda_pre: write failed (doc.dat)");
    if (fm && fclose(fm) != 0 && rc == ONI_OK) rc = oni_set_error(ONI_ERR_IO, "lda_pre: write failed (model.dat)");
    free(words.slot); free(docs.slot); free(row_doc); free(row_word); free(row_cnt); free(row_cnt_len);
    free(doc_ptr); free(cursor); free(order); free(buf);
    return rc;
}
