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
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

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
static char *slurp(const char *path, size_t *len)
{
    FILE *f = fopen(path, "rb");
    if (!f) return NULL;
    if (fseek(f, 0, SEEK_END) != 0) { fclose(f); return NULL; }
    long n = ftell(f);
    if (n < 0) { fclose(f); return NULL; }
    rewind(f);
    char *buf = malloc((size_t)n + 1);
    if (!buf) { fclose(f); errno = ENOMEM; return NULL; }
    size_t got = fread(buf, 1, (size_t)n, f);
    fclose(f);
    buf[got] = 0;
    *len = got;
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

int oni_lda_read_corpus(const char *path, oni_lda_corpus *out)
{
    if (!path || !out) return oni_set_error(ONI_ERR_ARG, "read_corpus: null argument");
    memset(out, 0, sizeof *out);
    size_t len = 0;
    char *buf = slurp(path, &len);
    if (!buf) return oni_set_error(ONI_ERR_IO, "cannot read corpus %s: %s", path, strerror(errno));

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
static FILE *open_out(const char *path)
{
    FILE *f = fopen(path, "w");
    if (f) setvbuf(f, NULL, _IOFBF, 1 << 20);
    return f;
}

int oni_lda_save_model(const char *root, int32_t k, int32_t v, const double *logbeta, double alpha)
{
    char name[4096];
    if (!root) return oni_set_error(ONI_ERR_ARG, "save_model: null root");
    if (logbeta) {
        snprintf(name, sizeof name, "%s.beta", root);
        FILE *f = open_out(name);
        if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", name, strerror(errno));
        for (int32_t i = 0; i < k; i++) {
            const double *row = logbeta + (size_t)i * v;
            for (int32_t j = 0; j < v; j++) fprintf(f, " %5.10f", row[j]); /* lda-model.c:205 */
            fputc('\n', f);
        }
        if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %s", name);
    }
    snprintf(name, sizeof name, "%s.other", root);
    FILE *f = open_out(name);
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
    FILE *f = open_out(path);
    if (!f) return oni_set_error(ONI_ERR_IO, "cannot write %s: %s", path, strerror(errno));
    for (int32_t i = 0; i < d; i++) {
        const double *row = gamma + (size_t)i * k;
        fprintf(f, "%5.10f", row[0]);                                  /* lda-estimate.c:97 */
        for (int32_t j = 1; j < k; j++) fprintf(f, " %5.10f", row[j]); /* :100 */
        fputc('\n', f);
    }
    if (fclose(f) != 0) return oni_set_error(ONI_ERR_IO, "write failed on %s", path);
    return ONI_OK;
}
