/*
 * score_oracle.c -- CPU restatement of the post-LDA result shaping and event
 * scoring (lda_post.py, flow_post_lda.scala, dns_post_lda.scala).
 *
 * TEST INFRASTRUCTURE ONLY (see lda_oracle.c header for who may call this).
 *
 * PARITY UNPINNED for the text conventions: lda_post.py is Python-2-only and
 * the scoring loop runs inside spark-shell; neither interpreter exists in this
 * image, and the reference holds no fixture for either.  What is restated is
 * the arithmetic that IS in the reference sources, in its written order:
 *   - 10-decimal text of final.gamma / final.beta ("%5.10f",
 *     oni-lda-c/lda-estimate.c:97-100, lda-model.c:205) read back as double;
 *   - theta = gamma_i / sum(gamma) with a left-to-right sum and the all-zero
 *     row rule (lda_post.py:44-54);
 *   - p(w|z) = exp(beta_kw) / sum_w exp(beta_kw), left-to-right sum
 *     (lda_post.py:88-96);
 *   - both tables pass through 12-significant-digit decimal text, which is what
 *     Python-2 str(float) and numpy<1.14 str(float64) print (lda_post.py:50,
 *     :119-121) -- this is the convention the build FIXES, not one the
 *     reference pins;
 *   - score = sum_{i=0..K-1} theta[i]*p[i], one rounded multiply and one
 *     rounded add per term, no FMA (JVM semantics; flow_post_lda.scala:232-237,
 *     dns_post_lda.scala:315-318); a missing document or word scores with a
 *     constant vector (0.05 flow, 0.1 DNS; flow_post_lda.scala:228-231,
 *     dns_post_lda.scala:313-314);
 *   - flow key = min(src, dst) (flow_post_lda.scala:238); keep key < tol
 *     (strict; :245), ascending by key.  Spark's sortByKey leaves the order of
 *     equal keys unspecified; ties here are by ascending event index.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* value -> printf(fmt) -> strtod */
double orc_text_roundtrip(double x, const char *fmt)
{
    char buf[512];
    snprintf(buf, sizeof buf, fmt, x);
    return strtod(buf, NULL);
}

/* lda_post.py:35-59.  gamma [D][K] (already 10-decimal text values).
 * theta [D][K] receives the 12-digit values the Scala stage would parse. */
void orc_theta_from_gamma(int D, int K, const double *gamma, double *theta)
{
    for (int d = 0; d < D; d++) {
        const double *g = gamma + (size_t)d * K;
        double total = 0; /* python sum(): starts from int 0, adds left to right */
        for (int k = 0; k < K; k++) total = total + g[k];
        for (int k = 0; k < K; k++)
            theta[(size_t)d * K + k] =
                (total > 0) ? orc_text_roundtrip(g[k] / total, "%.12g") : 0.0;
    }
}

/* lda_post.py:70-96,115-122.  beta [K][V] (10-decimal text values) ->
 * pwz [V][K] word-major, 12-digit values. */
void orc_pwz_from_beta(int K, int V, const double *beta, double *pwz)
{
    double *raw = malloc(sizeof(double) * (size_t)V);
    for (int k = 0; k < K; k++) {
        double total = 0;
        for (int w = 0; w < V; w++) {
            raw[w] = exp(beta[(size_t)k * V + w]);
            total = total + raw[w];
        }
        for (int w = 0; w < V; w++)
            pwz[(size_t)w * K + k] = orc_text_roundtrip(raw[w] / total, "%.12g");
    }
    free(raw);
}

/* The same two shapings WITHOUT the 12-digit text round trip: the arithmetic of lda_post.py:45-50
 * and :88-96 alone, which tests/golden/post pins (values produced by those source lines
 * themselves under Python 3, oracle/pin_lda_post.py). */
void orc_theta_from_gamma_raw(int D, int K, const double *gamma, double *theta)
{
    for (int d = 0; d < D; d++) {
        const double *g = gamma + (size_t)d * K;
        double total = 0;
        for (int k = 0; k < K; k++) total = total + g[k];
        for (int k = 0; k < K; k++) theta[(size_t)d * K + k] = (total > 0) ? g[k] / total : 0.0;
    }
}

void orc_pwz_from_beta_raw(int K, int V, const double *beta, double *pwz)
{
    double *raw = malloc(sizeof(double) * (size_t)V);
    for (int k = 0; k < K; k++) {
        double total = 0;
        for (int w = 0; w < V; w++) {
            raw[w] = exp(beta[(size_t)k * V + w]);
            total = total + raw[w];
        }
        for (int w = 0; w < V; w++) pwz[(size_t)w * K + k] = raw[w] / total;
    }
    free(raw);
}

static inline double dot_seq(const double *a, const double *b, int K)
{
    volatile double acc = 0.0; /* volatile: forbid contraction/reassociation */
    for (int i = 0; i < K; i++) {
        volatile double prod = a[i] * b[i];
        acc = acc + prod;
    }
    return acc;
}

/* One side of an event: theta row (or default) . pwz row (or default). */
static double side_score(const double *theta, const double *pwz, int K, int32_t doc,
                         int32_t word, double dflt, double *tmp_a, double *tmp_b)
{
    const double *a, *b;
    if (doc >= 0) a = theta + (size_t)doc * K;
    else { for (int i = 0; i < K; i++) tmp_a[i] = dflt; a = tmp_a; }
    if (word >= 0) b = pwz + (size_t)word * K;
    else { for (int i = 0; i < K; i++) tmp_b[i] = dflt; b = tmp_b; }
    return dot_seq(a, b, K);
}

/* flow_post_lda.scala:227-238.  src/dst score per event; index < 0 = lookup miss. */
void orc_score_flow(int64_t n, int K, const double *theta, const double *pwz,
                    const int32_t *src_doc, const int32_t *dst_doc, const int32_t *src_word,
                    const int32_t *dst_word, double *src_score, double *dst_score)
{
    double *ta = malloc(sizeof(double) * (size_t)K * 2), *tb = ta + K;
    for (int64_t e = 0; e < n; e++) {
        src_score[e] = side_score(theta, pwz, K, src_doc[e], src_word[e], 0.05, ta, tb);
        dst_score[e] = side_score(theta, pwz, K, dst_doc[e], dst_word[e], 0.05, ta, tb);
    }
    free(ta);
}

/* dns_post_lda.scala:312-321 */
void orc_score_dns(int64_t n, int K, const double *theta, const double *pwz,
                   const int32_t *doc, const int32_t *word, double *score)
{
    double *ta = malloc(sizeof(double) * (size_t)K * 2), *tb = ta + K;
    for (int64_t e = 0; e < n; e++)
        score[e] = side_score(theta, pwz, K, doc[e], word[e], 0.1, ta, tb);
    free(ta);
}

typedef struct { double key; int64_t idx; } ranked;

static int ranked_cmp(const void *pa, const void *pb)
{
    const ranked *a = pa, *b = pb;
    if (a->key < b->key) return -1;
    if (a->key > b->key) return 1;
    return (a->idx > b->idx) - (a->idx < b->idx);
}

/* flow_post_lda.scala:245 / dns_post_lda.scala:326: filter(key < tol).sortByKey().
 * key[e] = min(a[e], b[e]) when b != NULL (flow), else a[e] (DNS).
 * Writes up to cap (idx, key) pairs in rank order; returns the number kept. */
int64_t orc_rank_below(int64_t n, const double *a, const double *b, double tol, int64_t cap,
                       int64_t *out_idx, double *out_key)
{
    ranked *r = malloc(sizeof(ranked) * (size_t)(n > 0 ? n : 1));
    int64_t m = 0;
    for (int64_t e = 0; e < n; e++) {
        /* breeze.linalg.min(x, y) on two doubles = math.min */
        double key = b ? (a[e] < b[e] ? a[e] : b[e]) : a[e];
        if (key < tol) { r[m].key = key; r[m].idx = e; m++; }
    }
    qsort(r, (size_t)m, sizeof(ranked), ranked_cmp);
    for (int64_t i = 0; i < m && i < cap; i++) { out_idx[i] = r[i].idx; out_key[i] = r[i].key; }
    free(r);
    return m;
}
