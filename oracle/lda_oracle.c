/*
 * lda_oracle.c -- CPU restatement of the oni-lda-c variational-EM hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  Nothing in the product library, the `lda`
 * executable or the Python host package may call into this file; only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline leg load it, and there
 * only as the checker.
 *
 * Parity pin: this restatement is checked (tests/test_oracle_pin.py) against
 * the outputs of the reference itself, compiled unmodified by
 * oracle/ref_build/Makefile into oracle/_ref/lda_ref and run on the committed
 * corpus tests/golden/c0_model.dat -- 000.beta byte-for-byte, likelihood.dat,
 * final.beta, final.gamma, final.other byte-for-byte.  The reference ships no
 * tests or golden vectors of its own.
 *
 * Every function names the reference lines it restates (paths relative to
 * /root/reference/oni-lda-c/).  The data layout is ours (flat CSR corpus,
 * flat row-major matrices); the arithmetic, its order of evaluation and the
 * control flow are the reference's, so that a build with the same compiler
 * flags (-O3 -ffp-contract=off) produces the same bits.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

/* ---------------------------------------------------------------- RNG ---- */
/* MT19937 as recoded in cokus.c:52-144.  The reference keeps the state in
 * `unsigned long` (64-bit here) but every stored word is masked or built from
 * 32-bit pieces, so uint32_t state gives the same stream. */
#define MT_N 624
#define MT_M 397
#define MT_MAGIC 0x9908B0DFu

typedef struct {
    uint32_t s[MT_N + 1];
    int next; /* index of the next untempered word */
    int left; /* words left before the next reload; -1 = unseeded */
} orc_mt;

static orc_mt g_mt = { .left = -1 };

/* cokus.c:52-105 -- seed forced odd, LCG 69069 fill */
void orc_mt_seed(uint32_t seed)
{
    uint32_t x = seed | 1u;
    g_mt.s[0] = x;
    for (int j = 1; j < MT_N; j++) {
        x *= 69069u;
        g_mt.s[j] = x;
    }
    g_mt.left = 0;
    g_mt.next = 0;
}

static inline uint32_t mt_twist(uint32_t u, uint32_t v)
{
    /* cokus.h:21-24 mixBits/loBit and cokus.c:119 */
    uint32_t mixed = (u & 0x80000000u) | (v & 0x7FFFFFFFu);
    return (mixed >> 1) ^ ((v & 1u) ? MT_MAGIC : 0u);
}

static inline uint32_t mt_temper(uint32_t y)
{
    /* cokus.c:125-128 and :139-142 */
    y ^= (y >> 11);
    y ^= (y << 7) & 0x9D2C5680u;
    y ^= (y << 15) & 0xEFC60000u;
    return y ^ (y >> 18);
}

/* cokus.c:108-129 */
static uint32_t mt_reload(void)
{
    uint32_t *s = g_mt.s;
    if (g_mt.left < -1) orc_mt_seed(4357u);
    g_mt.left = MT_N - 1;
    g_mt.next = 1;
    int i;
    for (i = 0; i < MT_N - MT_M; i++) s[i] = s[i + MT_M] ^ mt_twist(s[i], s[i + 1]);
    for (; i < MT_N - 1; i++) s[i] = s[i + MT_M - MT_N] ^ mt_twist(s[i], s[i + 1]);
    s[MT_N - 1] = s[MT_M - 1] ^ mt_twist(s[MT_N - 1], s[0]);
    return mt_temper(s[0]);
}

/* cokus.c:131-144 */
uint32_t orc_mt_next(void)
{
    if (--g_mt.left < 0) return mt_reload();
    return mt_temper(g_mt.s[g_mt.next++]);
}

/* lda-model.h:11  myrand() = randomMT() / 2^32 */
double orc_myrand(void) { return (double)orc_mt_next() / 4294967296.; }

/* ------------------------------------------------- special functions ---- */
/* utils.c:8-21 */
double orc_log_sum(double log_a, double log_b)
{
    if (log_a < log_b) return log_b + log(1 + exp(log_a - log_b));
    return log_a + log(1 + exp(log_b - log_a));
}

/* utils.c:55-64 -- coefficients are the reference's literals, including the
 * -0.003968253986254 that is not quite 1/252 */
double orc_digamma(double x)
{
    double p;
    x = x + 6;
    p = 1 / (x * x);
    p = (((0.004166666666667 * p - 0.003968253986254) * p + 0.008333333333333) * p -
         0.083333333333333) * p;
    p = p + log(x) - 0.5 / x - 1 / (x - 1) - 1 / (x - 2) - 1 / (x - 3) - 1 / (x - 4) -
        1 / (x - 5) - 1 / (x - 6);
    return p;
}

/* utils.c:32-47 */
double orc_trigamma(double x)
{
    double p;
    x = x + 6;
    p = 1 / (x * x);
    p = (((((0.075757575757576 * p - 0.033333333333333) * p + 0.0238095238095238) * p -
           0.033333333333333) * p + 0.166666666666667) * p + 1) / x + 0.5 * p;
    for (int i = 0; i < 6; i++) {
        x = x - 1;
        p = 1 / (x * x) + p;
    }
    return p;
}

/* ------------------------------------------------------- alpha Newton ---- */
/* lda-alpha.c:30-34 */
static double d_alhood(double a, double ss, int D, int K)
{
    return D * (K * orc_digamma(K * a) - K * orc_digamma(a)) + ss;
}
static double d2_alhood(double a, int D, int K)
{
    return D * (K * K * orc_trigamma(K * a) - K * orc_trigamma(a));
}

/* lda-alpha.c:42-68 -- Newton on log(alpha), always from 100; NaN restarts
 * at ten times the previous start; stop at |df| <= 1e-5 or 1000 iterations
 * (lda-alpha.h:11-12).  alhood() is evaluated by the reference but unused. */
double orc_opt_alpha(double ss, int D, int K)
{
    double a, log_a, init_a = 100;
    double df, d2f;
    int iter = 0;
    log_a = log(init_a);
    do {
        iter++;
        a = exp(log_a);
        if (isnan(a)) {
            init_a = init_a * 10;
            a = init_a;
            log_a = log(a);
        }
        df = d_alhood(a, ss, D, K);
        d2f = d2_alhood(a, D, K);
        log_a = log_a - df / (d2f * a + df);
    } while ((fabs(df) > 1e-5) && (iter < 1000));
    return exp(log_a);
}

/* ------------------------------------------------------------ M-step ---- */
/* lda-model.c:27-54 (beta part).  logbeta, class_word: [K][V] row-major. */
void orc_mle_beta(int K, int V, const double *class_word, const double *class_total,
                  double *logbeta)
{
    for (int k = 0; k < K; k++)
        for (int w = 0; w < V; w++) {
            double cw = class_word[(size_t)k * V + w];
            logbeta[(size_t)k * V + w] = (cw > 0) ? log(cw) - log(class_total[k]) : -100;
        }
}

/* lda-model.c:104-117, on zeroed statistics (lda-model.c:61-80) */
void orc_random_init_ss(int K, int V, double *class_word, double *class_total)
{
    for (int k = 0; k < K; k++) {
        class_total[k] = 0;
        for (int n = 0; n < V; n++) {
            double v = 0;
            v += 1.0 / V + orc_myrand();
            class_word[(size_t)k * V + n] = v;
            class_total[k] += v;
        }
    }
}

/* ------------------------------------------------------------ E-step ---- */
/* lda-inference.c:93-127.  phi: [len][K]. */
static double doc_bound(int len, const int *words, const int *counts, int K, int V,
                        double alpha, const double *logbeta, const double *phi,
                        const double *gamma, double *dig /* K scratch */)
{
    double likelihood = 0, digsum = 0, gsum = 0;
    for (int k = 0; k < K; k++) {
        dig[k] = orc_digamma(gamma[k]);
        gsum += gamma[k];
    }
    digsum = orc_digamma(gsum);
    likelihood = lgamma(alpha * K) - K * lgamma(alpha) - (lgamma(gsum));
    for (int k = 0; k < K; k++) {
        likelihood += (alpha - 1) * (dig[k] - digsum) + lgamma(gamma[k]) -
                      (gamma[k] - 1) * (dig[k] - digsum);
        for (int n = 0; n < len; n++) {
            double p = phi[(size_t)n * K + k];
            if (p > 0)
                likelihood += counts[n] *
                              (p * ((dig[k] - digsum) - log(p) +
                                    logbeta[(size_t)k * V + words[n]]));
        }
    }
    return likelihood;
}

/* lda-inference.c:27-85.  Returns the document's bound; writes gamma[K] and
 * phi[len][K]; *iters_out (optional) receives the variational passes run.
 * var_converged is the reference's float threshold promoted to double
 * (lda-inference.h:10). */
double orc_doc_inference(int len, const int *words, const int *counts, int total, int K,
                         int V, double alpha, const double *logbeta, int var_max_iter,
                         double var_converged, double *gamma, double *phi, int *iters_out)
{
    double converged = 1, phisum = 0, likelihood = 0, likelihood_old = 0;
    double *oldphi = malloc(sizeof(double) * (size_t)K * 3);
    double *dg = oldphi + K, *scratch = dg + K;
    int var_iter = 0;

    for (int k = 0; k < K; k++) {
        gamma[k] = alpha + (total / ((double)K));
        dg[k] = orc_digamma(gamma[k]);
        for (int n = 0; n < len; n++) phi[(size_t)n * K + k] = 1.0 / K;
    }
    while ((converged > var_converged) && ((var_iter < var_max_iter) || (var_max_iter == -1))) {
        var_iter++;
        for (int n = 0; n < len; n++) {
            double *ph = phi + (size_t)n * K;
            phisum = 0;
            for (int k = 0; k < K; k++) {
                oldphi[k] = ph[k];
                ph[k] = dg[k] + logbeta[(size_t)k * V + words[n]];
                phisum = (k > 0) ? orc_log_sum(phisum, ph[k]) : ph[k];
            }
            for (int k = 0; k < K; k++) {
                ph[k] = exp(ph[k] - phisum);
                gamma[k] = gamma[k] + counts[n] * (ph[k] - oldphi[k]);
                dg[k] = orc_digamma(gamma[k]);
            }
        }
        likelihood = doc_bound(len, words, counts, K, V, alpha, logbeta, phi, gamma, scratch);
        if (isnan(likelihood)) { /* lda-inference.c:78 assert(!isnan) */
            if (iters_out) *iters_out = -var_iter;
            free(oldphi);
            return likelihood;
        }
        converged = (likelihood_old - likelihood) / likelihood_old;
        likelihood_old = likelihood;
    }
    if (iters_out) *iters_out = var_iter;
    free(oldphi);
    return likelihood;
}

/* One E-step over documents d = first, first+stride, ... (lda-estimate.c:201)
 * accumulating the sufficient statistics the way doc_e_step does
 * (lda-estimate.c:28-60).  class_word [K][V], class_total [K] are ADDED to.
 * gamma: [D][K] (only the visited rows are written).  Returns the summed bound;
 * *alpha_ss and *ndocs are added to; *word_visits (optional) accumulates
 * sum over docs of len * passes. */
double orc_e_step(int D, const int64_t *doc_ptr, const int *words, const int *counts,
                  int first, int stride, int K, int V, double alpha, const double *logbeta,
                  int var_max_iter, double var_converged, double *gamma, double *class_word,
                  double *class_total, double *alpha_ss, int *ndocs, int64_t *word_visits)
{
    int maxlen = 0;
    for (int d = 0; d < D; d++) {
        int len = (int)(doc_ptr[d + 1] - doc_ptr[d]);
        if (len > maxlen) maxlen = len;
    }
    double *phi = malloc(sizeof(double) * (size_t)(maxlen > 0 ? maxlen : 1) * K);
    double lik = 0;
    for (int d = first; d < D; d += stride) {
        int len = (int)(doc_ptr[d + 1] - doc_ptr[d]);
        const int *w = words + doc_ptr[d];
        const int *c = counts + doc_ptr[d];
        int total = 0;
        for (int n = 0; n < len; n++) total += c[n]; /* lda-data.c:47 */
        double *g = gamma + (size_t)d * K;
        int iters = 0;
        lik += orc_doc_inference(len, w, c, total, K, V, alpha, logbeta, var_max_iter,
                                 var_converged, g, phi, &iters);
        if (word_visits) *word_visits += (int64_t)len * (iters < 0 ? -iters : iters);
        double gsum = 0;
        for (int k = 0; k < K; k++) {
            gsum += g[k];
            *alpha_ss += orc_digamma(g[k]);
        }
        *alpha_ss -= K * orc_digamma(gsum);
        for (int n = 0; n < len; n++)
            for (int k = 0; k < K; k++) {
                class_word[(size_t)k * V + w[n]] += c[n] * phi[(size_t)n * K + k];
                class_total[k] += c[n] * phi[(size_t)n * K + k];
            }
        *ndocs = *ndocs + 1;
    }
    free(phi);
    return lik;
}

/* ---------------------------------------------------------- EM driver ---- */
typedef struct {
    int var_max_iter;    /* settings.txt line 1 */
    float var_converged; /* line 2, stored as float (lda-inference.h:10)  */
    int em_max_iter;     /* line 3 */
    float em_converged;  /* line 4, stored as float (lda-estimate.h:21)   */
    int estimate_alpha;  /* line 5: 0 only for the literal "fixed"        */
} orc_settings;

/* lda-estimate.c:112-321 for start == "random", one rank (or `nproc` ranks
 * emulated in-process: partials are formed per rank in the reference's strided
 * order and added in rank order, which is what MPI_SUM over our shim does).
 *
 * Outputs: logbeta0 (model written as 000.beta), logbeta (final.beta), gamma
 * [D][K] (final.gamma), *alpha_out (final.other), lik[i]/conv[i] (the two
 * columns of likelihood.dat line i).  Returns the number of EM iterations. */
int orc_run_em(int D, const int64_t *doc_ptr, const int *words, const int *counts, int K,
               int V, double initial_alpha, const orc_settings *st, uint32_t seed, int nproc,
               double *logbeta0, double *logbeta, double *gamma, double *alpha_out,
               double *lik, double *conv, int lik_cap)
{
    size_t kv = (size_t)K * V;
    double *cw = calloc(kv, sizeof(double)), *ct = calloc((size_t)K, sizeof(double));
    double *cw_r = calloc(kv, sizeof(double)), *ct_r = calloc((size_t)K, sizeof(double));
    int var_max_iter = st->var_max_iter;
    double alpha;

    orc_mt_seed(seed);                      /* lda-estimate.c:414-416 */
    orc_random_init_ss(K, V, cw, ct);       /* :157 */
    orc_mle_beta(K, V, cw, ct, logbeta);    /* :158 lda_mle(model, ss, 0) */
    alpha = initial_alpha;                  /* :159 */
    memcpy(logbeta0, logbeta, kv * sizeof(double));

    int i = 0;
    double likelihood, likelihood_old = 0, converged = 1;
    while (((converged < 0) || (converged > st->em_converged) || (i <= 2)) &&
           (i <= st->em_max_iter)) {
        i++;
        likelihood = 0;
        double alpha_ss = 0;
        int ndocs = 0;
        memset(cw, 0, kv * sizeof(double));
        memset(ct, 0, (size_t)K * sizeof(double));
        for (int r = 0; r < nproc; r++) {
            double a_r = 0, l_r;
            int n_r = 0;
            if (nproc == 1) {
                l_r = orc_e_step(D, doc_ptr, words, counts, 0, 1, K, V, alpha, logbeta,
                                 var_max_iter, st->var_converged, gamma, cw, ct, &a_r, &n_r, NULL);
            } else {
                memset(cw_r, 0, kv * sizeof(double));
                memset(ct_r, 0, (size_t)K * sizeof(double));
                l_r = orc_e_step(D, doc_ptr, words, counts, r, nproc, K, V, alpha, logbeta,
                                 var_max_iter, st->var_converged, gamma, cw_r, ct_r, &a_r, &n_r,
                                 NULL);
                for (size_t j = 0; j < kv; j++) cw[j] = (r == 0) ? cw_r[j] : cw[j] + cw_r[j];
                for (int k = 0; k < K; k++) ct[k] = (r == 0) ? ct_r[k] : ct[k] + ct_r[k];
            }
            alpha_ss = (r == 0) ? a_r : alpha_ss + a_r;
            ndocs = (r == 0) ? n_r : ndocs + n_r;
            likelihood = (r == 0) ? l_r : likelihood + l_r;
        }
        orc_mle_beta(K, V, cw, ct, logbeta);                               /* :256 */
        if (st->estimate_alpha == 1) alpha = orc_opt_alpha(alpha_ss, ndocs, K);
        converged = (likelihood_old - likelihood) / (likelihood_old);      /* :260 */
        if (converged < 0) var_max_iter = var_max_iter * 2;                /* :261 */
        likelihood_old = likelihood;
        if (i - 1 < lik_cap) {
            lik[i - 1] = likelihood;
            conv[i - 1] = converged;
        }
    }
    *alpha_out = alpha;
    free(cw); free(ct); free(cw_r); free(ct_r);
    return i;
}
