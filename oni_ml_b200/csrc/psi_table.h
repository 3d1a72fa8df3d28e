// psi_table.h -- table-driven digamma and a short-chain exp for the long-document E-step.
//
// A long document is a serial chain of word visits (lda-inference.c:50-75 updates gamma and
// digamma(gamma) after every word), so what sets its pace is the LENGTH of the dependency chain
// gamma -> digamma -> exp -> sum -> reciprocal -> gamma, not the instruction count.  The
// reference's digamma (oni-lda-c/utils.c:55-64) as restated in devmath.cuh is 19 dependent fp64
// operations deep; here the same FUNCTION (that series with its literal coefficients and the six
// recurrence terms, not the exact psi) is tabulated once on the host in long double:
//
//   psi_ref(g), g in [2^PSI_EXP_LO, 2^PSI_EXP_HI):  the top 6 mantissa bits pick one of 64
//   intervals per octave; on each a degree-5 polynomial in r = g - centre (Chebyshev
//   interpolation, error < 1e-14 * max(1, |psi|)), evaluated Estrin-style: 4 dependent operations
//   after one 48-byte table read.  Arguments outside the table make the caller fall back to
//   digamma_ref.
//
//   exp_k(x): like exp_tab but with a 2048-entry table of 2^(i/2048), so the polynomial is a
//   cubic evaluated in two levels: 6 dependent operations instead of 9.
//
// The tables (93 KB + 16 KB) are built at context creation (build_long_tables) and copied to
// shared memory by the long-document kernel; the main kernel does not use them (its load/store
// pipe has no room for three more random 16-byte reads per topic).
#pragma once
#include "devmath.cuh"

namespace oni {

constexpr int PSI_EXP_LO = -8;                       // table covers [2^-8, 2^23)
constexpr int PSI_EXP_HI = 23;
constexpr int PSI_SUB_BITS = 6;                      // 64 intervals per octave
constexpr int PSI_N = (PSI_EXP_HI - PSI_EXP_LO) << PSI_SUB_BITS;
constexpr int PSI_BASE = (1023 + PSI_EXP_LO) << PSI_SUB_BITS;
constexpr int PSI_COEF = 6;
constexpr int PSI_SHIFT = 20 - PSI_SUB_BITS;         // hi word >> PSI_SHIFT = (exponent, sub-interval)
constexpr int EXK_BITS = 11;
constexpr int EXK_N = 1 << EXK_BITS;

#define ONI_INVLN2K 0x1.71547652b82fep+11 /* 2048 / ln 2 */
#define ONI_LN2K 0x1.62e42fefa39efp-12    /* ln 2 / 2048 */

// Index of g's interval, or something outside [0, PSI_N) when g is not covered (also NaN, <= 0).
ONI_HD int psi_index(double g) { return (hi_word(g) >> PSI_SHIFT) - PSI_BASE; }

// Centre of the interval that holds g (top bits of g, the bit below them set).
ONI_HD double psi_centre(double g)
{
    const int m = (1 << PSI_SHIFT) - 1;
    return make_double((hi_word(g) & ~m) | (1 << (PSI_SHIFT - 1)), 0);
}

// psi_ref(g) from the six coefficients c[0..5] of g's interval; `add` is folded into the constant
// term (the caller passes log beta to get psi + log beta out of the same chain, or 0).
ONI_HD double psi_poly(double g, const double *c, double add)
{
    const double r = g - psi_centre(g);
    const double r2 = r * r;
    const double p01 = fma_(c[1], r, c[0] + add);
    const double p23 = fma_(c[3], r, c[2]);
    const double p45 = fma_(c[5], r, c[4]);
    return fma_(fma_(p45, r2, p23), r2, p01);
}

// exp(x) for -7e5 < x <= 709 (~3 ulp + 0.3 ulp per unit of |x|); like exp_tab it stays at ~1e-306
// instead of underflowing.  The caller clamps x from below.
ONI_HD double exp_k(double x, const double *tab)
{
    const double SHIFT = 6755399441055744.0;  // 1.5 * 2^52
    const double z = fma_(x, ONI_INVLN2K, SHIFT);
    const int ki = lo_word(z);
    const double kd = z - SHIFT;
    const double r = fma_(kd, -ONI_LN2K, x);      // |r| <= ln2/4096
    const double t = tab[ki & (EXK_N - 1)];      // 2^(i/2048)
    int top = ki >> EXK_BITS;
    top = top < -1018 ? -1018 : top;
    const double scale = make_double(hi_word(t) + top * 1048576, lo_word(t));
    // 1 + r + r^2/2 + r^3/6 in two levels; the r^4/24 term is < 4e-17
    const double a = 1.0 + r;
    const double b = fma_(r, 1.0 / 6.0, 0.5);
    return scale * fma_(r * r, b, a);
}

// ---- host: build the tables (long double; x87 extended is ample for a 1e-15 fit)
static inline long double psi_ref_ld(long double g)
{
    // oni-lda-c/utils.c:55-64, term for term
    long double x = g + 6, p = 1 / (x * x);
    p = (((0.004166666666667L * p - 0.003968253986254L) * p + 0.008333333333333L) * p - 0.083333333333333L) * p;
    return p + logl(x) - 0.5L / x - 1 / (x - 1) - 1 / (x - 2) - 1 / (x - 3) - 1 / (x - 4) - 1 / (x - 5) - 1 / (x - 6);
}

// psi[PSI_N * PSI_COEF], exk[EXK_N]
static inline void build_long_tables(double *psi, double *exk)
{
    const long double pi = 3.14159265358979323846264338327950288L;
    long double node[6];
    for (int j = 0; j < 6; j++) node[j] = cosl(pi * (2 * j + 1) / 12.0L);
    for (int i = 0; i < PSI_N; i++) {
        const int e = PSI_EXP_LO + (i >> PSI_SUB_BITS), sub = i & ((1 << PSI_SUB_BITS) - 1);
        const long double h = ldexpl(1.0L, e - PSI_SUB_BITS - 1);          // half width
        const long double centre = ldexpl(1.0L, e) + (2 * sub + 1) * h;
        long double f[6], b[6];
        for (int j = 0; j < 6; j++) f[j] = psi_ref_ld(centre + h * node[j]);
        for (int m = 0; m < 6; m++) {
            long double s = 0;
            for (int j = 0; j < 6; j++) s += f[j] * cosl(pi * m * (2 * j + 1) / 12.0L);
            b[m] = s * (m == 0 ? 1.0L : 2.0L) / 6.0L;
        }
        // Chebyshev -> monomial in u = r/h, then r = h u (h is a power of two: exact)
        const long double a[6] = {b[0] - b[2] + b[4], b[1] - 3 * b[3] + 5 * b[5], 2 * b[2] - 8 * b[4],
                                  4 * b[3] - 20 * b[5], 8 * b[4], 16 * b[5]};
        long double scale = 1;
        for (int m = 0; m < 6; m++) {
            psi[(size_t)i * PSI_COEF + m] = (double)(a[m] * scale);
            scale /= h;
        }
    }
    for (int i = 0; i < EXK_N; i++) exk[i] = (double)exp2l((long double)i / EXK_N);
}

}  // namespace oni
