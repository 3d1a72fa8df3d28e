// Host harness for csrc/devmath.cuh: the table-driven exp/log and the folded
// digamma are compiled for the CPU (same tables, same coefficients; only the
// reciprocal seed differs) and compared with long-double evaluations.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "../oni_ml_b200/csrc/devmath.cuh"
#include "../oni_ml_b200/csrc/psi_table.h"
#include <vector>

static const oni::MathTables T = ONI_MATH_TABLES_INIT;

static double ulp_err(double got, long double want)
{
    if (want == 0) return got == 0 ? 0 : 1e30;
    int e;
    frexpl(want, &e);
    long double ulp = ldexpl(1.0L, e - 53);
    return (double)(fabsl((long double)got - want) / ulp);
}

// the reference's digamma (utils.c:55-64) in long double
static long double digamma_ld(long double x)
{
    long double p;
    x = x + 6;
    p = 1 / (x * x);
    p = (((0.004166666666667L * p - 0.003968253986254L) * p + 0.008333333333333L) * p - 0.083333333333333L) * p;
    p = p + logl(x) - 0.5L / x - 1 / (x - 1) - 1 / (x - 2) - 1 / (x - 3) - 1 / (x - 4) - 1 / (x - 5) - 1 / (x - 6);
    return p;
}

static double urand(void) { return (rand() + 0.5) / ((double)RAND_MAX + 1.0); }

int main(void)
{
    srand(12345);
    double worst_exp = 0, worst_log = 0, worst_dg = 0, worst_dg_abs = 0;
    for (int i = 0; i < 2000000; i++) {
        double x = -700.0 + 1400.0 * urand();
        if (i & 1) x = -40.0 + 60.0 * urand();
        // the error model of exp_tab: ~2 ulp from the table and the polynomial, plus |x| * 5e-17
        // relative from the one-constant argument reduction (0.45 ulp per unit of |x|)
        double e = ulp_err(oni::exp_tab(x, T.ex), expl((long double)x)) - 0.45 * fabs(x);
        if (e > worst_exp) worst_exp = e;
        double y = exp(-690.0 + 1390.0 * urand());
        if (i % 3 == 0) y = 0.5 + 20.0 * urand();
        if (i % 3 == 1) y = 1.0 + 1e-3 * (urand() - 0.5);
        double l = fabs((double)((long double)oni::log_tab(y, T.lg) - logl((long double)y)));
        double lrel = l / fmax(fabs(log(y)), 1e-3);
        if (lrel > worst_log) worst_log = lrel;
        double g = exp(-7.0 + 25.0 * urand());  // 1e-3 .. 6e7
        long double want = digamma_ld((long double)g);
        double d = fabs((double)((long double)oni::digamma_ref(g, T.lg) - want));
        double drel = d / fmax(fabsl(want), 1.0L);
        if (drel > worst_dg) worst_dg = drel;
        if (d > worst_dg_abs && g > 0.05) worst_dg_abs = d;
    }
    double worst_lg = 0;
    for (int i = 0; i < 1000000; i++) {
        double g = exp(-7.0 + 30.0 * urand());  // 1e-3 .. 1e10
        long double want = lgammal((long double)g);
        double d = fabs((double)((long double)oni::lgamma_pos(g, T.lg) - want));
        // error relative to the size of the Stirling intermediates (what the bound sums)
        double scale = fmax(fabs((double)want), fmax(g * log(g + 12.0), 20.0));
        if (d / scale > worst_lg) worst_lg = d / scale;
    }
    printf("lgamma_max_rel %.3e\n", worst_lg);
    // the long-document kernel's tables (psi_table.h): digamma by interval polynomial, short-chain exp
    {
        std::vector<double> psi((size_t)oni::PSI_N * oni::PSI_COEF), exk(oni::EXK_N);
        oni::build_long_tables(psi.data(), exk.data());
        double worst_psi = 0, worst_expk = 0;
        int index_ok = oni::psi_index(ldexp(1.0, oni::PSI_EXP_LO)) == 0 &&
                       oni::psi_index(nextafter(ldexp(1.0, oni::PSI_EXP_HI), 0.0)) == oni::PSI_N - 1 &&
                       (unsigned)oni::psi_index(nextafter(ldexp(1.0, oni::PSI_EXP_LO), 0.0)) >= (unsigned)oni::PSI_N &&
                       oni::psi_index(ldexp(1.0, oni::PSI_EXP_HI)) >= oni::PSI_N &&
                       (unsigned)oni::psi_index(NAN) >= (unsigned)oni::PSI_N && (unsigned)oni::psi_index(-1.0) >= (unsigned)oni::PSI_N &&
                       (unsigned)oni::psi_index(0.0) >= (unsigned)oni::PSI_N && (unsigned)oni::psi_index(INFINITY) >= (unsigned)oni::PSI_N;
        for (int i = 0; i < 4000000; i++) {
            double g = ldexp(1.0 + urand(), oni::PSI_EXP_LO + (int)(urand() * (oni::PSI_EXP_HI - oni::PSI_EXP_LO)));
            if (i % 5 == 0) {  // interval edges
                const int idx = (int)(urand() * oni::PSI_N);
                g = ldexp(1.0 + (double)(idx & 63) / 64.0, oni::PSI_EXP_LO + (idx >> 6));
                if (i % 10 == 0) g = nextafter(g, 0.0);
                if (g < ldexp(1.0, oni::PSI_EXP_LO)) g = ldexp(1.0, oni::PSI_EXP_LO);
            }
            const int idx = oni::psi_index(g);
            if (idx < 0 || idx >= oni::PSI_N) { index_ok = 0; continue; }
            const long double want = oni::psi_ref_ld((long double)g);
            const double got = oni::psi_poly(g, &psi[(size_t)idx * oni::PSI_COEF], 0.0);
            const double d = fabs((double)((long double)got - want)) / fmax(fabsl(want), 1.0L);
            if (d > worst_psi) worst_psi = d;
            double x = -700.0 + 1400.0 * urand();
            if (i & 1) x = -40.0 + 60.0 * urand();
            double e = ulp_err(oni::exp_k(x, exk.data()), expl((long double)x)) - 0.6 * fabs(x);
            if (e > worst_expk) worst_expk = e;
        }
        int edge = oni::exp_k(-5e5, exk.data()) < 1e-300 && oni::exp_k(-5e5, exk.data()) > 0 &&
                   fabs(oni::exp_k(0.0, exk.data()) - 1.0) < 5e-16;
        printf("psi_table_max_rel %.3e\nexpk_max_ulp_beyond_model %.3f\npsi_index_ok %d\nexpk_edge_ok %d\n",
               worst_psi, worst_expk, index_ok, edge);
    }
    // flush and edge behaviour
    int edge_ok = oni::exp_tab(-708.5, T.ex) < 1e-306 && oni::exp_tab(-1e6, T.ex) < 1e-306 && oni::exp_tab(-1e6, T.ex) > 0 &&
                  fabs(oni::exp_tab(0.0, T.ex) - 1.0) < 5e-16 && fabs(oni::log_tab(1.0, T.lg)) < 1e-16;
    printf("exp_max_ulp_beyond_model %.3f\nlog_max_rel %.3e\ndigamma_max_rel %.3e\ndigamma_max_abs %.3e\nedge_ok %d\n",
           worst_exp, worst_log, worst_dg, worst_dg_abs, edge_ok);
    return 0;
}
