// Host harness for csrc/devmath.cuh: the table-driven exp/log and the folded
// digamma are compiled for the CPU (same tables, same coefficients; only the
// reciprocal seed differs) and compared with long-double evaluations.
#include <math.h>
#include <stdio.h>
#include <stdlib.h>

#include "../oni_ml_b200/csrc/devmath.cuh"

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
    // flush and edge behaviour
    int edge_ok = oni::exp_tab(-708.5, T.ex) < 1e-306 && oni::exp_tab(-1e6, T.ex) < 1e-306 && oni::exp_tab(-1e6, T.ex) > 0 &&
                  fabs(oni::exp_tab(0.0, T.ex) - 1.0) < 5e-16 && fabs(oni::log_tab(1.0, T.lg)) < 1e-16;
    printf("exp_max_ulp_beyond_model %.3f\nlog_max_rel %.3e\ndigamma_max_rel %.3e\ndigamma_max_abs %.3e\nedge_ok %d\n",
           worst_exp, worst_log, worst_dg, worst_dg_abs, edge_ok);
    return 0;
}
