// log of Student's t distribution function, both tails, for non-integer degrees of freedom:
// what stats::pt(q, df, lower.tail, log.p = TRUE) gives the reference at
// R/pairwiseTTests.R:439-461 (.run_t_test), where df is Welch's (R/pairwiseTTests.R:311-332).
//
// P(T < -|t|) = 1/2 I_x(df / 2, 1 / 2), x = df / (df + t^2), with the regularised incomplete beta
// function evaluated by its continued fraction (modified Lentz), directly when x lies below
// (a + 1) / (a + b + 2) and through I_x(a, b) = 1 - I_{1 - x}(b, a) otherwise (|t| below ~1.7,
// where neither tail is small). Everything stays in log space, so tails far below the smallest
// double (log p < -1000) keep full relative accuracy; nmath/pt.c takes the same two branches.
//
// The header compiles as plain C++ too (tests/test_tdist_cpu.py builds it with g++ and checks it
// against mpmath), so the numerics can be pinned without a GPU; the product only ever runs it in
// device code (pvalues.cu).
#pragma once

#include <math.h>

#ifdef __CUDACC__
#define SB_HD __host__ __device__ inline
#else
#define SB_HD inline
#endif

namespace sb {

// log B(a, 1/2) = lgamma(a) + lgamma(1/2) - lgamma(a + 1/2). For large a the two lgamma values are
// ~ a log a and their difference ~ log a / 2: subtracting them would lose a log a * 1e-16 absolute;
// Stirling's series is subtracted term by term instead.
SB_HD double log_beta_half(double a) {
    const double half_log_pi = 0.572364942924700087071713675676529356;
    if (a < 12.0) return lgamma(a) + half_log_pi - lgamma(a + 0.5);
    auto stirling = [](double z) {
        const double r = 1.0 / z, r2 = r * r;
        return r * (1.0 / 12.0 + r2 * (-1.0 / 360.0 + r2 * (1.0 / 1260.0 + r2 * (-1.0 / 1680.0 + r2 * (1.0 / 1188.0)))));
    };
    // lgamma(a + 1/2) - lgamma(a) = log(a) / 2 + a log1p(1 / 2a) - 1/2 + S(a + 1/2) - S(a)
    const double d = 0.5 * log(a) + (a * log1p(0.5 / a) - 0.5) + (stirling(a + 0.5) - stirling(a));
    return half_log_pi - d;
}

// Continued fraction of the incomplete beta function: I_x(a, b) = front * cf(a, b, x) / a with
// front = x^a (1 - x)^b / B(a, b); converges quickly for x < (a + 1) / (a + b + 2).
SB_HD double beta_cf(double a, double b, double x) {
    const double tiny = 1e-300, eps = 1e-16;
    const double qab = a + b, qap = a + 1.0, qam = a - 1.0;
    double c = 1.0, d = 1.0 - qab * x / qap;
    if (fabs(d) < tiny) d = tiny;
    d = 1.0 / d;
    double h = d;
    for (int m = 1; m <= 20000; ++m) {
        const double m2 = 2.0 * m;
        double aa = m * (b - m) * x / ((qam + m2) * (a + m2));
        d = 1.0 + aa * d;
        if (fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c;
        if (fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        h *= d * c;
        aa = -(a + m) * (qab + m) * x / ((a + m2) * (qap + m2));
        d = 1.0 + aa * d;
        if (fabs(d) < tiny) d = tiny;
        c = 1.0 + aa / c;
        if (fabs(c) < tiny) c = tiny;
        d = 1.0 / d;
        const double del = d * c;
        h *= del;
        if (fabs(del - 1.0) < eps) break;
    }
    return h;
}

// log P(T <= t) (lower) or log P(T > t) for T ~ Student(df). NaN in, or df <= 0: NaN (R warns and
// returns NaN); t = +-Inf: the limits.
SB_HD double log_pt(double t, double df, bool lower) {
    const double nan = NAN, ninf = -INFINITY;
    if (isnan(t) || isnan(df) || !(df > 0.0)) return nan;
    if (isinf(t)) return ((t < 0.0) == lower) ? ninf : 0.0;
    const double a = 0.5 * df;
    const double t2 = t * t;
    // log of the smaller tail, P(T < -|t|)
    double log_tail;
    if (isinf(df)) {  // the normal limit
        const double h = 0.5 * erfc(fabs(t) * 0.707106781186547524400844362104849039);
        if (h > 1e-290) {
            log_tail = log(h);
        } else {
            const double r = 1.0 / t2;
            const double s = 1.0 + r * (-1.0 + r * (3.0 + r * (-15.0 + r * (105.0 + r * (-945.0 + r * 10395.0)))));
            log_tail = -0.5 * t2 - log(fabs(t)) - 0.918938533204672741780329736405617639 + log(s);
        }
    } else {
        const double u = t2 / df;  // x = 1 / (1 + u), 1 - x = u / (1 + u)
        const double x = 1.0 / (1.0 + u), y = u / (1.0 + u);
        if (x < (a + 1.0) / (a + 2.5)) {
            // log front = a log x + log(1 - x) / 2 - log B(a, 1/2); log x = -log1p(u)
            const double l1pu = log1p(u);
            const double log_y = isinf(u) ? 0.0 : log(u) - l1pu;
            const double log_front = -a * l1pu + 0.5 * log_y - log_beta_half(a);
            log_tail = log_front + log(beta_cf(a, 0.5, x) / a) - 0.693147180559945309417232121458176568;
        } else {
            // central region: I_y(1/2, a) is the mass between -|t| and |t|
            double mid = 0.0;
            if (y > 0.0) {
                const double log_front = 0.5 * log(y) - a * log1p(u) - log_beta_half(a);
                mid = exp(log_front) * beta_cf(0.5, a, y) / 0.5;
            }
            log_tail = log(0.5 - 0.5 * mid);
        }
    }
    const bool want_small = (t <= 0.0) == lower;  // the requested tail is the one beyond -|t|
    if (want_small) return log_tail;
    return log1p(-exp(log_tail));
}

}  // namespace sb
