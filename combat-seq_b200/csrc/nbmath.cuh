// nbmath.cuh — FP64 device functions for the negative-binomial distribution.
//
// Replaces, on the device, the R nmath routines that the reference calls as
//   dnbinom(x, mu=, size=)  /root/reference/helper_seq.R:28
//   pnbinom(q, mu=, size=)  /root/reference/helper_seq.R:53
//   qnbinom(p, mu=, size=)  /root/reference/helper_seq.R:58
// R/nmath is a third-party dependency that is not under /root/reference; the algorithms follow its published
// form: Loader's saddle-point pmf (stirlerr / bd0), the TOMS 708 continued fraction for the regularised
// incomplete beta function, Wichura's AS 241 inverse normal and qnbinom's Cornish-Fisher start + discrete
// search with the p*(1-64*eps) fuzz.  Everything is evaluated on the FP64 pipe.
#pragma once
#include <cuda_runtime.h>
#include <cfloat>
#include <cmath>

#define CBS_HD __host__ __device__ __forceinline__

namespace cbs {

#define CBS_LN_SQRT_2PI 0.918938533204672741780329736406
#define CBS_LN_2PI      1.837877066409345483560659472811
#define CBS_2PI         6.283185307179586476925286766559

// stirlerr(k/2), k = 0..30:  log(n!) - log(sqrt(2 pi n) (n/e)^n), generated with mpmath at 40 digits
#define CBS_SFERR_HALVES_INIT { \
    0.0, \
    0.1534264097200273452913839, 0.08106146679532725821967026, 0.0548141210519176538961387, \
    0.04134069595540929409382208, 0.03316287351993628748511051, 0.02767792568499833914878929, \
    0.02374616365629749597133028, 0.02079067210376509311152277, 0.01848845053267318523077936, \
    0.01664469118982119216319487, 0.01513497322191737887351384, 0.01387612882307074799874573, \
    0.01281046524292022692425066, 0.01189670994589177009505572, 0.01110455975820691732663076, \
    0.01041126526197209649747857, 0.009799416126158803298390373, 0.009255462182712732917728637, \
    0.008768700134139385462955047, 0.008330563433362871256469319, 0.007934114564314020547249562, \
    0.007573675487951840794972024, 0.007244554301320383179546197, 0.006942840107209529865664153, \
    0.006665247032707682442356181, 0.006408994188004207068439631, 0.006171712263039457647534605, \
    0.005951370112758847735624416, 0.005746216513010115682026102, 0.00555473355196280137103869}
__device__ __constant__ double c_sferr_halves[31] = CBS_SFERR_HALVES_INIT;
static const double h_sferr_halves[31] = CBS_SFERR_HALVES_INIT;   // host copy (tests/hostmath compiles this header for the CPU)
CBS_HD double sferr_half(int k) {
#ifdef __CUDA_ARCH__
    return c_sferr_halves[k];
#else
    return h_sferr_halves[k];
#endif
}

// Stirling-series remainder of log(n!)
CBS_HD double stirlerr(double n) {
    const double S0 = 0.083333333333333333333, S1 = 0.00277777777777777777778, S2 = 0.00079365079365079365079365,
                 S3 = 0.000595238095238095238095238, S4 = 0.0008417508417508417508417508;
    if (n <= 15.0) {
        const double nn = n + n;
        if (nn == (double)(int)nn) return sferr_half((int)nn);
        return lgamma(n + 1.0) - (n + 0.5) * log(n) + n - CBS_LN_SQRT_2PI;
    }
    const double nn = n * n;
    if (n > 500) return (S0 - S1 / nn) / n;
    if (n > 80) return (S0 - (S1 - S2 / nn) / nn) / n;
    if (n > 35) return (S0 - (S1 - (S2 - S3 / nn) / nn) / nn) / n;
    return (S0 - (S1 - (S2 - (S3 - S4 / nn) / nn) / nn) / nn) / n;
}

// bd0(x, np) = x log(x/np) + np - x, without cancellation when x ~ np
CBS_HD double bd0(double x, double np) {
    if (!isfinite(x) || !isfinite(np) || np == 0.0) return NAN;
    if (fabs(x - np) < 0.1 * (x + np)) {
        double v = (x - np) / (x + np);
        double s = (x - np) * v;
        if (fabs(s) < DBL_MIN) return s;
        double ej = 2 * x * v;
        v = v * v;
        for (int j = 1; j < 1000; j++) {
            ej *= v;
            const double s1 = s + ej / ((j << 1) + 1);
            if (s1 == s) return s1;
            s = s1;
        }
    }
    return x * log(x / np) + np - x;
}

// binomial pmf kernel of Loader's algorithm, linear (log_p = false) or log scale
template <bool LOG>
CBS_HD double dbinom_raw(double x, double n, double p, double q) {
    if (p == 0) return (x == 0) ? (LOG ? 0.0 : 1.0) : (LOG ? -INFINITY : 0.0);
    if (q == 0) return (x == n) ? (LOG ? 0.0 : 1.0) : (LOG ? -INFINITY : 0.0);
    double lc;
    if (x == 0) {
        if (n == 0) return LOG ? 0.0 : 1.0;
        lc = (p < 0.1) ? -bd0(n, n * q) - n * p : n * log(q);
        return LOG ? lc : exp(lc);
    }
    if (x == n) {
        lc = (q < 0.1) ? -bd0(n, n * p) - n * q : n * log(p);
        return LOG ? lc : exp(lc);
    }
    if (x < 0 || x > n) return LOG ? -INFINITY : 0.0;
    lc = stirlerr(n) - stirlerr(x) - stirlerr(n - x) - bd0(x, n * p) - bd0(n - x, n * q);
    const double lf = CBS_LN_2PI + log(x) + log1p(-x / n);
    return LOG ? (lc - 0.5 * lf) : exp(lc - 0.5 * lf);
}

// dnbinom(x; mu, size) for integer-valued x >= 0 (helper_seq.R:28), linear or log scale
template <bool LOG>
CBS_HD double dnbinom_mu(double x, double size, double mu) {
    if (isnan(x) || isnan(size) || isnan(mu)) return x + size + mu;
    if (mu < 0 || size < 0) return NAN;
    if (x < 0 || !isfinite(x)) return LOG ? -INFINITY : 0.0;
    if (x == 0 && size == 0) return LOG ? 0.0 : 1.0;
    if (!isfinite(size)) {  // Poisson limit
        if (mu == 0) return (x == 0) ? (LOG ? 0.0 : 1.0) : (LOG ? -INFINITY : 0.0);
        if (x == 0) return LOG ? -mu : exp(-mu);
        const double l = -stirlerr(x) - bd0(x, mu) - 0.5 * log(CBS_2PI * x);
        return LOG ? l : exp(l);
    }
    if (x == 0) {
        const double l = size * (size < mu ? log(size / (size + mu)) : log1p(-mu / (size + mu)));
        return LOG ? l : exp(l);
    }
    if (x < 1e-10 * size) {
        const double p = (size < mu ? log(size / (1 + size / mu)) : log(mu / (1 + mu / size)));
        const double l = x * p - mu - lgamma(x + 1) + log1p(x * (x - 1) / (2 * size));
        return LOG ? l : exp(l);
    }
    const double p = size / (size + x);
    const double ans = dbinom_raw<LOG>(size, x + size, size / (size + mu), mu / (size + mu));
    return LOG ? (log(p) + ans) : (p * ans);
}

// x^a y^b / B(a,b) in saddle-point form (valid for all a, b > 0)
CBS_HD double brcomp(double a, double b, double x, double y) {
    if (x == 0 || y == 0) return 0.0;
    const double n = a + b;
    const double lc = stirlerr(n) - stirlerr(a) - stirlerr(b) - bd0(a, n * x) - bd0(b, n * y);
    return sqrt(a * b / (n * CBS_2PI)) * exp(lc);
}

// TOMS 708 continued fraction for I_x(a,b), lambda = (a+b)y - b >= 0
CBS_HD double bfrac(double a, double b, double x, double y, double lambda, double eps) {
    const double brc = brcomp(a, b, x, y);
    if (brc == 0.) return 0.;
    const double c = lambda + 1., c0 = b / a, c1 = 1. / a + 1., yp1 = y + 1.;
    double n = 0., p = 1., s = a + 1., an = 0., bn = 1., anp1 = 1., bnp1 = c / c1, r = c1 / c;
    do {
        n += 1.;
        double t = n / a;
        const double w = n * (b - n) * x;
        double e = a / s;
        const double alpha = p * (p + c0) * e * e * (w * x);
        e = (t + 1.) / (c1 + t + t);
        const double beta = n + w / s + e * (c + n * yp1);
        p = t + 1.;
        s += 2.;
        t = alpha * an + beta * anp1; an = anp1; anp1 = t;
        t = alpha * bn + beta * bnp1; bn = bnp1; bnp1 = t;
        const double r0 = r;
        r = anp1 / bnp1;
        if (fabs(r - r0) <= eps * r) break;
        an /= bnp1; bn /= bnp1; anp1 = r; bnp1 = 1.;
    } while (n < 10000);
    return brc * r;
}

// regularised incomplete beta I_x(a,b) (w) and its complement (w1); y = 1 - x is passed separately
CBS_HD void bratio(double a, double b, double x, double y, double* w, double* w1) {
    if (isnan(a) || isnan(b) || isnan(x) || isnan(y) || a < 0 || b < 0 || x < 0 || x > 1) { *w = *w1 = NAN; return; }
    if (x == 0) { *w = 0; *w1 = 1; return; }
    if (y == 0) { *w = 1; *w1 = 0; return; }
    if (a == 0) { *w = 1; *w1 = 0; return; }
    if (b == 0) { *w = 0; *w1 = 1; return; }
    const double eps = 2 * DBL_EPSILON;
    const double lambda = (a > b) ? (a + b) * y - b : a - (a + b) * x;
    if (lambda >= 0) {
        *w = bfrac(a, b, x, y, lambda, eps);
        *w1 = 0.5 - *w + 0.5;
    } else {
        *w1 = bfrac(b, a, y, x, -lambda, eps);
        *w = 0.5 - *w1 + 0.5;
    }
}

// pnbinom(q; mu, size), lower tail (helper_seq.R:53)
CBS_HD double pnbinom_mu(double x, double size, double mu) {
    if (isnan(x) || isnan(size) || isnan(mu)) return x + size + mu;
    if (!isfinite(mu) || size < 0 || mu < 0) return NAN;
    if (size == 0) return (x >= 0) ? 1.0 : 0.0;
    if (x < 0) return 0.0;
    if (!isfinite(x)) return 1.0;
    x = floor(x + 1e-7);
    double w, wc;
    bratio(size, x + 1, size / (size + mu), mu / (size + mu), &w, &wc);
    return w;
}

// pnbinom(q; size, prob), lower tail
CBS_HD double pnbinom_prob(double x, double size, double prob) {
    if (isnan(x) || isnan(size) || isnan(prob)) return x + size + prob;
    if (!isfinite(size) || !isfinite(prob) || size < 0 || prob <= 0 || prob > 1) return NAN;
    if (size == 0) return (x >= 0) ? 1.0 : 0.0;
    if (x < 0) return 0.0;
    if (!isfinite(x)) return 1.0;
    x = floor(x + 1e-7);
    double w, wc;
    bratio(size, x + 1, prob, 0.5 - prob + 0.5, &w, &wc);
    return w;
}

// standard normal quantile, AS 241 (PPND16)
CBS_HD double qnorm01(double p) {
    if (isnan(p)) return p;
    if (p <= 0) return (p == 0) ? -INFINITY : nan("");
    if (p >= 1) return (p == 1) ? INFINITY : nan("");
    const double q = p - 0.5;
    double r, val;
    if (fabs(q) <= .425) {
        r = .180625 - q * q;
        val = q * (((((((r * 2509.0809287301226727 + 33430.575583588128105) * r + 67265.770927008700853) * r
                        + 45921.953931549871457) * r + 13731.693765509461125) * r + 1971.5909503065514427) * r
                        + 133.14166789178437745) * r + 3.387132872796366608)
            / (((((((r * 5226.495278852854561 + 28729.085735721942674) * r + 39307.89580009271061) * r
                        + 21213.794301586595867) * r + 5394.1960214247511077) * r + 687.1870074920579083) * r
                        + 42.313330701600911252) * r + 1.);
        return val;
    }
    r = (q < 0) ? p : (0.5 - p + 0.5);
    r = sqrt(-log(r));
    if (r <= 5.) {
        r += -1.6;
        val = (((((((r * 7.7454501427834140764e-4 + .0227238449892691845833) * r + .24178072517745061177) * r
                      + 1.27045825245236838258) * r + 3.64784832476320460504) * r + 5.7694972214606914055) * r
                      + 4.6303378461565452959) * r + 1.42343711074968357734)
            / (((((((r * 1.05075007164441684324e-9 + 5.475938084995344946e-4) * r + .0151986665636164571966) * r
                      + .14810397642748007459) * r + .68976733498510000455) * r + 1.6763848301838038494) * r
                      + 2.05319162663775882187) * r + 1.);
    } else {
        r += -5.;
        val = (((((((r * 2.01033439929228813265e-7 + 2.71155556874348757815e-5) * r + .0012426609473880784386) * r
                      + .026532189526576123093) * r + .29656057182850489123) * r + 1.7848265399172913358) * r
                      + 5.4637849111641143699) * r + 6.6579046435011037772)
            / (((((((r * 2.04426310338993978564e-15 + 1.4215117583164458887e-7) * r + 1.8463183175100546818e-5) * r
                      + 7.868691311456132591e-4) * r + .0148753612908506148525) * r + .13692988092273580531) * r
                      + .59983220655588793769) * r + 1.);
    }
    return (q < 0.0) ? -val : val;
}

// discrete search on the incr-grid: returns y with F(y) >= p and F(y - incr) < p (or y == 0); *z = F(y)
CBS_HD double qnb_search(double y, double* z, double p, double n, double pr, double incr) {
    if (*z >= p) {
        for (;;) {
            double newz;
            if (y == 0 || (newz = pnbinom_prob(y - incr, n, pr)) < p) return y;
            y = fmax(0.0, y - incr);
            *z = newz;
        }
    } else {
        for (;;) {
            y = y + incr;
            if ((*z = pnbinom_prob(y, n, pr)) >= p) return y;
        }
    }
}

// qnbinom(p; size, prob): smallest integer y >= 0 with pnbinom(y) >= p*(1-64*DBL_EPSILON)
CBS_HD double qnbinom_prob(double p, double size, double prob) {
    if (isnan(p) || isnan(size) || isnan(prob)) return p + size + prob;
    if (prob == 0 && size == 0) return 0;
    if (prob <= 0 || prob > 1 || size < 0) return NAN;
    if (prob == 1 || size == 0) return 0;
    if (p < 0 || p > 1) return NAN;
    if (p == 0) return 0;
    if (p == 1) return INFINITY;
    const double Q = 1.0 / prob, P = (1.0 - prob) * Q, mu = size * P, sigma = sqrt(size * P * Q);
    const double gamma = (Q + P) / sigma;
    if (p + 1.01 * DBL_EPSILON >= 1.) return INFINITY;
    double z = qnorm01(p);
    double y = nearbyint(mu + sigma * (z + gamma * (z * z - 1) / 6));
    if (!(y >= 0)) y = 0;      // Cornish-Fisher can undershoot 0 for very skewed distributions; F(y<0) = 0 anyway
    z = pnbinom_prob(y, size, prob);
    p *= 1 - 64 * DBL_EPSILON;
    if (y < 1e5) return qnb_search(y, &z, p, size, prob, 1);
    double incr = floor(y * 0.001), oldincr;
    do {
        oldincr = incr;
        y = qnb_search(y, &z, p, size, prob, incr);
        incr = fmax(1.0, floor(incr / 100));
    } while (oldincr > 1 && incr > y * 1e-15);
    return y;
}

// qnbinom(p; mu, size) (helper_seq.R:58)
CBS_HD double qnbinom_mu(double p, double size, double mu) {
    if (!isfinite(size)) return NAN;
    return qnbinom_prob(p, size, size / (size + mu));
}

// ---- fast path used by match_quantiles -----------------------------------------------------------------------
// Same mathematics as pnbinom/qnbinom above (TOMS 708 continued fraction, Cornish-Fisher start, smallest y with
// F(y) >= p(1 - 64 eps)), reorganised for the GPU: (i) the continued fraction uses three reciprocals per term
// instead of seven divisions; (ii) the pmf comes for free from the fraction's prefactor,
// x^a y^b / B(a, k+1) = pmf(k) * y * (a + k); (iii) qnbinom's unit-step search, which costs one continued fraction
// per step and up to hundreds of steps for skewed distributions, becomes a walk along the pmf recurrence
// F(y+1) = F(y) + pmf(y+1), pmf(y+1) = pmf(y) (y + size)/(y + 1) q.  The walk's rounding error is bounded; whenever
// the decision F(y) >= p is closer than that bound the result is re-derived with continued fractions only, so the
// answer is the one the unit-step search gives.
struct NbCdf { double F, pmf; };

CBS_HD double bfrac_fast(double a, double b, double x, double y, double lambda) {
    const double eps = 2 * DBL_EPSILON;
    const double c = lambda + 1., c0 = b / a, inva = 1. / a, c1 = inva + 1., yp1 = y + 1.;
    double n = 0., p = 1., s = a + 1., an = 0., bn = 1., anp1 = 1., bnp1 = c / c1, r = c1 / c;
    do {
        n += 1.;
        const double t = n * inva, w = n * (b - n) * x, rs = 1. / s;
        double e = a * rs;
        const double alpha = p * (p + c0) * e * e * (w * x);
        e = (t + 1.) / (c1 + t + t);
        const double beta = n + w * rs + e * (c + n * yp1);
        p = t + 1.;
        s += 2.;
        const double ta = alpha * an + beta * anp1, tb = alpha * bn + beta * bnp1;
        const double inv = 1. / tb, r0 = r;
        r = ta * inv;
        if (fabs(r - r0) <= eps * r) return r;
        an = anp1 * inv; bn = bnp1 * inv; anp1 = r; bnp1 = 1.;
    } while (n < 10000);
    return NAN;                                               // not converged: the callers take the general bratio routine
}

// F = P(X <= k), pmf = P(X = k) for X ~ NB(size, prob), q = 1 - prob, k a non-negative integer, 0 < prob < 1
CBS_HD NbCdf nb_cdf_pmf(double k, double size, double prob, double q) {
    NbCdf o;
    const double a = size, b = k + 1.;
    const double brc = brcomp(a, b, prob, q);
    o.pmf = brc / (q * (a + k));
    const double lambda = (a > b) ? (a + b) * q - b : a - (a + b) * prob;
    if (brc == 0.) { o.F = (lambda >= 0) ? 0.0 : 1.0; return o; }      // prefactor underflow: far tail
    if (lambda >= 0) o.F = brc * bfrac_fast(a, b, prob, q, lambda);
    else o.F = 0.5 - brc * bfrac_fast(b, a, q, prob, -lambda) + 0.5;
    if (!(o.F == o.F)) o.F = pnbinom_mu(k, size, size * q / prob);      // (the fraction hit its term cap: R's bratio picks another expansion there)
    return o;
}

// Out-of-line copies (NI = true in the templates below).  Fully inlined, match_one_fast is ~100 000 SASS instructions
// (1.6 MB: seven copies of the continued fraction, two of the general nmath routines), far beyond the instruction
// cache, and its warps diverge onto different copies (ncu, round 1: "no instruction" = 24 % of the stall cycles).
// Same arithmetic, same results; only the call structure differs.
static __host__ __device__ __noinline__ double bfrac_fast_ni(double a, double b, double x, double y, double lambda) { return bfrac_fast(a, b, x, y, lambda); }
static __host__ __device__ __noinline__ double pnbinom_mu_ni(double x, double size, double mu) { return pnbinom_mu(x, size, mu); }
static __host__ __device__ __noinline__ NbCdf nb_cdf_pmf_ni(double k, double size, double prob, double q) {
    NbCdf o;                                                  // nb_cdf_pmf with ONE copy of the continued fraction
    const double a = size, b = k + 1.;
    const double brc = brcomp(a, b, prob, q);
    o.pmf = brc / (q * (a + k));
    const double lambda = (a > b) ? (a + b) * q - b : a - (a + b) * prob;
    if (brc == 0.) { o.F = (lambda >= 0) ? 0.0 : 1.0; return o; }
    const bool fwd = lambda >= 0;
    const double cf = bfrac_fast_ni(fwd ? a : b, fwd ? b : a, fwd ? prob : q, fwd ? q : prob, fwd ? lambda : -lambda);
    o.F = fwd ? brc * cf : 0.5 - brc * cf + 0.5;
    if (!(cf == cf)) o.F = pnbinom_mu_ni(k, size, size * q / prob);     // (the fraction hit its term cap: R's bratio picks another expansion there)
    return o;
}
static __host__ __device__ __noinline__ double qnbinom_mu_ni(double p, double size, double mu) { return qnbinom_mu(p, size, mu); }
template <bool NI> CBS_HD NbCdf nb_cdf_pmf_t(double k, double size, double prob, double q) {
    if (NI) return nb_cdf_pmf_ni(k, size, prob, q);
    return nb_cdf_pmf(k, size, prob, q);
}

// unit-step search with continued fractions only (qnbinom's do_search with incr = 1), from y with z = F(y)
template <bool NI = false>
CBS_HD double nb_search_exact(double y, double z, double p, double size, double prob, double q) {
    if (z >= p) {
        for (;;) {
            if (y == 0) return y;
            if (nb_cdf_pmf_t<NI>(y - 1, size, prob, q).F < p) return y;
            y -= 1;
        }
    }
    for (;;) {
        y += 1;
        if (nb_cdf_pmf_t<NI>(y, size, prob, q).F >= p) return y;
    }
}

// qnbinom(p; size, prob) for 0 < p < 1, 0 < prob < 1, size > 0 (the callers handle the boundaries)
template <bool NI = false>
CBS_HD double nb_quantile_fast(double p, double size, double prob) {
    const double q = 0.5 - prob + 0.5;
    const double Q = 1.0 / prob, P = (1.0 - prob) * Q, m = size * P, sigma = sqrt(size * P * Q);
    const double gam = (Q + P) / sigma;
    const double zn = qnorm01(p);
    double y = nearbyint(m + sigma * (zn + gam * (zn * zn - 1) / 6));
    if (!(y >= 0)) y = 0;
    const double pp = p * (1 - 64 * DBL_EPSILON);
    NbCdf c = nb_cdf_pmf_t<NI>(y, size, prob, q);
    // up to two Newton refinements on the continuous extension when the start is far off (pmf ~ density)
    for (int it = 0; it < 2; ++it) {
        if (!(c.pmf > 0)) break;
        double d = (pp - c.F) / c.pmf;
        if (fabs(d) < 12.0) break;
        d = fmin(fmax(d, -0.5 * y), fmax(y, 16.0));
        const double y2 = fmax(0.0, nearbyint(y + d));
        if (y2 == y) break;
        y = y2;
        c = nb_cdf_pmf_t<NI>(y, size, prob, q);
    }
    const double F0 = c.F;
    double F = c.F, f = c.pmf, Fm, steps = 0;
    bool bad = !(f > 0) || !isfinite(f);
    if (!bad) {
        if (F >= pp) {                                         // walk left to the smallest y with F(y) >= pp
            for (;;) {
                Fm = F - f;                                    // F(y - 1)
                if (y == 0 || Fm < pp) break;
                f = f * y / ((y - 1 + size) * q);
                y -= 1; F = Fm; steps += 1;
                if (!(f > 0)) { bad = true; break; }
            }
        } else {                                               // walk right until F(y) >= pp
            for (;;) {
                f = f * ((y + size) / (y + 1)) * q;
                y += 1; F += f; steps += 1;
                if (F >= pp) break;
                if (!(f > 1e-300) || steps > 2e6) { bad = true; break; }
            }
            Fm = F - f;
        }
    }
    if (!bad) {
        const double tol = (steps + 16.0) * 4e-15 * fmax(F0, fmax(F, pp));
        const bool tie = fabs(F - pp) <= tol || (y > 0 && fabs(Fm - pp) <= tol);
        if (!tie) return y;
    }
    return nb_search_exact<NI>(y, nb_cdf_pmf_t<NI>(y, size, prob, q).F, pp, size, prob, q);
}

// Small counts: both distribution functions by direct summation of the pmf recurrence from 0,
//   pmf(0) = prob^size,  pmf(j+1) = pmf(j) (j + size)/(j + 1) q,   F(k) = sum_{j<=k} pmf(j),
// three FP64 instructions per term with the reciprocals 1/(j+1) from a table, against ~1500 for one continued fraction:
// p = F_old(x - 1) costs x terms and the quantile is found by summing the NEW distribution until it reaches
// p (1 - 64 eps) (no Cornish-Fisher start, no refinement).  Each sum carries a relative error below 1e-13 (term
// recurrence + exp(size log prob)); whenever a decision (the |p - 1| < 1e-4 pass-through, F_new(y) >= p') is closer than
// 1e-12 the element is handed to the continued-fraction routines below, so the answer is the one they give.
// inv[j] = 1/j for 1 <= j < MQ_INV_N.  Returns false when the element was not decided here.
constexpr int MQ_INV_N = 1024;
constexpr double MQ_SMALL_X = 320.0;
CBS_HD bool match_small(double x, double old_mu, double so, double new_mu, double sn, const double* __restrict__ inv, double* out) {
    const int xi = (int)x;                                       // 2 <= x <= MQ_SMALL_X
    if (so < 0.05 || sn < 0.05) return false;                    // (the DFMA form of the term ratio cancels for sizes near 0)
    // prob = size / (size + mu) rounds to 1 for dispersions below ~1e-16 mu^-1 (what Brent's lower end gives on Poisson-like
    // data): R 3.5's qnbinom returns 0 there ("if (prob == 1 || size == 0) return 0"), and so do the general routines below
    if (!(sn / (sn + new_mu) < 1)) return false;
    const double qo = old_mu / (so + old_mu);
    double f = exp(-so * log1p(old_mu / so));                    // pmf_old(0) = (so / (so + mu))^so
    if (!(f > 1e-250)) return false;
    // ratio of consecutive terms: (j + s) q / (j + 1) = q + q (s - 1) / (j + 1): one DFMA on the table entry
    const double ao = qo * (so - 1);
    double F = f;
    for (int j = 1; j < xi; ++j) { f *= fma(ao, inv[j], qo); F += f; }      // F = P_old(X <= x - 1)
    const double d1 = fabs(F - 1);
    if (fabs(d1 - 1e-4) < 1e-11 || !(F > 0)) return false;
    if (d1 < 1e-4) { *out = x; return true; }                    // helper_seq.R:54-56
    const double qn = new_mu / (sn + new_mu);
    double g = exp(-sn * log1p(new_mu / sn));
    if (!(g > 1e-250)) return false;
    const double an = qn * (sn - 1);
    const double pp = F * (1 - 64 * DBL_EPSILON);
    double Fn = g, Fm = 0;
    int y = 0;
    while (Fn < pp) {
        if (y >= MQ_INV_N - 2) return false;
        ++y;
        g *= fma(an, inv[y], qn);
        Fm = Fn; Fn += g;
    }
    const double tol = 1e-12 * fmax(Fn, pp);
    if (fabs(Fn - pp) <= tol || (y > 0 && fabs(Fm - pp) <= tol)) return false;
    *out = 1 + (double)y;                                        // helper_seq.R:58
    return true;
}

// match_quantiles for one element (helper_seq.R:46-64); inv (optional): the reciprocal table of match_small
template <bool NI = false>
CBS_HD double match_one_fast(double x, double old_mu, double old_phi, double new_mu, double new_phi,
                             const double* __restrict__ inv = nullptr) {
    if (x <= 1) return x;                                      // helper_seq.R:50-51
    const double so = 1.0 / old_phi, sn = 1.0 / new_phi;
    if (!(old_mu > 0) || !(new_mu > 0) || !isfinite(old_mu) || !isfinite(new_mu) || !(so > 0) || !(sn > 0) ||
        !isfinite(so) || !isfinite(sn)) {                      // degenerate parameters: the general routines
        const double tmp_p = NI ? pnbinom_mu_ni(x - 1, so, old_mu) : pnbinom_mu(x - 1, so, old_mu);
        if (fabs(tmp_p - 1) < 1e-4) return x;
        return 1 + (NI ? qnbinom_mu_ni(tmp_p, sn, new_mu) : qnbinom_mu(tmp_p, sn, new_mu));
    }
    if (inv != nullptr && x <= MQ_SMALL_X) {
        double r;
        if (match_small(x, old_mu, so, new_mu, sn, inv, &r)) return r;
    }
    const double tmp_p = nb_cdf_pmf_t<NI>(x - 1, so, so / (so + old_mu), old_mu / (so + old_mu)).F;     // helper_seq.R:53
    if (fabs(tmp_p - 1) < 1e-4) return x;                      // helper_seq.R:54-56
    const double prob = sn / (sn + new_mu);
    if (!(tmp_p > 0) || !(prob > 0) || !(prob < 1)) return 1 + (NI ? qnbinom_mu_ni(tmp_p, sn, new_mu) : qnbinom_mu(tmp_p, sn, new_mu));
    return 1 + nb_quantile_fast<NI>(tmp_p, sn, prob);          // helper_seq.R:58
}

}  // namespace cbs
