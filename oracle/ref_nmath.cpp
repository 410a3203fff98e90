// ORACLE — TEST INFRASTRUCTURE ONLY (see ref_nmath.hpp).
#include "ref_nmath.hpp"
#include <algorithm>
#include <limits>

namespace ref {

static const double kLnSqrt2Pi = 0.918938533204672741780329736406;  // log(sqrt(2*pi))
static const double kLn2Pi     = 1.837877066409345483560659472811;  // log(2*pi)

// nmath/stirlerr.c: stirlerr(n) = log(n!) - log( sqrt(2*pi*n)*(n/e)^n ).
// R keeps a 31-entry table for n = 0, 0.5, ..., 15 and uses lgammafn for other n <= 15; both are the
// same mathematical function, evaluated here in 80-bit long double so no table has to be transcribed.
double stirlerr(double n) {
    const double S0 = 0.083333333333333333333;        /* 1/12 */
    const double S1 = 0.00277777777777777777778;      /* 1/360 */
    const double S2 = 0.00079365079365079365079365;   /* 1/1260 */
    const double S3 = 0.000595238095238095238095238;  /* 1/1680 */
    const double S4 = 0.0008417508417508417508417508; /* 1/1188 */
    if (n <= 15.0) {
        if (n == 0.0) return 0.0;
        long double ln = n;
        return (double)(lgammal(ln + 1.0L) - (ln + 0.5L) * logl(ln) + ln - 0.918938533204672741780329736406L);
    }
    double nn = n * n;
    if (n > 500) return (S0 - S1 / nn) / n;
    if (n > 80) return (S0 - (S1 - S2 / nn) / nn) / n;
    if (n > 35) return (S0 - (S1 - (S2 - S3 / nn) / nn) / nn) / n;
    return (S0 - (S1 - (S2 - (S3 - S4 / nn) / nn) / nn) / nn) / n;
}

// nmath/bd0.c: bd0(x,np) = x log(x/np) + np - x, stable for x ~ np.
double bd0(double x, double np) {
    if (!std::isfinite(x) || !std::isfinite(np) || np == 0.0) return std::numeric_limits<double>::quiet_NaN();
    if (std::fabs(x - np) < 0.1 * (x + np)) {
        double v = (x - np) / (x + np);
        double s = (x - np) * v;
        if (std::fabs(s) < DBL_MIN) return s;
        double ej = 2 * x * v;
        v = v * v;
        for (int j = 1; j < 1000; j++) {
            ej *= v;
            double s1 = s + ej / ((j << 1) + 1);
            if (s1 == s) return s1;
            s = s1;
        }
    }
    return x * std::log(x / np) + np - x;
}

// nmath/dbinom.c dbinom_raw, give_log = FALSE
double dbinom_raw(double x, double n, double p, double q) {
    if (p == 0) return (x == 0) ? 1.0 : 0.0;
    if (q == 0) return (x == n) ? 1.0 : 0.0;
    double lc;
    if (x == 0) {
        if (n == 0) return 1.0;
        lc = (p < 0.1) ? -bd0(n, n * q) - n * p : n * std::log(q);
        return std::exp(lc);
    }
    if (x == n) {
        lc = (q < 0.1) ? -bd0(n, n * p) - n * q : n * std::log(p);
        return std::exp(lc);
    }
    if (x < 0 || x > n) return 0.0;
    lc = stirlerr(n) - stirlerr(x) - stirlerr(n - x) - bd0(x, n * p) - bd0(n - x, n * q);
    double lf = kLn2Pi + std::log(x) + std::log1p(-x / n);
    return std::exp(lc - 0.5 * lf);
}

// nmath/dnbinom.c dnbinom_mu, give_log = FALSE
double dnbinom_mu(double x, double size, double mu) {
    if (std::isnan(x) || std::isnan(size) || std::isnan(mu)) return x + size + mu;
    if (mu < 0 || size < 0) return std::numeric_limits<double>::quiet_NaN();
    if (std::fabs(x - std::nearbyint(x)) > 1e-7 * std::max(1.0, std::fabs(x))) return 0.0;  // R_D_nonint_check
    if (x < 0 || !std::isfinite(x)) return 0.0;
    if (x == 0 && size == 0) return 1.0;
    x = std::nearbyint(x);
    if (!std::isfinite(size)) {  // limit case: Poisson
        if (mu == 0) return (x == 0) ? 1.0 : 0.0;
        if (x == 0) return std::exp(-mu);
        return std::exp(-stirlerr(x) - bd0(x, mu)) / std::sqrt(2 * M_PI * x);
    }
    if (x == 0)
        return std::exp(size * (size < mu ? std::log(size / (size + mu)) : std::log1p(-mu / (size + mu))));
    if (x < 1e-10 * size) {
        double p = (size < mu ? std::log(size / (1 + size / mu)) : std::log(mu / (1 + mu / size)));
        return std::exp(x * p - mu - std::lgamma(x + 1) + std::log1p(x * (x - 1) / (2 * size)));
    }
    double p = size / (size + x);
    double ans = dbinom_raw(size, x + size, size / (size + mu), mu / (size + mu));
    return p * ans;
}

// x^a y^b / B(a,b).  toms708.c brcomp() in its a0,b0 >= 8 form (saddle point: rlog1 + bcorr) is the same
// quantity as the bd0/stirlerr form used here, which is valid for all a,b > 0.
static double brcomp(double a, double b, double x, double y) {
    if (x == 0 || y == 0) return 0.0;
    double n = a + b;
    double lc = stirlerr(n) - stirlerr(a) - stirlerr(b) - bd0(a, n * x) - bd0(b, n * y);
    return std::sqrt(a * b / (n * 2 * M_PI)) * std::exp(lc);
}

// toms708.c bfrac(): continued fraction expansion for I_x(a,b), lambda = (a+b)y - b = a - (a+b)x >= 0.
static double bfrac(double a, double b, double x, double y, double lambda, double eps) {
    double brc = brcomp(a, b, x, y);
    if (brc == 0.) return 0.;
    double c = lambda + 1., c0 = b / a, c1 = 1. / a + 1., yp1 = y + 1.;
    double n = 0., p = 1., s = a + 1., an = 0., bn = 1., anp1 = 1., bnp1 = c / c1, r = c1 / c;
    do {
        n += 1.;
        double t = n / a, w = n * (b - n) * x, e = a / s;
        double alpha = p * (p + c0) * e * e * (w * x);
        e = (t + 1.) / (c1 + t + t);
        double beta = n + w / s + e * (c + n * yp1);
        p = t + 1.;
        s += 2.;
        t = alpha * an + beta * anp1; an = anp1; anp1 = t;
        t = alpha * bn + beta * bnp1; bn = bnp1; bnp1 = t;
        double r0 = r;
        r = anp1 / bnp1;
        if (std::fabs(r - r0) <= eps * r) break;
        an /= bnp1; bn /= bnp1; anp1 = r; bnp1 = 1.;
    } while (n < 10000);
    return brc * r;
}

// toms708.c bratio(): R dispatches over bpser/bup/bgrat/bfrac/basym by regime.  This restatement routes
// every (a,b,x) through the bfrac continued fraction on whichever side has lambda >= 0 (the same
// side-selection rule as toms708.c), which the golden vectors show is accurate to ~1e-14 over the NB
// domain (a = size in [1e-4,1e4], b = count+1 integer).
void bratio(double a, double b, double x, double y, double* w, double* w1) {
    if (std::isnan(a) || std::isnan(b) || std::isnan(x) || std::isnan(y) || a < 0 || b < 0 || x < 0 || x > 1) {
        *w = *w1 = std::numeric_limits<double>::quiet_NaN(); return;
    }
    if (x == 0) { *w = 0; *w1 = 1; return; }
    if (y == 0) { *w = 1; *w1 = 0; return; }
    if (a == 0) { *w = 1; *w1 = 0; return; }
    if (b == 0) { *w = 0; *w1 = 1; return; }
    const double eps = 2 * DBL_EPSILON;
    double lambda = (a > b) ? (a + b) * y - b : a - (a + b) * x;
    if (lambda >= 0) {
        *w = bfrac(a, b, x, y, lambda, eps);
        *w1 = 0.5 - *w + 0.5;
    } else {
        *w1 = bfrac(b, a, y, x, -lambda, eps);
        *w = 0.5 - *w1 + 0.5;
    }
}

// nmath/pnbinom.c pnbinom_mu(x, size, mu, lower_tail = TRUE, log_p = FALSE)
double pnbinom_mu(double x, double size, double mu) {
    if (std::isnan(x) || std::isnan(size) || std::isnan(mu)) return x + size + mu;
    if (!std::isfinite(mu) || size < 0 || mu < 0) return std::numeric_limits<double>::quiet_NaN();
    if (size == 0) return (x >= 0) ? 1.0 : 0.0;
    if (x < 0) return 0.0;
    if (!std::isfinite(x)) return 1.0;
    x = std::floor(x + 1e-7);
    double w, wc;
    bratio(size, x + 1, size / (size + mu), mu / (size + mu), &w, &wc);
    return w;
}

// nmath/pnbinom.c pnbinom(x, size, prob) -> pbeta(prob, size, x+1) -> bratio(size, x+1, prob, 0.5-prob+0.5)
double pnbinom_prob(double x, double size, double prob) {
    if (std::isnan(x) || std::isnan(size) || std::isnan(prob)) return x + size + prob;
    if (!std::isfinite(size) || !std::isfinite(prob) || size < 0 || prob <= 0 || prob > 1)
        return std::numeric_limits<double>::quiet_NaN();
    if (size == 0) return (x >= 0) ? 1.0 : 0.0;
    if (x < 0) return 0.0;
    if (!std::isfinite(x)) return 1.0;
    x = std::floor(x + 1e-7);
    double w, wc;
    bratio(size, x + 1, prob, 0.5 - prob + 0.5, &w, &wc);
    return w;
}

// nmath/qnorm.c: Wichura (1988) AS 241, PPND16; lower tail, mu = 0, sigma = 1.
double qnorm01(double p) {
    if (std::isnan(p)) return p;
    if (p <= 0) return (p == 0) ? -INFINITY : std::numeric_limits<double>::quiet_NaN();
    if (p >= 1) return (p == 1) ? INFINITY : std::numeric_limits<double>::quiet_NaN();
    double q = p - 0.5, r, val;
    if (std::fabs(q) <= .425) {
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
    r = std::sqrt(-std::log(r));
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
    if (q < 0.0) val = -val;
    return val;
}

// nmath/qnbinom.c do_search().
// R < 4.2 quirk: the left search overwrites *z with pnbinom(y - incr) (< p) before returning y, so for
// y >= 1e5 (multi-pass search with shrinking incr) the next pass believes F(y) < p, steps RIGHT and the
// function returns a y that is up to floor(y*0.001) above the smallest y with F(y) >= p.  R >= 4.2 rewrote
// the search (qDiscrete_search.h) and returns the smallest y.  The oracle's contract is the documented one
// (smallest y with pnbinom(y) >= p*(1-64*DBL_EPSILON)); g_legacy_search = 1 reproduces the R < 4.2 quirk.
static int g_legacy_search = 0;
void set_legacy_qnbinom_search(int on) { g_legacy_search = on; }

static double do_search(double y, double* z, double p, double n, double pr, double incr) {
    if (*z >= p) {  /* search to the left */
        for (;;) {
            double newz;
            if (y == 0 || (newz = pnbinom_prob(y - incr, n, pr)) < p) {
                if (g_legacy_search && y != 0) *z = newz;
                return y;
            }
            y = std::max(0.0, y - incr);
            *z = newz;
        }
    } else {        /* search to the right */
        for (;;) {
            y = y + incr;
            if ((*z = pnbinom_prob(y, n, pr)) >= p) return y;
        }
    }
}

// nmath/qnbinom.c qnbinom(p, size, prob, lower_tail = TRUE, log_p = FALSE), R 3.5.x
double qnbinom_prob(double p, double size, double prob) {
    if (std::isnan(p) || std::isnan(size) || std::isnan(prob)) return p + size + prob;
    if (prob == 0 && size == 0) return 0;
    if (prob <= 0 || prob > 1 || size < 0) return std::numeric_limits<double>::quiet_NaN();
    if (prob == 1 || size == 0) return 0;
    if (p < 0 || p > 1) return std::numeric_limits<double>::quiet_NaN();   // R_Q_P01_boundaries
    if (p == 0) return 0;
    if (p == 1) return INFINITY;
    double Q = 1.0 / prob, P = (1.0 - prob) * Q, mu = size * P, sigma = std::sqrt(size * P * Q);
    double gamma = (Q + P) / sigma;
    if (p + 1.01 * DBL_EPSILON >= 1.) return INFINITY;
    double z = qnorm01(p);
    double y = std::nearbyint(mu + sigma * (z + gamma * (z * z - 1) / 6));
    z = pnbinom_prob(y, size, prob);
    p *= 1 - 64 * DBL_EPSILON;   /* fuzz to ensure left continuity */
    if (y < 1e5) return do_search(y, &z, p, size, prob, 1);
    double incr = std::floor(y * 0.001), oldincr;
    do {
        oldincr = incr;
        y = do_search(y, &z, p, size, prob, incr);
        incr = std::max(1.0, std::floor(incr / 100));
    } while (oldincr > 1 && incr > y * 1e-15);
    return y;
}

// nmath/qnbinom.c qnbinom_mu (R 3.5.x): prob = size/(size+mu), then qnbinom()
double qnbinom_mu(double p, double size, double mu) {
    if (!std::isfinite(size)) return std::numeric_limits<double>::quiet_NaN();  // Poisson limit not on the path
    return qnbinom_prob(p, size, size / (size + mu));
}

// Cross-check only: F(k) by summing the pmf recurrence in long double from the mode outwards is not needed;
// a plain ascending sum has only positive terms and is accurate to ~1e-18 relative.
double pnbinom_mu_bysum(double q, double size, double mu) {
    if (q < 0) return 0.0;
    long double r = size, m = mu;
    long double lp = (size < mu) ? logl(r / (r + m)) : log1pl(-m / (r + m));
    long double qq = m / (r + m);
    long double term = expl(r * lp), tot = term;
    long long k = (long long)std::floor(q + 1e-7);
    for (long long j = 0; j < k; ++j) {
        term = term * ((j + r) / (j + 1)) * qq;
        tot += term;
    }
    return (double)tot;
}

}  // namespace ref
