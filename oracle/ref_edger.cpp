// ORACLE — TEST INFRASTRUCTURE ONLY (see ref_edger.hpp).
#include "ref_edger.hpp"
#include <cmath>
#include <cfloat>
#include <algorithm>
#include <limits>
#include <numeric>
#include <map>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace ref {

static const double low_value = 1e-10, supremely_low_value = 1e-13, ridiculously_low_value = 1e-100,
                    mildly_low_value = 1e-8;
static const double kNaN = std::numeric_limits<double>::quiet_NaN();

// ---------------------------------------------------------------------------------------------------
// LINPACK dqrdc2 (R's default qr()): Householder QR with limited column pivoting; returns the rank k.
int qr_rank(const double* Xin, int n, int p, double tol) {
    if (p == 0) return 0;
    std::vector<double> x(Xin, Xin + (size_t)n * p), qraux(p), w1(p), w2(p);
    auto X = [&](int i, int j) -> double& { return x[(size_t)j * n + i]; };
    auto nrm2 = [&](int i0, int j) { double s = 0; for (int i = i0; i < n; ++i) s += X(i, j) * X(i, j); return std::sqrt(s); };
    for (int j = 0; j < p; ++j) { qraux[j] = nrm2(0, j); w1[j] = qraux[j]; w2[j] = qraux[j]; if (w2[j] == 0) w2[j] = 1; }
    int lup = std::min(n, p), k = p + 1;   // k is 1-based count as in the Fortran
    for (int l = 1; l <= lup; ++l) {
        int l0 = l - 1;
        while (l < k && qraux[l0] < w2[l0] * tol) {
            for (int i = 0; i < n; ++i) { double t = X(i, l0); for (int j = l0 + 1; j < p; ++j) X(i, j - 1) = X(i, j); X(i, p - 1) = t; }
            double t = qraux[l0], tt = w1[l0], ttt = w2[l0];
            for (int j = l0 + 1; j < p; ++j) { qraux[j - 1] = qraux[j]; w1[j - 1] = w1[j]; w2[j - 1] = w2[j]; }
            qraux[p - 1] = t; w1[p - 1] = tt; w2[p - 1] = ttt;
            --k;
        }
        if (l == n) break;
        double nrmxl = nrm2(l0, l0);
        if (nrmxl != 0) {
            if (X(l0, l0) != 0) nrmxl = std::copysign(nrmxl, X(l0, l0));
            for (int i = l0; i < n; ++i) X(i, l0) /= nrmxl;
            X(l0, l0) = 1 + X(l0, l0);
            for (int j = l0 + 1; j < p; ++j) {
                double dot = 0; for (int i = l0; i < n; ++i) dot += X(i, l0) * X(i, j);
                double t = -dot / X(l0, l0);
                for (int i = l0; i < n; ++i) X(i, j) += t * X(i, l0);
                if (qraux[j] != 0) {
                    double tt = 1.0 - std::pow(std::fabs(X(l0, j)) / qraux[j], 2);
                    tt = std::max(tt, 0.0);
                    t = tt;
                    if (std::fabs(t) < 1e-6) { qraux[j] = nrm2(l0 + 1, j); w1[j] = qraux[j]; }
                    else qraux[j] = qraux[j] * std::sqrt(t);
                }
            }
            qraux[l0] = X(l0, l0);
            X(l0, l0) = -nrmxl;
        }
    }
    return std::min(k - 1, n);
}

// R/designAsFactor.R: hash each row with powers of z = (e+pi)/5, factor of the distinct values (ascending).
int design_as_factor(const double* design, int n, int p, std::vector<int>& group) {
    const double z = (std::exp(1.0) + M_PI) / 5;
    std::vector<double> h(n);
    for (int s = 0; s < n; ++s) {
        double acc = 0, zp = 1;
        for (int k = 0; k < p; ++k) { acc += design[(size_t)k * n + s] * zp; zp *= z; }
        h[s] = acc / p;
    }
    std::vector<double> lev(h);
    std::sort(lev.begin(), lev.end());
    lev.erase(std::unique(lev.begin(), lev.end()), lev.end());
    group.resize(n);
    for (int s = 0; s < n; ++s) group[s] = (int)(std::lower_bound(lev.begin(), lev.end(), h[s]) - lev.begin());
    return (int)lev.size();
}

// ---------------------------------------------------------------------------------------------------
// src/glm_one_group.cpp (weights == 1, start = NA)
double glm_one_group(int n, const double* y, const double* offset, const double* disp, int maxit, double tol) {
    bool nonzero = false;
    double cur_beta = 0, totweight = 0;
    for (int j = 0; j < n; ++j) {
        if (y[j] > low_value) { cur_beta += y[j] / std::exp(offset[j]); nonzero = true; }
        totweight += 1;
    }
    cur_beta = std::log(cur_beta / totweight);
    if (!nonzero) return -INFINITY;
    for (int i = 0; i < maxit; ++i) {
        double dl = 0, info = 0;
        for (int j = 0; j < n; ++j) {
            const double mu = std::exp(cur_beta + offset[j]), denominator = 1 + mu * disp[j];
            dl += (y[j] - mu) / denominator;
            info += mu / denominator;
        }
        const double step = dl / info;
        cur_beta += step;
        if (std::fabs(step) < tol) break;
    }
    return cur_beta;
}

// src/nbdev.cpp compute_unit_nb_deviance.  "2/3*resid" is C++ integer division (== 0) in edgeR 3.22.
double unit_nb_deviance(double y, double mu, double phi) {
    y += mildly_low_value; mu += mildly_low_value;
    if (phi < 1e-4) {
        const double resid = y - mu;
        return 2 * (y * std::log(y / mu) - resid - 0.5 * resid * resid * phi * (1 + phi * ((2 / 3) * resid - y)));
    }
    const double product = mu * phi;
    if (product > 1e6) return 2 * ((y - mu) / mu - std::log(y / mu)) * mu / (1 + product);
    const double invphi = 1 / phi;
    return 2 * (y * std::log(y / mu) + (y + invphi) * std::log((mu + invphi) / (y + invphi)));
}

// least-squares solution of design * b ~= rhs via Householder QR (get_levenberg_start's QRdecomposition)
static void qr_lstsq(const double* design, int n, int p, const double* rhs, double* sol) {
    std::vector<double> a(design, design + (size_t)n * p), b(rhs, rhs + n);
    auto A = [&](int i, int j) -> double& { return a[(size_t)j * n + i]; };
    for (int k = 0; k < p; ++k) {
        double nrm = 0; for (int i = k; i < n; ++i) nrm += A(i, k) * A(i, k);
        nrm = std::sqrt(nrm);
        if (nrm == 0) continue;
        double alpha = (A(k, k) > 0) ? -nrm : nrm;
        std::vector<double> v(n, 0.0);
        for (int i = k; i < n; ++i) v[i] = A(i, k);
        v[k] -= alpha;
        double vnorm2 = 0; for (int i = k; i < n; ++i) vnorm2 += v[i] * v[i];
        if (vnorm2 == 0) continue;
        for (int j = k; j < p; ++j) {
            double dot = 0; for (int i = k; i < n; ++i) dot += v[i] * A(i, j);
            double f = 2 * dot / vnorm2;
            for (int i = k; i < n; ++i) A(i, j) -= f * v[i];
        }
        double dot = 0; for (int i = k; i < n; ++i) dot += v[i] * b[i];
        double f = 2 * dot / vnorm2;
        for (int i = k; i < n; ++i) b[i] -= f * v[i];
    }
    for (int k = p - 1; k >= 0; --k) {
        double s = b[k];
        for (int j = k + 1; j < p; ++j) s -= A(k, j) * sol[j];
        sol[k] = s / A(k, k);
    }
}

// LAPACK dpotf2 (uplo = 'U') on a dense p x p column-major matrix; returns 0 on success
static int chol_upper(double* a, int p) {
    for (int j = 0; j < p; ++j) {
        double ajj = a[j * p + j];
        for (int k = 0; k < j; ++k) ajj -= a[j * p + k] * a[j * p + k];
        if (ajj <= 0 || std::isnan(ajj)) return j + 1;
        ajj = std::sqrt(ajj);
        a[j * p + j] = ajj;
        for (int c = j + 1; c < p; ++c) {
            double v = a[c * p + j];
            for (int k = 0; k < j; ++k) v -= a[j * p + k] * a[c * p + k];
            a[c * p + j] = v / ajj;
        }
    }
    return 0;
}
// dpotrs with the factor from chol_upper: solve U'U x = b
static void chol_solve_upper(const double* a, int p, double* b) {
    for (int i = 0; i < p; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= a[i * p + k] * b[k]; b[i] = s / a[i * p + i]; }
    for (int i = p - 1; i >= 0; --i) { double s = b[i]; for (int k = i + 1; k < p; ++k) s -= a[k * p + i] * b[k]; b[i] = s / a[i * p + i]; }
}

// src/glm_levenberg.cpp glm_levenberg::fit (weights == 1), maxit = 250, tolerance = 1e-6
struct LevWork { std::vector<double> W, z, xtwx, xtwx_copy, dl, dbeta, beta_new, mu_new; };
static void levenberg_fit(int n, int p, const double* design, const double* y, const double* offset, const double* disp,
                          double* mu, double* beta, LevWork& wk, int maxit = 250, double tolerance = 1e-6) {
    double ymax = 0;
    for (int s = 0; s < n; ++s) if (y[s] > ymax) ymax = y[s];
    if (ymax < low_value) { std::fill(beta, beta + p, kNaN); std::fill(mu, mu + n, 0.0); return; }
    auto autofill = [&](const double* b, double* m) {
        for (int s = 0; s < n; ++s) {
            double eta = offset[s];
            for (int k = 0; k < p; ++k) eta += design[(size_t)k * n + s] * b[k];
            m[s] = std::exp(eta);
        }
    };
    auto nb_dev = [&](const double* m) { double d = 0; for (int s = 0; s < n; ++s) d += unit_nb_deviance(y[s], m[s], disp[s]); return d; };
    autofill(beta, mu);
    double dev = nb_dev(mu);
    double max_info = -1, lambda = 0;
    wk.W.resize(n); wk.z.resize(n); wk.xtwx.assign((size_t)p * p, 0.0); wk.xtwx_copy.resize((size_t)p * p);
    wk.dl.resize(p); wk.dbeta.resize(p); wk.beta_new.resize(p); wk.mu_new.resize(n);
    int iter = 0; bool failed = false;
    while ((++iter) <= maxit) {
        for (int s = 0; s < n; ++s) { const double denom = 1 + mu[s] * disp[s]; wk.W[s] = mu[s] / denom; wk.z[s] = (y[s] - mu[s]) / denom; }
        for (int c1 = 0; c1 < p; ++c1)
            for (int c2 = 0; c2 <= c1; ++c2) {
                double acc = 0;
                for (int s = 0; s < n; ++s) acc += design[(size_t)c1 * n + s] * design[(size_t)c2 * n + s] * wk.W[s];
                wk.xtwx[c1 * p + c2] = acc;
            }
        for (int c = 0; c < p; ++c) { double acc = 0; for (int s = 0; s < n; ++s) acc += design[(size_t)c * n + s] * wk.z[s]; wk.dl[c] = acc; }
        for (int i = 0; i < p; ++i) if (wk.xtwx[i * p + i] > max_info) max_info = wk.xtwx[i * p + i];
        if (iter == 1) { lambda = max_info * 1e-6; if (lambda < 1e-13) lambda = 1e-13; }
        int lev = 0; bool low_dev = false;
        while (++lev) {
            do {
                std::copy(wk.dl.begin(), wk.dl.end(), wk.dbeta.begin());
                std::copy(wk.xtwx.begin(), wk.xtwx.end(), wk.xtwx_copy.begin());
                for (int i = 0; i < p; ++i) wk.xtwx_copy[i * p + i] += lambda;
                if (chol_upper(wk.xtwx_copy.data(), p) != 0) { lambda *= 10; if (lambda <= 0) lambda = ridiculously_low_value; }
                else break;
            } while (1);
            chol_solve_upper(wk.xtwx_copy.data(), p, wk.dbeta.data());
            for (int i = 0; i < p; ++i) wk.beta_new[i] = beta[i] + wk.dbeta[i];
            autofill(wk.beta_new.data(), wk.mu_new.data());
            const double dev_new = nb_dev(wk.mu_new.data());
            if (dev_new / ymax < supremely_low_value) low_dev = true;
            if (dev_new <= dev || low_dev) {
                std::copy(wk.beta_new.begin(), wk.beta_new.end(), beta);
                std::copy(wk.mu_new.begin(), wk.mu_new.end(), mu);
                dev = dev_new;
                break;
            }
            lambda *= 2; if (lambda <= 0) lambda = ridiculously_low_value;
            if (lambda / max_info > 1 / supremely_low_value) { failed = true; break; }
        }
        if (failed) break;
        if (low_dev) break;
        double divergence = 0;
        for (int i = 0; i < p; ++i) divergence += wk.dl[i] * wk.dbeta[i];
        if (divergence < tolerance) break;
        if (lev == 1) lambda /= 10;
    }
}

// solve a small dense system A X = B (A p x p column-major, B p x m) by LU with partial pivoting: R's solve()
static bool lu_solve(std::vector<double> A, int p, std::vector<double>& B, int m) {
    std::vector<int> piv(p);
    for (int k = 0; k < p; ++k) {
        int best = k; double bv = std::fabs(A[(size_t)k * p + k]);
        for (int i = k + 1; i < p; ++i) if (std::fabs(A[(size_t)k * p + i]) > bv) { bv = std::fabs(A[(size_t)k * p + i]); best = i; }
        if (bv == 0) return false;
        if (best != k) {
            for (int j = 0; j < p; ++j) std::swap(A[(size_t)j * p + k], A[(size_t)j * p + best]);
            for (int j = 0; j < m; ++j) std::swap(B[(size_t)j * p + k], B[(size_t)j * p + best]);
        }
        for (int i = k + 1; i < p; ++i) {
            double f = A[(size_t)k * p + i] / A[(size_t)k * p + k];
            A[(size_t)k * p + i] = f;
            for (int j = k + 1; j < p; ++j) A[(size_t)j * p + i] -= f * A[(size_t)j * p + k];
            for (int j = 0; j < m; ++j) B[(size_t)j * p + i] -= f * B[(size_t)j * p + k];
        }
    }
    for (int j = 0; j < m; ++j)
        for (int i = p - 1; i >= 0; --i) {
            double s = B[(size_t)j * p + i];
            for (int k = i + 1; k < p; ++k) s -= A[(size_t)k * p + i] * B[(size_t)j * p + k];
            B[(size_t)j * p + i] = s / A[(size_t)i * p + i];
        }
    return true;
}

// One plain fit (prior.count = 0) of all genes: R/glmfit.R dispatch between mglmOneWay and mglmLevenberg.
// yfun/offfun give the (possibly prior-augmented) response and offset.
template <class YF, class OF>
static void fit_all(long G, int n, const double* design, int p, const Bcast& disp, YF yfun, OF offfun,
                    double* coef, double* fitted, int nthreads) {
    std::vector<int> group;
    int nlev = design_as_factor(design, n, p, group);
    const bool oneway = (nlev == p);
    std::vector<std::vector<int>> members(nlev);
    for (int s = 0; s < n; ++s) members[group[s]].push_back(s);
    std::vector<double> designunique;          // p x p, rows = first sample of each level
    std::vector<double> lev_start(p, 0.0);     // least-squares solution of design*b ~= 1
    if (oneway) {
        designunique.resize((size_t)p * p);
        for (int l = 0; l < nlev; ++l)
            for (int k = 0; k < p; ++k) designunique[(size_t)k * p + l] = design[(size_t)k * n + members[l][0]];
    } else {
        std::vector<double> ones(n, 1.0);
        qr_lstsq(design, n, p, ones.data(), lev_start.data());
    }
#ifdef _OPENMP
#pragma omp parallel num_threads(nthreads > 0 ? nthreads : 1)
#endif
    {
        std::vector<double> yy(n), off(n), dd(n), mu(n), beta(p), ysub(n), osub(n), dsub(n);
        LevWork wk;
#ifdef _OPENMP
#pragma omp for schedule(dynamic, 64)
#endif
        for (long g = 0; g < G; ++g) {
            for (int s = 0; s < n; ++s) { yy[s] = yfun(g, s); off[s] = offfun(g, s); dd[s] = disp.at(g, s); }
            if (oneway) {
                std::vector<double> gb(p);
                for (int l = 0; l < nlev; ++l) {
                    int m = (int)members[l].size();
                    for (int t = 0; t < m; ++t) { int s = members[l][t]; ysub[t] = yy[s]; osub[t] = off[s]; dsub[t] = dd[s]; }
                    double b = glm_one_group(m, ysub.data(), osub.data(), dsub.data(), 50, 1e-10);
                    gb[l] = std::max(b, -1e8);                          // mglmOneWay: pmax(groupbeta, -1e8)
                }
                for (int s = 0; s < n; ++s) mu[s] = std::exp(gb[group[s]] + off[s]);
                std::vector<double> rhs(gb);                             // beta = solve(designunique, groupbeta)
                lu_solve(designunique, p, rhs, 1);
                for (int k = 0; k < p; ++k) beta[k] = rhs[k];
            } else {
                // get_levenberg_start, "null" method
                double sum_exprs = 0, sum_weight = 0;
                for (int s = 0; s < n; ++s) {
                    const double curN = std::exp(off[s]);
                    const double curweight = curN / (1 + dd[s] * curN);
                    sum_exprs += yy[s] * curweight / curN;
                    sum_weight += curweight;
                }
                const double c = std::log(sum_exprs / sum_weight);
                for (int k = 0; k < p; ++k) beta[k] = lev_start[k] * c;
                levenberg_fit(n, p, design, yy.data(), off.data(), dd.data(), mu.data(), beta.data(), wk);
            }
            if (coef) for (int k = 0; k < p; ++k) coef[g + (size_t)k * G] = beta[k];
            if (fitted) for (int s = 0; s < n; ++s) fitted[g + (size_t)s * G] = mu[s];
        }
    }
}

void glm_fit(const double* y, long G, int n, const double* design, int p, const Bcast& disp, const Bcast& offset,
             double prior_count, double* coef, double* fitted, int nthreads) {
    auto yf = [&](long g, int s) { return y[g + (size_t)s * G]; };
    auto of = [&](long g, int s) { return offset.at(g, s); };
    fit_all(G, n, design, p, disp, yf, of, coef, fitted, nthreads);
    if (prior_count > 0 && coef) {
        // R/predFC.R + src/add_prior.cpp: y' = y + prior*lib/mean(lib), offset' = log(lib + 2*prior*lib/mean(lib))
        std::vector<double> yp((size_t)G * n), op((size_t)G * n);
        for (long g = 0; g < G; ++g) {
            std::vector<double> lib(n);
            double ave = 0;
            for (int s = 0; s < n; ++s) { lib[s] = std::exp(offset.at(g, s)); ave += lib[s]; }
            ave /= n;
            for (int s = 0; s < n; ++s) {
                double pr = prior_count * lib[s] / ave;
                yp[g + (size_t)s * G] = y[g + (size_t)s * G] + pr;
                op[g + (size_t)s * G] = std::log(lib[s] + 2 * pr);
            }
        }
        auto yf2 = [&](long g, int s) { return yp[g + (size_t)s * G]; };
        auto of2 = [&](long g, int s) { return op[g + (size_t)s * G]; };
        fit_all(G, n, design, p, disp, yf2, of2, coef, (double*)nullptr, nthreads);
        for (size_t i = 0; i < (size_t)G * p; ++i) coef[i] = coef[i] / M_LN2 * M_LN2;   // predFC returns log2; glmFit rescales
    }
}

// LAPACK dsytf2 (uplo = 'L'): Bunch-Kaufman LDL'; a is p x p column-major, lower triangle referenced.
static void sytf2_lower(double* a, int n) {
    const double alpha = (1. + std::sqrt(17.)) / 8.;
    auto A = [&](int i, int j) -> double& { return a[(size_t)j * n + i]; };
    int k = 0;
    while (k < n) {
        int kstep = 1, kp = k, imax = k;
        double absakk = std::fabs(A(k, k)), colmax = 0;
        if (k < n - 1) { for (int i = k + 1; i < n; ++i) if (std::fabs(A(i, k)) > colmax) { colmax = std::fabs(A(i, k)); imax = i; } }
        if (std::max(absakk, colmax) == 0 || std::isnan(absakk)) { kp = k; }
        else {
            if (absakk >= alpha * colmax) kp = k;
            else {
                double rowmax = 0;
                for (int j = k; j < imax; ++j) rowmax = std::max(rowmax, std::fabs(A(imax, j)));
                for (int i = imax + 1; i < n; ++i) rowmax = std::max(rowmax, std::fabs(A(i, imax)));
                if (absakk >= alpha * colmax * (colmax / rowmax)) kp = k;
                else if (std::fabs(A(imax, imax)) >= alpha * rowmax) kp = imax;
                else { kp = imax; kstep = 2; }
            }
            int kk = k + kstep - 1;
            if (kp != kk) {
                for (int i = kp + 1; i < n; ++i) std::swap(A(i, kk), A(i, kp));
                for (int j = kk + 1; j < kp; ++j) std::swap(A(j, kk), A(kp, j));
                std::swap(A(kk, kk), A(kp, kp));
                if (kstep == 2) std::swap(A(k + 1, k), A(kp, k));
            }
            if (kstep == 1) {
                if (k < n - 1) {
                    double d11 = 1. / A(k, k);
                    for (int j = k + 1; j < n; ++j) {
                        double t = -d11 * A(j, k);
                        for (int i = j; i < n; ++i) A(i, j) += A(i, k) * t;
                    }
                    for (int i = k + 1; i < n; ++i) A(i, k) *= d11;
                }
            } else if (k < n - 2) {
                double d21 = A(k + 1, k), d11 = A(k + 1, k + 1) / d21, d22 = A(k, k) / d21;
                double t = 1. / (d11 * d22 - 1.);
                d21 = t / d21;
                for (int j = k + 2; j < n; ++j) {
                    double wk = d21 * (d11 * A(j, k) - A(j, k + 1));
                    double wkp1 = d21 * (d22 * A(j, k + 1) - A(j, k));
                    for (int i = j; i < n; ++i) A(i, j) -= A(i, k) * wk + A(i, k + 1) * wkp1;
                    A(j, k) = wk; A(j, k + 1) = wkp1;
                }
            }
        }
        k += kstep;
    }
}

// R/adjustedProfileLik.R + src/R_compute_apl.cpp + src/adj_coxreid.cpp (weights == 1, adjust = TRUE)
void adjusted_profile_lik(const double* y, long G, int n, const double* design, int p, const Bcast& disp,
                          const Bcast& offset, double* apl, int nthreads) {
    std::vector<double> mu((size_t)G * n);
    glm_fit(y, G, n, design, p, disp, offset, 0.0, nullptr, mu.data(), nthreads);
    const double log_low_value = std::log(low_value);
#ifdef _OPENMP
#pragma omp parallel for schedule(static) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (long g = 0; g < G; ++g) {
        std::vector<double> W(n, 0.0), wm((size_t)p * p, 0.0);
        double sum_loglik = 0;
        for (int s = 0; s < n; ++s) {
            const double cury = y[g + (size_t)s * G], curmu = mu[g + (size_t)s * G], curd = disp.at(g, s);
            if (curmu == 0) continue;
            double loglik;
            if (curd > 0) {
                const double r = 1 / curd, logmur = std::log(curmu + r);
                loglik = cury * std::log(curmu) - cury * logmur + r * std::log(r) - r * logmur
                       + std::lgamma(cury + r) - std::lgamma(cury + 1) - std::lgamma(r);
            } else {
                loglik = cury * std::log(curmu) - curmu - std::lgamma(cury + 1);
            }
            W[s] = curmu / (1 + curd * curmu);
            sum_loglik += loglik;
        }
        double adj;
        if (p == 1) {
            double sw = 0; for (int s = 0; s < n; ++s) sw += W[s];
            adj = std::log(std::fabs(sw)) / 2;
        } else {
            for (int row = 0; row < p; ++row)
                for (int col = 0; col <= row; ++col) {
                    double acc = 0;
                    for (int s = 0; s < n; ++s) acc += design[(size_t)row * n + s] * design[(size_t)col * n + s] * W[s];
                    wm[(size_t)col * p + row] = acc;
                }
            sytf2_lower(wm.data(), p);
            double sld = 0;
            for (int i = 0; i < p; ++i) {
                const double v = wm[(size_t)i * p + i];
                if (v < low_value || !std::isfinite(v)) sld += log_low_value; else sld += std::log(v);
            }
            adj = sld * 0.5;
        }
        apl[g] = sum_loglik - adj;
    }
}

// R src/library/stats/src/optimize.c Brent_fmin
double brent_fmin(double ax, double bx, double (*f)(double, void*), void* info, double tol) {
    const double c = (3. - std::sqrt(5.)) * .5;
    double a, b, d, e, p, q, r, u, v, w, x, t2, fu, fv, fw, fx, xm, eps, tol1, tol3;
    eps = std::sqrt(DBL_EPSILON);
    a = ax; b = bx; v = a + c * (b - a); w = v; x = v;
    d = 0.; e = 0.;
    fx = (*f)(x, info); fv = fx; fw = fx; tol3 = tol / 3.;
    for (;;) {
        xm = (a + b) * .5; tol1 = eps * std::fabs(x) + tol3; t2 = tol1 * 2.;
        if (std::fabs(x - xm) <= t2 - (b - a) * .5) break;
        p = 0.; q = 0.; r = 0.;
        if (std::fabs(e) > tol1) {
            r = (x - w) * (fx - fv); q = (x - v) * (fx - fw); p = (x - v) * q - (x - w) * r; q = (q - r) * 2.;
            if (q > 0.) p = -p; else q = -q;
            r = e; e = d;
        }
        if (std::fabs(p) >= std::fabs(q * .5 * r) || p <= q * (a - x) || p >= q * (b - x)) {
            if (x < xm) e = b - x; else e = a - x;
            d = c * e;
        } else {
            d = p / q; u = x + d;
            if (u - a < t2 || b - u < t2) { d = tol1; if (x >= xm) d = -d; }
        }
        if (std::fabs(d) >= tol1) u = x + d; else if (d > 0.) u = x + tol1; else u = x - tol1;
        fu = (*f)(u, info);
        if (fu <= fx) {
            if (u < x) b = x; else a = x;
            v = w; w = x; x = u; fv = fw; fw = fx; fx = fu;
        } else {
            if (u < x) a = u; else b = u;
            if (fu <= fw || w == x) { v = w; fv = fw; w = u; fw = fu; }
            else if (fu <= fv || v == x || v == w) { v = u; fv = fu; }
        }
    }
    return x;
}

// R/estimateGLMCommonDisp.R + R/dispCoxReid.R (interval = c(0,4), tol = 1e-5, min.row.sum = 5, no subsetting
// because ComBat_seq.R:103,105 passes subset = nrow(counts))
struct CommonCtx { const double* y; long G; int n; const double* design; int p; const double* off; int nthreads; int neval; };
static double neg_sum_apl(double par, void* vp) {
    CommonCtx* c = (CommonCtx*)vp;
    double disp = par * par * par * par;
    std::vector<double> apl(c->G);
    Bcast d{&disp, c->G, c->n, 0}, o{c->off, c->G, c->n, 2};
    adjusted_profile_lik(c->y, c->G, c->n, c->design, c->p, d, o, apl.data(), c->nthreads);
    long double s = 0;
    for (long g = 0; g < c->G; ++g) s += apl[g];
    c->neval++;
    return -(double)s;
}
double estimate_common_disp(const double* y, long G, int n, const double* design_in, int p_in, int nthreads, int* n_eval) {
    std::vector<double> ones(n, 1.0);
    const double* design = design_in ? design_in : ones.data();
    int p = design_in ? p_in : 1;
    if (p >= n) return kNaN;
    std::vector<double> off(n);
    for (int s = 0; s < n; ++s) { double cs = 0; for (long g = 0; g < G; ++g) cs += y[g + (size_t)s * G]; off[s] = std::log(cs); }
    std::vector<long> rows;
    for (long g = 0; g < G; ++g) { double rs = 0; for (int s = 0; s < n; ++s) rs += y[g + (size_t)s * G]; if (!(rs < 5)) rows.push_back(g); }
    if (rows.empty()) return kNaN;
    std::vector<double> ysub((size_t)rows.size() * n);
    for (size_t i = 0; i < rows.size(); ++i) for (int s = 0; s < n; ++s) ysub[i + (size_t)s * rows.size()] = y[rows[i] + (size_t)s * G];
    CommonCtx ctx{ysub.data(), (long)rows.size(), n, design, p, off.data(), nthreads, 0};
    double x = brent_fmin(0.0, std::pow(4.0, 0.25), neg_sum_apl, &ctx, 1e-5);
    if (n_eval) *n_eval = ctx.neval;
    return x * x * x * x;
}

// src/fmm_spline.c (R's fmm_spline), 0-based
void fmm_spline(int n, const double* x, const double* y, double* b, double* c, double* d) {
    if (n < 2) return;
    if (n < 3) { double t = (y[1] - y[0]); b[0] = t / (x[1] - x[0]); b[1] = b[0]; c[0] = c[1] = d[0] = d[1] = 0.0; return; }
    const int nm1 = n - 1;
    d[0] = x[1] - x[0];
    c[1] = (y[1] - y[0]) / d[0];
    for (int i = 1; i < nm1; i++) {
        d[i] = x[i + 1] - x[i];
        b[i] = 2.0 * (d[i - 1] + d[i]);
        c[i + 1] = (y[i + 1] - y[i]) / d[i];
        c[i] = c[i + 1] - c[i];
    }
    b[0] = -d[0]; b[nm1] = -d[n - 2]; c[0] = c[nm1] = 0.0;
    if (n > 3) {
        c[0] = c[2] / (x[3] - x[1]) - c[1] / (x[2] - x[0]);
        c[nm1] = c[n - 2] / (x[nm1] - x[n - 3]) - c[n - 3] / (x[n - 2] - x[n - 4]);
        c[0] = c[0] * d[0] * d[0] / (x[3] - x[0]);
        c[nm1] = -c[nm1] * d[n - 2] * d[n - 2] / (x[nm1] - x[n - 4]);
    }
    for (int i = 1; i <= nm1; i++) { double t = d[i - 1] / b[i - 1]; b[i] = b[i] - t * d[i - 1]; c[i] = c[i] - t * c[i - 1]; }
    c[nm1] = c[nm1] / b[nm1];
    for (int i = n - 2; i >= 0; i--) c[i] = (c[i] - d[i] * c[i + 1]) / b[i];
    b[nm1] = (y[nm1] - y[n - 2]) / d[n - 2] + d[n - 2] * (c[n - 2] + 2.0 * c[nm1]);
    for (int i = 0; i < nm1; i++) {
        b[i] = (y[i + 1] - y[i]) / d[i] - d[i] * (c[i + 1] + 2.0 * c[i]);
        d[i] = (c[i + 1] - c[i]) / d[i];
        c[i] = 3.0 * c[i];
    }
    c[nm1] = 3.0 * c[nm1];
    d[nm1] = d[n - 2];
}

// src/interpolator.cpp interpolator::find_max (+ quad_solver, always sol1)
double maximize_interpolant(int npts, const double* x, const double* y) {
    double maxed = -1; int maxed_at = -1;
    for (int i = 0; i < npts; ++i) if (maxed_at < 0 || y[i] > maxed) { maxed = y[i]; maxed_at = i; }
    double x_max = x[maxed_at];
    std::vector<double> b(npts), c(npts), d(npts);
    fmm_spline(npts, x, y, b.data(), c.data(), d.data());
    auto try_seg = [&](int k, double width) {
        const double A = 3 * d[k], B = 2 * c[k], C = b[k];
        double back = B * B - 4 * A * C;
        if (back < 0) return;
        double front = -B / (2 * A);
        back = std::sqrt(back) / (2 * A);
        double sol = front - back;
        if (sol > 0 && sol < width) {
            double temp = ((d[k] * sol + c[k]) * sol + b[k]) * sol + y[k];
            if (temp > maxed) { maxed = temp; x_max = sol + x[k]; }
        }
    };
    if (maxed_at > 0) try_seg(maxed_at - 1, x[maxed_at] - x[maxed_at - 1]);
    if (maxed_at < npts - 1) try_seg(maxed_at, x[maxed_at + 1] - x[maxed_at]);
    return x_max;
}

// R/estimateGLMTagwiseDisp.R + R/dispCoxReidInterpolateTagwise.R with prior.df = 0 (ComBat_seq.R:116):
// prior.n = 0 => apl.smooth == apl, so the abundance trend / moving average never contributes.
// R/aveLogCPM.R (aveLogCPM.default): prior.count = 2 scaled by lib.size / mean(lib.size) (src/add_prior.cpp), one-group NB fit
// with dispersion 0.05, (beta + log(1e6)) / log(2)                                                        [edgeR-recall, A.9]
static double ave_log_cpm(const double* yrow, int n, const double* lib, double mean_lib) {
    std::vector<double> yy(n), off(n), disp(n, 0.05);
    for (int s = 0; s < n; ++s) {
        const double prior = 2.0 * lib[s] / mean_lib;
        yy[s] = yrow[s] + prior;
        off[s] = std::log(lib[s] + 2 * prior);
    }
    return (glm_one_group(n, yy.data(), off.data(), disp.data(), 50, 1e-10) + std::log(1e6)) / std::log(2.0);
}

// R/movingAverageByCol.R with full.length = TRUE: zero-padded cumulative sums; every row becomes the mean of the rows that
// exist inside its window [r - ceil(width/2) + 1, r + floor(width/2)]                                        [edgeR-recall]
static void moving_average_rows(const std::vector<double>& x, long n, int width, std::vector<double>& out) {
    out = x;
    if (width <= 1) return;
    if (width > n) width = (int)n;
    const long half1 = (width + 1) / 2, half2 = width / 2;
    std::vector<long double> cs(n + 1, 0.0L);                        // R's cumsum accumulates in long double
    for (long i = 0; i < n; ++i) cs[i + 1] = cs[i] + (long double)x[i];
    for (long r = 0; r < n; ++r) {
        const long lo = std::max(0L, r - half1 + 1), hi = std::min(n - 1, r + half2);
        out[r] = (double)((double)(cs[hi + 1]) - (double)(cs[lo])) / (double)(hi - lo + 1);
    }
}

// R/estimateGLMTagwiseDisp.R + R/dispCoxReidInterpolateTagwise.R.  prior_df = 0 is what ComBat-seq passes
// (/root/reference/ComBat_seq.R:116): no moderation.  prior_df > 0: the weighted-likelihood empirical-Bayes moderation,
// apl.smooth = (apl + prior.n * trend) / (1 + prior.n) with prior.n = prior.df / (nlibs - ncoefs) and trend = moving average
// (width floor(span * ntags), span = (10 / ntags)^0.23) of the APL rows ordered by AveLogCPM.               [edgeR-recall, A.7]
void estimate_tagwise_disp(const double* y, long G, int n, const double* design, int p, double common,
                           double* tagwise, int nthreads, double prior_df) {
    for (long g = 0; g < G; ++g) tagwise[g] = common;
    if (p >= n) { for (long g = 0; g < G; ++g) tagwise[g] = kNaN; return; }
    std::vector<double> off(n), lib(n);
    for (int s = 0; s < n; ++s) { double cs = 0; for (long g = 0; g < G; ++g) cs += y[g + (size_t)s * G]; lib[s] = cs; off[s] = std::log(cs); }
    std::vector<long> rows;
    for (long g = 0; g < G; ++g) { double rs = 0; for (int s = 0; s < n; ++s) rs += y[g + (size_t)s * G]; if (rs >= 5) rows.push_back(g); }
    if (rows.empty()) return;
    const long Gs = (long)rows.size();
    std::vector<double> ysub((size_t)Gs * n);
    for (long i = 0; i < Gs; ++i) for (int s = 0; s < n; ++s) ysub[i + (size_t)s * Gs] = y[rows[i] + (size_t)s * G];
    const int npts = 11;
    double pts[npts];
    for (int k = 0; k < npts; ++k) pts[k] = -6.0 + k * (12.0 / (npts - 1));     // seq(-6, 6, length = 11)
    std::vector<double> apl((size_t)Gs * npts), sd(Gs);
    for (int k = 0; k < npts; ++k) {
        for (long i = 0; i < Gs; ++i) sd[i] = common * std::pow(2.0, pts[k]);
        Bcast d{sd.data(), Gs, n, 1}, o{off.data(), Gs, n, 2};
        adjusted_profile_lik(ysub.data(), Gs, n, design, p, d, o, apl.data() + (size_t)k * Gs, nthreads);
    }
    if (prior_df > 0) {
        const double prior_n = prior_df / (double)(n - p);
        const double span = G > 10 ? std::pow(10.0 / (double)G, 0.23) : 1.0;       // ntags of estimateGLMTagwiseDisp: all rows of y
        const int width = (int)std::floor(span * (double)Gs);                      // ntags of the filtered recursive call
        double mean_lib = 0;
        for (int s = 0; s < n; ++s) mean_lib += lib[s];
        mean_lib /= n;
        std::vector<double> alc(Gs), row(n);
        for (long i = 0; i < Gs; ++i) {
            for (int s = 0; s < n; ++s) row[s] = ysub[i + (size_t)s * Gs];
            alc[i] = ave_log_cpm(row.data(), n, lib.data(), mean_lib);
        }
        std::vector<long> o(Gs);
        for (long i = 0; i < Gs; ++i) o[i] = i;
        std::stable_sort(o.begin(), o.end(), [&](long a, long b) { return alc[a] < alc[b]; });     // order(): ties by index
        std::vector<double> col(Gs), sm;
        for (int k = 0; k < npts; ++k) {
            for (long r = 0; r < Gs; ++r) col[r] = apl[o[r] + (size_t)k * Gs];
            moving_average_rows(col, Gs, width, sm);
            for (long r = 0; r < Gs; ++r) {
                double& a = apl[o[r] + (size_t)k * Gs];
                a = (a + prior_n * sm[r]) / (1 + prior_n);
            }
        }
    }
    for (long i = 0; i < Gs; ++i) {
        double yv[npts];
        for (int k = 0; k < npts; ++k) yv[k] = apl[i + (size_t)k * Gs];
        double dmax = maximize_interpolant(npts, pts, yv);
        tagwise[rows[i]] = common * std::pow(2.0, dmax);
    }
}

}  // namespace ref
