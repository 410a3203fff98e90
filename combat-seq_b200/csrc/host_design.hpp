// host_design.hpp — host-side preparation of the (tiny) design matrices for the device fitters.
//
// Covers the base-R / edgeR bookkeeping that surrounds the per-gene arithmetic on the reference path:
//   qr(design)$rank checks                      /root/reference/ComBat_seq.R:84-90,101,111  (LINPACK dqrdc2, tol 1e-7)
//   edgeR designAsFactor / one-way detection    glmFit.default, called at ComBat_seq.R:130,134
//   mglmLevenberg's "null" start (QR least squares of design * b = 1)
// Everything here is O(N p^2) on the host; no per-gene work.
#pragma once
#include <vector>
#include <cmath>
#include <algorithm>
#include <cstddef>

namespace cbs {

// rank of an n x p column-major matrix as R's qr()$rank (Householder QR with limited column pivoting)
inline int qr_rank(const double* Xin, int n, int p, double tol = 1e-7) {
    if (p == 0) return 0;
    std::vector<double> x(Xin, Xin + (size_t)n * p), qraux(p), work(p), work2(p);
    auto X = [&](int i, int j) -> double& { return x[(size_t)j * n + i]; };
    auto colnorm = [&](int i0, int j) { double s = 0; for (int i = i0; i < n; ++i) s += X(i, j) * X(i, j); return std::sqrt(s); };
    for (int j = 0; j < p; ++j) { qraux[j] = work[j] = work2[j] = colnorm(0, j); if (work2[j] == 0) work2[j] = 1; }
    const int lup = std::min(n, p);
    int k = p + 1;
    for (int l = 1; l <= lup; ++l) {
        const int c = l - 1;
        while (l < k && qraux[c] < work2[c] * tol) {      // move a negligible column to the end
            for (int i = 0; i < n; ++i) { const double t = X(i, c); for (int j = c + 1; j < p; ++j) X(i, j - 1) = X(i, j); X(i, p - 1) = t; }
            std::rotate(qraux.begin() + c, qraux.begin() + c + 1, qraux.end());
            std::rotate(work.begin() + c, work.begin() + c + 1, work.end());
            std::rotate(work2.begin() + c, work2.begin() + c + 1, work2.end());
            --k;
        }
        if (l == n) break;
        double nrm = colnorm(c, c);
        if (nrm == 0) continue;
        if (X(c, c) != 0) nrm = std::copysign(nrm, X(c, c));
        for (int i = c; i < n; ++i) X(i, c) /= nrm;
        X(c, c) += 1;
        for (int j = c + 1; j < p; ++j) {
            double dot = 0;
            for (int i = c; i < n; ++i) dot += X(i, c) * X(i, j);
            const double t = -dot / X(c, c);
            for (int i = c; i < n; ++i) X(i, j) += t * X(i, c);
            if (qraux[j] != 0) {
                const double ratio = std::fabs(X(c, j)) / qraux[j];
                const double tt = std::max(1.0 - ratio * ratio, 0.0);
                if (std::fabs(tt) < 1e-6) { qraux[j] = colnorm(c + 1, j); work[j] = qraux[j]; }
                else qraux[j] *= std::sqrt(tt);
            }
        }
        qraux[c] = X(c, c);
        X(c, c) = -nrm;
    }
    return std::min(k - 1, n);
}

struct HostDesign {
    int n = 0, p = 0, oneway = 0;
    int ncls = 0;                  // > 0: Levenberg design whose samples are grouped by distinct design row (cell_start has ncls+1 entries)
    std::vector<int> col;          // global column of each sample slot (cell-grouped when oneway)
    std::vector<int> cell_start;   // p+1
    std::vector<double> minv;      // p x p
    std::vector<int> rowptr, cidx;
    std::vector<double> cval;
    std::vector<double> start;     // p
};

// run_glm_fit's class_form argument: bit 0 = class form (ncls > 0), bit 1 = every class lies inside one dispersion column
// (phi_col: dispersion column of each global sample column; nullptr = one column)
inline int glm_class_flags(const HostDesign& d, const int* phi_col) {
    if (d.ncls <= 0) return 0;
    int flags = 3;
    if (phi_col)
        for (int c = 0; c < d.ncls && flags == 3; ++c)
            for (int t = d.cell_start[c] + 1; t < d.cell_start[c + 1]; ++t)
                if (phi_col[d.col[t]] != phi_col[d.col[d.cell_start[c]]]) { flags = 1; break; }
    return flags;
}

// least-squares solution of X b = rhs by Householder QR (X n x p column-major, full column rank)
inline void lstsq(const double* Xin, int n, int p, const double* rhs, double* sol) {
    std::vector<double> a(Xin, Xin + (size_t)n * p), b(rhs, rhs + n), v(n);
    auto A = [&](int i, int j) -> double& { return a[(size_t)j * n + i]; };
    for (int k = 0; k < p; ++k) {
        double nrm = 0;
        for (int i = k; i < n; ++i) nrm += A(i, k) * A(i, k);
        nrm = std::sqrt(nrm);
        if (nrm == 0) continue;
        const double alpha = (A(k, k) > 0) ? -nrm : nrm;
        double vn2 = 0;
        for (int i = k; i < n; ++i) { v[i] = A(i, k); if (i == k) v[i] -= alpha; vn2 += v[i] * v[i]; }
        if (vn2 == 0) continue;
        for (int j = k; j < p; ++j) {
            double dot = 0;
            for (int i = k; i < n; ++i) dot += v[i] * A(i, j);
            const double f = 2 * dot / vn2;
            for (int i = k; i < n; ++i) A(i, j) -= f * v[i];
        }
        double dot = 0;
        for (int i = k; i < n; ++i) dot += v[i] * b[i];
        const double f = 2 * dot / vn2;
        for (int i = k; i < n; ++i) b[i] -= f * v[i];
    }
    for (int k = p - 1; k >= 0; --k) {
        double s = b[k];
        for (int j = k + 1; j < p; ++j) s -= A(k, j) * sol[j];
        sol[k] = s / A(k, k);
    }
}

// inverse of a p x p column-major matrix by Gauss-Jordan with partial pivoting; false if singular
inline bool invert(std::vector<double> A, int p, std::vector<double>& inv) {
    inv.assign((size_t)p * p, 0.0);
    for (int i = 0; i < p; ++i) inv[(size_t)i * p + i] = 1.0;
    for (int k = 0; k < p; ++k) {
        int best = k;
        for (int i = k + 1; i < p; ++i) if (std::fabs(A[(size_t)k * p + i]) > std::fabs(A[(size_t)k * p + best])) best = i;
        if (A[(size_t)k * p + best] == 0) return false;
        if (best != k)
            for (int j = 0; j < p; ++j) { std::swap(A[(size_t)j * p + k], A[(size_t)j * p + best]); std::swap(inv[(size_t)j * p + k], inv[(size_t)j * p + best]); }
        const double piv = A[(size_t)k * p + k];
        for (int j = 0; j < p; ++j) { A[(size_t)j * p + k] /= piv; inv[(size_t)j * p + k] /= piv; }
        for (int i = 0; i < p; ++i) {
            if (i == k) continue;
            const double f = A[(size_t)k * p + i];
            if (f == 0) continue;
            for (int j = 0; j < p; ++j) { A[(size_t)j * p + i] -= f * A[(size_t)j * p + k]; inv[(size_t)j * p + i] -= f * inv[(size_t)j * p + k]; }
        }
    }
    return true;
}

// X: n x p column-major design of the sub-problem; cols[t] = global column of sample t.
inline HostDesign make_design(const double* X, int n, int p, const int* cols) {
    HostDesign d;
    d.n = n; d.p = p;
    // edgeR designAsFactor: hash each row with powers of (e + pi)/5; levels in ascending hash order
    const double z = (std::exp(1.0) + M_PI) / 5;
    std::vector<double> h(n);
    for (int s = 0; s < n; ++s) {
        double acc = 0, zp = 1;
        for (int k = 0; k < p; ++k) { acc += X[(size_t)k * n + s] * zp; zp *= z; }
        h[s] = acc / p;
    }
    std::vector<double> lev(h);
    std::sort(lev.begin(), lev.end());
    lev.erase(std::unique(lev.begin(), lev.end()), lev.end());
    const int nlev = (int)lev.size();
    std::vector<int> order(n);
    for (int s = 0; s < n; ++s) order[s] = s;
    d.oneway = (nlev == p);
    d.cell_start.assign(p + 1, 0);
    d.minv.assign((size_t)p * p, 0.0);
    d.start.assign(p, 0.0);
    if (d.oneway) {
        std::vector<int> grp(n);
        for (int s = 0; s < n; ++s) grp[s] = (int)(std::lower_bound(lev.begin(), lev.end(), h[s]) - lev.begin());
        std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return grp[a] < grp[b]; });
        for (int s = 0; s < n; ++s) d.cell_start[grp[s] + 1]++;
        for (int l = 0; l < p; ++l) d.cell_start[l + 1] += d.cell_start[l];
        std::vector<double> du((size_t)p * p);
        for (int l = 0; l < p; ++l)
            for (int k = 0; k < p; ++k) du[(size_t)k * p + l] = X[(size_t)k * n + order[d.cell_start[l]]];
        if (!invert(du, p, d.minv)) d.oneway = 0;      // cannot happen for distinct rows of a full-rank design
    }
    if (!d.oneway) {
        for (int s = 0; s < n; ++s) order[s] = s;
        std::vector<double> ones(n, 1.0);
        lstsq(X, n, p, ones.data(), d.start.data());
        // few distinct design rows (categorical covariates): group the samples by row so that the device sweeps can share
        // exp(x'beta) and the X'WX update inside a class (glm.cuh lev_sweep)
        if (nlev <= 32 && nlev < n) {
            std::vector<int> grp(n);
            for (int s = 0; s < n; ++s) grp[s] = (int)(std::lower_bound(lev.begin(), lev.end(), h[s]) - lev.begin());
            std::stable_sort(order.begin(), order.end(), [&](int a, int b) { return grp[a] < grp[b]; });
            d.ncls = nlev;
            d.cell_start.assign(nlev + 1, 0);
            for (int s = 0; s < n; ++s) d.cell_start[grp[s] + 1]++;
            for (int l = 0; l < nlev; ++l) d.cell_start[l + 1] += d.cell_start[l];
        }
    }
    d.col.resize(n);
    d.rowptr.assign(n + 1, 0);
    for (int t = 0; t < n; ++t) {
        const int s = order[t];
        d.col[t] = cols ? cols[s] : s;
        for (int k = 0; k < p; ++k) {
            const double v = X[(size_t)k * n + s];
            if (v != 0) { d.cidx.push_back(k); d.cval.push_back(v); }
        }
        d.rowptr[t + 1] = (int)d.cidx.size();
    }
    return d;
}

// R optimize() (Brent's fmin) as a resumable state machine, so that several one-dimensional searches can share
// one device launch per step.  Usage: init(); while (!done) { feed(f(query())); }  -> result()
struct BrentFmin {
    double a, b, d, e, p, q, r, u, v, w, x, t2, fu, fv, fw, fx, xm, eps, tol1, tol3, c;
    bool first, done;
    void init(double ax, double bx, double tol) {
        c = (3. - std::sqrt(5.)) * .5;
        eps = std::sqrt(2.220446049250313e-16);
        a = ax; b = bx; v = a + c * (b - a); w = v; x = v;
        d = 0.; e = 0.; tol3 = tol / 3.;
        first = true; done = false; u = x;
    }
    double query() const { return first ? x : u; }
    double result() const { return x; }
    void feed(double f) {
        if (first) { fx = fv = fw = f; first = false; }
        else {
            fu = f;
            if (fu <= fx) {
                if (u < x) b = x; else a = x;
                v = w; w = x; x = u; fv = fw; fw = fx; fx = fu;
            } else {
                if (u < x) a = u; else b = u;
                if (fu <= fw || w == x) { v = w; fv = fw; w = u; fw = fu; }
                else if (fu <= fv || v == x || v == w) { v = u; fv = fu; }
            }
        }
        xm = (a + b) * .5; tol1 = eps * std::fabs(x) + tol3; t2 = tol1 * 2.;
        if (std::fabs(x - xm) <= t2 - (b - a) * .5) { done = true; return; }
        p = 0.; q = 0.; r = 0.;
        if (std::fabs(e) > tol1) {
            r = (x - w) * (fx - fv); q = (x - v) * (fx - fw); p = (x - v) * q - (x - w) * r; q = (q - r) * 2.;
            if (q > 0.) p = -p; else q = -q;
            r = e; e = d;
        }
        if (std::fabs(p) >= std::fabs(q * .5 * r) || p <= q * (a - x) || p >= q * (b - x)) {
            e = (x < xm) ? b - x : a - x;
            d = c * e;
        } else {
            d = p / q; u = x + d;
            if (u - a < t2 || b - u < t2) { d = tol1; if (x >= xm) d = -d; }
        }
        if (std::fabs(d) >= tol1) u = x + d; else if (d > 0.) u = x + tol1; else u = x - tol1;
    }
};

}  // namespace cbs
