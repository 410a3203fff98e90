// glm.cuh — per-gene NB log-link GLM fitters, Cox-Reid adjusted profile likelihood and the spline maximiser,
// as FP64 device functions (one gene per thread; every reduction over samples is sequential in sample order).
//
// Replaces, on the device, the edgeR routines the reference calls at
//   /root/reference/ComBat_seq.R:103,105  estimateGLMCommonDisp   -> adjustedProfileLik  (apl_gene)
//   /root/reference/ComBat_seq.R:115-116  estimateGLMTagwiseDisp  -> adjustedProfileLik + maximizeInterpolant
//   /root/reference/ComBat_seq.R:130,134  glmFit / glmFit.default -> mglmOneWay | mglmLevenberg (+ predFC refit)
// edgeR (pinned "edgeR 3.22.1", ComBat_seq.R:42) is a third-party dependency that is not under /root/reference;
// the algorithms follow its published form: glm_one_group Newton iteration (maxit 50, tol 1e-10), the one-way
// shortcut when #distinct design rows == ncol(design), glm_levenberg (null start, maxit 250, tol 1e-6, lambda
// schedule x10 on Cholesky failure / x2 on deviance increase / /10 after a first-try success), the unit NB
// deviance of nbdev.cpp, compute_apl + adj_coxreid (LDL' log-determinant), and interpolator::find_max on
// the Forsythe-Malcolm-Moler spline.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cmath>
#include "fastmath.cuh"

#ifndef CBS_GLM_UNROLL      // samples of a class sweep in flight per lane (development builds: tools/build_variant.sh)
#define CBS_GLM_UNROLL 1
#endif

namespace cbs {

constexpr int kGlmUnroll = CBS_GLM_UNROLL;
constexpr int PMAX = 32;                     // max design columns handled by the per-thread solvers
constexpr int PTRI = PMAX * (PMAX + 1) / 2;  // packed lower triangle

#define CBS_LOW_VALUE 1e-10
#define CBS_SUPREMELY_LOW 1e-13
#define CBS_RIDICULOUSLY_LOW 1e-100
#define CBS_MILDLY_LOW 1e-8

// A (sub)problem design: n samples, p columns.  Arrays live in global memory (read-only, tiny, L1-resident).
struct DesignDev {
    int n, p, oneway;
    int ncls;              // > 0 (Levenberg designs only): samples grouped by distinct design row, cell_start[ncls+1]
    const int* col;        // [n] column of each sample in the G x N matrices; grouped by cell when oneway
    const int* cell_start; // [p+1] sample ranges of the cells (oneway)
    const double* minv;    // [p*p] column-major inverse of the distinct design rows: coef = minv * cellbeta
    const int* rowptr;     // [n+1] CSR of the design rows (non-zeros, ascending column)
    const int* cidx;
    const double* cval;
    const double* start;   // [p] least-squares solution of design * b = 1 (Levenberg null start direction)
};

// How y, offset and dispersion of gene g at sample-slot column c are obtained.
struct RowSrc {
    const int32_t* cnt;    // G x N column-major counts of the kept genes
    long G;
    const double* logoff;  // [2N] interleaved pairs (log(lib.size[c]), lib.size[c])
    const double* yadd;    // [N] prior count added to y (predFC), or nullptr
    const double* phi;     // per-gene dispersion columns phi[g + phi_col[c]*G], or nullptr => phi_scalar
    const int* phi_col;    // [N] dispersion column of each sample (batch index), nullptr => column 0
    double phi_scalar;
    const int32_t* cnt_row = nullptr;   // optional: THIS gene's counts indexed by column (a shared-memory copy), instead of cnt
    bool phi_by_class = false;          // the samples of a design class share one dispersion column (true for every fit of the
                                        // pipeline: a class lies inside one batch): one dispersion load per class, not per sample
};

struct Row {
    const RowSrc& s; long g; double goff;
    __device__ Row(const RowSrc& s_, long g_, double goff_) : s(s_), g(g_), goff(goff_) {}
    __device__ __forceinline__ double y(int c) const {
        // counts are non-negative and below 2^31: 2^52 + x is exact in the low mantissa word, one DADD instead of I2F.F64
        // (the conversion sits in every sweep of the fitters; I2F.F64 was the top stall of the GLM kernel, ncu round 1)
        const double v = __hiloint2double(0x43300000, s.cnt_row ? s.cnt_row[c] : s.cnt[g + (long)c * s.G]) - 4503599627370496.0;
        return s.yadd ? v + s.yadd[c] : v;
    }
    // y(c) in two halves, so that the load can be issued one sample ahead of its use (lev_sweep)
    __device__ __forceinline__ int yraw(int c) const { return s.cnt_row ? s.cnt_row[c] : s.cnt[g + (long)c * s.G]; }
    __device__ __forceinline__ double ycvt(int raw, int c) const {
        const double v = __hiloint2double(0x43300000, raw) - 4503599627370496.0;
        return s.yadd ? v + s.yadd[c] : v;
    }
    __device__ __forceinline__ double off(int c) const { return s.logoff[2 * c] + goff; }
    __device__ __forceinline__ double lib(int c) const { return s.logoff[2 * c + 1]; }     // exp(logoff) without the exp
    __device__ __forceinline__ double phi(int c) const {
        return s.phi ? s.phi[g + (long)(s.phi_col ? s.phi_col[c] : 0) * s.G] : s.phi_scalar;
    }
};

__device__ __forceinline__ int tri(int i, int j) { return i * (i + 1) / 2 + j; }   // i >= j

// ---- lane groups --------------------------------------------------------------------------------------------------
// The fitters are templates on LANES, the number of consecutive lanes of a warp that share one gene.  LANES = 1 is the
// plain one-thread-per-gene form.  With LANES > 1 lane `ln` of the group takes the samples t = ln, ln + LANES, ... of
// every sweep and the partial sums are combined with an XOR butterfly over the group's lanes (mask `gm`): a + b and
// b + a are the same double, so every lane of the group ends up with bit-identical sums and the (data-dependent)
// control flow of the iteration stays uniform inside the group.  Everything that is not a sweep over the samples
// (Cholesky, lambda schedule, convergence tests) is done redundantly by all lanes of the group.
template <int LANES>
__device__ __forceinline__ double lane_sum(double v, unsigned gm) {
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) v += __shfl_xor_sync(gm, v, o);
    return v;
}
template <int LANES>
__device__ __forceinline__ double lane_max(double v, unsigned gm) {
#pragma unroll
    for (int o = LANES / 2; o > 0; o >>= 1) v = fmax(v, __shfl_xor_sync(gm, v, o));
    return v;
}

// edgeR glm_one_group on samples [t0,t1) of the design's sample list
template <int LANES = 1>
__device__ inline double fit_one_group(const DesignDev& d, const Row& r, int t0, int t1, int maxit, double tol,
                                       int ln = 0, unsigned gm = 0xffffffffu) {
    // exp(beta + offset_s) = exp(beta + gene offset) * lib.size_s: one exp per Newton iteration instead of one per sample
    // (rounding differs from exp(beta + offset_s) in the last place; the iteration converges to |step| < 1e-10 either way)
    const double eg = exp(r.goff);
    double nz = 0;
    double cur = 0;
    for (int t = t0 + ln; t < t1; t += LANES) {
        const int c = d.col[t];
        const double y = r.y(c);
        if (y > CBS_LOW_VALUE) { cur += y / (r.lib(c) * eg); nz = 1; }
    }
    if (LANES > 1) { cur = lane_sum<LANES>(cur, gm); nz = lane_max<LANES>(nz, gm); }
    const bool nonzero = nz > 0;
    cur = log(cur / (double)(t1 - t0));
    if (!nonzero) return -INFINITY;
    for (int i = 0; i < maxit; ++i) {
        double dl = 0, info = 0;
        const double ec = exp(cur) * eg;
        for (int t = t0 + ln; t < t1; t += LANES) {
            const int c = d.col[t];
            const double mu = ec * r.lib(c), inv = rcp_pos(1 + mu * r.phi(c));
            dl += (r.y(c) - mu) * inv;
            info += mu * inv;
        }
        if (LANES > 1) { dl = lane_sum<LANES>(dl, gm); info = lane_sum<LANES>(info, gm); }
        const double step = dl / info;
        cur += step;
        if (fabs(step) < tol) break;
    }
    return cur;
}

// edgeR compute_unit_nb_deviance ("2/3" is an integer division, i.e. 0, in edgeR 3.22)
__device__ __forceinline__ double unit_nb_deviance(double y, double mu, double phi) {
    y += CBS_MILDLY_LOW; mu += CBS_MILDLY_LOW;
    if (phi < 1e-4) {
        const double resid = y - mu;
        return 2 * (y * log(y / mu) - resid - 0.5 * resid * resid * phi * (1 + phi * (0.0 * resid - y)));
    }
    const double product = mu * phi;
    if (product > 1e6) return 2 * ((y - mu) / mu - log(y / mu)) * mu / (1 + product);
    // main branch: 1e-8 <= y < 2^31, 1e-8 <= mu <= 1e10, 1e-4 <= phi: every quotient is a positive normal number
    const double invphi = rcp_pos(phi), ypr = y + invphi;
    return 2 * (y * log_pos(y * rcp_pos(mu)) + ypr * log_pos((mu + invphi) * rcp_pos(ypr)));
}

// the same with 1 / phi handed in (invphi = rcp_pos(phi), only read when phi >= 1e-4): identical arithmetic, identical bits
__device__ __forceinline__ double unit_nb_deviance_inv(double y, double mu, double phi, double invphi) {
    y += CBS_MILDLY_LOW; mu += CBS_MILDLY_LOW;
    if (phi < 1e-4) {
        const double resid = y - mu;
        return 2 * (y * log(y / mu) - resid - 0.5 * resid * resid * phi * (1 + phi * (0.0 * resid - y)));
    }
    const double product = mu * phi;
    if (product > 1e6) return 2 * ((y - mu) / mu - log(y / mu)) * mu / (1 + product);
    const double ypr = y + invphi;
    return 2 * (y * log_pos(y * rcp_pos(mu)) + ypr * log_pos((mu + invphi) * rcp_pos(ypr)));
}

__device__ __forceinline__ double lin_pred(const DesignDev& d, const Row& r, int t, const double* beta) {
    double eta = r.off(d.col[t]);
    for (int k = d.rowptr[t]; k < d.rowptr[t + 1]; ++k) eta += d.cval[k] * beta[d.cidx[k]];
    return eta;
}

// One sweep over the samples at coefficients beta: deviance, X'WX (packed lower) and X'z.
// SHARED: xtwx and dl are ONE shared-memory copy owned by the lane group (class form only): lane 0 of the group is the only
// writer, and the sweep ends with a __syncwarp(gm) that publishes the result to the other lanes.
template <int LANES = 1, bool SHARED = false>
__device__ inline double lev_sweep(const DesignDev& d, const Row& r, const double* beta, double* xtwx, double* dl,
                                   int ln = 0, unsigned gm = 0xffffffffu) {
    const int p = d.p;
    const bool writer = !SHARED || ln == 0;
    if (writer) {
        for (int i = 0; i < p * (p + 1) / 2; ++i) xtwx[i] = 0;
        for (int i = 0; i < p; ++i) dl[i] = 0;
    }
    double dev = 0;
    if (d.ncls > 0) {
        // Samples grouped by distinct design row: mu_s = exp(gene offset + x_c'beta) * lib.size_s needs one exp per class,
        // and X'WX, X'z only need the class sums of W and z (the packed matrix is touched once per class, not per sample).
        for (int c = 0; c < d.ncls; ++c) {
            const int t0 = d.cell_start[c], t1 = d.cell_start[c + 1], k0 = d.rowptr[t0], k1 = d.rowptr[t0 + 1];
            double eta = r.goff;
            for (int k = k0; k < k1; ++k) eta += d.cval[k] * beta[d.cidx[k]];
            const double E = exp(eta);
            double sw = 0, sz = 0;
            if (r.s.phi_by_class) {
                // one dispersion (and one reciprocal of it) per class; the count and the library size of the lane's next sample
                // are requested before the current one is worked on (ncu: a quarter of the kernel's stall samples sat on the
                // first use of these loads, and the dispersion added a third dependent load: column -> batch -> phi)
                const double phi = r.phi(d.col[t0]);
                const double invphi = phi >= 1e-4 ? rcp_pos(phi) : 0.0;
                int t = t0 + ln, col = 0, yr = 0;
                double lib = 0;
                if (t < t1) { col = d.col[t]; lib = r.lib(col); yr = r.yraw(col); }
                for (; t < t1; t += LANES) {
                    const double mu = E * lib, y = r.ycvt(yr, col);
                    if (t + LANES < t1) { col = d.col[t + LANES]; lib = r.lib(col); yr = r.yraw(col); }
                    dev += unit_nb_deviance_inv(y, mu, phi, invphi);
                    const double inv = rcp_pos(1 + mu * phi);
                    sw += mu * inv; sz += (y - mu) * inv;
                }
            } else {
#pragma unroll kGlmUnroll
                for (int t = t0 + ln; t < t1; t += LANES) {
                    const int col = d.col[t];
                    const double mu = E * r.lib(col), y = r.y(col), phi = r.phi(col);
                    dev += unit_nb_deviance(y, mu, phi);
                    const double inv = rcp_pos(1 + mu * phi);
                    sw += mu * inv; sz += (y - mu) * inv;
                }
            }
            if (LANES > 1) { sw = lane_sum<LANES>(sw, gm); sz = lane_sum<LANES>(sz, gm); }
            if (writer)
                for (int k = k0; k < k1; ++k) {
                    const int ik = d.cidx[k];
                    const double vk = d.cval[k];
                    dl[ik] += vk * sz;
                    for (int k2 = k0; k2 <= k; ++k2) xtwx[tri(ik, d.cidx[k2])] += vk * d.cval[k2] * sw;
                }
        }
        if (LANES > 1) dev = lane_sum<LANES>(dev, gm);
        if (SHARED) __syncwarp(gm);
        return dev;
    }
    for (int t = ln; t < d.n; t += LANES) {
        const int c = d.col[t];
        const double mu = exp(lin_pred(d, r, t, beta)), y = r.y(c), phi = r.phi(c);
        dev += unit_nb_deviance(y, mu, phi);
        const double inv = rcp_pos(1 + mu * phi), W = mu * inv, z = (y - mu) * inv;
        for (int k = d.rowptr[t]; k < d.rowptr[t + 1]; ++k) {
            const int ik = d.cidx[k];
            const double vk = d.cval[k];
            dl[ik] += vk * z;
            for (int k2 = d.rowptr[t]; k2 <= k; ++k2) xtwx[tri(ik, d.cidx[k2])] += vk * d.cval[k2] * W;
        }
    }
    if (LANES > 1) {
        for (int i = 0; i < p * (p + 1) / 2; ++i) xtwx[i] = lane_sum<LANES>(xtwx[i], gm);
        for (int i = 0; i < p; ++i) dl[i] = lane_sum<LANES>(dl[i], gm);
        dev = lane_sum<LANES>(dev, gm);
    }
    return dev;
}

// Cholesky of the packed lower triangle (in place); returns false when a pivot is <= 0 or NaN
__device__ inline bool chol_packed(double* a, int p) {
    for (int j = 0; j < p; ++j) {
        double ajj = a[tri(j, j)];
        for (int k = 0; k < j; ++k) ajj -= a[tri(j, k)] * a[tri(j, k)];
        if (!(ajj > 0)) return false;
        ajj = sqrt(ajj);
        a[tri(j, j)] = ajj;
        for (int c = j + 1; c < p; ++c) {
            double v = a[tri(c, j)];
            for (int k = 0; k < j; ++k) v -= a[tri(j, k)] * a[tri(c, k)];
            a[tri(c, j)] = v / ajj;
        }
    }
    return true;
}
__device__ inline void chol_solve_packed(const double* a, int p, double* b) {
    for (int i = 0; i < p; ++i) { double s = b[i]; for (int k = 0; k < i; ++k) s -= a[tri(i, k)] * b[k]; b[i] = s / a[tri(i, i)]; }
    for (int i = p - 1; i >= 0; --i) { double s = b[i]; for (int k = i + 1; k < p; ++k) s -= a[tri(k, i)] * b[k]; b[i] = s / a[tri(i, i)]; }
}

// edgeR glm_levenberg::fit with the "null" start.  beta (out, p).  Returns the number of sweeps over the samples (1 + one per
// Levenberg trial step), or -1 when y is all zero (beta = NaN, mu = 0).
// Work arrays of one Levenberg fit: 3 packed p x p triangles and 4 p-vectors.  One thread per gene keeps them in local
// memory.  A lane group can share ONE copy in shared memory (SCR = true, `scr` = lev_scratch_doubles<PM>() doubles owned
// by the group): outside the sample sweeps every lane of the group follows the same control flow on bit-identical values; lane 0
// does the small dense algebra and is the only writer, the others read after a __syncwarp.  That needs the class form of the sweep
// (d.ncls > 0), where the lanes' partial sums are combined BEFORE X'WX is touched; the per-sample form accumulates a
// private X'WX per lane and keeps the local arrays.
template <int PM>
__host__ __device__ constexpr int lev_scratch_doubles() {
    // 3 PT + 4 PM, padded so that the 4 groups of a warp fall on disjoint shared-memory banks (stride = 4 mod 8)
    return (3 * (PM * (PM + 1) / 2) + 4 * PM + 3) / 8 * 8 + 4;
}
template <int LANES = 1, int PM = PMAX, bool SCR = false>
__device__ inline int fit_levenberg(const DesignDev& d, const Row& r, double* beta, int ln = 0, unsigned gm = 0xffffffffu,
                                    int maxit = 250, double tolerance = 1e-6, double* scr = nullptr, int max_sweeps = 0x7fffffff) {
    constexpr int PT = PM * (PM + 1) / 2;
    constexpr int LT = SCR ? 1 : PT, LV = SCR ? 1 : PM;
    const int p = d.p, pt = p * (p + 1) / 2;
    double l_xtwx[LT], l_xtwx_new[LT], l_work[LT], l_dl[LV], l_dl_new[LV], l_dbeta[LV], l_beta_new[LV];
    double* const xtwx = SCR ? scr : l_xtwx;
    double* const xtwx_new = SCR ? scr + PT : l_xtwx_new;
    double* const work = SCR ? scr + 2 * PT : l_work;
    double* const dl = SCR ? scr + 3 * PT : l_dl;
    double* const dl_new = SCR ? scr + 3 * PT + PM : l_dl_new;
    double* const dbeta = SCR ? scr + 3 * PT + 2 * PM : l_dbeta;
    double* const beta_new = SCR ? scr + 3 * PT + 3 * PM : l_beta_new;
    double ymax = 0, sum_exprs = 0, sum_weight = 0;
    const double eg0 = exp(r.goff);
    for (int t = ln; t < d.n; t += LANES) {
        const int c = d.col[t];
        const double y = r.y(c);
        if (y > ymax) ymax = y;
        const double curN = r.lib(c) * eg0, w = curN / (1 + r.phi(c) * curN);
        sum_exprs += y * w / curN;
        sum_weight += w;
    }
    if (LANES > 1) { ymax = lane_max<LANES>(ymax, gm); sum_exprs = lane_sum<LANES>(sum_exprs, gm); sum_weight = lane_sum<LANES>(sum_weight, gm); }
    if (ymax < CBS_LOW_VALUE) { for (int i = 0; i < p; ++i) beta[i] = nan(""); return -1; }
    const double c0 = log(sum_exprs / sum_weight);
    for (int i = 0; i < p; ++i) beta[i] = d.start[i] * c0;
    // Shared work arrays (SCR): lane 0 of the group is the only writer of the shared words; every hand-over between its writes and
    // the other lanes' reads goes through a __syncwarp(gm) (the shuffles below are such barriers too).  All lanes of the group
    // take the same branches: the values they branch on are bit-identical.
    const bool writer = !SCR || ln == 0;
    const int leader = __ffs(gm) - 1;
    double dev = lev_sweep<LANES, SCR>(d, r, beta, xtwx, dl, ln, gm);
    double max_info = -1, lambda = 0;
    int iter = 0, sweeps = 1;
    while ((++iter) <= maxit) {
        for (int i = 0; i < p; ++i) if (xtwx[tri(i, i)] > max_info) max_info = xtwx[tri(i, i)];
        if (iter == 1) { lambda = max_info * 1e-6; if (lambda < 1e-13) lambda = 1e-13; }
        int lev = 0;
        bool low_dev = false, failed = false;
        while (++lev) {
            do {
                int ok = 0;
                if (writer) {
                    for (int i = 0; i < pt; ++i) work[i] = xtwx[i];
                    for (int i = 0; i < p; ++i) { work[tri(i, i)] += lambda; dbeta[i] = dl[i]; }
                    ok = chol_packed(work, p) ? 1 : 0;
                }
                if (SCR) ok = __shfl_sync(gm, ok, leader);
                if (ok) break;
                lambda *= 10;
                if (lambda <= 0) lambda = CBS_RIDICULOUSLY_LOW;
            } while (1);
            if (writer) {
                chol_solve_packed(work, p, dbeta);
                for (int i = 0; i < p; ++i) beta_new[i] = beta[i] + dbeta[i];
            }
            if (SCR) __syncwarp(gm);
            if (sweeps >= max_sweeps) return -2;                   // handed to the wide second pass (tu_glm.cu): nothing is written
            const double dev_new = lev_sweep<LANES, SCR>(d, r, beta_new, xtwx_new, dl_new, ln, gm);
            ++sweeps;
            if (dev_new / ymax < CBS_SUPREMELY_LOW) low_dev = true;
            if (dev_new <= dev || low_dev) {
                for (int i = 0; i < p; ++i) beta[i] = beta_new[i];
                dev = dev_new;
                break;
            }
            lambda *= 2;
            if (lambda <= 0) lambda = CBS_RIDICULOUSLY_LOW;
            if (lambda / max_info > 1 / CBS_SUPREMELY_LOW) { failed = true; break; }
            if (SCR) __syncwarp(gm);                              // (nothing shared is read on this path; keeps the group together)
        }
        if (failed || low_dev) break;
        double divergence = 0;
        for (int i = 0; i < p; ++i) divergence += dl[i] * dbeta[i];
        if (divergence < tolerance) break;
        if (lev == 1) lambda /= 10;
        if (SCR) __syncwarp(gm);                                  // every lane has read beta_new, dl, dbeta, xtwx
        if (writer) {
            for (int i = 0; i < pt; ++i) xtwx[i] = xtwx_new[i];     // information at the accepted coefficients
            for (int i = 0; i < p; ++i) dl[i] = dl_new[i];
        }
        if (SCR) __syncwarp(gm);
    }
    return sweeps;
}

// glmFit.default's plain fit of one gene: coefficients beta[p]; cellbeta[p] additionally for one-way designs.
// Returns the Levenberg sweep count (0 for one-way designs, -1 for an all-zero row).
template <int LANES = 1, int PM = PMAX>
__device__ inline int fit_gene(const DesignDev& d, const Row& r, double* beta, double* cellbeta, int ln = 0,
                               unsigned gm = 0xffffffffu, double* scr = nullptr, int max_sweeps = 0x7fffffff) {
    if (d.oneway) {
        for (int l = 0; l < d.p; ++l) {
            const double b = fit_one_group<LANES>(d, r, d.cell_start[l], d.cell_start[l + 1], 50, 1e-10, ln, gm);
            cellbeta[l] = fmax(b, -1e8);                        // mglmOneWay: pmax(beta, -1e8)
        }
        for (int k = 0; k < d.p; ++k) {
            double s = 0;
            for (int l = 0; l < d.p; ++l) s += d.minv[k + l * d.p] * cellbeta[l];
            beta[k] = s;
        }
    } else {
        if (LANES > 1 && scr != nullptr && d.ncls > 0) return fit_levenberg<LANES, PM, true>(d, r, beta, ln, gm, 250, 1e-6, scr, max_sweeps);
        return fit_levenberg<LANES, PM, false>(d, r, beta, ln, gm, 250, 1e-6, nullptr, max_sweeps);
    }
    return 0;
}

// LAPACK dsytf2 (uplo = 'L', Bunch-Kaufman) on a packed lower triangle; the diagonal of D is left on the diagonal
__device__ inline void sytf2_packed(double* a, int n) {
    const double alpha = (1. + sqrt(17.)) / 8.;
    int k = 0;
    while (k < n) {
        int kstep = 1, kp = k, imax = k;
        const double absakk = fabs(a[tri(k, k)]);
        double colmax = 0;
        for (int i = k + 1; i < n; ++i) if (fabs(a[tri(i, k)]) > colmax) { colmax = fabs(a[tri(i, k)]); imax = i; }
        if (fmax(absakk, colmax) == 0 || isnan(absakk)) { kp = k; }
        else {
            if (absakk >= alpha * colmax) kp = k;
            else {
                double rowmax = 0;
                for (int j = k; j < imax; ++j) rowmax = fmax(rowmax, fabs(a[tri(imax, j)]));
                for (int i = imax + 1; i < n; ++i) rowmax = fmax(rowmax, fabs(a[tri(i, imax)]));
                if (absakk >= alpha * colmax * (colmax / rowmax)) kp = k;
                else if (fabs(a[tri(imax, imax)]) >= alpha * rowmax) kp = imax;
                else { kp = imax; kstep = 2; }
            }
            const int kk = k + kstep - 1;
            if (kp != kk) {
                for (int i = kp + 1; i < n; ++i) { const double t = a[tri(i, kk)]; a[tri(i, kk)] = a[tri(i, kp)]; a[tri(i, kp)] = t; }
                for (int j = kk + 1; j < kp; ++j) { const double t = a[tri(j, kk)]; a[tri(j, kk)] = a[tri(kp, j)]; a[tri(kp, j)] = t; }
                { const double t = a[tri(kk, kk)]; a[tri(kk, kk)] = a[tri(kp, kp)]; a[tri(kp, kp)] = t; }
                if (kstep == 2) { const double t = a[tri(k + 1, k)]; a[tri(k + 1, k)] = a[tri(kp, k)]; a[tri(kp, k)] = t; }
            }
            if (kstep == 1) {
                if (k < n - 1) {
                    const double d11 = 1. / a[tri(k, k)];
                    for (int j = k + 1; j < n; ++j) {
                        const double t = -d11 * a[tri(j, k)];
                        for (int i = j; i < n; ++i) a[tri(i, j)] += a[tri(i, k)] * t;
                    }
                    for (int i = k + 1; i < n; ++i) a[tri(i, k)] *= d11;
                }
            } else if (k < n - 2) {
                double d21 = a[tri(k + 1, k)];
                const double d11 = a[tri(k + 1, k + 1)] / d21, d22 = a[tri(k, k)] / d21;
                const double t = 1. / (d11 * d22 - 1.);
                d21 = t / d21;
                for (int j = k + 2; j < n; ++j) {
                    const double wk = d21 * (d11 * a[tri(j, k)] - a[tri(j, k + 1)]);
                    const double wkp1 = d21 * (d22 * a[tri(j, k + 1)] - a[tri(j, k)]);
                    for (int i = j; i < n; ++i) a[tri(i, j)] -= a[tri(i, k)] * wk + a[tri(i, k + 1)] * wkp1;
                    a[tri(j, k)] = wk; a[tri(j, k + 1)] = wkp1;
                }
            }
        }
        k += kstep;
    }
}

// edgeR adjustedProfileLik for one gene at the dispersion carried by `r`: fit, NB log-likelihood, Cox-Reid term.
// lgfact (optional): table of lgamma(y + 1), y = 0..LGFACT_N-1, filled on the device by the same lgamma() (k_lgfact_init),
// so a lookup returns the bits the call would.  With one dispersion per evaluation (the only form the pipeline uses)
// log(r) and lgamma(r) are loop invariants and are evaluated once per gene instead of once per sample.
constexpr int LGFACT_N = 2048;
__device__ inline double apl_gene(const DesignDev& d, const Row& r, const double* __restrict__ lgfact = nullptr,
                                  const double* __restrict__ lgr = nullptr) {      // lgr[y] = lgamma(y + 1/phi), y < LGFACT_N (scalar phi)
    double beta[PMAX], cellbeta[PMAX], wm[PTRI];
    const int p = d.p;
    fit_gene(d, r, beta, cellbeta);
    const bool fast = lgfact != nullptr;                    // the "apl_fast" knob: table + hoisting + product form
    const bool scalar_phi = fast && (r.s.phi == nullptr) && r.s.phi_scalar > 0;
    double rr_c = 0, logrr_c = 0, lgrr_c = 0;
    if (scalar_phi) { rr_c = 1 / r.s.phi_scalar; logrr_c = log(rr_c); lgrr_c = lgamma(rr_c); }
    for (int i = 0; i < p * (p + 1) / 2; ++i) wm[i] = 0;
    double sum_loglik = 0, sumw = 0;
    int l = -1;
    double ecell = 0;
    for (int t = 0; t < d.n; ++t) {
        const int c = d.col[t];
        double mu, logmu;
        if (d.oneway) {
            // mu = exp(cell coefficient + gene offset) * lib.size and log(mu) = the linear predictor itself: one exp per
            // cell instead of an exp and a log per sample (last-place differences against exp/log of the sum)
            if (fast) {
                if (l < 0 || t >= d.cell_start[l + 1]) { if (l < 0) l = 0; while (t >= d.cell_start[l + 1]) ++l; ecell = exp(cellbeta[l] + r.goff); }
                mu = ecell * r.lib(c);
                logmu = cellbeta[l] + r.off(c);
            } else {
                if (l < 0) l = 0;
                while (t >= d.cell_start[l + 1]) ++l;
                mu = exp(cellbeta[l] + r.off(c));
                logmu = log(mu);
            }
        } else {
            mu = isnan(beta[0]) ? 0.0 : exp(lin_pred(d, r, t, beta));
            logmu = log(mu);
        }
        if (mu == 0) continue;
        const double y = r.y(c), phi = r.phi(c);
        double loglik;
        const double lgy1 = (lgfact && y < (double)LGFACT_N && !r.s.yadd) ? lgfact[(int)y] : lgamma(y + 1);
        if (phi > 0) {
            const double rr = scalar_phi ? rr_c : 1 / phi, logmur = log_pos(mu + rr);
            const double logrr = scalar_phi ? logrr_c : log(rr), lgrr = scalar_phi ? lgrr_c : lgamma(rr);
            const double lgyr = (scalar_phi && lgr && y < (double)LGFACT_N && !r.s.yadd) ? lgr[(int)y] : lgamma(y + rr);
            loglik = y * logmu - y * logmur + rr * logrr - rr * logmur + lgyr - lgy1 - lgrr;
        } else {
            loglik = y * logmu - mu - lgy1;
        }
        const double W = mu * rcp_pos(1 + phi * mu);
        sum_loglik += loglik;
        sumw += W;
        if (p > 1)
            for (int k = d.rowptr[t]; k < d.rowptr[t + 1]; ++k)
                for (int k2 = d.rowptr[t]; k2 <= k; ++k2) wm[tri(d.cidx[k], d.cidx[k2])] += d.cval[k] * d.cval[k2] * W;
    }
    double adj;
    if (p == 1) adj = log(fabs(sumw)) / 2;
    else {
        sytf2_packed(wm, p);
        const double log_low = log(CBS_LOW_VALUE);
        double sld = 0;
        for (int i = 0; i < p; ++i) {
            const double v = wm[tri(i, i)];
            sld += (v < CBS_LOW_VALUE || !isfinite(v)) ? log_low : log(v);
        }
        adj = sld * 0.5;
    }
    return sum_loglik - adj;
}

// R's fmm_spline (n >= 4 here) and edgeR interpolator::find_max: abscissa of the spline maximum
template <int NP>
__device__ inline double maximize_interpolant(const double* x, const double* y) {
    double b[NP], c[NP], dd[NP];
    double maxed = -1; int at = -1;
    for (int i = 0; i < NP; ++i) if (at < 0 || y[i] > maxed) { maxed = y[i]; at = i; }
    double x_max = x[at];
    const int nm1 = NP - 1;
    dd[0] = x[1] - x[0];
    c[1] = (y[1] - y[0]) / dd[0];
    for (int i = 1; i < nm1; i++) {
        dd[i] = x[i + 1] - x[i];
        b[i] = 2.0 * (dd[i - 1] + dd[i]);
        c[i + 1] = (y[i + 1] - y[i]) / dd[i];
        c[i] = c[i + 1] - c[i];
    }
    b[0] = -dd[0]; b[nm1] = -dd[NP - 2];
    c[0] = c[2] / (x[3] - x[1]) - c[1] / (x[2] - x[0]);
    c[nm1] = c[NP - 2] / (x[nm1] - x[NP - 3]) - c[NP - 3] / (x[NP - 2] - x[NP - 4]);
    c[0] = c[0] * dd[0] * dd[0] / (x[3] - x[0]);
    c[nm1] = -c[nm1] * dd[NP - 2] * dd[NP - 2] / (x[nm1] - x[NP - 4]);
    for (int i = 1; i <= nm1; i++) { const double t = dd[i - 1] / b[i - 1]; b[i] = b[i] - t * dd[i - 1]; c[i] = c[i] - t * c[i - 1]; }
    c[nm1] = c[nm1] / b[nm1];
    for (int i = NP - 2; i >= 0; i--) c[i] = (c[i] - dd[i] * c[i + 1]) / b[i];
    b[nm1] = (y[nm1] - y[NP - 2]) / dd[NP - 2] + dd[NP - 2] * (c[NP - 2] + 2.0 * c[nm1]);
    for (int i = 0; i < nm1; i++) {
        b[i] = (y[i + 1] - y[i]) / dd[i] - dd[i] * (c[i + 1] + 2.0 * c[i]);
        dd[i] = (c[i + 1] - c[i]) / dd[i];
        c[i] = 3.0 * c[i];
    }
    c[nm1] = 3.0 * c[nm1];
    dd[nm1] = dd[NP - 2];
#pragma unroll
    for (int side = 0; side < 2; ++side) {
        const int k = (side == 0) ? at - 1 : at;
        if (k < 0 || k >= NP - 1) continue;
        const double width = x[k + 1] - x[k];
        const double A = 3 * dd[k], B = 2 * c[k], C = b[k];
        double back = B * B - 4 * A * C;
        if (back < 0) continue;
        const double front = -B / (2 * A);
        back = sqrt(back) / (2 * A);
        const double sol = front - back;
        if (sol > 0 && sol < width) {
            const double temp = ((dd[k] * sol + c[k]) * sol + b[k]) * sol + y[k];
            if (temp > maxed) { maxed = temp; x_max = sol + x[k]; }
        }
    }
    return x_max;
}

}  // namespace cbs
