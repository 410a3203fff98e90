// ORACLE — TEST INFRASTRUCTURE ONLY.  Never linked into, imported by, or executed from the product
// (combat-seq_b200/).  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may load liboracle.so.
//
// CPU restatement of /root/reference/ComBat_seq.R:22-204 and /root/reference/helper_seq.R:2-64, stage by
// stage, on top of ref_edger.cpp (edgeR 3.22 restated) and ref_nmath.cpp (R 3.5 nmath restated).
// PARITY UNPINNED against real R/edgeR: neither is installed in this image.
#include "../include/combatseq.h"
#include "ref_nmath.hpp"
#include "ref_edger.hpp"
#include <vector>
#include <cmath>
#include <cstring>
#include <cstdio>
#include <limits>
#include <algorithm>
#include <random>
#include <chrono>
#ifdef _OPENMP
#include <omp.h>
#endif

using namespace ref;
static const double kNaN = std::numeric_limits<double>::quiet_NaN();

// The engine's gene.subset.n sampler (include/combatseq.h, cbs_input.seed): a keyed 4-round Feistel permutation of
// [0, G-1) with cycle walking; source k of target i is perm(k), shifted past i.  Same distribution as R's
// sample(1:(G-1), n) (a uniformly random n-subset without replacement, redrawn per target gene, helper_seq.R:17-21);
// R's Mersenne-Twister stream itself is not reproduced (PARITY UNPINNED for this option).
static inline uint32_t feistel_mix(uint32_t x, uint32_t k) {
    x ^= k; x *= 0x9E3779B1u; x ^= x >> 15; x *= 0x85EBCA77u; x ^= x >> 13; x *= 0xC2B2AE3Du; x ^= x >> 16;
    return x;
}
static inline uint32_t feistel_perm(uint32_t idx, uint32_t m, int hb, uint64_t key) {
    const uint32_t mask = (1u << hb) - 1u;
    uint32_t v = idx;
    do {
        uint32_t l = v >> hb, r = v & mask;
        for (int rd = 0; rd < 4; ++rd) {
            const uint32_t f = feistel_mix(r, (uint32_t)(key >> (16 * rd)) + 0x632BE5ABu * (rd + 1) + (uint32_t)(key >> 32)) & mask;
            const uint32_t nl = r; r = l ^ f; l = nl;
        }
        v = (l << hb) | r;
    } while (v >= m);
    return v;
}

// helper_seq.R:8-42 monte_carlo_int_NB for one batch.  dat G x n, mu G x n, gamma/phi G.
// `targets` (optional) restricts the outer loop to a subset of target genes (for bounded CPU timing).
static void mc_int_ref(long G, int n, const double* dat, const double* mu, const double* gamma, const double* phi,
                       int mode, int gene_subset_n, uint64_t seed, double* gamma_star, double* phi_star,
                       int64_t* n_fallback, const long* targets, long n_targets, int nthreads) {
    long nt = targets ? n_targets : G;
    int64_t nfb = 0;
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 4) reduction(+ : nfb) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (long ti = 0; ti < nt; ++ti) {
        const long i = targets ? targets[ti] : ti;
        // subset of source genes (helper_seq.R:17-21), drawn with the engine's sampler (see feistel_perm above)
        std::vector<long> src;
        if (gene_subset_n > 0 && gene_subset_n < G - 1) {
            int hb = 1;
            while ((1ull << (2 * hb)) < (unsigned long long)(G - 1)) ++hb;
            const uint64_t key = seed * 0x9E3779B97F4A7C15ULL + (uint64_t)i * 0xD1B54A32D192ED03ULL + 0x2545F4914F6CDD1DULL;
            src.resize(gene_subset_n);
            for (int k = 0; k < gene_subset_n; ++k) {
                long j = feistel_perm((uint32_t)k, (uint32_t)(G - 1), hb, key);
                if (j >= i) ++j;
                src[k] = j;
            }
        }
        const long ns = src.empty() ? G : (long)src.size();
        long double sumLH = 0, sumG = 0, sumP = 0;      // R's sum() accumulates in long double
        double smax = -INFINITY;                        // CBS_MC_STABLE: two-pass log-sum-exp
        std::vector<double> logw;
        if (mode == CBS_MC_STABLE) logw.assign(ns, -INFINITY);
        for (long jj = 0; jj < ns; ++jj) {
            const long j = src.empty() ? jj : src[jj];
            if (j == i) continue;                       // mu[-i,], gamma[-i], phi[-i]  helper_seq.R:11-14
            const double size = 1.0 / phi[j];
            if (mode == CBS_MC_REFERENCE) {
                long double LH = 1.0L;                  // prod() accumulates in long double, result is double
                for (int s = 0; s < n; ++s) {
                    const double x = dat[i + (size_t)s * G];
                    if (std::isnan(x)) continue;        // !is.na(dat[i,])  helper_seq.R:10-11
                    LH *= dnbinom_mu(x, size, mu[j + (size_t)s * G]);
                }
                double lh = (double)LH;
                if (std::isnan(lh)) lh = 0;             // LH[is.nan(LH)] = 0  helper_seq.R:29
                // gamma_sub*LH and phi_sub*LH are double vector products in R (they round in the denormal range
                // when LH has underflowed gradually); only sum() accumulates in long double  (helper_seq.R:33)
                const double gl = gamma[j] * lh, pl = phi[j] * lh;
                sumLH += lh; sumG += gl; sumP += pl;
            } else {
                double L = 0;
                for (int s = 0; s < n; ++s) {
                    const double x = dat[i + (size_t)s * G];
                    if (std::isnan(x)) continue;
                    L += std::log(dnbinom_mu(x, size, mu[j + (size_t)s * G]));
                }
                if (std::isnan(L)) L = -INFINITY;
                logw[jj] = L; if (L > smax) smax = L;
            }
        }
        if (mode == CBS_MC_STABLE) {
            if (smax > -INFINITY)
                for (long jj = 0; jj < ns; ++jj) {
                    const long j = src.empty() ? jj : src[jj];
                    if (j == i || logw[jj] == -INFINITY) continue;
                    double w = std::exp(logw[jj] - smax);
                    const double gl = gamma[j] * w, pl = phi[j] * w;
                    sumLH += w; sumG += gl; sumP += pl;
                }
        }
        double s = (double)sumLH;
        if (s == 0 || std::isnan(s)) {                  // helper_seq.R:30-31
            gamma_star[i] = gamma[i]; phi_star[i] = phi[i]; nfb++;
        } else {
            gamma_star[i] = (double)sumG / s;           // sum(gamma_sub*LH)/sum(LH)  helper_seq.R:33
            phi_star[i] = (double)sumP / s;
        }
    }
    if (n_fallback) *n_fallback = nfb;
}

// helper_seq.R:46-64 match_quantiles for one batch; old_phi/new_phi per gene.
static void match_quantiles_ref(long G, int n, const double* counts, const double* old_mu, const double* old_phi,
                                const double* new_mu, const double* new_phi, double* out, int nthreads) {
#ifdef _OPENMP
#pragma omp parallel for schedule(dynamic, 16) num_threads(nthreads > 0 ? nthreads : 1)
#endif
    for (long a = 0; a < G; ++a)
        for (int b = 0; b < n; ++b) {
            const size_t k = a + (size_t)b * G;
            const double x = counts[k];
            if (x <= 1) { out[k] = x; continue; }
            const double tmp_p = pnbinom_mu(x - 1, 1.0 / old_phi[a], old_mu[k]);
            if (std::fabs(tmp_p - 1) < 1e-4) out[k] = x;
            else out[k] = 1 + qnbinom_mu(tmp_p, 1.0 / new_phi[a], new_mu[k]);
        }
}

static thread_local char g_err[512];
static int fail(int code, const char* msg) { snprintf(g_err, sizeof g_err, "%s", msg); return code; }

extern "C" {

const char* ref_last_error(void) { return g_err; }

// Mirrors cbs_combat_seq() (same structs) so tests can diff every intermediate.  nthreads = OpenMP threads.
// mc_target_stride > 1 evaluates the Monte-Carlo stage only for every stride-th target gene (others keep the
// unshrunk estimate): used ONLY by bench.py's bounded CPU baseline, never by parity tests.
int ref_combat_seq(const cbs_input* in, cbs_output* out, int nthreads, int mc_target_stride) {
    const long G0 = in->G0; const int N = in->N, B = in->B, p_mod = in->p_mod;
    double stage[CBS_N_STAGES] = {0};
    auto clk = [] { return std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t_begin = clk();
    double t_last = t_begin;
    auto lap = [&](int which) { const double t = clk(); stage[which] += t - t_last; t_last = t; };
    if (G0 <= 0 || N <= 0 || B <= 0 || p_mod < 1) return fail(CBS_EINVAL, "bad dimensions");
    if (B < 2) return fail(CBS_EINVAL, "ComBat-seq needs at least 2 batches");
    // ---- Preparation (ComBat_seq.R:25-51) ----
    std::vector<double> countsOri((size_t)G0 * N);
    for (size_t k = 0; k < (size_t)G0 * N; ++k) countsOri[k] = in->counts_f64 ? in->counts_f64[k] : (double)in->counts_i32[k];
    std::vector<std::vector<int>> batches_ind(B);
    for (int s = 0; s < N; ++s) {
        if (in->batch_idx[s] < 0 || in->batch_idx[s] >= B) return fail(CBS_EINVAL, "batch_idx out of range");
        batches_ind[in->batch_idx[s]].push_back(s);
    }
    for (int b = 0; b < B; ++b)
        if (batches_ind[b].size() <= 1) return fail(CBS_ESINGLE_SAMPLE_BATCH, "ComBat-seq doesn't support 1 sample per batch yet");
    // keep: genes that are not all-zero within any batch (ComBat_seq.R:33-37)
    std::vector<long> keep;
    for (long g = 0; g < G0; ++g) {
        bool ok = true;
        for (int b = 0; b < B && ok; ++b) {
            bool allzero = true;
            for (int s : batches_ind[b]) if (countsOri[g + (size_t)s * G0] != 0) { allzero = false; break; }
            if (allzero) ok = false;
        }
        if (ok) keep.push_back(g);
    }
    const long G = (long)keep.size();
    if (out->keep) { std::memset(out->keep, 0, G0); for (long g : keep) out->keep[g] = 1; }
    if (G == 0) return fail(CBS_EINVAL, "no genes left after the zero filter");
    std::vector<double> counts((size_t)G * N);
    for (long i = 0; i < G; ++i) for (int s = 0; s < N; ++s) counts[i + (size_t)s * G] = countsOri[keep[i] + (size_t)s * G0];
    // DGEList: lib.size = colSums(counts) (ComBat_seq.R:43)
    std::vector<double> lib(N), offset(N);
    for (int s = 0; s < N; ++s) { double cs = 0; for (long i = 0; i < G; ++i) cs += counts[i + (size_t)s * G]; lib[s] = cs; offset[s] = std::log(cs); }
    if (out->lib_size) std::copy(lib.begin(), lib.end(), out->lib_size);
    // ---- design = cbind(batchmod, mod)[, !all-ones]  (ComBat_seq.R:53-81) ----
    const int p = B + p_mod - 1;
    std::vector<double> design((size_t)N * p, 0.0);
    for (int s = 0; s < N; ++s) design[(size_t)in->batch_idx[s] * N + s] = 1.0;
    for (int k = 1; k < p_mod; ++k) for (int s = 0; s < N; ++s) design[(size_t)(B + k - 1) * N + s] = in->mod[(size_t)k * N + s];
    // confounding checks (ComBat_seq.R:84-90)
    if (qr_rank(design.data(), N, p) < p) {
        if (p == B + 1) return fail(CBS_ECONF_COVAR_BATCH, "The covariate is confounded with batch! Remove the covariate and rerun ComBat-Seq");
        if (p > B + 1) {
            if (qr_rank(design.data() + (size_t)B * N, N, p - B) < p - B)
                return fail(CBS_ECONF_COVARS, "The covariates are confounded! Please remove one or more of the covariates so the design is not confounded");
            return fail(CBS_ECONF_COVAR_WITH_BATCH, "At least one covariate is confounded with batch! Please remove confounded covariates and rerun ComBat-Seq");
        }
    }
    lap(CBS_ST_PREP);
    // ---- dispersions within each batch (ComBat_seq.R:97-125) ----
    std::vector<double> disp_common(B), phi_hat((size_t)G * B);
    for (int b = 0; b < B; ++b) {
        const int nb = (int)batches_ind[b].size();
        std::vector<double> yb((size_t)G * nb), modb((size_t)nb * p_mod);
        for (int t = 0; t < nb; ++t) {
            const int s = batches_ind[b][t];
            for (long i = 0; i < G; ++i) yb[i + (size_t)t * G] = counts[i + (size_t)s * G];
            for (int k = 0; k < p_mod; ++k) modb[(size_t)k * nb + t] = in->mod[(size_t)k * N + s];
        }
        bool fallback;
        if (in->use_fallback) fallback = in->use_fallback[b] != 0;
        else fallback = (nb <= p - B + 1) || (qr_rank(modb.data(), nb, p_mod) < p_mod);     // ComBat_seq.R:101
        if (fallback) {
            disp_common[b] = estimate_common_disp(yb.data(), G, nb, nullptr, 1, nthreads);
            for (long i = 0; i < G; ++i) phi_hat[i + (size_t)b * G] = disp_common[b];        // ComBat_seq.R:113
        } else {
            disp_common[b] = estimate_common_disp(yb.data(), G, nb, modb.data(), p_mod, nthreads);
            estimate_tagwise_disp(yb.data(), G, nb, modb.data(), p_mod, disp_common[b], phi_hat.data() + (size_t)b * G, nthreads);
        }
    }
    if (out->disp_common) std::copy(disp_common.begin(), disp_common.end(), out->disp_common);
    lap(CBS_ST_TAGWISE_DISP);      // common + tagwise together
    // phi_matrix (ComBat_seq.R:122-125)
    std::vector<double> phi_matrix((size_t)G * N);
    for (int s = 0; s < N; ++s) for (long i = 0; i < G; ++i) phi_matrix[i + (size_t)s * G] = phi_hat[i + (size_t)in->batch_idx[s] * G];
    // ---- NB GLM (ComBat_seq.R:128-138) ----
    std::vector<double> coef1((size_t)G * p), coef2((size_t)G * p), mu_hat((size_t)G * N);
    Bcast dm{phi_matrix.data(), G, N, 3}, o1{offset.data(), G, N, 2};
    glm_fit(counts.data(), G, N, design.data(), p, dm, o1, 1e-4, coef1.data(), nullptr, nthreads);
    std::vector<double> new_offset((size_t)G * N);
    for (long i = 0; i < G; ++i) {
        double alpha = 0;
        for (int b = 0; b < B; ++b) alpha += coef1[i + (size_t)b * G] * ((double)batches_ind[b].size() / N);
        for (int s = 0; s < N; ++s) new_offset[i + (size_t)s * G] = offset[s] + alpha;
    }
    Bcast o2{new_offset.data(), G, N, 3};
    glm_fit(counts.data(), G, N, design.data(), p, dm, o2, 1e-4, coef2.data(), mu_hat.data(), nthreads);
    const double* gamma_hat = coef2.data();   // first B columns
    lap(CBS_ST_GLM);
    // ---- Monte-Carlo EB (ComBat_seq.R:141-170) ----
    std::vector<double> gamma_star(gamma_hat, gamma_hat + (size_t)G * B), phi_star_mat(phi_hat);
    if (in->shrink) {
        for (int b = 0; b < B; ++b) {
            const int nb = (int)batches_ind[b].size();
            std::vector<double> yb((size_t)G * nb), mub((size_t)G * nb);
            for (int t = 0; t < nb; ++t) {
                const int s = batches_ind[b][t];
                for (long i = 0; i < G; ++i) { yb[i + (size_t)t * G] = counts[i + (size_t)s * G]; mub[i + (size_t)t * G] = mu_hat[i + (size_t)s * G]; }
            }
            std::vector<long> targets;
            if (mc_target_stride > 1) for (long i = 0; i < G; i += mc_target_stride) targets.push_back(i);
            int64_t nfb = 0;
            mc_int_ref(G, nb, yb.data(), mub.data(), gamma_hat + (size_t)b * G, phi_hat.data() + (size_t)b * G,
                       in->mc_mode, in->gene_subset_n, in->seed + (uint64_t)b, gamma_star.data() + (size_t)b * G,
                       phi_star_mat.data() + (size_t)b * G, &nfb, targets.empty() ? nullptr : targets.data(),
                       (long)targets.size(), nthreads);
            if (out->n_mc_fallback) out->n_mc_fallback[b] = nfb;
        }
        if (!in->shrink_disp) phi_star_mat = phi_hat;     // ComBat_seq.R:162-165
    }
    lap(CBS_ST_MC);
    // ---- batch-free distribution + match_quantiles (ComBat_seq.R:173-193) ----
    std::vector<double> phi_star(G), adjust((size_t)G * N);
    for (long i = 0; i < G; ++i) { double s = 0; for (int b = 0; b < B; ++b) s += phi_star_mat[i + (size_t)b * G]; phi_star[i] = s / B; }
    for (int b = 0; b < B; ++b) {
        const int nb = (int)batches_ind[b].size();
        std::vector<double> yb((size_t)G * nb), omu((size_t)G * nb), nmu((size_t)G * nb), res((size_t)G * nb);
        for (int t = 0; t < nb; ++t) {
            const int s = batches_ind[b][t];
            for (long i = 0; i < G; ++i) {
                yb[i + (size_t)t * G] = counts[i + (size_t)s * G];
                omu[i + (size_t)t * G] = mu_hat[i + (size_t)s * G];
                nmu[i + (size_t)t * G] = std::exp(std::log(mu_hat[i + (size_t)s * G]) - gamma_star[i + (size_t)b * G]);
            }
        }
        match_quantiles_ref(G, nb, yb.data(), omu.data(), phi_hat.data() + (size_t)b * G, nmu.data(), phi_star.data(), res.data(), nthreads);
        for (int t = 0; t < nb; ++t) { const int s = batches_ind[b][t]; for (long i = 0; i < G; ++i) adjust[i + (size_t)s * G] = res[i + (size_t)t * G]; }
    }
    lap(CBS_ST_MATCH_QUANTILES);
    // ---- outputs (ComBat_seq.R:198-203) ----
    if (out->adjusted) {
        std::copy(countsOri.begin(), countsOri.end(), out->adjusted);
        for (long i = 0; i < G; ++i) for (int s = 0; s < N; ++s) out->adjusted[keep[i] + (size_t)s * G0] = adjust[i + (size_t)s * G];
    }
    auto scatter = [&](double* dst, const double* src, int ncol) {
        if (!dst) return;
        std::fill(dst, dst + (size_t)G0 * ncol, kNaN);
        for (long i = 0; i < G; ++i) for (int c = 0; c < ncol; ++c) dst[keep[i] + (size_t)c * G0] = src[i + (size_t)c * G];
    };
    scatter(out->phi_hat, phi_hat.data(), B);
    scatter(out->gamma_hat, gamma_hat, B);
    scatter(out->gamma_star, gamma_star.data(), B);
    scatter(out->phi_star, phi_star_mat.data(), B);
    scatter(out->mu_hat, mu_hat.data(), N);
    out->row_begin = 0; out->row_end = G0;
    lap(CBS_ST_D2H);
    stage[CBS_ST_TOTAL] = clk() - t_begin;
    if (out->stage_ms) std::copy(stage, stage + CBS_N_STAGES, out->stage_ms);
    return CBS_OK;
}

// ---- elementwise / per-subsystem exports for the tests --------------------------------------------
void ref_nb_pmf(int64_t n, const double* x, const double* size, const double* mu, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = dnbinom_mu(x[i], size[i], mu[i]); }
void ref_nb_cdf(int64_t n, const double* x, const double* size, const double* mu, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = pnbinom_mu(x[i], size[i], mu[i]); }
void ref_nb_cdf_bysum(int64_t n, const double* x, const double* size, const double* mu, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = pnbinom_mu_bysum(x[i], size[i], mu[i]); }
void ref_nb_quantile(int64_t n, const double* p, const double* size, const double* mu, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = qnbinom_mu(p[i], size[i], mu[i]); }
void ref_qnorm(int64_t n, const double* p, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = qnorm01(p[i]); }
void ref_stirlerr(int64_t n, const double* x, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = stirlerr(x[i]); }
void ref_bd0(int64_t n, const double* x, const double* np, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = bd0(x[i], np[i]); }
void ref_pbeta(int64_t n, const double* x, const double* a, const double* b, double* out) { for (int64_t i = 0; i < n; ++i) { double w, w1; bratio(a[i], b[i], x[i], 0.5 - x[i] + 0.5, &w, &w1); out[i] = w; } }
void ref_set_legacy_qnbinom_search(int on) { set_legacy_qnbinom_search(on); }
int  ref_qr_rank(const double* X, int n, int p) { return qr_rank(X, n, p); }
double ref_unit_nb_deviance(double y, double mu, double phi) { return unit_nb_deviance(y, mu, phi); }
double ref_maximize_interpolant(int npts, const double* x, const double* y) { return maximize_interpolant(npts, x, y); }
void ref_fmm_spline(int n, const double* x, const double* y, double* b, double* c, double* d) { fmm_spline(n, x, y, b, c, d); }

// glmFit.default with phi_matrix given as disp G x n_disp_cols + column selector; offset = log(lib[s]) + gene_offset[g]
void ref_glm_fit(int64_t G, int32_t n, const double* y, const double* design, int32_t p, const double* disp,
                 int32_t n_disp_cols, const int32_t* disp_cols, const double* lib, const double* gene_offset,
                 double prior_count, double* coef, double* fitted, int nthreads) {
    std::vector<double> dm((size_t)G * n), om((size_t)G * n);
    for (int s = 0; s < n; ++s) for (int64_t g = 0; g < G; ++g) {
        dm[g + (size_t)s * G] = disp[g + (size_t)(disp_cols ? disp_cols[s] : 0) * G];
        om[g + (size_t)s * G] = std::log(lib[s]) + (gene_offset ? gene_offset[g] : 0.0);
    }
    (void)n_disp_cols;
    Bcast d{dm.data(), (long)G, n, 3}, o{om.data(), (long)G, n, 3};
    glm_fit(y, G, n, design, p, d, o, prior_count, coef, fitted, nthreads);
}
void ref_apl(int64_t G, int32_t n, const double* y, const double* design, int32_t p, const double* lib,
             const double* disp, int32_t D, int32_t disp_per_gene, double* apl, int nthreads) {
    std::vector<double> off(n), ones(n, 1.0);
    for (int s = 0; s < n; ++s) off[s] = std::log(lib[s]);
    const double* des = design ? design : ones.data(); int pp = design ? p : 1;
    for (int d = 0; d < D; ++d) {
        Bcast dd = disp_per_gene ? Bcast{disp + (size_t)d * G, (long)G, n, 1} : Bcast{disp + d, (long)G, n, 0};
        Bcast o{off.data(), (long)G, n, 2};
        adjusted_profile_lik(y, G, n, des, pp, dd, o, apl + (size_t)d * G, nthreads);
    }
}
double ref_common_disp(int64_t G, int32_t n, const double* y, const double* design, int32_t p, int nthreads, int* n_eval) {
    return estimate_common_disp(y, G, n, design, p, nthreads, n_eval);
}
void ref_tagwise_disp_prior(int64_t G, int32_t n, const double* y, const double* design, int32_t p, double common, double prior_df,
                            double* tagwise, int nthreads) {
    std::vector<double> ones(n, 1.0);
    estimate_tagwise_disp(y, G, n, design ? design : ones.data(), design ? p : 1, common, tagwise, nthreads, prior_df);
}
void ref_tagwise_disp(int64_t G, int32_t n, const double* y, const double* design, int32_t p, double common, double* tagwise, int nthreads) {
    std::vector<double> ones(n, 1.0);
    estimate_tagwise_disp(y, G, n, design ? design : ones.data(), design ? p : 1, common, tagwise, nthreads);
}
void ref_mc_int(int64_t G, int32_t n, const double* dat, const double* mu, const double* gamma, const double* phi,
                int32_t mode, int32_t gene_subset_n, uint64_t seed, double* gamma_star, double* phi_star,
                int64_t* n_fallback, const int64_t* targets, int64_t n_targets, int nthreads) {
    std::vector<long> t; if (targets) t.assign(targets, targets + n_targets);
    if (targets) { for (int64_t g = 0; g < G; ++g) { gamma_star[g] = gamma[g]; phi_star[g] = phi[g]; } }
    mc_int_ref(G, n, dat, mu, gamma, phi, mode, gene_subset_n, seed, gamma_star, phi_star, n_fallback,
               targets ? t.data() : nullptr, (long)t.size(), nthreads);
}
void ref_match_quantiles(int64_t G, int32_t n, const double* counts, const double* old_mu, const double* old_phi,
                         const double* new_mu, const double* new_phi, double* out, int nthreads) {
    match_quantiles_ref(G, n, counts, old_mu, old_phi, new_mu, new_phi, out, nthreads);
}
int ref_max_threads(void) {
#ifdef _OPENMP
    return omp_get_max_threads();
#else
    return 1;
#endif
}

}  // extern "C"
