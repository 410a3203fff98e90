// kernels.cuh — sm_100a kernels of the ComBat-seq engine other than the Monte-Carlo integration (mc.cuh).
// Data layout: every per-gene matrix is column-major with the gene index fastest (R layout), so one thread per
// gene (consecutive threads = consecutive genes) gives fully coalesced loads and stores.
#pragma once
#include "nbmath.cuh"
#include "glm.cuh"

namespace cbs {

// ---- a1: zero filter (ComBat_seq.R:33-37): keep[g] = every batch has a non-zero count in row g ----------
template <class T>
__global__ void k_keep_mask(const T* __restrict__ counts, long G0, int N, const int* __restrict__ batch_idx, int B,
                            unsigned char* __restrict__ keep) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G0) return;
    unsigned long long seen_lo = 0, seen_hi = 0;   // up to 128 batches as a bit set
    for (int s = 0; s < N; ++s) {
        const T v = counts[g + (long)s * G0];
        if (v != (T)0) { const int b = batch_idx[s]; if (b < 64) seen_lo |= 1ull << b; else seen_hi |= 1ull << (b - 64); }
    }
    const int cnt = __popcll(seen_lo) + __popcll(seen_hi);
    keep[g] = (cnt == B) ? 1 : 0;
}

// ---- a1/a2: compaction to int32 + exact integer column sums (DGEList lib.size, ComBat_seq.R:43) -----------
// grid: (ceil(G/256), N).  cnt is G x N column-major over the kept genes.
template <class T>
__global__ void k_compact_colsum(const T* __restrict__ counts, long G0, const int* __restrict__ keep_idx, long G,
                                 int32_t* __restrict__ cnt, unsigned long long* __restrict__ colsum) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    long long v = 0;
    if (i < G) {
        v = (long long)counts[keep_idx[i] + (long)s * G0];
        cnt[i + (long)s * G] = (int32_t)v;
    }
    // integer sums are exact and order-independent, so atomics keep the result deterministic
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __shared__ long long part[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) part[wid] = v;
    __syncthreads();
    if (wid == 0) {
        v = (lane < (int)(blockDim.x >> 5)) ? part[lane] : 0;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0 && v != 0) atomicAdd(&colsum[s], (unsigned long long)v);
    }
}

// per-batch row sums of the kept genes: sel[b*G+g] = rowSums(counts[g, batch b]) >= 5  (edgeR min.row.sum)
__global__ void k_rowsum_sel(const int32_t* __restrict__ cnt, long G, int N, const int* __restrict__ batch_idx, int B,
                             unsigned char* __restrict__ sel) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (g >= G) return;
    long long rs = 0;
    for (int s = 0; s < N; ++s) if (batch_idx[s] == b) rs += cnt[g + (long)s * G];
    sel[(long)b * G + g] = (rs >= 5) ? 1 : 0;
}

// ---- a5/a6: adjusted profile likelihood, one thread per (gene, evaluation) -------------------------------
constexpr int APL_MAX_EVALS = 96;
struct AplLaunch {
    int n_evals;
    int design[APL_MAX_EVALS];   // which DesignDev (batch) each evaluation uses
    double phi[APL_MAX_EVALS];   // scalar dispersion of the evaluation (ignored when phi_gene != nullptr)
};
// apl[e*G + g]; sel[design*G + g] == 0 genes are skipped (apl = 0).  phi_gene (optional): G x n_evals per-gene dispersions.
__global__ void __launch_bounds__(128)
k_apl(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ logoff,
      const DesignDev* __restrict__ designs, const unsigned char* __restrict__ sel, const double* __restrict__ phi_gene,
      AplLaunch L, double* __restrict__ apl, const double* __restrict__ lgfact, const double* __restrict__ lgr) {
    const long g = g0 + (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int e = blockIdx.y;
    if (g >= g1) return;
    const int di = L.design[e];
    if (sel && !sel[(long)di * G + g]) { apl[(long)e * G + g] = 0.0; return; }
    const DesignDev d = designs[di];
    RowSrc src{cnt, G, logoff, nullptr, nullptr, nullptr, phi_gene ? phi_gene[(long)e * G + g] : L.phi[e]};
    Row r(src, g, 0.0);
    apl[(long)e * G + g] = apl_gene(d, r, lgfact, (lgr && !phi_gene) ? lgr + (long)e * LGFACT_N : nullptr);
}

// lgamma(y + 1) for y = 0..LGFACT_N-1 by the device lgamma() itself (see apl_gene)
__global__ void k_lgfact_init(double* __restrict__ tab) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < LGFACT_N) tab[i] = lgamma((double)i + 1.0);
}

// lgamma(y + 1/phi_e) for y = 0..LGFACT_N-1 and every evaluation e of a launch: with one dispersion per evaluation the
// term depends on the count only, so 2048 calls per evaluation replace one call per gene and sample (same bits).
__global__ void k_lgr_init(AplLaunch L, double* __restrict__ tab) {      // grid (LGFACT_N / 256, n_evals)
    const int y = blockIdx.x * blockDim.x + threadIdx.x, e = blockIdx.y;
    const double rr = 1 / L.phi[e];
    if (y < LGFACT_N) tab[(long)e * LGFACT_N + y] = lgamma((double)y + rr);
}

// deterministic chunked sums: partial[e*nchunk + chunk] = sum of apl[e*G + g] over the selected genes of the chunk
constexpr int APL_CHUNK = 1024;
__global__ void __launch_bounds__(256)
k_chunk_sum(const double* __restrict__ apl, long G, const unsigned char* __restrict__ sel, AplLaunch L,
            long chunk0, double* __restrict__ partial, long nchunk_total) {
    const long chunk = chunk0 + blockIdx.x;
    const int e = blockIdx.y;
    const long base = chunk * APL_CHUNK;
    const unsigned char* s = sel ? sel + (long)L.design[e] * G : nullptr;
    double v = 0;
    for (int k = 0; k < APL_CHUNK / 256; ++k) {
        const long g = base + threadIdx.x + k * 256;
        if (g < G && (!s || s[g])) v += apl[(long)e * G + g];
    }
    __shared__ double sh[256];
    sh[threadIdx.x] = v;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) sh[threadIdx.x] += sh[threadIdx.x + o];
        __syncthreads();
    }
    if (threadIdx.x == 0) partial[(long)e * nchunk_total + chunk] = sh[0];
}

// a6: spline maximisation over the 11-point grid; apl is laid out [k*G + g] for this batch
__global__ void k_tagwise_max(const double* __restrict__ apl, long G, long g0, long g1, const unsigned char* __restrict__ sel,
                              double common, double* __restrict__ phi_out) {
    const long g = g0 + (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= g1) return;
    if (sel && !sel[g]) { phi_out[g] = common; return; }
    double x[11], y[11];
#pragma unroll
    for (int k = 0; k < 11; ++k) { x[k] = -6.0 + k * (12.0 / 10); y[k] = apl[(long)k * G + g]; }
    const double dmax = maximize_interpolant<11>(x, y);
    phi_out[g] = common * pow(2.0, dmax);
}

__global__ void k_fill(double* __restrict__ p, long n, double v) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ---- a8-a10: glmFit.default, one thread per gene --------------------------------------------------------
// mode 0: plain fit, gene offset = goff_in[g] (or 0)            -> coef (optional), fitted (optional)
// mode 1: predFC fit (y + yadd, gene offset = log(exp(goff_in[g]) + 2*prior/mean(lib))) -> coef
// alpha_w (optional, length n_alpha): additionally write alpha[g] = sum_b alpha_w[b] * coef[g,b]   (ComBat_seq.R:131)
__global__ void __launch_bounds__(128)
k_glm_fit(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ logoff,
          const double* __restrict__ yadd, double two_prior_over_meanlib, const double* __restrict__ phi,
          const int* __restrict__ phi_col, const DesignDev* __restrict__ design, const double* __restrict__ goff_in,
          int mode, double* __restrict__ coef, double* __restrict__ fitted, const double* __restrict__ alpha_w,
          int n_alpha, double* __restrict__ alpha_out) {
    const long g = g0 + (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= g1) return;
    const DesignDev d = *design;
    double goff = goff_in ? goff_in[g] : 0.0;
    if (mode == 1) goff = log(exp(goff) + two_prior_over_meanlib);
    RowSrc src{cnt, G, logoff, mode == 1 ? yadd : nullptr, phi, phi_col, 0.0};
    Row r(src, g, goff);
    double beta[PMAX], cellbeta[PMAX];
    fit_gene(d, r, beta, cellbeta);
    if (coef) for (int k = 0; k < d.p; ++k) coef[g + (long)k * G] = beta[k];
    if (alpha_out) {
        double a = 0;
        for (int b = 0; b < n_alpha; ++b) a += beta[b] * alpha_w[b];
        alpha_out[g] = a;
    }
    if (fitted) {
        if (d.oneway) {
            for (int l = 0; l < d.p; ++l)
                for (int t = d.cell_start[l]; t < d.cell_start[l + 1]; ++t) {
                    const int c = d.col[t];
                    fitted[g + (long)c * G] = exp(cellbeta[l] + r.off(c));
                }
        } else {
            const bool dead = isnan(beta[0]);
            for (int t = 0; t < d.n; ++t) fitted[g + (long)d.col[t] * G] = dead ? 0.0 : exp(lin_pred(d, r, t, beta));
        }
    }
}

// Same contract as k_glm_fit with GLM_LANES consecutive lanes per gene (glm.cuh "lane groups"): the sweeps over the
// samples are split over the lanes, so a 60k-gene x 1000-sample fit exposes 480k threads instead of 60k and the
// per-thread dependent chain is 125 samples per sweep instead of 1000.  A warp holds 32 / GLM_LANES consecutive genes:
// each load instruction touches GLM_LANES columns x (32 / GLM_LANES) consecutive rows of the column-major matrices.
// Used for designs with p <= GLM_LANES_PMAX columns.  use_scr: the Levenberg work arrays (3 packed X'WX + 4 vectors) are
// one shared-memory copy per gene instead of a local-memory copy per lane (ncu, round 1: the local copies made 10.9 GB of
// DRAM traffic per launch, 15x the algorithmic bytes); dynamic shared memory = (128 / GLM_LANES) * lev_scratch_doubles<PM>().
constexpr int GLM_LANES = 8;
constexpr int GLM_LANES_PMAX = 16;
template <int PM>                                            // PM >= design columns: sizes the per-thread packed X'WX (local memory)
__global__ void __launch_bounds__(128, 5)
k_glm_fit_lanes(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ logoff,
                const double* __restrict__ yadd, double two_prior_over_meanlib, const double* __restrict__ phi,
                const int* __restrict__ phi_col, const DesignDev* __restrict__ design, const double* __restrict__ goff_in,
                int mode, double* __restrict__ coef, double* __restrict__ fitted, const double* __restrict__ alpha_w,
                int n_alpha, double* __restrict__ alpha_out, int use_scr) {
    const long slot = ((long)blockIdx.x * blockDim.x + threadIdx.x) / GLM_LANES;
    const long g = g0 + slot;
    const int ln = threadIdx.x % GLM_LANES;
    if (g >= g1) return;                                    // whole groups leave together
    const unsigned gm = ((1u << GLM_LANES) - 1u) << ((threadIdx.x & 31) - ln);
    const DesignDev d = *design;
    double goff = goff_in ? goff_in[g] : 0.0;
    if (mode == 1) goff = log(exp(goff) + two_prior_over_meanlib);
    RowSrc src{cnt, G, logoff, mode == 1 ? yadd : nullptr, phi, phi_col, 0.0};
    Row r(src, g, goff);
    double beta[PM], cellbeta[PM];
    extern __shared__ __align__(16) double glm_scr[];       // lev_scratch_doubles<PM>() per lane group, or nothing
    double* scr = use_scr ? glm_scr + (threadIdx.x / GLM_LANES) * lev_scratch_doubles<PM>() : nullptr;
    fit_gene<GLM_LANES, PM>(d, r, beta, cellbeta, ln, gm, scr);
    if (coef) for (int k = ln; k < d.p; k += GLM_LANES) coef[g + (long)k * G] = beta[k];
    if (alpha_out && ln == 0) {
        double a = 0;
        for (int b = 0; b < n_alpha; ++b) a += beta[b] * alpha_w[b];
        alpha_out[g] = a;
    }
    if (fitted) {
        if (d.oneway) {
            for (int l = 0; l < d.p; ++l)
                for (int t = d.cell_start[l] + ln; t < d.cell_start[l + 1]; t += GLM_LANES) {
                    const int c = d.col[t];
                    fitted[g + (long)c * G] = exp(cellbeta[l] + r.off(c));
                }
        } else {
            const bool dead = isnan(beta[0]);
            for (int t = ln; t < d.n; t += GLM_LANES) fitted[g + (long)d.col[t] * G] = dead ? 0.0 : exp(lin_pred(d, r, t, beta));
        }
    }
}

// ---- a12/a13: batch-free parameters + match_quantiles (helper_seq.R:46-64), one thread per element -------
__global__ void k_phi_star_mean(const double* __restrict__ phi_star_mat, long G, int B, double* __restrict__ phi_star) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    double s = 0;
    for (int b = 0; b < B; ++b) s += phi_star_mat[g + (long)b * G];
    phi_star[g] = s / B;                                    // rowMeans  ComBat_seq.R:178
}

// grid: (ceil(G/128), N).  out is G0-row column-major; row of kept gene i is keep_idx[i] (or i when keep_idx == nullptr)
__global__ void __launch_bounds__(128)
k_match_quantiles(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ mu_hat,
                  const int* __restrict__ batch_idx, const double* __restrict__ phi_hat,
                  const double* __restrict__ gamma_star, const double* __restrict__ phi_star,
                  const int* __restrict__ keep_idx, long G0, double* __restrict__ out) {
    const long g = g0 + (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (g >= g1) return;
    const int b = batch_idx[s];
    const double x = (double)cnt[g + (long)s * G];
    const double old_mu = mu_hat[g + (long)s * G];
    const double new_mu = exp(log(old_mu) - gamma_star[g + (long)b * G]);     // ComBat_seq.R:176
    const double v = match_one_fast(x, old_mu, phi_hat[g + (long)b * G], new_mu, phi_star[g]);
    out[(keep_idx ? (long)keep_idx[g] : g) + (long)s * G0] = v;
}

// Same contract, different thread mapping: one warp = one gene x 32 samples of (mostly) one batch.  The nmath routines
// branch on the magnitudes of x, mu and size (continued-fraction length, pmf-walk length, pass-throughs), and those are
// similar along a gene's row and wildly different down a column: with one warp per 32 consecutive genes only 10 of 32
// lanes were active per instruction (ncu, round 1).  A block owns an 8-gene x 32-sample tile: it is loaded with 8
// consecutive genes per sample (whole 32-byte sectors), transposed through shared memory, computed with warp w = gene
// w, lane l = sample l, and stored the way it was loaded.  order[] lists the samples sorted by batch.
// grid: (ceil((g1-g0)/8), ceil(N/32)), block 256.
template <bool NI>                                           // NI: out-of-line continued fraction (nbmath.cuh)
__global__ void __launch_bounds__(256)
k_match_quantiles_rows(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ mu_hat,
                       const int* __restrict__ batch_idx, const int* __restrict__ order, int N,
                       const double* __restrict__ phi_hat, const double* __restrict__ gamma_star,
                       const double* __restrict__ phi_star, const int* __restrict__ keep_idx, long G0,
                       double* __restrict__ out) {
    __shared__ double s_mu[32][9], s_out[32][9];
    __shared__ int s_x[32][9];
    const int t = threadIdx.x, gi = t & 7, si = t >> 3, w = t >> 5, l = t & 31;
    const long gl = g0 + (long)blockIdx.x * 8;
    const int sbase = blockIdx.y * 32;
    const bool io_ok = (gl + gi < g1) && (sbase + si < N);
    const int s_io = io_ok ? order[sbase + si] : 0;
    if (io_ok) {
        s_x[si][gi] = cnt[gl + gi + (long)s_io * G];
        s_mu[si][gi] = mu_hat[gl + gi + (long)s_io * G];
    }
    __syncthreads();
    if (gl + w < g1 && sbase + l < N) {
        const long g = gl + w;
        const int b = batch_idx[order[sbase + l]];
        const double old_mu = s_mu[l][w];
        const double new_mu = exp(log(old_mu) - gamma_star[g + (long)b * G]);     // ComBat_seq.R:176
        s_out[l][w] = match_one_fast<NI>((double)s_x[l][w], old_mu, phi_hat[g + (long)b * G], new_mu, phi_star[g]);
    }
    __syncthreads();
    if (io_ok) out[(keep_idx ? (long)keep_idx[gl + gi] : gl + gi) + (long)s_io * G0] = s_out[si][gi];
}

// Warp = gene, lane = sample like k_match_quantiles_rows, but without the shared-memory tile: the kernel is bound by
// FP64 arithmetic and divergence (ncu, round 1: 34 % FP64 pipe, 14 of 32 lanes active, DRAM traffic 0.3 % of peak), so
// the strided loads and stores (one 32-byte sector per element, mostly L2 hits because the 4 warps of a block and the
// neighbouring blocks share the sectors) cost nothing, while the two block barriers of the tiled form kept finished
// warps waiting for the slowest gene of the tile (23 % of the stall samples).  grid: (ceil((g1-g0)/4), ceil(N/32)), block 128.
template <bool NI>                                           // NI: out-of-line continued fraction (nbmath.cuh)
__global__ void __launch_bounds__(128, 5)
k_match_quantiles_warp(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ mu_hat,
                       const int* __restrict__ batch_idx, const int* __restrict__ order, int N,
                       const double* __restrict__ phi_hat, const double* __restrict__ gamma_star,
                       const double* __restrict__ phi_star, const int* __restrict__ keep_idx, long G0,
                       double* __restrict__ out) {
    const long g = g0 + (long)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int si = blockIdx.y * 32 + (threadIdx.x & 31);
    if (g >= g1 || si >= N) return;
    const int s = order[si], b = batch_idx[s];
    const double x = (double)cnt[g + (long)s * G];
    const double old_mu = mu_hat[g + (long)s * G];
    const double new_mu = exp(log(old_mu) - gamma_star[g + (long)b * G]);     // ComBat_seq.R:176
    out[(keep_idx ? (long)keep_idx[g] : g) + (long)s * G0] = match_one_fast<NI>(x, old_mu, phi_hat[g + (long)b * G], new_mu, phi_star[g]);
}

// generic per-batch form used by cbs_match_quantiles (explicit new_mu)
__global__ void __launch_bounds__(128)
k_match_quantiles_plain(const int32_t* __restrict__ cnt, long G, int n, const double* __restrict__ old_mu,
                        const double* __restrict__ old_phi, const double* __restrict__ new_mu,
                        const double* __restrict__ new_phi, double* __restrict__ out) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (g >= G) return;
    const long k = g + (long)s * G;
    out[k] = match_one_fast((double)cnt[k], old_mu[k], old_phi[g], new_mu[k], new_phi[g]);
}

// a14: output pre-fill with the original counts as doubles (rows of filtered genes pass through; ComBat_seq.R:202)
template <class T>
__global__ void k_counts_to_f64(const T* __restrict__ counts, long n, double* __restrict__ out) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (double)counts[i];
}

// ---- elementwise NB primitives (cbs_nb_logpmf / cbs_nb_cdf / cbs_nb_quantile) ----------------------------
__global__ void k_nb_elementwise(int what, long n, const double* __restrict__ a, const double* __restrict__ size,
                                 const double* __restrict__ mu, double* __restrict__ out) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    if (what == 0) out[i] = dnbinom_mu<true>(a[i], size[i], mu[i]);
    else if (what == 1) out[i] = pnbinom_mu(a[i], size[i], mu[i]);
    else out[i] = qnbinom_mu(a[i], size[i], mu[i]);
}

}  // namespace cbs
