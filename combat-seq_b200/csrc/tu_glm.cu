// tu_glm.cu — the two glmFit calls (a8-a10): per-gene NB log-link GLM kernels and their stage driver.
//
// Replaces glmFit / glmFit.default of edgeR 3.22 as called at /root/reference/ComBat_seq.R:130,134 (plain fit for the
// fitted values, prior.count refit for the coefficients, alpha_g of ComBat_seq.R:131 fused).
#include "engine.hpp"

namespace cbs {

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
#ifndef CBS_GLM_MINBLOCKS
#define CBS_GLM_MINBLOCKS 5
#endif
// what a fit leaves behind: coefficients, alpha_g (ComBat_seq.R:131), fitted values; LANES lanes of the gene share the stores
template <int LANES, int PM>
__device__ __forceinline__ void glm_store(const DesignDev& d, const Row& r, long g, long G, int ln, const double* beta,
                                          const double* cellbeta, double* __restrict__ coef, double* __restrict__ fitted,
                                          const double* __restrict__ alpha_w, int n_alpha, double* __restrict__ alpha_out) {
    if (coef) for (int k = ln; k < d.p; k += LANES) coef[g + (long)k * G] = beta[k];
    if (alpha_out && ln == 0) {
        double a = 0;
        for (int b = 0; b < n_alpha; ++b) a += beta[b] * alpha_w[b];
        alpha_out[g] = a;
    }
    if (fitted) {
        if (d.oneway) {
            for (int l = 0; l < d.p; ++l)
                for (int t = d.cell_start[l] + ln; t < d.cell_start[l + 1]; t += LANES) {
                    const int c = d.col[t];
                    fitted[g + (long)c * G] = exp(cellbeta[l] + r.off(c));
                }
        } else {
            const bool dead = isnan(beta[0]);
            for (int t = ln; t < d.n; t += LANES) fitted[g + (long)d.col[t] * G] = dead ? 0.0 : exp(lin_pred(d, r, t, beta));
        }
    }
}

// A fit that needs more than GLM_SWEEP_CAP sweeps over the samples (Levenberg trial steps that keep being rejected: 0.05 % of
// C5's fits, up to 64 sweeps against the usual 5-7) is abandoned here and its gene appended to `strag`; k_glm_fit_wide redoes
// it with a whole warp and the gene's counts in shared memory.  On 8 lanes such a fit is a serial chain ten times longer
// than its neighbours': it alone set the duration of a launch once the genes were spread over 8 GPUs (7 ms per launch
// whatever the shard).  The decision depends on the gene only, so it is the same for every sharding.
constexpr int GLM_SWEEP_CAP = 9;
template <int PM>                                            // PM >= design columns: sizes the per-thread packed X'WX (local memory)
__global__ void __launch_bounds__(128, CBS_GLM_MINBLOCKS)
k_glm_fit_lanes(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ logoff,
                const double* __restrict__ yadd, double two_prior_over_meanlib, const double* __restrict__ phi,
                const int* __restrict__ phi_col, const DesignDev* __restrict__ design, const double* __restrict__ goff_in,
                int mode, double* __restrict__ coef, double* __restrict__ fitted, const double* __restrict__ alpha_w,
                int n_alpha, double* __restrict__ alpha_out, int use_scr, int* __restrict__ iter_hist,
                int* __restrict__ strag, int* __restrict__ n_strag) {
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
    const int n_sweeps = fit_gene<GLM_LANES, PM>(d, r, beta, cellbeta, ln, gm, scr, strag ? GLM_SWEEP_CAP : 0x7fffffff);
    if (n_sweeps == -2) {                                    // too long for 8 lanes: the wide pass fits it
        if (ln == 0) strag[atomicAdd(n_strag, 1)] = (int)(g - g0);
        return;
    }
    if (iter_hist && ln == 0) atomicAdd(&iter_hist[max(0, min(n_sweeps, 255))], 1);      // diagnostics: sweeps per fit
    glm_store<GLM_LANES, PM>(d, r, g, G, ln, beta, cellbeta, coef, fitted, alpha_w, n_alpha, alpha_out);
}

// Second pass: one warp per abandoned gene, the gene's counts staged in shared memory (the sweeps then run at shared-memory
// instead of DRAM latency), the Levenberg work arrays in shared memory as in the first pass.  The fit restarts from the null
// start; with 32 lanes the sums associate differently than with 8, so the coefficients of these genes differ from an 8-lane
// fit in the last places: what matters is that the route a gene takes does not depend on the sharding.
// grid: any; warps stride over the list.  Dynamic shared memory: 4 x (scratch + N ints).
template <int PM>
__global__ void __launch_bounds__(128)
k_glm_fit_wide(const int32_t* __restrict__ cnt, long G, long g0, int N, const double* __restrict__ logoff,
               const double* __restrict__ yadd, double two_prior_over_meanlib, const double* __restrict__ phi,
               const int* __restrict__ phi_col, const DesignDev* __restrict__ design, const double* __restrict__ goff_in,
               int mode, double* __restrict__ coef, double* __restrict__ fitted, const double* __restrict__ alpha_w,
               int n_alpha, double* __restrict__ alpha_out, int* __restrict__ iter_hist, const int* __restrict__ strag,
               const int* __restrict__ n_strag, int stage_counts) {
    extern __shared__ __align__(16) double glm_scr[];
    const int warp = threadIdx.x >> 5, ln = threadIdx.x & 31;
    double* scr = glm_scr + warp * lev_scratch_doubles<PM>();
    int32_t* xs = reinterpret_cast<int32_t*>(glm_scr + 4 * lev_scratch_doubles<PM>()) + (long)warp * N;
    const DesignDev d = *design;
    const int count = *n_strag;
    for (int i = blockIdx.x * 4 + warp; i < count; i += gridDim.x * 4) {
        const long g = g0 + strag[i];
        if (stage_counts) {
            __syncwarp();
            for (int c = ln; c < N; c += 32) xs[c] = cnt[g + (long)c * G];
            __syncwarp();
        }
        double goff = goff_in ? goff_in[g] : 0.0;
        if (mode == 1) goff = log(exp(goff) + two_prior_over_meanlib);
        RowSrc src{cnt, G, logoff, mode == 1 ? yadd : nullptr, phi, phi_col, 0.0, stage_counts ? xs : nullptr};
        Row r(src, g, goff);
        double beta[PM], cellbeta[PM];
        const int n_sweeps = fit_gene<32, PM>(d, r, beta, cellbeta, ln, 0xffffffffu, scr);
        if (iter_hist && ln == 0) atomicAdd(&iter_hist[max(0, min(n_sweeps, 255))], 1);
        glm_store<32, PM>(d, r, g, G, ln, beta, cellbeta, coef, fitted, alpha_w, n_alpha, alpha_out);
    }
}

// glmFit.default(prior.count) on genes [g0,g1): coef (G x p, may be null) from the prior-count refit when
// prior_count > 0 else from the plain fit; fitted (G x N, may be null) from the plain fit; alpha (G, may be null)
// = sum_b alpha_w[b] * coef[g,b].  d_lib is the per-sample library size on the device (N); h_lib the host copy.
void run_glm_fit(const int32_t* cnt, long G, long g0, long g1, int N, const double* d_logoff,
                        const std::vector<double>& h_lib, const double* d_phi, const int* d_phi_col,
                        const DesignDev* d_design, int h_p, const double* d_goff, double prior_count, double* d_coef,
                        double* d_fitted, const double* d_alpha_w, int n_alpha, double* d_alpha, cudaStream_t st) {
    if (g1 <= g0) return;
    const dim3 grid = grid1(g1 - g0, 128);
    double two_prior = 0;
    const double* d_yadd = nullptr;
    if (prior_count > 0) {
        double meanlib = 0;
        for (int s = 0; s < N; ++s) meanlib += h_lib[s];
        meanlib /= N;
        std::vector<double> yadd(N);
        for (int s = 0; s < N; ++s) yadd[s] = prior_count * h_lib[s] / meanlib;     // add_prior.cpp
        d_yadd = upload(g_dev, "glm_yadd", yadd.data(), (size_t)N, st);
        two_prior = 2 * prior_count / meanlib;
    }
    // designs with few columns go to the lane-group kernel (GLM_LANES lanes per gene); wider ones stay one thread per gene
    const bool lanes = h_p <= GLM_LANES_PMAX;
    int* d_hist = g_tune.glm_hist ? g_dev.get<int>("glm_iter_hist", 256) : nullptr;     // zeroed by cbs_get_diag("reset")
    const dim3 lgrid = grid1((g1 - g0) * GLM_LANES, 128);
    // the wide second pass needs the class form of the sweep with shared scratch (every Levenberg design of the pipeline)
    const bool two_pass = g_tune.glm_scr && g_tune.glm_two_pass;
    int* d_strag = g_dev.get<int>("glm_strag", (size_t)(g1 - g0));
    int* d_nstrag = g_dev.get<int>("glm_nstrag", 1);
    auto launch = [&](const double* yadd, double tp, int mode, double* coef, double* fitted, double* alpha) {
        // one shared-memory copy of the Levenberg work arrays per gene (16 genes per block)
#define CBS_GLM_LANES(PM_) do { \
            const size_t smem = g_tune.glm_scr ? (size_t)(128 / GLM_LANES) * lev_scratch_doubles<PM_>() * sizeof(double) : 0; \
            t_ctx->set_func_attr((const void*)k_glm_fit_lanes<PM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                 (int)((128 / GLM_LANES) * lev_scratch_doubles<PM_>() * sizeof(double))); \
            CBS_CUDA(cudaMemsetAsync(d_nstrag, 0, sizeof(int), st)); \
            CBS_LAUNCH(k_glm_fit_lanes<PM_>, lgrid, 128, smem, st, cnt, G, g0, g1, d_logoff, yadd, tp, d_phi, \
                       d_phi_col, d_design, d_goff, mode, coef, fitted, d_alpha_w, n_alpha, alpha, g_tune.glm_scr ? 1 : 0, d_hist, \
                       two_pass ? d_strag : (int*)nullptr, d_nstrag); \
            if (two_pass) { \
                const int stage = N <= 8192; \
                const size_t wsmem = 4 * lev_scratch_doubles<PM_>() * sizeof(double) + (stage ? (size_t)4 * N * sizeof(int32_t) : 0); \
                t_ctx->set_func_attr((const void*)k_glm_fit_wide<PM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024); \
                CBS_LAUNCH(k_glm_fit_wide<PM_>, dim3(64), 128, wsmem, st, cnt, G, g0, N, d_logoff, yadd, tp, d_phi, d_phi_col, d_design, \
                           d_goff, mode, coef, fitted, d_alpha_w, n_alpha, alpha, d_hist, (const int*)d_strag, (const int*)d_nstrag, stage); \
            } } while (0)
        if (lanes) {
            if (h_p <= 4) CBS_GLM_LANES(4); else if (h_p <= 8) CBS_GLM_LANES(8); else if (h_p <= 12) CBS_GLM_LANES(12); else CBS_GLM_LANES(16);
        } else
#undef CBS_GLM_LANES
            CBS_LAUNCH(k_glm_fit, grid, 128, 0, st, cnt, G, g0, g1, d_logoff, yadd, tp, d_phi, d_phi_col, d_design,
                       d_goff, mode, coef, fitted, d_alpha_w, n_alpha, alpha);
    };
    if (d_fitted || (prior_count <= 0 && (d_coef || d_alpha)))
        launch(nullptr, 0.0, 0, prior_count > 0 ? (double*)nullptr : d_coef, d_fitted, prior_count > 0 ? (double*)nullptr : d_alpha);
    if (prior_count > 0 && (d_coef || d_alpha))
        launch(d_yadd, two_prior, 1, d_coef, (double*)nullptr, d_alpha);
}

}  // namespace cbs
