// stages.inl — internal device-level stage drivers shared by the per-subsystem entry points and cbs_combat_seq().
// Included by combatseq.cu inside no namespace; everything here is static.

namespace cbs {

struct AplJob { int design; double phi; };

// gene shard [g0,g1) of rank r: contiguous blocks rounded to the APL reduction chunk so that the chunk partial
// sums (and therefore disp_common and everything downstream) are identical for every GPU count.
static void shard_range(long G, int rank, int world, long* g0, long* g1) {
    long per = (G + world - 1) / world;
    per = (per + APL_CHUNK - 1) / APL_CHUNK * APL_CHUNK;
    *g0 = std::min(G, (long)rank * per);
    *g1 = std::min(G, (long)(rank + 1) * per);
}

// APL of genes [g0,g1) for every job -> d_apl[job*G + g]
static void launch_apl(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignDev* designs,
                       const unsigned char* sel, const double* phi_gene, const std::vector<AplJob>& jobs, double* d_apl,
                       cudaStream_t st) {
    if (g1 <= g0) return;
    double* lgfact = g_dev.get<double>("lgfact", (size_t)LGFACT_N);
    CBS_LAUNCH(k_lgfact_init, grid1(LGFACT_N, 256), 256, 0, st, lgfact);
    for (size_t e0 = 0; e0 < jobs.size(); e0 += APL_MAX_EVALS) {
        AplLaunch L;
        L.n_evals = (int)std::min<size_t>(APL_MAX_EVALS, jobs.size() - e0);
        for (int e = 0; e < L.n_evals; ++e) { L.design[e] = jobs[e0 + e].design; L.phi[e] = jobs[e0 + e].phi; }
        dim3 grid((unsigned)((g1 - g0 + 127) / 128), (unsigned)L.n_evals);
        const double* lgr = nullptr;
        if (g_apl_fast && !phi_gene) {
            double* t = g_dev.get<double>("lgr", (size_t)APL_MAX_EVALS * LGFACT_N);
            CBS_LAUNCH(k_lgr_init, dim3(LGFACT_N / 256, (unsigned)L.n_evals), 256, 0, st, L, t);
            lgr = t;
        }
        CBS_LAUNCH(k_apl, grid, 128, 0, st, cnt, G, g0, g1, logoff, designs, sel,
                   phi_gene ? phi_gene + e0 * G : nullptr, L, d_apl + e0 * G, g_apl_fast ? (const double*)lgfact : (const double*)nullptr, lgr);
    }
}

// per-job sum over the selected genes, deterministic: chunk partials on the device (tree order inside a chunk),
// all-reduced across ranks (each rank owns disjoint chunks, the others contribute exact zeros) and added on the
// host in chunk order.
static void apl_sums(const double* d_apl, long G, long g0, long g1, const unsigned char* sel,
                     const std::vector<AplJob>& jobs, cbs_comm* comm, cudaStream_t st, std::vector<double>& sums) {
    const long nchunk = (G + APL_CHUNK - 1) / APL_CHUNK;
    const size_t nj = jobs.size();
    double* d_part = g_dev.get<double>("apl_partial", nj * nchunk);
    CBS_CUDA(cudaMemsetAsync(d_part, 0, nj * nchunk * sizeof(double), st));
    const long c0 = g0 / APL_CHUNK, c1 = (g1 + APL_CHUNK - 1) / APL_CHUNK;
    if (g1 > g0)                              // a rank without genes contributes exact zeros (its g0 == G need not be chunk-aligned)
        for (size_t e0 = 0; e0 < nj; e0 += APL_MAX_EVALS) {
            AplLaunch L;
            L.n_evals = (int)std::min<size_t>(APL_MAX_EVALS, nj - e0);
            for (int e = 0; e < L.n_evals; ++e) { L.design[e] = jobs[e0 + e].design; L.phi[e] = jobs[e0 + e].phi; }
            dim3 grid((unsigned)(c1 - c0), (unsigned)L.n_evals);
            CBS_LAUNCH(k_chunk_sum, grid, 256, 0, st, d_apl + e0 * G, G, sel, L, c0, d_part + e0 * nchunk, nchunk);
        }
    if (comm && comm->world > 1)
        CBS_NCCL(nccl_api()->AllReduce(d_part, d_part, nj * nchunk, NCCL_FLOAT64, NCCL_SUM, comm->nccl, st));
    std::vector<double> h(nj * nchunk);
    CBS_CUDA(cudaMemcpyAsync(h.data(), d_part, h.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaStreamSynchronize(st));
    sums.assign(nj, 0.0);
    for (size_t e = 0; e < nj; ++e) { double s = 0; for (long c = 0; c < nchunk; ++c) s += h[e * nchunk + c]; sums[e] = s; }
}

// estimateGLMCommonDisp for several (batch) designs at once: the Brent searches advance in lock-step, one APL
// launch per step evaluates every unfinished batch at its current delta.  designs[b] is batch b's DesignDev,
// sel[b*G+g] its rowSums>=5 mask.  Returns NaN for a batch whose design has p >= n (edgeR: warning + NA).
static void run_common_disp(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignSet& ds,
                            const unsigned char* sel, cbs_comm* comm, cudaStream_t st, std::vector<double>& common,
                            int* n_steps) {
    const int nb = (int)ds.h_array.size();
    common.assign(nb, std::nan(""));
    std::vector<BrentFmin> br(nb);
    std::vector<char> active(nb, 0);
    for (int b = 0; b < nb; ++b)
        if (ds.h_array[b].p < ds.h_array[b].n) { br[b].init(0.0, std::pow(4.0, 0.25), 1e-5); active[b] = 1; }
    double* d_apl = g_dev.get<double>("apl_common", (size_t)nb * G);
    int steps = 0;
    for (;;) {
        std::vector<AplJob> jobs;
        std::vector<int> who;
        for (int b = 0; b < nb; ++b)
            if (active[b] && !br[b].done) { const double x = br[b].query(); jobs.push_back({b, x * x * x * x}); who.push_back(b); }
        if (jobs.empty()) break;
        launch_apl(cnt, G, g0, g1, logoff, ds.d_array, sel, nullptr, jobs, d_apl, st);
        std::vector<double> sums;
        apl_sums(d_apl, G, g0, g1, sel, jobs, comm, st, sums);
        for (size_t k = 0; k < who.size(); ++k) br[who[k]].feed(-sums[k]);     // optimize(maximum = TRUE)
        ++steps;
    }
    for (int b = 0; b < nb; ++b) if (active[b]) { const double x = br[b].result(); common[b] = x * x * x * x; }
    if (n_steps) *n_steps = steps;
}

// estimateGLMTagwiseDisp(prior.df = 0) for the batches with do_batch[b] != 0: phi_hat[g + b*G] for g in [g0,g1).
// Batches with do_batch[b] == 0 get rep(common[b]) (ComBat_seq.R:113).
static void run_tagwise_disp(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignSet& ds,
                             const unsigned char* sel, const std::vector<double>& common, const std::vector<char>& do_batch,
                             cudaStream_t st, double* d_phi_hat) {
    const int nb = (int)ds.h_array.size();
    if (g1 <= g0) return;
    std::vector<AplJob> jobs;
    std::vector<int> job0(nb, -1);
    for (int b = 0; b < nb; ++b) {
        if (!do_batch[b]) continue;
        job0[b] = (int)jobs.size();
        for (int k = 0; k < 11; ++k) jobs.push_back({b, common[b] * std::pow(2.0, -6.0 + k * (12.0 / 10))});
    }
    double* d_apl = jobs.empty() ? nullptr : g_dev.get<double>("apl_grid", jobs.size() * (size_t)G);
    if (!jobs.empty()) launch_apl(cnt, G, g0, g1, logoff, ds.d_array, sel, nullptr, jobs, d_apl, st);
    for (int b = 0; b < nb; ++b) {
        if (do_batch[b]) {
            CBS_LAUNCH(k_tagwise_max, grid1(g1 - g0, 128), 128, 0, st, d_apl + (size_t)job0[b] * G, G, g0, g1,
                       sel ? sel + (size_t)b * G : nullptr, common[b], d_phi_hat + (size_t)b * G);
        } else {
            CBS_LAUNCH(k_fill, grid1(g1 - g0, 256), 256, 0, st, d_phi_hat + (size_t)b * G + g0, g1 - g0, common[b]);
        }
    }
}

// glmFit.default(prior.count) on genes [g0,g1): coef (G x p, may be null) from the prior-count refit when
// prior_count > 0 else from the plain fit; fitted (G x N, may be null) from the plain fit; alpha (G, may be null)
// = sum_b alpha_w[b] * coef[g,b].  d_lib is the per-sample library size on the device (N); h_lib the host copy.
static void run_glm_fit(const int32_t* cnt, long G, long g0, long g1, int N, const double* d_logoff,
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
    const dim3 lgrid = grid1((g1 - g0) * GLM_LANES, 128);
    auto launch = [&](const double* yadd, double tp, int mode, double* coef, double* fitted, double* alpha) {
        // one shared-memory copy of the Levenberg work arrays per gene (16 genes per block)
#define CBS_GLM_LANES(PM_) do { \
            const size_t smem = g_glm_scr ? (size_t)(128 / GLM_LANES) * lev_scratch_doubles<PM_>() * sizeof(double) : 0; \
            static bool attr_set = false; \
            if (!attr_set) { CBS_CUDA(cudaFuncSetAttribute(k_glm_fit_lanes<PM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                                                           (int)((128 / GLM_LANES) * lev_scratch_doubles<PM_>() * sizeof(double)))); attr_set = true; } \
            CBS_LAUNCH(k_glm_fit_lanes<PM_>, lgrid, 128, smem, st, cnt, G, g0, g1, d_logoff, yadd, tp, d_phi, \
                       d_phi_col, d_design, d_goff, mode, coef, fitted, d_alpha_w, n_alpha, alpha, g_glm_scr ? 1 : 0); } while (0)
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

// number of Chebyshev terms for the range h = (uhi - ulo)/2 of u = log(r + c): the series of lgamma(x + e^u - c)
// converges like rho^-K with rho fixed by the singularities at Im u = +-pi above the end points of the interval.
static int mcg_terms(double h) {
    if (h < 1e-3) return 8;
    const double wr = 1.0, wi = M_PI / h;                     // w = 1 + i pi/h
    const double ar = wr * wr - wi * wi - 1.0, ai = 2 * wr * wi;   // w^2 - 1
    const double mod = std::sqrt(std::hypot(ar, ai)), arg = 0.5 * std::atan2(ai, ar);
    const double rho = std::hypot(wr + mod * std::cos(arg), wi + mod * std::sin(arg));
    int K = (int)std::ceil(std::log(1e14) / std::log(rho)) + 4;
    K = (K + 1) & ~1;
    return std::max(8, K);
}

// monte_carlo_int_NB for one batch by the direct pairwise kernel (every lgamma evaluated)
static void run_mc_batch_direct(const int32_t* cnt, long G, int n, const int* d_cols, long i0, long i1, const double* d_mu_hat,
                         const double* d_gamma, const double* d_phi, int mc_mode, int subset_n, uint64_t seed,
                         double* d_gamma_star, double* d_phi_star, unsigned long long* d_nfb, cudaStream_t st) {
    double* A = g_dev.get<double>("mc_A", (size_t)n * G);
    double* cj = g_dev.get<double>("mc_cj", (size_t)G);
    double* rj = g_dev.get<double>("mc_rj", (size_t)G);
    double* lgr = g_dev.get<double>("mc_lgr", (size_t)G);
    double* di = g_dev.get<double>("mc_di", (size_t)G);
    CBS_LAUNCH(k_mc_prep_src, grid1(G, 128), 128, 0, st, d_mu_hat, G, n, d_cols, d_phi, A, cj, rj, lgr);
    CBS_LAUNCH(k_mc_prep_tgt, grid1(G, 128), 128, 0, st, cnt, G, n, d_cols, di);
    if (i1 <= i0) return;
    if (subset_n >= G - 1) subset_n = 0;      // sample(1:(G-1), G-1) is a permutation of all other genes
    const size_t smem = (size_t)MC_TI * n * sizeof(double);
    const unsigned nblk = (unsigned)((i1 - i0 + MC_TI - 1) / MC_TI);
    CBS_LAUNCH(k_mc_direct, dim3(nblk), 256, smem, st, cnt, G, n, d_cols, i0, i1, A, cj, rj, lgr, di, d_gamma, d_phi,
               mc_mode == CBS_MC_STABLE ? 1 : 0, subset_n, seed, d_gamma_star, d_phi_star, d_nfb);
}

// monte_carlo_int_NB for one batch as P Q^T on the FP64 tensor pipe (mc_gemm.cuh).  Returns false (nothing
// written) when the batch's dispersion range needs more Chebyshev terms than the kernel supports.
// *tail_out receives the series' tail estimate (synchronises the stream).
static bool run_mc_batch_gemm(const int32_t* cnt, long G, int n, const int* d_cols, long i0, long i1, const double* d_mu_hat,
                              const double* d_gamma, const double* d_phi, int mc_mode, double* d_gamma_star,
                              double* d_phi_star, unsigned long long* d_nfb, cudaStream_t st, double* tail_out,
                              cbs_comm* comm = nullptr) {
    double* d_rng = g_dev.get<double>("mcg_range", 2);
    CBS_LAUNCH(k_mcg_phi_range, dim3(1), 256, 0, st, d_phi, G, d_rng);
    double rng[2];
    CBS_CUDA(cudaMemcpyAsync(rng, d_rng, sizeof rng, cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaStreamSynchronize(st));
    if (!(rng[0] > 0) || !std::isfinite(rng[1])) return false;
    McgShape sh;
    sh.n = n;
    sh.ulo = std::log(1.0 / rng[1] + MCG_SHIFT);
    sh.uhi = std::log(1.0 / rng[0] + MCG_SHIFT);
    if (sh.uhi - sh.ulo < 0.02) { const double mid = 0.5 * (sh.ulo + sh.uhi); sh.ulo = mid - 0.01; sh.uhi = mid + 0.01; }
    sh.Kc = mcg_terms(0.5 * (sh.uhi - sh.ulo));
    if (sh.Kc > MCG_MAX_KC) return false;
    sh.KT = n + sh.Kc + MCG_SHIFT + 1;
    // K is padded to whole 8-wide k-chunks only; at least two full hand-overs per source tile (gamma/phi buffering)
    sh.KTP = std::max((sh.KT + MCG_KC - 1) / MCG_KC * MCG_KC, 2 * MCG_KC * MCG_SUB);
    static_assert((MCG_GPBUF - 1) * 2 >= MCG_STAGES + 1, "gamma/phi buffers must outlive the stage ring");
    sh.rows = (G + MCG_TILE - 1) / MCG_TILE * MCG_TILE;
    g_last_mc_kc = sh.Kc;
    std::vector<double> dct((size_t)sh.Kc * sh.Kc);
    for (int k = 0; k < sh.Kc; ++k)
        for (int m = 0; m < sh.Kc; ++m) dct[(size_t)k * sh.Kc + m] = (k ? 2.0 : 1.0) / sh.Kc * std::cos(M_PI * k * (m + 0.5) / sh.Kc);
    double* d_dct = upload(g_dev, "mcg_dct", dct.data(), dct.size(), st);
    double* P = g_dev.get<double>("mcg_P", (size_t)sh.KTP * sh.rows);
    double* Q = g_dev.get<double>("mcg_Q", (size_t)sh.KTP * sh.rows);
    double* Fn = g_dev.get<double>("mcg_Fnode", (size_t)sh.Kc * G);
    double2* gp = (double2*)g_dev.get<double>("mcg_gp", (size_t)2 * sh.rows);
    unsigned long long* d_tail = g_dev.get<unsigned long long>("mcg_tail", 1);
    CBS_CUDA(cudaMemsetAsync(d_tail, 0, sizeof(unsigned long long), st));
    CBS_LAUNCH(k_mcg_prep_src, grid1(sh.rows, 128), 128, 0, st, d_mu_hat, G, d_cols, d_phi, sh, Q);
    // P is only needed for this rank's target tiles (Q for all sources)
    const long ip0 = i1 > i0 ? i0 / MCG_TI * MCG_TI : 0, ip1 = i1 > i0 ? std::min(sh.rows, (i1 + MCG_TI - 1) / MCG_TI * MCG_TI) : 0;
    if (ip1 > ip0) {
        double* d_lgtab = g_dev.get<double>("mcg_lgtab", (size_t)sh.Kc * MCG_LGTAB);
        CBS_LAUNCH(k_mcg_node_table, dim3(MCG_LGTAB / 256, (unsigned)sh.Kc), 256, 0, st, sh, d_lgtab);
        CBS_LAUNCH(k_mcg_prep_nodes, dim3((unsigned)((ip1 - ip0 + 127) / 128), (unsigned)sh.Kc), 128, 0, st, cnt, G, d_cols, sh, ip0, ip1, d_lgtab, Fn);
        CBS_LAUNCH(k_mcg_prep_tgt, grid1(ip1 - ip0, 128), 128, 0, st, cnt, G, d_cols, sh, ip0, ip1, Fn, d_dct, P, d_tail);
    }
    CBS_LAUNCH(k_mcg_pack_gp, grid1(sh.rows, 256), 256, 0, st, d_gamma, d_phi, G, sh.rows, gp);
    if (i1 > i0) {
        const long tile0 = i0 / MCG_TI, ntiles = (i1 - i0 + MCG_TI - 1) / MCG_TI, NJ = sh.rows / MCG_TJ;
        int sms = 148;
        { int dev = 0; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
        // source-tile splits: at least ~8 waves of CTAs (two CTAs per SM), and among the candidates the split count whose
        // last wave is fullest (2811 CTAs on 296 slots waste half a wave; 4690 waste 1 %)
        const long slots = (long)MCG_CTAS * sms;
        const long lo = std::max(1L, (8 * slots + ntiles - 1) / ntiles), hi = std::min(32L, std::max(1L, NJ / 4));
        long nsplit = std::min(lo, hi);
        double best = 0;
        for (long c = std::min(lo, hi); c <= std::min(hi, lo + 4); ++c) {
            const long ctas = ntiles * c;
            const double eff = (double)ctas / (double)(((ctas + slots - 1) / slots) * slots);
            if (eff > best + 1e-9) { best = eff; nsplit = c; }
        }
        McAcc* part = (McAcc*)g_dev.get<double>("mcg_part", (size_t)nsplit * sh.rows * (sizeof(McAcc) / sizeof(double)));
        static bool attr_set = false;
        if (!attr_set) {
            CBS_CUDA(cudaFuncSetAttribute(k_mc_gemm<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MCG_SMEM_BYTES));
            CBS_CUDA(cudaFuncSetAttribute(k_mc_gemm<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MCG_SMEM_BYTES));
            CBS_CUDA(cudaFuncSetAttribute(k_mc_gemm<0>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));   // two CTAs per SM
            CBS_CUDA(cudaFuncSetAttribute(k_mc_gemm<1>, cudaFuncAttributePreferredSharedMemoryCarveout, 100));
            attr_set = true;
        }
        const dim3 grid((unsigned)ntiles, (unsigned)nsplit);
        if (mc_mode == CBS_MC_STABLE)
            CBS_LAUNCH(k_mc_gemm<1>, grid, MCG_BLOCK, MCG_SMEM_BYTES, st, P, Q, gp, sh, tile0, part);
        else
            CBS_LAUNCH(k_mc_gemm<0>, grid, MCG_BLOCK, MCG_SMEM_BYTES, st, P, Q, gp, sh, tile0, part);
        CBS_LAUNCH(k_mcg_finalize, grid1(i1 - i0, 128), 128, 0, st, part, sh.rows, (int)nsplit, i0, i1,
                   mc_mode == CBS_MC_STABLE ? 1 : 0, d_gamma, d_phi, d_gamma_star, d_phi_star, d_nfb);
    }
    // the series' tail is judged over ALL targets, so that every rank takes the same decision a single GPU would
    if (comm && comm->world > 1) CBS_NCCL(nccl_api()->AllReduce(d_tail, d_tail, 1, NCCL_UINT64, NCCL_MAX, comm->nccl, st));
    unsigned long long tb = 0;
    CBS_CUDA(cudaMemcpyAsync(&tb, d_tail, sizeof tb, cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaStreamSynchronize(st));
    double tail; memcpy(&tail, &tb, sizeof tail);
    *tail_out = tail;
    return true;
}

// monte_carlo_int_NB for one batch, targets [i0,i1); sources are all G genes.  d_cols: the batch's columns (n).
// d_nfb must be zero on entry.
static void run_mc_batch(const int32_t* cnt, long G, int n, const int* d_cols, long i0, long i1, const double* d_mu_hat,
                         const double* d_gamma, const double* d_phi, int mc_mode, int subset_n, uint64_t seed,
                         double* d_gamma_star, double* d_phi_star, unsigned long long* d_nfb, cudaStream_t st,
                         cbs_comm* comm = nullptr) {
    if (subset_n >= G - 1) subset_n = 0;
    const bool want_gemm = subset_n <= 0 && g_mc_algo != 1 && G >= 2;
    if (want_gemm) {
        double tail = 0;
        if (run_mc_batch_gemm(cnt, G, n, d_cols, i0, i1, d_mu_hat, d_gamma, d_phi, mc_mode, d_gamma_star, d_phi_star, d_nfb, st, &tail, comm)) {
            g_last_mc_tail = std::max(g_last_mc_tail, tail);
            if (tail <= g_mc_tail_tol || g_mc_algo == 2) return;
            CBS_CUDA(cudaMemsetAsync(d_nfb, 0, sizeof(unsigned long long), st));     // series not converged: redo exactly
        }
    }
    run_mc_batch_direct(cnt, G, n, d_cols, i0, i1, d_mu_hat, d_gamma, d_phi, mc_mode, subset_n, seed, d_gamma_star, d_phi_star, d_nfb, st);
}

}  // namespace cbs
