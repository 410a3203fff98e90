// api_subsystems.inl — per-subsystem C-ABI entry points (host buffers in/out).  Included by combatseq.cu.

namespace cbs {

// design of a stand-alone sub-problem: X (n x p, column-major) or nullptr => intercept only
static HostDesign subproblem_design(const double* X, int n, int p) {
    std::vector<double> ones;
    if (!X) { ones.assign(n, 1.0); X = ones.data(); p = 1; }
    return make_design(X, n, p, nullptr);
}

static int check_design_size(int n, int p) {
    if (n < 1) return fail(CBS_EINVAL, "need at least one sample");
    if (p < 1 || p > PMAX) return fail(CBS_EINVAL, "design has %d columns; this build supports 1..%d", p, PMAX);
    return CBS_OK;
}

static int nb_elementwise(int what, int64_t n, const double* a, const double* size, const double* mu, double* out) {
    CBS_NEED_GPU();
    if (n < 0 || (n > 0 && (!a || !size || !mu || !out))) return fail(CBS_EINVAL, "null pointer");
    if (n == 0) return CBS_OK;
    CBS_ENTRY_BEGIN
    double* da = upload(g_dev, "ew_a", a, (size_t)n);
    double* ds = upload(g_dev, "ew_size", size, (size_t)n);
    double* dm = upload(g_dev, "ew_mu", mu, (size_t)n);
    double* dout = g_dev.get<double>("ew_out", (size_t)n);
    launch_nb_elementwise(what, (long)n, da, ds, dm, dout, 0);
    CBS_CUDA(cudaMemcpy(out, dout, (size_t)n * sizeof(double), cudaMemcpyDeviceToHost));
    CBS_ENTRY_END
}

}  // namespace cbs

extern "C" {

int cbs_nb_logpmf(int64_t n, const double* x, const double* size, const double* mu, double* out) { return nb_elementwise(0, n, x, size, mu, out); }
int cbs_nb_cdf(int64_t n, const double* x, const double* size, const double* mu, double* out) { return nb_elementwise(1, n, x, size, mu, out); }
int cbs_nb_quantile(int64_t n, const double* p, const double* size, const double* mu, double* out) { return nb_elementwise(2, n, p, size, mu, out); }

int cbs_simulate_counts(int64_t G, int32_t N, const double* base_mu, const double* col_factor, const int32_t* level,
                        int32_t n_levels, const double* fold, const double* phi_col, uint64_t seed, int32_t* counts,
                        void* counts_dev) {
    CBS_NEED_GPU();
    if (G < 0 || N < 0 || (G > 0 && N > 0 && (!base_mu || !col_factor || !phi_col))) return fail(CBS_EINVAL, "null pointer / bad size");
    if (!counts && !counts_dev) return fail(CBS_EINVAL, "no output buffer");
    if (fold && (!level || n_levels < 1)) return fail(CBS_EINVAL, "fold needs level[] and n_levels >= 1");
    if (fold) for (int s = 0; s < N; ++s) if (level[s] < 0 || level[s] >= n_levels) return fail(CBS_EINVAL, "level out of range");
    if (G == 0 || N == 0) return CBS_OK;
    CBS_ENTRY_BEGIN
    double* d_base = upload(g_dev, "sim_base", base_mu, (size_t)G);
    double* d_cf = upload(g_dev, "sim_colfac", col_factor, (size_t)N);
    double* d_phi = upload(g_dev, "sim_phi", phi_col, (size_t)N);
    int* d_level = fold ? upload(g_dev, "sim_level", level, (size_t)N) : nullptr;
    double* d_fold = fold ? upload(g_dev, "sim_fold", fold, (size_t)G * n_levels) : nullptr;
    int32_t* d_out = counts_dev ? (int32_t*)counts_dev : g_dev.get<int32_t>("sim_out", (size_t)G * N);
    launch_simulate_counts((long)G, N, d_base, d_cf, d_level, d_fold, d_phi, seed, d_out, 0);
    if (counts) CBS_CUDA(cudaMemcpy(counts, d_out, (size_t)G * N * sizeof(int32_t), cudaMemcpyDeviceToHost));
    CBS_CUDA(cudaDeviceSynchronize());
    CBS_ENTRY_END
}

int cbs_glm_fit(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p, const double* disp,
                int32_t n_disp_cols, const int32_t* disp_cols, const double* lib, const double* gene_offset,
                double prior_count, double* coef, double* fitted) {
    CBS_NEED_GPU();
    if (G < 0 || !counts || !design || !disp || !lib || n_disp_cols < 1) return fail(CBS_EINVAL, "null pointer / bad size");
    if (int rc = check_design_size(n, p)) return rc;
    if (disp_cols) for (int s = 0; s < n; ++s) if (disp_cols[s] < 0 || disp_cols[s] >= n_disp_cols) return fail(CBS_EINVAL, "disp_cols out of range");
    if (G == 0) return CBS_OK;
    CBS_ENTRY_BEGIN
    DesignSet ds("sub");
    ds.add(make_design(design, n, p, nullptr));
    ds.finish();
    int32_t* d_cnt = upload(g_dev, "cnt", counts, (size_t)G * n);
    double* d_phi = upload(g_dev, "phi_hat", disp, (size_t)G * n_disp_cols);
    std::vector<int> pc(n, 0);
    if (disp_cols) for (int s = 0; s < n; ++s) pc[s] = disp_cols[s];
    int* d_pc = upload(g_dev, "phi_col", pc.data(), (size_t)n);
    std::vector<double> hlib(lib, lib + n), logoff((size_t)2 * n);
    for (int s = 0; s < n; ++s) { logoff[2 * s] = std::log(hlib[s]); logoff[2 * s + 1] = hlib[s]; }
    double* d_logoff = upload(g_dev, "logoff", logoff.data(), (size_t)2 * n);
    double* d_goff = gene_offset ? upload(g_dev, "goff", gene_offset, (size_t)G) : nullptr;
    double* d_coef = coef ? g_dev.get<double>("coef", (size_t)G * p) : nullptr;
    double* d_fit = fitted ? g_dev.get<double>("mu_hat", (size_t)G * n) : nullptr;
    run_glm_fit(d_cnt, G, 0, G, n, d_logoff, hlib, d_phi, d_pc, ds.d_array, p, d_goff, prior_count, d_coef, d_fit,
                nullptr, 0, nullptr, 0, glm_class_flags(ds.host[0], pc.data()));
    if (coef) CBS_CUDA(cudaMemcpy(coef, d_coef, (size_t)G * p * sizeof(double), cudaMemcpyDeviceToHost));
    if (fitted) CBS_CUDA(cudaMemcpy(fitted, d_fit, (size_t)G * n * sizeof(double), cudaMemcpyDeviceToHost));
    CBS_CUDA(cudaDeviceSynchronize());
    CBS_ENTRY_END
}

int cbs_apl(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p, const double* lib,
            const double* disp, int32_t D, int32_t disp_per_gene, double* apl) {
    CBS_NEED_GPU();
    if (G < 0 || !counts || !lib || !disp || !apl || D < 1) return fail(CBS_EINVAL, "null pointer / bad size");
    if (int rc = check_design_size(n, design ? p : 1)) return rc;
    if (G == 0) return CBS_OK;
    CBS_ENTRY_BEGIN
    DesignSet ds("sub");
    ds.add(subproblem_design(design, n, p));
    ds.finish();
    int32_t* d_cnt = upload(g_dev, "cnt", counts, (size_t)G * n);
    std::vector<double> logoff((size_t)2 * n);
    for (int s = 0; s < n; ++s) { logoff[2 * s] = std::log(lib[s]); logoff[2 * s + 1] = lib[s]; }
    double* d_logoff = upload(g_dev, "logoff", logoff.data(), (size_t)2 * n);
    std::vector<AplJob> jobs(D);
    for (int d = 0; d < D; ++d) jobs[d] = {0, disp_per_gene ? 0.0 : disp[d]};
    double* d_pg = disp_per_gene ? upload(g_dev, "apl_phi_gene", disp, (size_t)G * D) : nullptr;
    double* d_apl = g_dev.get<double>("apl_grid", (size_t)G * D);
    launch_apl(d_cnt, G, 0, G, d_logoff, ds.d_array, nullptr, d_pg, jobs, d_apl, 0, &ds.h_array);
    CBS_CUDA(cudaMemcpy(apl, d_apl, (size_t)G * D * sizeof(double), cudaMemcpyDeviceToHost));
    CBS_ENTRY_END
}

}  // extern "C"

namespace cbs {
// shared set-up of cbs_common_disp / cbs_tagwise_disp: one batch whose columns are the whole matrix
struct OneBatch {
    DesignSet ds{"sub"};
    int32_t* d_cnt; double* d_logoff; unsigned char* d_sel; int* d_bidx;
    std::vector<double> h_lib;
    void init(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p) {
        ds.add(subproblem_design(design, n, p));
        ds.finish();
        d_cnt = upload(g_dev, "cnt", counts, (size_t)G * n);
        // offset = log(colSums(y))   (estimateGLMCommonDisp.default / estimateGLMTagwiseDisp.default)
        unsigned long long* d_cs = g_dev.get<unsigned long long>("colsum", (size_t)n);
        CBS_CUDA(cudaMemset(d_cs, 0, (size_t)n * sizeof(unsigned long long)));
        std::vector<int> ident(G);
        std::iota(ident.begin(), ident.end(), 0);
        int* d_ident = upload(g_dev, "keep_idx", ident.data(), (size_t)G);
        int32_t* d_tmp = g_dev.get<int32_t>("cnt_tmp", (size_t)G * n);
        CBS_LAUNCH(k_compact_colsum<int32_t>, dim3((unsigned)((G + 255) / 256), (unsigned)n), 256, 0, 0, d_cnt, (long)G,
                   d_ident, (long)G, d_tmp, d_cs);
        std::vector<unsigned long long> cs(n);
        CBS_CUDA(cudaMemcpy(cs.data(), d_cs, (size_t)n * sizeof(unsigned long long), cudaMemcpyDeviceToHost));
        std::vector<double> logoff((size_t)2 * n);
        for (int s = 0; s < n; ++s) { logoff[2 * s] = std::log((double)cs[s]); logoff[2 * s + 1] = (double)cs[s]; }
        h_lib.resize(n);
        for (int s = 0; s < n; ++s) h_lib[s] = (double)cs[s];
        d_logoff = upload(g_dev, "logoff", logoff.data(), (size_t)2 * n);
        std::vector<int> bidx(n, 0);
        d_bidx = upload(g_dev, "batch_idx", bidx.data(), (size_t)n);
        d_sel = g_dev.get<unsigned char>("sel", (size_t)G);
        CBS_LAUNCH(k_rowsum_sel, dim3((unsigned)((G + 127) / 128), 1), 128, 0, 0, d_cnt, (long)G, (int)n, d_bidx, 1, d_sel);
    }
};
}  // namespace cbs

extern "C" {

int cbs_common_disp(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p, double* common) {
    CBS_NEED_GPU();
    if (G < 1 || !counts || !common) return fail(CBS_EINVAL, "null pointer / bad size");
    if (int rc = check_design_size(n, design ? p : 1)) return rc;
    CBS_ENTRY_BEGIN
    OneBatch ob;
    ob.init(G, n, counts, design, p);
    std::vector<double> c;
    run_common_disp(ob.d_cnt, G, 0, G, ob.d_logoff, ob.ds, ob.d_sel, nullptr, 0, c, nullptr);
    *common = c[0];
    CBS_ENTRY_END
}

int cbs_tagwise_disp_prior(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p, double common,
                           double prior_df, double* tagwise) {
    CBS_NEED_GPU();
    if (G < 1 || !counts || !tagwise || !(prior_df >= 0)) return fail(CBS_EINVAL, "null pointer / bad size / negative prior.df");
    if (int rc = check_design_size(n, design ? p : 1)) return rc;
    CBS_ENTRY_BEGIN
    OneBatch ob;
    ob.init(G, n, counts, design, p);
    if (ob.ds.h_array[0].p >= n) {      // estimateGLMTagwiseDisp.default: ncol(design) >= ncol(y) -> NA
        for (int64_t g = 0; g < G; ++g) tagwise[g] = std::nan("");
    } else {
        DesignSet one("onecell");
        one.add(subproblem_design(nullptr, n, 1));
        one.finish();
        double* d_phi = g_dev.get<double>("phi_hat", (size_t)G);
        run_tagwise_disp_prior(ob.d_cnt, G, ob.d_logoff, ob.h_lib, ob.ds, one, ob.d_sel, common, prior_df, 0, d_phi);
        CBS_CUDA(cudaMemcpy(tagwise, d_phi, (size_t)G * sizeof(double), cudaMemcpyDeviceToHost));
    }
    CBS_ENTRY_END
}

int cbs_tagwise_disp(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p, double common,
                     double* tagwise) {
    CBS_NEED_GPU();
    if (G < 1 || !counts || !tagwise) return fail(CBS_EINVAL, "null pointer / bad size");
    if (int rc = check_design_size(n, design ? p : 1)) return rc;
    CBS_ENTRY_BEGIN
    OneBatch ob;
    ob.init(G, n, counts, design, p);
    double* d_phi = g_dev.get<double>("phi_hat", (size_t)G);
    if (ob.ds.h_array[0].p >= n) {      // estimateGLMTagwiseDisp.default: ncol(design) >= ncol(y) -> NA
        for (int64_t g = 0; g < G; ++g) tagwise[g] = std::nan("");
    } else {
        run_tagwise_disp(ob.d_cnt, G, 0, G, ob.d_logoff, ob.ds, ob.d_sel, std::vector<double>{common}, std::vector<char>{1}, 0, d_phi);
        CBS_CUDA(cudaMemcpy(tagwise, d_phi, (size_t)G * sizeof(double), cudaMemcpyDeviceToHost));
    }
    CBS_ENTRY_END
}

int cbs_mc_int(int64_t G, int32_t n, const int32_t* counts, const double* mu, const double* gamma, const double* phi,
               int32_t mc_mode, int32_t gene_subset_n, uint64_t seed, double* gamma_star, double* phi_star,
               int64_t* n_fallback) {
    CBS_NEED_GPU();
    if (G < 1 || n < 1 || !counts || !mu || !gamma || !phi || !gamma_star || !phi_star) return fail(CBS_EINVAL, "null pointer / bad size");
    if (gene_subset_n > 0 && gene_subset_n > G - 1) return fail(CBS_EINVAL, "cannot take a sample larger than the population when 'replace = FALSE'");
    CBS_ENTRY_BEGIN
    int32_t* d_cnt = upload(g_dev, "cnt", counts, (size_t)G * n);
    double* d_mu = upload(g_dev, "mu_hat", mu, (size_t)G * n);
    double* d_g = upload(g_dev, "gamma_hat", gamma, (size_t)G);
    double* d_p = upload(g_dev, "phi_hat", phi, (size_t)G);
    std::vector<std::vector<int>> bcols(1, std::vector<int>(n));
    std::iota(bcols[0].begin(), bcols[0].end(), 0);
    double* d_gs = g_dev.get<double>("gamma_star", (size_t)G);
    double* d_ps = g_dev.get<double>("phi_star_mat", (size_t)G);
    unsigned long long nfb = 0;
    run_mc_stage(d_cnt, G, 1, bcols, 0, G, d_mu, d_g, d_p, mc_mode, gene_subset_n > 0 ? gene_subset_n : 0, seed, d_gs, d_ps, &nfb, 0, nullptr);
    CBS_CUDA(cudaMemcpy(gamma_star, d_gs, (size_t)G * sizeof(double), cudaMemcpyDeviceToHost));
    CBS_CUDA(cudaMemcpy(phi_star, d_ps, (size_t)G * sizeof(double), cudaMemcpyDeviceToHost));
    if (n_fallback) *n_fallback = (int64_t)nfb;
    CBS_ENTRY_END
}

int cbs_match_quantiles(int64_t G, int32_t n, const int32_t* counts, const double* old_mu, const double* old_phi,
                        const double* new_mu, const double* new_phi, double* adjusted) {
    CBS_NEED_GPU();
    if (G < 0 || n < 0 || ((G > 0 && n > 0) && (!counts || !old_mu || !old_phi || !new_mu || !new_phi || !adjusted)))
        return fail(CBS_EINVAL, "null pointer / bad size");
    if (G == 0 || n == 0) return CBS_OK;
    CBS_ENTRY_BEGIN
    int32_t* d_cnt = upload(g_dev, "cnt", counts, (size_t)G * n);
    double* d_omu = upload(g_dev, "mu_hat", old_mu, (size_t)G * n);
    double* d_nmu = upload(g_dev, "mq_new_mu", new_mu, (size_t)G * n);
    double* d_ophi = upload(g_dev, "phi_hat", old_phi, (size_t)G);
    double* d_nphi = upload(g_dev, "phi_star", new_phi, (size_t)G);
    double* d_out = g_dev.get<double>("adjusted", (size_t)G * n);
    launch_match_quantiles_plain(d_cnt, (long)G, (int)n, d_omu, d_ophi, d_nmu, d_nphi, d_out, 0);
    CBS_CUDA(cudaMemcpy(adjusted, d_out, (size_t)G * n * sizeof(double), cudaMemcpyDeviceToHost));
    CBS_ENTRY_END
}

}  // extern "C"
