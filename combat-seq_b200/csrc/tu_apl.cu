// tu_apl.cu — dispersion estimation (a5/a6): Cox-Reid adjusted profile likelihood kernels and their stage drivers.
//
// Replaces, per batch, estimateGLMCommonDisp (/root/reference/ComBat_seq.R:100-107) and estimateGLMTagwiseDisp
// (/root/reference/ComBat_seq.R:110-119) of edgeR 3.22: APL evaluation + Brent's optimize() on delta = phi^(1/4),
// then the 11-point grid + FMM spline maximiser.
#include "engine.hpp"

namespace cbs {

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
// (APL_CHUNK = 128 genes, one thread each, fixed tree order).
__global__ void __launch_bounds__(APL_CHUNK)
k_chunk_sum(const double* __restrict__ apl, long G, const unsigned char* __restrict__ sel, AplLaunch L,
            long chunk0, double* __restrict__ partial, long nchunk_total) {
    const long chunk = chunk0 + blockIdx.x;
    const int e = blockIdx.y;
    const long g = chunk * APL_CHUNK + threadIdx.x;
    const unsigned char* s = sel ? sel + (long)L.design[e] * G : nullptr;
    __shared__ double sh[APL_CHUNK];
    sh[threadIdx.x] = (g < G && (!s || s[g])) ? apl[(long)e * G + g] : 0.0;
    __syncthreads();
    for (int o = APL_CHUNK / 2; o > 0; o >>= 1) {
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

__global__ void k_fill_apl(double* __restrict__ p, long n, double v) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// APL of genes [g0,g1) for every job -> d_apl[job*G + g]
void launch_apl(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignDev* designs,
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
        if (g_tune.apl_fast && !phi_gene) {
            double* t = g_dev.get<double>("lgr", (size_t)APL_MAX_EVALS * LGFACT_N);
            CBS_LAUNCH(k_lgr_init, dim3(LGFACT_N / 256, (unsigned)L.n_evals), 256, 0, st, L, t);
            lgr = t;
        }
        CBS_LAUNCH(k_apl, grid, 128, 0, st, cnt, G, g0, g1, logoff, designs, sel,
                   phi_gene ? phi_gene + e0 * G : nullptr, L, d_apl + e0 * G, g_tune.apl_fast ? (const double*)lgfact : (const double*)nullptr, lgr);
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
            CBS_LAUNCH(k_chunk_sum, grid, APL_CHUNK, 0, st, d_apl + e0 * G, G, sel, L, c0, d_part + e0 * nchunk, nchunk);
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
void run_common_disp(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignSet& ds,
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
void run_tagwise_disp(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignSet& ds,
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
            CBS_LAUNCH(k_fill_apl, grid1(g1 - g0, 256), 256, 0, st, d_phi_hat + (size_t)b * G + g0, g1 - g0, common[b]);
        }
    }
}

}  // namespace cbs
