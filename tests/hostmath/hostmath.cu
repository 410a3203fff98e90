// TEST INFRASTRUCTURE: compiles the product's device math header (combat-seq_b200/csrc/nbmath.cuh) for the HOST so that
// the not-gpu test tier can check the exact routines the kernels run (continued fraction, pmf walk, quantile search)
// against the oracle and the golden vectors without a GPU.  Nothing in the product links this.
#include "../../combat-seq_b200/csrc/nbmath.cuh"
#include <cstdint>
using namespace cbs;
extern "C" {
void hm_nb_logpmf(int64_t n, const double* x, const double* size, const double* mu, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = dnbinom_mu<true>(x[i], size[i], mu[i]); }
void hm_nb_cdf(int64_t n, const double* x, const double* size, const double* mu, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = pnbinom_mu(x[i], size[i], mu[i]); }
void hm_nb_cdf_fast(int64_t n, const double* x, const double* size, const double* mu, double* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = nb_cdf_pmf(x[i], size[i], size[i] / (size[i] + mu[i]), mu[i] / (size[i] + mu[i])).F;
}
void hm_nb_pmf_fast(int64_t n, const double* x, const double* size, const double* mu, double* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = nb_cdf_pmf(x[i], size[i], size[i] / (size[i] + mu[i]), mu[i] / (size[i] + mu[i])).pmf;
}
void hm_nb_quantile(int64_t n, const double* p, const double* size, const double* mu, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = qnbinom_mu(p[i], size[i], mu[i]); }
void hm_nb_quantile_fast(int64_t n, const double* p, const double* size, const double* mu, double* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = nb_quantile_fast(p[i], size[i], size[i] / (size[i] + mu[i]));
}
void hm_match_quantiles(int64_t n, const double* x, const double* omu, const double* ophi, const double* nmu, const double* nphi, double* out, int fast) {
    static double inv[MQ_INV_N];
    for (int j = 1; j < MQ_INV_N; ++j) inv[j] = 1.0 / (double)j;
    for (int64_t i = 0; i < n; ++i) {
        if (fast == 2) out[i] = match_one_fast(x[i], omu[i], ophi[i], nmu[i], nphi[i], inv);      // with the small-count path
        else if (fast) out[i] = match_one_fast(x[i], omu[i], ophi[i], nmu[i], nphi[i]);
        else {
            if (x[i] <= 1) { out[i] = x[i]; continue; }
            const double tp = pnbinom_mu(x[i] - 1, 1.0 / ophi[i], omu[i]);
            out[i] = (fabs(tp - 1) < 1e-4) ? x[i] : 1 + qnbinom_mu(tp, 1.0 / nphi[i], nmu[i]);
        }
    }
}
}
#include "../../combat-seq_b200/csrc/fastmath.cuh"
extern "C" {
void hm_log_pos(int64_t n, const double* x, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = cbs::log_pos(x[i]); }
void hm_rcp_pos(int64_t n, const double* x, double* out) { for (int64_t i = 0; i < n; ++i) out[i] = cbs::rcp_pos(x[i]); }
}
#include "../../combat-seq_b200/csrc/rng.cuh"
extern "C" {
// the synthetic-count sampler (rng.cuh) for elements index0 .. index0 + n - 1 of a matrix
void hm_sample_nb(int64_t n, uint64_t seed, uint64_t index0, const double* mu, const double* phi, int32_t* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = cbs::sample_nb(seed, index0 + (uint64_t)i, mu[i], phi[i]);
}
void hm_philox(uint32_t* c, uint32_t k0, uint32_t k1) { cbs::philox4x32_10(c, k0, k1); }
}
