// tu_mq.cu — match_quantiles (a13) and the elementwise NB primitives.
//
// Replaces match_quantiles (/root/reference/helper_seq.R:46-64) with the batch-free parameters of
// /root/reference/ComBat_seq.R:174-178 fused in (mu_star is never stored), and R's dnbinom/pnbinom/qnbinom as called at
// helper_seq.R:28,53,58 (cbs_nb_logpmf / cbs_nb_cdf / cbs_nb_quantile).
#include "engine.hpp"
#include "nbmath.cuh"

namespace cbs {

// Warp = gene, lane = sample (samples visited in batch order, order[]): the nmath routines branch on the magnitudes of
// x, mu and size (continued-fraction length, pmf-walk length, pass-throughs), and those are similar along a gene's row
// and wildly different down a column.  The kernel is bound by FP64 arithmetic and divergence (ncu: DRAM traffic = the
// algorithmic 20 B per element, 0.6 % of HBM peak), so the strided loads and stores (one 32-byte sector per element,
// mostly L2 hits because the 4 warps of a block and the neighbouring blocks share the sectors) cost nothing.
// out is G0-row column-major; row of kept gene i is keep_idx[i] (or i when keep_idx == nullptr).
// grid: (ceil((g1-g0)/4), ceil(N/32)), block 128.
#ifndef CBS_MQ_MINBLOCKS
#define CBS_MQ_MINBLOCKS 6      // 80 registers: measured 25.3 -> 23.4 ms on C5 against 5 blocks of 96 registers
#endif
__global__ void __launch_bounds__(128, CBS_MQ_MINBLOCKS)
k_match_quantiles_warp(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ mu_hat,
                       const int* __restrict__ batch_idx, const int* __restrict__ order, int N,
                       const double* __restrict__ phi_hat, const double* __restrict__ gamma_star,
                       const double* __restrict__ phi_star, const int* __restrict__ keep_idx, long G0,
                       double* __restrict__ out, int si_begin, int si_end) {
    __shared__ double inv_s[MQ_INV_N];                          // 1/j for the small-count path (nbmath.cuh match_small)
    for (int j = threadIdx.x; j < MQ_INV_N; j += blockDim.x) inv_s[j] = j ? 1.0 / (double)j : 0.0;
    __syncthreads();
    const long g = g0 + (long)blockIdx.x * 4 + (threadIdx.x >> 5);
    const int si = si_begin + blockIdx.y * 32 + (threadIdx.x & 31);     // position in batch order; [si_begin, si_end): this launch's share
    if (g >= g1 || si >= si_end) return;
    const int s = order[si], b = batch_idx[s];
    const double x = (double)cnt[g + (long)s * G];
    const double old_mu = mu_hat[g + (long)s * G];
    const double new_mu = exp(log(old_mu) - gamma_star[g + (long)b * G]);     // ComBat_seq.R:176
    out[(keep_idx ? (long)keep_idx[g] : g) + (long)s * G0] = match_one_fast<true>(x, old_mu, phi_hat[g + (long)b * G], new_mu, phi_star[g], inv_s);
}

// generic per-batch form used by cbs_match_quantiles (explicit new_mu)
__global__ void __launch_bounds__(128)
k_match_quantiles_plain(const int32_t* __restrict__ cnt, long G, int n, const double* __restrict__ old_mu,
                        const double* __restrict__ old_phi, const double* __restrict__ new_mu,
                        const double* __restrict__ new_phi, double* __restrict__ out) {
    __shared__ double inv_s[MQ_INV_N];
    for (int j = threadIdx.x; j < MQ_INV_N; j += blockDim.x) inv_s[j] = j ? 1.0 / (double)j : 0.0;
    __syncthreads();
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (g >= G) return;
    const long k = g + (long)s * G;
    out[k] = match_one_fast<true>((double)cnt[k], old_mu[k], old_phi[g], new_mu[k], new_phi[g], inv_s);
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

void launch_match_quantiles(const int32_t* cnt, long G, long g0, long g1, const double* mu_hat, const int* batch_idx,
                            const int* order, int N, const double* phi_hat, const double* gamma_star,
                            const double* phi_star, const int* keep_idx, long G0, double* out, cudaStream_t st,
                            int si_begin, int si_end) {
    if (si_end < 0) si_end = N;
    if (g1 <= g0 || si_end <= si_begin) return;
    CBS_LAUNCH(k_match_quantiles_warp, dim3((unsigned)((g1 - g0 + 3) / 4), (unsigned)((si_end - si_begin + 31) / 32)), 128, 0, st, cnt, G,
               g0, g1, mu_hat, batch_idx, order, N, phi_hat, gamma_star, phi_star, keep_idx, G0, out, si_begin, si_end);
}

void launch_match_quantiles_plain(const int32_t* cnt, long G, int n, const double* old_mu, const double* old_phi,
                                  const double* new_mu, const double* new_phi, double* out, cudaStream_t st) {
    CBS_LAUNCH(k_match_quantiles_plain, dim3((unsigned)((G + 127) / 128), (unsigned)n), 128, 0, st, cnt, G, n,
               old_mu, old_phi, new_mu, new_phi, out);
}

void launch_nb_elementwise(int what, long n, const double* a, const double* size, const double* mu, double* out, cudaStream_t st) {
    CBS_LAUNCH(k_nb_elementwise, grid1(n, 128), 128, 0, st, what, n, a, size, mu, out);
}

}  // namespace cbs
