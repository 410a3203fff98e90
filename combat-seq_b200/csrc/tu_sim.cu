// tu_sim.cu — synthetic negative-binomial count matrices on the device (generator half).
//
// Replaces the count generation of the reference's simulation study (/root/reference/simulations/sim_DEpipe.R:98-113:
// polyester::simulate_experiment(reads_per_transcript, fold_changes, size), i.e. rnbinom(mu = reads * fold change, size))
// for benchmark-sized matrices: 6e7 draws in about a millisecond-scale launch instead of seconds of host RNG.
#include "engine.hpp"
#include "rng.cuh"

namespace cbs {

// out[g + s*G] ~ NB(mean = base_mu[g] * col_factor[s] * fold[g + level[s]*G], dispersion phi_col[s]); gene index fastest, so a
// warp writes 128 contiguous bytes.  fold may be null (no fold changes).  One Philox stream per element (rng.cuh).
__global__ void __launch_bounds__(256)
k_simulate_counts(long G, int N, const double* __restrict__ base_mu, const double* __restrict__ col_factor,
                  const int* __restrict__ level, const double* __restrict__ fold, const double* __restrict__ phi_col,
                  uint64_t seed, int32_t* __restrict__ out) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    if (g >= G) return;
    double mu = base_mu[g] * col_factor[s];
    if (fold) mu *= fold[g + (long)level[s] * G];
    out[g + (long)s * G] = sample_nb(seed, (uint64_t)g + (uint64_t)s * (uint64_t)G, mu, phi_col[s]);
}

void launch_simulate_counts(long G, int N, const double* base_mu, const double* col_factor, const int* level,
                            const double* fold, const double* phi_col, uint64_t seed, int32_t* out, cudaStream_t st) {
    if (G <= 0 || N <= 0) return;
    CBS_LAUNCH(k_simulate_counts, dim3((unsigned)((G + 255) / 256), (unsigned)N), 256, 0, st, G, N, base_mu, col_factor, level,
               fold, phi_col, seed, out);
}

}  // namespace cbs
