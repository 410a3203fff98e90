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
#define CBS_MQ_MINBLOCKS 8      // 64 registers; binned kernel on C5 (ms per step): 5 -> 9.75, 6 -> 8.9, 8 -> 8.15, 10 -> 8.13, 12 -> 8.33
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

// ---- binned form: a warp's 32 lanes get elements with the same amount of work ------------------------------------------
// The cost of an element is its count: the small-count path sums x terms of the old and about x * new_mu / old_mu terms of
// the new distribution, the continued fractions need O(sqrt x) terms.  With lane = sample the 32 counts of a warp are NB draws
// around the gene's mean: the longest lane ran about twice the average (ncu, round 2: half of the lanes active on average).
// Here the elements with x >= 2 (the others are pass-throughs, already in the pre-filled output) are put into 256 bins by a
// counting sort on the device (histogram, scan, scatter of element indices: two streaming passes over the counts), the key
// being the quarter-octave of x (1 + new_mu / old_mu), small-count and continued-fraction elements in separate halves, and
// k_match_quantiles_binned gives consecutive index entries to consecutive lanes.  One thread per element and the same
// routine as before: the results are the same bits (test_kernel_variants_give_the_same_bits, "mq_bins").
constexpr int MQ_BINS = 256;
__device__ __forceinline__ int mq_bin_key(int x, double gamma_star) {
    // (single precision: the key only balances the work; both passes compute it with the same instructions)
    float ratio = __expf(-(float)gamma_star);                          // new_mu / old_mu  (ComBat_seq.R:176)
    ratio = fminf(fmaxf(ratio, 0.f), 64.f);                            // (NaN -> 64)
    const float w = (float)x * (1.f + ratio);                          // >= 2
    int k = ((__float_as_int(w) >> 21) & 0x3FF) - (127 << 2);          // exponent and the top two mantissa bits
    k = min(max(k, 0), MQ_BINS / 2 - 1);
    return k + (x > (int)MQ_SMALL_X ? MQ_BINS / 2 : 0);
}

// grid (ceil(ng / 256), ceil(nsi / MQ_BIN_SPB)), block 256: a block takes 256 genes x MQ_BIN_SPB sample positions (one sample
// per block made the two passes block-turnover bound: 235 000 blocks of a microsecond each per pass on C5).
// pass 0: hist[key] += count.  pass 1: ids[cursor[key]++] = element index (gene - g0) + (sample position - si_begin) * ng,
// the block reserving one range per bin.
constexpr int MQ_BIN_SPB = 8;
template <int PASS>
__global__ void __launch_bounds__(256)
k_mq_bin(const int32_t* __restrict__ cnt, long G, long g0, long g1, const int* __restrict__ batch_idx,
         const int* __restrict__ order, int si_begin, int nsi, const double* __restrict__ gamma_star, int* __restrict__ hist,
         int* __restrict__ cursor, int* __restrict__ ids) {
    __shared__ int cnt_s[MQ_BINS];
    cnt_s[threadIdx.x] = 0;
    __syncthreads();
    const long ng = g1 - g0, gl = (long)blockIdx.x * 256 + threadIdx.x;
    const int sl0 = blockIdx.y * MQ_BIN_SPB, sl1 = min(nsi, sl0 + MQ_BIN_SPB);
    int keys[MQ_BIN_SPB];
#pragma unroll
    for (int u = 0; u < MQ_BIN_SPB; ++u) {
        keys[u] = -1;
        if (gl < ng && sl0 + u < sl1) {
            const int s = order[si_begin + sl0 + u];
            const long g = g0 + gl;
            const int x = cnt[g + (long)s * G];
            if (x >= 2) {
                keys[u] = mq_bin_key(x, gamma_star[g + (long)batch_idx[s] * G]);
                atomicAdd(&cnt_s[keys[u]], 1);
            }
        }
    }
    __syncthreads();
    const int c = cnt_s[threadIdx.x];
    if (PASS == 0) { if (c) atomicAdd(&hist[threadIdx.x], c); return; }
    __syncthreads();
    cnt_s[threadIdx.x] = c ? atomicAdd(&cursor[threadIdx.x], c) : 0;       // now the block's write position in every bin
    __syncthreads();
#pragma unroll
    for (int u = 0; u < MQ_BIN_SPB; ++u)
        if (keys[u] >= 0) ids[atomicAdd(&cnt_s[keys[u]], 1)] = (int)(gl + (long)(sl0 + u) * ng);
}
// cursor = exclusive scan of hist; cursor[MQ_BINS] = number of elements
__global__ void __launch_bounds__(MQ_BINS) k_mq_bin_scan(const int* __restrict__ hist, int* __restrict__ cursor) {
    __shared__ int sc[MQ_BINS];
    const int v = hist[threadIdx.x];
    sc[threadIdx.x] = v;
    __syncthreads();
    for (int o = 1; o < MQ_BINS; o <<= 1) {
        const int a = (int)threadIdx.x >= o ? sc[threadIdx.x - o] : 0;
        __syncthreads();
        sc[threadIdx.x] += a;
        __syncthreads();
    }
    cursor[threadIdx.x] = sc[threadIdx.x] - v;
    if (threadIdx.x == MQ_BINS - 1) cursor[MQ_BINS] = sc[threadIdx.x];
}

__global__ void __launch_bounds__(128, CBS_MQ_MINBLOCKS)
k_match_quantiles_binned(const int32_t* __restrict__ cnt, long G, long g0, long ng, const double* __restrict__ mu_hat,
                         const int* __restrict__ batch_idx, const int* __restrict__ order, int si_begin,
                         const double* __restrict__ phi_hat, const double* __restrict__ gamma_star,
                         const double* __restrict__ phi_star, const int* __restrict__ keep_idx, long G0,
                         double* __restrict__ out, const int* __restrict__ ids, const int* __restrict__ total,
                         const double* __restrict__ inv_tab) {
    __shared__ double inv_s[MQ_INV_N];
    const long k0 = (long)blockIdx.x * blockDim.x;
    const int n_el = *total;
    if (k0 >= n_el) return;                                     // (the grid covers the chunk; the list is shorter)
    for (int j = threadIdx.x; j < MQ_INV_N; j += blockDim.x) inv_s[j] = inv_tab[j];
    __syncthreads();
    const long k = k0 + threadIdx.x;
    if (k >= n_el) return;
    const int e = ids[k];
    const long g = g0 + e % ng;
    const int s = order[si_begin + (int)(e / ng)], b = batch_idx[s];
    const double x = (double)cnt[g + (long)s * G];
    const double old_mu = mu_hat[g + (long)s * G];
    const double new_mu = exp(log(old_mu) - gamma_star[g + (long)b * G]);     // ComBat_seq.R:176
    out[(keep_idx ? (long)keep_idx[g] : g) + (long)s * G0] = match_one_fast<true>(x, old_mu, phi_hat[g + (long)b * G], new_mu, phi_star[g], inv_s);
}
__global__ void k_mq_inv_table(double* __restrict__ t) {
    const int j = blockIdx.x * blockDim.x + threadIdx.x;
    if (j < MQ_INV_N) t[j] = j ? 1.0 / (double)j : 0.0;
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
    const long ng = g1 - g0, nsi = si_end - si_begin;
    // binned form: needs the pre-filled output for the pass-through elements (the pipeline's k_counts_to_f64) and element
    // indices that fit an int
    if (g_tune.mq_bins && ng * nsi < 2000000000L && nsi <= 65535) {
        int* hist = g_dev.get<int>("mq_hist", 2 * MQ_BINS + 2);
        int* cursor = hist + MQ_BINS;
        int* ids = g_dev.get<int>("mq_ids", (size_t)(ng * nsi));
        double* inv_tab = g_dev.get<double>("mq_inv", MQ_INV_N);
        CBS_CUDA(cudaMemsetAsync(hist, 0, (2 * MQ_BINS + 2) * sizeof(int), st));
        CBS_LAUNCH(k_mq_inv_table, dim3(MQ_INV_N / 256), 256, 0, st, inv_tab);
        const dim3 bgrid((unsigned)((ng + 255) / 256), (unsigned)((nsi + MQ_BIN_SPB - 1) / MQ_BIN_SPB));
        CBS_LAUNCH(k_mq_bin<0>, bgrid, 256, 0, st, cnt, G, g0, g1, batch_idx, order, si_begin, (int)nsi, gamma_star, hist, cursor, ids);
        CBS_LAUNCH(k_mq_bin_scan, dim3(1), MQ_BINS, 0, st, (const int*)hist, cursor);
        CBS_LAUNCH(k_mq_bin<1>, bgrid, 256, 0, st, cnt, G, g0, g1, batch_idx, order, si_begin, (int)nsi, gamma_star, hist, cursor, ids);
        // after the scatter cursor[k] is the END of bin k: cursor[MQ_BINS - 1] = cursor[MQ_BINS] = number of elements
        CBS_LAUNCH(k_match_quantiles_binned, dim3((unsigned)((ng * nsi + 127) / 128)), 128, 0, st, cnt, G, g0, ng, mu_hat, batch_idx,
                   order, si_begin, phi_hat, gamma_star, phi_star, keep_idx, G0, out, (const int*)ids, (const int*)(cursor + MQ_BINS),
                   (const double*)inv_tab);
        return;
    }
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
