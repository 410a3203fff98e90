// mc.cuh — Monte-Carlo empirical-Bayes integration, device side.
//
// Replaces monte_carlo_int_NB (/root/reference/helper_seq.R:8-42): for every target gene i of a batch the
// posterior means of (gamma, phi) over all other genes j, weighted by  LH_ij = prod_s dnbinom(x_is; mu_js, 1/phi_j).
// The reference evaluates G^2 * n_b scalar dnbinom calls and multiplies them in linear space.  Here
//     log LH_ij = d_i + c_j + sum_s x_is * A_js + sum_s lgamma(x_is + r_j),        r_j = 1/phi_j
//     A_js = log(mu_js / (r_j + mu_js)),  c_j = r_j * sum_s log(r_j / (r_j + mu_js)) - n_b * lgamma(r_j),
//     d_i  = -sum_s lgamma(x_is + 1)
// which is the same NB log-pmf split into a separable part (per source j, per target i), a bilinear part
// (a dense contraction over the samples) and the one non-separable term lgamma(x_is + r_j).
// Weight modes: CBS_MC_REFERENCE reproduces the reference's linear-space product (w = exp(log LH) with no shift, so
// weights underflow to 0 where R's product does and sum(LH)==0 falls back to the gene's own estimate,
// helper_seq.R:29-31); CBS_MC_STABLE uses an online log-sum-exp.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cmath>

namespace cbs {

constexpr double MC_LOG_ZERO = -1e30;   // stands for log(0) in A: x*A stays finite, exp() of the sum is exactly 0

// lgamma(z) for z >= 16 by the Stirling series (truncation error < 2e-16 at z = 16)
__device__ __forceinline__ double lgamma_stirling(double z) {
    const double zi = 1.0 / z, z2 = zi * zi;
    const double ser = zi * (0.083333333333333333 + z2 * (-0.0027777777777777778 + z2 * (0.00079365079365079365 +
                       z2 * (-0.00059523809523809524 + z2 * 0.00084175084175084175))));
    return (z - 0.5) * log(z) - z + 0.918938533204672741780329736406 + ser;
}
__device__ __forceinline__ double lgamma_pos(double z) { return z >= 16.0 ? lgamma_stirling(z) : lgamma(z); }

// per-source quantities for one batch.  cols[t] = global column of the batch's t-th sample.
// A is [t*G + j]; c, r, lgr are [j].
__global__ void k_mc_prep_src(const double* __restrict__ mu_hat, long G, int n, const int* __restrict__ cols,
                              const double* __restrict__ phi, double* __restrict__ A, double* __restrict__ cj,
                              double* __restrict__ rj, double* __restrict__ lgr) {
    const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= G) return;
    const double r = 1.0 / phi[j];
    double c = 0;
    for (int t = 0; t < n; ++t) {
        const double mu = mu_hat[j + (long)cols[t] * G];
        double a, l;
        if (mu == 0) { a = MC_LOG_ZERO; l = 0; }
        else {
            a = (mu < r) ? log(mu / (r + mu)) : log1p(-r / (r + mu));
            l = (r < mu) ? log(r / (r + mu)) : log1p(-mu / (r + mu));
        }
        A[(long)t * G + j] = a;
        c += r * l;
    }
    const double lg = lgamma(r);
    cj[j] = c - n * lg;
    rj[j] = r;
    lgr[j] = lg;
}

// per-target constant d_i = -sum_s lgamma(x_is + 1)
__global__ void k_mc_prep_tgt(const int32_t* __restrict__ cnt, long G, int n, const int* __restrict__ cols,
                              double* __restrict__ di) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= G) return;
    double d = 0;
    for (int t = 0; t < n; ++t) d -= lgamma_pos((double)cnt[i + (long)cols[t] * G] + 1.0);
    di[i] = d;
}

struct McAcc {          // running (max, sum w, sum w*gamma, sum w*phi); max stays 0 in reference mode
    double m, sw, sg, sp;
};
__device__ __forceinline__ void mc_add(McAcc& a, double L, double g, double p, int stable) {
    if (!(L == L)) return;                                   // LH[is.nan(LH)] = 0   helper_seq.R:29
    if (!stable) {
        const double w = exp(L);
        a.sw += w; a.sg += w * g; a.sp += w * p;
    } else {
        if (L == -INFINITY) return;
        if (L > a.m) { const double sc = exp(a.m - L); a.sw *= sc; a.sg *= sc; a.sp *= sc; a.m = L; }
        const double w = exp(L - a.m);
        a.sw += w; a.sg += w * g; a.sp += w * p;
    }
}
__device__ __forceinline__ void mc_merge(McAcc& a, const McAcc& b, int stable) {
    if (!stable) { a.sw += b.sw; a.sg += b.sg; a.sp += b.sp; return; }
    const double M = fmax(a.m, b.m);
    if (M == -INFINITY) return;
    const double sa = exp(a.m - M), sb = exp(b.m - M);
    a.sw = a.sw * sa + b.sw * sb; a.sg = a.sg * sa + b.sg * sb; a.sp = a.sp * sa + b.sp * sb; a.m = M;
}
__device__ __forceinline__ McAcc mc_shfl_xor(const McAcc& a, int o) {
    McAcc b;
    b.m = __shfl_xor_sync(0xffffffffu, a.m, o); b.sw = __shfl_xor_sync(0xffffffffu, a.sw, o);
    b.sg = __shfl_xor_sync(0xffffffffu, a.sg, o); b.sp = __shfl_xor_sync(0xffffffffu, a.sp, o);
    return b;
}

// Exact pairwise kernel: every (i, j, s) term is evaluated, including lgamma(x_is + r_j).
// One CTA (256 threads) owns MC_TI consecutive targets; threads stride over the sources.
// subset_n > 0: each target uses only `subset_n` sources drawn without replacement from the other G-1 genes
// (helper_seq.R:17-21) through a keyed Feistel permutation of [0, G-1).
constexpr int MC_TI = 8;

__device__ __forceinline__ uint32_t mc_mix(uint32_t x, uint32_t k) {
    x ^= k; x *= 0x9E3779B1u; x ^= x >> 15; x *= 0x85EBCA77u; x ^= x >> 13; x *= 0xC2B2AE3Du; x ^= x >> 16;
    return x;
}
// pseudo-random permutation of [0, m): 4-round Feistel on 2*hb bits with cycle walking
__device__ __forceinline__ uint32_t mc_perm(uint32_t idx, uint32_t m, int hb, uint64_t key) {
    const uint32_t mask = (1u << hb) - 1u;
    uint32_t v = idx;
    do {
        uint32_t l = v >> hb, r = v & mask;
#pragma unroll
        for (int rd = 0; rd < 4; ++rd) {
            const uint32_t f = mc_mix(r, (uint32_t)(key >> (16 * rd)) + 0x632BE5ABu * (rd + 1) + (uint32_t)(key >> 32)) & mask;
            const uint32_t nl = r; r = l ^ f; l = nl;
        }
        v = (l << hb) | r;
    } while (v >= m);
    return v;
}

__global__ void __launch_bounds__(256)
k_mc_direct(const int32_t* __restrict__ cnt, long G, int n, const int* __restrict__ cols, long i0, long i1,
            const double* __restrict__ A, const double* __restrict__ cj, const double* __restrict__ rj,
            const double* __restrict__ lgr, const double* __restrict__ di, const double* __restrict__ gamma,
            const double* __restrict__ phi, int stable, int subset_n, uint64_t seed, int xs_shared,
            double* __restrict__ gamma_star, double* __restrict__ phi_star, unsigned long long* __restrict__ n_fallback) {
    extern __shared__ double xs[];                    // [MC_TI][n] when xs_shared (64 n bytes), else unused
    __shared__ McAcc red[8][MC_TI];
    const long ib = i0 + (long)blockIdx.x * MC_TI;
    if (xs_shared) {
        for (int k = threadIdx.x; k < MC_TI * n; k += blockDim.x) {
            const int ti = k / n, t = k - ti * n;
            const long i = ib + ti;
            xs[k] = (i < i1) ? (double)cnt[i + (long)cols[t] * G] : 0.0;
        }
    }
    __syncthreads();
    // count of target ti at the batch's t-th sample (batches too wide for shared memory: straight from L1/L2)
    auto xat = [&](int ti, int t) -> double {
        if (xs_shared) return xs[ti * n + t];
        const long i = ib + ti;
        return (i < i1) ? (double)cnt[i + (long)cols[t] * G] : 0.0;
    };
    McAcc acc[MC_TI];
#pragma unroll
    for (int ti = 0; ti < MC_TI; ++ti) { acc[ti].m = stable ? -INFINITY : 0.0; acc[ti].sw = acc[ti].sg = acc[ti].sp = 0.0; }

    if (subset_n <= 0) {
        for (long j = threadIdx.x; j < G; j += blockDim.x) {
            const double r = rj[j], lg0 = lgr[j], c = cj[j];
            double L[MC_TI];
#pragma unroll
            for (int ti = 0; ti < MC_TI; ++ti) L[ti] = 0.0;
            for (int t = 0; t < n; ++t) {
                const double a = A[(long)t * G + j];
#pragma unroll
                for (int ti = 0; ti < MC_TI; ++ti) {
                    const double x = xat(ti, t);
                    L[ti] += x * a + (x == 0.0 ? lg0 : lgamma_pos(x + r));
                }
            }
            const double gj = gamma[j], pj = phi[j];
#pragma unroll
            for (int ti = 0; ti < MC_TI; ++ti) {
                const long i = ib + ti;
                if (i < i1 && j != i) mc_add(acc[ti], L[ti] + c + di[i], gj, pj, stable);
            }
        }
    } else {
        int hb = 1;
        while ((1ull << (2 * hb)) < (unsigned long long)(G - 1)) ++hb;
        for (int ti = 0; ti < MC_TI; ++ti) {
            const long i = ib + ti;
            if (i >= i1) break;
            const uint64_t key = seed * 0x9E3779B97F4A7C15ull + (uint64_t)i * 0xD1B54A32D192ED03ull + 0x2545F4914F6CDD1Dull;
            for (int k = threadIdx.x; k < subset_n; k += blockDim.x) {
                long j = mc_perm((uint32_t)k, (uint32_t)(G - 1), hb, key);
                if (j >= i) ++j;                             // index into the G-1 genes other than i (mu[-i,])
                const double r = rj[j], lg0 = lgr[j];
                double L = 0.0;
                for (int t = 0; t < n; ++t) {
                    const double x = xat(ti, t);
                    L += x * A[(long)t * G + j] + (x == 0.0 ? lg0 : lgamma_pos(x + r));
                }
                mc_add(acc[ti], L + cj[j] + di[i], gamma[j], phi[j], stable);
            }
        }
    }
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
#pragma unroll
    for (int ti = 0; ti < MC_TI; ++ti) {
        for (int o = 16; o > 0; o >>= 1) { const McAcc b = mc_shfl_xor(acc[ti], o); mc_merge(acc[ti], b, stable); }
        if (lane == 0) red[wid][ti] = acc[ti];
    }
    __syncthreads();
    if (threadIdx.x < MC_TI) {
        const int ti = threadIdx.x;
        const long i = ib + ti;
        if (i < i1) {
            McAcc a = red[0][ti];
            for (int w = 1; w < 8; ++w) mc_merge(a, red[w][ti], stable);
            if (a.sw == 0 || !(a.sw == a.sw)) {              // sum(LH)==0 | is.na(sum(LH))   helper_seq.R:30-31
                gamma_star[i] = gamma[i]; phi_star[i] = phi[i];
                atomicAdd(n_fallback, 1ull);
            } else {
                gamma_star[i] = a.sg / a.sw; phi_star[i] = a.sp / a.sw;    // helper_seq.R:33
            }
        }
    }
}

}  // namespace cbs
