// mc_gemm.cuh — Monte-Carlo empirical-Bayes integration as one FP64 tensor-core contraction with a fused
// weighted-sum epilogue (the "flash-attention with an NB score" form of monte_carlo_int_NB).
//
// Replaces monte_carlo_int_NB (/root/reference/helper_seq.R:8-42) for gene.subset.n = NULL.  For target gene i and
// source gene j of one batch (n samples, r_j = 1/phi_j):
//
//   log LH_ij = sum_s log dnbinom(x_is; mu_js, r_j)                                              helper_seq.R:28
//             = sum_s x_is * A_js                      A_js = log(mu_js / (r_j + mu_js))          (bilinear)
//             + sum_s [lgamma(x_is + r_j) - lgamma(x_is + 1)]                                     (not separable)
//             + c_j                                    c_j  = r_j sum_s log(r_j/(r_j+mu_js)) - n lgamma(r_j)
//
// The only non-separable term depends on j through the scalar r_j alone, so for every target i the function
//   F_i(r) = sum_{s : x_is >= c} [lgamma(x_is + r) - lgamma(x_is + 1)]
// is a smooth function of one variable.  In u = log(r + c) its singularities sit at Im u = +-pi whatever x >= c
// is, so a Chebyshev series in u over the batch's range of r converges geometrically (about 24-30 terms for
// 1e-13 over a 4096-fold range of r): F_i(r_j) = sum_k C_ik T_k(t_j).  Counts below the shift c are handled
// exactly with a one-hot table, lgamma(v + r_j) - lgamma(v + 1) for v = 0..c-1 times the number of samples of
// gene i that equal v.  All of it is one inner product of length KT = n + Kc + c + 1:
//
//   P_i = [ x_i1 .. x_in | C_i0 .. C_i,Kc-1 | #{x_is = 0} .. #{x_is = c-1} | 1   ]
//   Q_j = [ A_j1 .. A_jn | T_0(t_j) ..      | lgamma(v + r_j) - lgamma(v+1) | c_j ]          log LH_ij = P_i . Q_j
//
// k_mc_gemm evaluates P Q^T tile by tile on the FP64 tensor pipe (mma.sync m8n8k4 f64 -> DMMA.8x8x4; tcgen05 has
// no FP64 path), streams the tiles into shared memory with 1-D TMA bulk copies (cp.async.bulk + mbarrier), and
// folds the weights w = exp(log LH) into running sums (sum w, sum w gamma_j, sum w phi_j) in registers, so the
// G x G likelihood matrix never exists.  Per (i, j) pair that is 2*KT flop instead of n lgamma evaluations.
// Truncation error of the series is monitored (tail coefficients) and reported; see DESIGN.md for the tolerance.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cmath>
#include "mc.cuh"
#include "mc_shape.hpp"

namespace cbs {

#ifndef MCG_SUB_N                    // (development builds override the ring geometry: tools/build_variants.sh)
#define MCG_SUB_N 2
#endif
#ifndef MCG_STAGES_N
#define MCG_STAGES_N 4
#endif
constexpr int MCG_SUB = MCG_SUB_N;   // k-chunks per pipeline stage (16 doubles of K per hand-over)
constexpr int MCG_STAGES = MCG_STAGES_N;
#ifndef MCG_GPBUF_N
#define MCG_GPBUF_N 4
#endif
constexpr int MCG_GPBUF = MCG_GPBUF_N;       // buffers of the per-source-tile (gamma, phi) pairs
constexpr int MCG_MAX_KC = 64;       // max Chebyshev terms
constexpr int MCG_MT = 2;            // 8-row DMMA tiles per warp along the targets (warp tile = 8*MCG_MT x 8*MCG_NT)
#ifndef MCG_NT_N                     // (development builds: 4 = 64 x 64 CTA tiles, 16 x 32 warp tiles, three CTAs per SM)
#define MCG_NT_N 8
#endif
constexpr int MCG_NT = MCG_NT_N;     // 8-column DMMA tiles per warp along the sources
constexpr int MCG_TJ = 2 * 8 * MCG_NT;   // source columns per CTA tile (two warps side by side): 128
constexpr int MCG_CTAS = MCG_NT == 8 ? 2 : 3;   // CTAs per SM the kernel is sized for
constexpr int MCG_THREADS = 32 * (MCG_TI / (8 * MCG_MT)) * 2;     // warps: (64 / warp rows) x 2 = 8

// lgamma(x + r) - lgamma(x + 1) for integer-valued x >= 0, r > 0, without cancellation for large x
__device__ __forceinline__ double lgamma_diff(double x, double r) {
    if (x < 15.0) return lgamma(x + r) - lgamma(x + 1.0);
    const double z1 = x + r, z2 = x + 1.0, d = r - 1.0;
    const double i1 = 1.0 / z1, i2 = 1.0 / z2, q1 = i1 * i1, q2 = i2 * i2;
    const double s1 = i1 * (0.083333333333333333 + q1 * (-0.0027777777777777778 + q1 * (0.00079365079365079365 +
                      q1 * (-0.00059523809523809524 + q1 * 0.00084175084175084175))));
    const double s2 = i2 * (0.083333333333333333 + q2 * (-0.0027777777777777778 + q2 * (0.00079365079365079365 +
                      q2 * (-0.00059523809523809524 + q2 * 0.00084175084175084175))));
    return (x + 0.5) * log1p(d / z2) + d * (log(z1) - 1.0) + (s1 - s2);
}

// ---- range of r = 1/phi over the genes, per batch: out[2b] = min phi, out[2b+1] = max phi ----------------
__global__ void __launch_bounds__(256)
k_mcg_phi_range(const double* __restrict__ phi, long G, double* __restrict__ out) {
    const int b = blockIdx.x;
    double lo = INFINITY, hi = -INFINITY;
    for (long g = threadIdx.x; g < G; g += blockDim.x) {
        const double v = phi[g + (long)b * G];
        if (v == v) { lo = fmin(lo, v); hi = fmax(hi, v); }
    }
    __shared__ double slo[256], shi[256];
    slo[threadIdx.x] = lo; shi[threadIdx.x] = hi;
    __syncthreads();
    for (int o = 128; o > 0; o >>= 1) {
        if ((int)threadIdx.x < o) { slo[threadIdx.x] = fmin(slo[threadIdx.x], slo[threadIdx.x + o]); shi[threadIdx.x] = fmax(shi[threadIdx.x], shi[threadIdx.x + o]); }
        __syncthreads();
    }
    if (threadIdx.x == 0) { out[2 * b] = slo[0]; out[2 * b + 1] = shi[0]; }
}

// ---- source side: row j of Q -------------------------------------------------------------------------------
// cols[t] = global column of the batch's t-th sample; phi, gamma = this batch's columns of phi_hat / gamma_hat.
// Row j of P and Q is gene perm[j] (the batch's abundance order, see k_mcg_sort_key): genes with similar expression sit
// in neighbouring rows, so the pairs that carry weight form a band around the diagonal.
// Rows j >= G are padding: zeros with c_j = MC_LOG_ZERO, so their weight is exactly 0.
// Two kernels.  k_mcg_prep_src: grid (rows / 128, KTP / 8), one thread per (row, k-chunk): the thread computes the chunk's eight
// columns and writes them as one 64-byte piece (consecutive threads = consecutive rows: whole lines); the chunks of the sample
// part also leave their partial sums of r l_s (for c_j) and of A_s^2 (for the screening bound) in part[].  One thread per row
// walking all KTP columns (round 1) gave 469 blocks of serial, 8-byte-strided stores per batch: a fifth of a wave, as long at
// 8 GPUs as at one.  k_mcg_prep_src_fin: one thread per row adds the partials in chunk order (a fixed order: the same bits for
// every GPU count), writes c_j and the row's summary for the screening pass, qsum[j] = (sqrt(sum A^2), max_v |lg_v|, |c_j|).
__device__ __forceinline__ double mcg_src_tail_col(int k, int n, int Kc, int KT, double r, double tt) {
    // column k >= n of row j, except c_j: Chebyshev T_{k-n}(tt), then lgamma(v + r) - lgamma(v + 1) for v < MCG_SHIFT, then zeros
    if (k < n + Kc) {
        double t0 = 1.0, t1 = tt;
        for (int i = 0; i < k - n; ++i) { const double t2 = 2.0 * tt * t1 - t0; t0 = t1; t1 = t2; }
        return t0;
    }
    if (k < n + Kc + MCG_SHIFT) { const int v = k - n - Kc; return lgamma(v + r) - lgamma(v + 1.0); }
    return 0.0;                                              // c_j (k = KT - 1) is written by k_mcg_prep_src_fin; padding
}
__global__ void __launch_bounds__(128)
k_mcg_prep_src(const double* __restrict__ mu_hat, long G, const int* __restrict__ cols, const double* __restrict__ phi,
               const int* __restrict__ perm, McgShape sh, double* __restrict__ Q, double2* __restrict__ part) {
    const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int kc = blockIdx.y;
    if (j >= sh.rows) return;
    const int n = sh.n, Kc = sh.Kc;
    double v[MCG_KC];
    double csum = 0, s2 = 0;
    if (j >= G) {
#pragma unroll
        for (int e = 0; e < MCG_KC; ++e) v[e] = 0.0;
    } else {
        const long gj = perm[j];
        const double r = 1.0 / phi[gj];
        double tt = 0.0;
        if (kc * MCG_KC + MCG_KC > n) {
            const double u = log(r + (double)MCG_SHIFT);
            tt = (sh.uhi > sh.ulo) ? (2.0 * u - (sh.ulo + sh.uhi)) / (sh.uhi - sh.ulo) : 0.0;
            tt = fmin(1.0, fmax(-1.0, tt));
        }
        double mu[MCG_KC];
#pragma unroll
        for (int e = 0; e < MCG_KC; ++e) { const int t = kc * MCG_KC + e; mu[e] = t < n ? mu_hat[gj + (long)cols[t] * G] : 1.0; }
#pragma unroll
        for (int e = 0; e < MCG_KC; ++e) {
            const int k = kc * MCG_KC + e;
            if (k < n) {
                double a, l;
                if (mu[e] == 0) { a = MC_LOG_ZERO; l = 0; }
                else {
                    a = (mu[e] < r) ? log(mu[e] / (r + mu[e])) : log1p(-r / (r + mu[e]));
                    l = (r < mu[e]) ? log(r / (r + mu[e])) : log1p(-mu[e] / (r + mu[e]));
                }
                v[e] = a;
                csum += r * l;
                s2 += a * a;
            } else v[e] = mcg_src_tail_col(k, n, Kc, sh.KT, r, tt);
        }
    }
    // the chunk's 64 bytes: column kk of the chunk sits at (kk ^ swz), swz as in mcg_index
    const int swz = (int)((j >> 1) & 1) << 2;
    static_assert(MCG_KC == 8, "chunk layout");
    double2* dst = reinterpret_cast<double2*>(Q + ((size_t)kc * sh.rows + j) * MCG_KC);
    const double2 a0 = make_double2(v[0], v[1]), a1 = make_double2(v[2], v[3]), b0 = make_double2(v[4], v[5]), b1 = make_double2(v[6], v[7]);
    dst[0] = swz ? b0 : a0; dst[1] = swz ? b1 : a1; dst[2] = swz ? a0 : b0; dst[3] = swz ? a1 : b1;
    if (kc * MCG_KC < n) part[(size_t)kc * sh.rows + j] = make_double2(csum, s2);
}
__global__ void __launch_bounds__(128)
k_mcg_prep_src_fin(long G, const double* __restrict__ phi, const int* __restrict__ perm, McgShape sh,
                   const double2* __restrict__ part, double* __restrict__ Q, float4* __restrict__ qsum) {
    const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j >= sh.rows) return;
    if (j >= G) {
        Q[mcg_index(sh.rows, j, sh.KT - 1)] = MC_LOG_ZERO;
        qsum[j] = make_float4(0.f, 0.f, -1.f, 0.f);          // (z < 0: an impossible source, left out of the tile maxima)
        return;
    }
    const double r = 1.0 / phi[perm[j]];
    double c = 0, s2 = 0;
    for (int kc = 0; kc * MCG_KC < sh.n; ++kc) { const double2 p = part[(size_t)kc * sh.rows + j]; c += p.x; s2 += p.y; }
    const double cj = c - sh.n * lgamma(r);
    Q[mcg_index(sh.rows, j, sh.KT - 1)] = cj;
    double mx = 0;
    for (int v = 0; v < MCG_SHIFT; ++v) mx = fmax(mx, fabs(lgamma(v + r) - lgamma(v + 1.0)));
    qsum[j] = make_float4((float)sqrt(s2), (float)mx, fabsf((float)cj), 0.f);
}

// ---- target side, step 0: the batch's counts of this rank's target rows in abundance order, xs[t * own_rows + local row] ----
// (one gathered read per count; the kernels below then read coalesced)
__global__ void __launch_bounds__(256)
k_mcg_gather_counts(const int32_t* __restrict__ cnt, long G, const int* __restrict__ cols, const int* __restrict__ perm,
                    McgShape sh, McgOwn own, long own_rows, int32_t* __restrict__ xs) {       // grid (own_rows/256, n)
    const long lr = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int t = blockIdx.y;
    if (lr >= own_rows) return;
    const long i = own.row(lr);
    xs[(long)t * own_rows + lr] = i < G ? cnt[perm[i] + (long)cols[t] * G] : 0;
}

// ---- target side, step 1: F_i at the Chebyshev nodes.  grid (own_rows/128, Kc); Fnode is [m*G + i] ----------
// The summand depends on the count and the node only, so it is tabulated once per batch for counts below MCG_LGTAB
// (k_mcg_node_table: Kc x MCG_LGTAB evaluations instead of Kc x G x n_b; the table holds the bits the direct call
// returns; one 64 KB table per node stays in L1 for the block that uses it) and the sweep over the samples becomes a
// lookup.  Larger counts take the direct call.
constexpr int MCG_LGTAB = 8192;
__device__ __forceinline__ double mcg_node_r(const McgShape& sh, int m) {
    const double tn = cospi((m + 0.5) / sh.Kc);
    return exp(0.5 * (sh.ulo + sh.uhi) + 0.5 * (sh.uhi - sh.ulo) * tn) - (double)MCG_SHIFT;
}
__global__ void __launch_bounds__(256)
k_mcg_node_table(McgShape sh, double* __restrict__ tab) {          // grid (MCG_LGTAB / 256, Kc)
    const int x = blockIdx.x * blockDim.x + threadIdx.x, m = blockIdx.y;
    tab[(long)m * MCG_LGTAB + x] = x >= MCG_SHIFT ? lgamma_diff((double)x, mcg_node_r(sh, m)) : 0.0;
}

__global__ void __launch_bounds__(128)
k_mcg_prep_nodes(const int32_t* __restrict__ xs, long G, McgShape sh, McgOwn own, long own_rows,
                 const double* __restrict__ tab, double* __restrict__ Fnode) {
    const long lr = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long i = own.row(lr);                                    // sorted row of one of this rank's targets
    const int m = blockIdx.y;
    if (lr >= own_rows || i >= G) return;
    const double r = mcg_node_r(sh, m);
    const double* __restrict__ t_m = tab + (long)m * MCG_LGTAB;
    double f = 0;
    for (int t = 0; t < sh.n; ++t) {
        const int x = xs[(long)t * own_rows + lr];
        if (x >= MCG_LGTAB) f += lgamma_diff((double)x, r);
        else if (x >= MCG_SHIFT) f += __ldg(t_m + x);
    }
    Fnode[(long)m * G + i] = f;
}

// ---- target side, step 2: row i of P (counts, Chebyshev coefficients by DCT, one-hot counts, 1) ------------
// dct is the Kc x Kc table dct[k*Kc + m] = (k ? 2 : 1)/Kc * cos(pi k (m + 1/2)/Kc).  tail_bits: atomicMax of the
// bit pattern of |C_{Kc-1}| + |C_{Kc-2}| (non-negative doubles order like their bit patterns).
// grid (own_rows / 128, KTP / 8): one thread per (row, k-chunk), the chunk written as one 64-byte piece (see k_mcg_prep_src).
__global__ void __launch_bounds__(128)
k_mcg_prep_tgt(const int32_t* __restrict__ xs, long G, McgShape sh, McgOwn own, long own_rows,
               const double* __restrict__ Fnode, const double* __restrict__ dct,
               double* __restrict__ P, unsigned long long* __restrict__ tail_bits) {
    const long lr = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int kc = blockIdx.y;
    if (lr >= own_rows) return;
    const long i = own.row(lr);
    if (i >= sh.rows) return;
    const int n = sh.n, Kc = sh.Kc, k0 = kc * MCG_KC;
    double v[MCG_KC];
#pragma unroll
    for (int e = 0; e < MCG_KC; ++e) v[e] = 0.0;
    if (i < G) {
#pragma unroll
        for (int e = 0; e < MCG_KC; ++e) if (k0 + e < n) v[e] = (double)xs[(long)(k0 + e) * own_rows + lr];
        if (k0 + MCG_KC > n && k0 < n + Kc) {                    // Chebyshev coefficients c_k = sum_m dct[k][m] f[m]
            double f[MCG_MAX_KC];
            for (int m = 0; m < Kc; ++m) f[m] = Fnode[(long)m * G + i];
#pragma unroll
            for (int e = 0; e < MCG_KC; ++e) {
                const int k = k0 + e - n;
                if (k < 0 || k >= Kc) continue;
                double c = 0;
                for (int m = 0; m < Kc; ++m) c += dct[k * Kc + m] * f[m];
                v[e] = c;
            }
            if (Kc >= 4 && k0 <= n + Kc - 1 && n + Kc - 1 < k0 + MCG_KC) {       // the thread that holds the last coefficient
                double c2 = 0;
                for (int m = 0; m < Kc; ++m) c2 += dct[(Kc - 2) * Kc + m] * f[m];
                double c1 = 0;
                for (int m = 0; m < Kc; ++m) c1 += dct[(Kc - 1) * Kc + m] * f[m];
                const double tail = fabs(c2) + fabs(c1);
                if (tail > 0) atomicMax(tail_bits, (unsigned long long)__double_as_longlong(tail));
            }
        }
        if (k0 + MCG_KC > n + Kc && k0 < n + Kc + MCG_SHIFT) {   // one-hot counts: how many samples have x == v
            int hot[MCG_SHIFT];
#pragma unroll
            for (int q = 0; q < MCG_SHIFT; ++q) hot[q] = 0;
            for (int t = 0; t < n; ++t) {
                const int x = xs[(long)t * own_rows + lr];
#pragma unroll
                for (int q = 0; q < MCG_SHIFT; ++q) hot[q] += (x == q);
            }
#pragma unroll
            for (int e = 0; e < MCG_KC; ++e)
#pragma unroll
                for (int q = 0; q < MCG_SHIFT; ++q) if (k0 + e == n + Kc + q) v[e] = (double)hot[q];
        }
#pragma unroll
        for (int e = 0; e < MCG_KC; ++e) if (k0 + e == sh.KT - 1) v[e] = 1.0;
    }
    const int swz = (int)((i >> 1) & 1) << 2;
    double2* dst = reinterpret_cast<double2*>(P + ((size_t)kc * sh.rows + i) * MCG_KC);
    const double2 a0 = make_double2(v[0], v[1]), a1 = make_double2(v[2], v[3]), b0 = make_double2(v[4], v[5]), b1 = make_double2(v[6], v[7]);
    dst[0] = swz ? b0 : a0; dst[1] = swz ? b1 : a1; dst[2] = swz ? a0 : b0; dst[3] = swz ? a1 : b1;
}

// ---- PTX helpers: mbarrier + 1-D TMA bulk copy (global -> shared) and the FP64 tensor-core MMA ---------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// A warp that finds its stage not landed yet backs off for ~200 ns instead of re-polling at once: the polls of the warps
// that run ahead of the ring (every warp but the slowest, in steady state) otherwise take issue slots from the warps that
// feed the FP64 units (measured on C5-sized batches: 34.1 -> 33.5 ms per launch; a try_wait suspend hint gave 33.9).
#ifndef MCG_WAIT_SLEEP
#define MCG_WAIT_SLEEP 200
#endif
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    for (;;) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
        if (ok) return;
        __nanosleep(MCG_WAIT_SLEEP);
    }
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void dmma884(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

// Pruning.  A source whose log-likelihood is more than MCG_PRUNE below the best one seen so far for that target
// carries a weight below e^-40 = 4e-18 of the largest weight; all G of them together change the weighted means by
// less than 3e-13 relative, so they are dropped before exp() is ever called (and they can never turn an all-zero
// row, the helper_seq.R:30-31 fallback, into a non-zero one: what is dropped is smaller than a weight already seen).
// In reference mode McAcc::m is only this running maximum; the weights themselves stay exp(L), unshifted, so that
// they underflow exactly where R's prod(dnbinom) does.  In stable mode the sums are kept relative to exp(m)
// (online log-sum-exp, rescaled once per source tile when the maximum moves).

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

// 2^(j/64), j = 0..63, correctly rounded (mpmath); copied to shared memory by every CTA of k_mc_gemm
__device__ __constant__ double c_exp2_64[64] = {
    1, 1.0108892860517005, 1.0218971486541166, 1.0330248790212284,
    1.0442737824274138, 1.0556451783605572, 1.0671404006768237, 1.0787607977571199,
    1.0905077326652577, 1.1023825833078409, 1.1143867425958924, 1.1265216186082418,
    1.1387886347566916, 1.1511892299529827, 1.1637248587775775, 1.1763969916502812,
    1.189207115002721, 1.2021567314527031, 1.215247359980469, 1.22848053610687,
    1.241857812073484, 1.2553807570246911, 1.2690509571917332, 1.2828700160787783,
    1.2968395546510096, 1.3109612115247644, 1.3252366431597413, 1.3396675240533029,
    1.3542555469368927, 1.3690024229745905, 1.383909881963832, 1.3989796725383112,
    1.4142135623730951, 1.42961333839197, 1.4451808069770467, 1.460917794180647,
    1.4768261459394993, 1.4929077282912648, 1.5091644275934228, 1.5255981507445384,
    1.5422108254079407, 1.5590044002378369, 1.5759808451078865, 1.593142151342267,
    1.6104903319492543, 1.6280274218573478, 1.6457554781539649, 1.6636765803267364,
    1.681792830507429, 1.7001063537185235, 1.7186192981224779, 1.7373338352737062,
    1.7562521603732995, 1.7753764925265212, 1.7947090750031072, 1.8142521755003989,
    1.8340080864093424, 1.8539791250833855, 1.8741676341103, 1.8945759815869656,
    1.9152065613971474, 1.9360617934922943, 1.9571441241754002, 1.9784560263879509
};

// exp(x) for x <= ~0.5 down to the underflow threshold (x < -745.14 gives 0; results in the denormal range are rounded
// once, like the libm exp).  DMMA and DFMA share the FP64 units on B200 (tools/dmma_bench.cu, "mix"), so every FP64
// instruction of the epilogue is taken from the contraction: x = (64 k + j) ln2/64 + r with |r| <= ln2/128, exp(x) =
// 2^k * T[j] * (1 + p(r)) with a 64-entry table in shared memory and a degree-5 polynomial: 11 FP64 instructions,
// the 2^k scaling is an integer add on the exponent field (plus one multiply that is 1.0 unless the result is
// subnormal).  Relative error < 3e-16.  Garbage (not a trap, no out-of-range table index) for x = -inf / NaN: callers mask.
__device__ __forceinline__ double mcg_exp(double x, const double* __restrict__ T) {
    const double t = fma(x, 92.33248261689366, 6755399441055744.0);            // 64/ln2; 1.5 * 2^52 rounds to an integer
    const int n = __double2loint(t);
    const double nf = t - 6755399441055744.0;
    double r = fma(nf, -0.010830424696249145, x);                              // ln2/64 = hi + lo
    r = fma(nf, -3.623510646634843e-19, r);
    double q = fma(r, 1.0 / 120, 1.0 / 24);
    q = fma(q, r, 1.0 / 6);
    q = fma(q, r, 0.5);
    const double p = fma(q, r * r, r);
    const double Tj = T[n & 63];
    const double v = fma(Tj, p, Tj);
    const int k = n >> 6, k1 = max(k, -1000);
    const double vs = __hiloint2double(__double2hiint(v) + (k1 << 20), __double2loint(v));
    return vs * __hiloint2double((k - k1 + 1023) << 20, 0);
}

constexpr int MCG_CONSUMER_WARPS = MCG_THREADS / 32;
constexpr int MCG_WI = MCG_CONSUMER_WARPS / 2;      // warps along the targets
constexpr int MCG_BLOCK = MCG_CONSUMER_WARPS * 32;
constexpr int MCG_STAGE_DBL = MCG_SUB * (MCG_TI + MCG_TJ) * MCG_KC;                             // doubles per stage
constexpr size_t MCG_SMEM_STAGE = (size_t)MCG_STAGE_DBL * sizeof(double);                         // SUB x (P chunk + Q chunk)
constexpr size_t MCG_SMEM_GP = (size_t)MCG_GPBUF * MCG_TJ * 2 * sizeof(double);
constexpr size_t MCG_SMEM_BYTES = MCG_STAGES * MCG_SMEM_STAGE + MCG_SMEM_GP + 2 * MCG_STAGES * sizeof(uint64_t) + 64   // full[] + done[]
                                  + 64 * sizeof(double);                                                                 // exp2 table
static_assert(MCG_THREADS == MCG_CONSUMER_WARPS * 32, "consumer thread count");

// One CTA = one 64-target tile x one split of the source tiles.  Eight MMA warps, arranged 4 (targets) x 2 (sources),
// warp tile 16 x 64 = 2 x 8 DMMA tiles (64 accumulator registers, 128 registers per thread); two CTAs share an SM, each
// with its own stage ring, so a CTA that waits on its ring or folds weights leaves the FP64 units to the other one.  The k-chunks stream through a ring of MCG_STAGES stages filled
// by TMA bulk copies that complete on a `full` mbarrier per stage.  There is no producer thread and no CTA-wide
// barrier in the main loop: every warp counts itself out of a stage (shared-memory atomic), and the LAST warp to
// leave a stage issues the copies that refill it.  A warp that is folding the weights of a finished source tile
// (FP64 pipe) therefore never holds up the refills the other warps are waiting for, and the other warp of its
// scheduler keeps the tensor pipe busy meanwhile.  part[(split*rows + i)] receives the partial sums of target row i.
// Work items: items[w] = (target tile, segment); segment s of target tile t covers entries [s*MCG_SEG, ...) of the tile's
// source-tile list srclist[t*NJ ..] (count[t] entries, ascending; built by the screening pass, or all NJ tiles).  The
// decomposition depends on the lists only, never on the number of GPUs, and the partials of a target tile are merged in
// segment order (k_mcg_finalize): results are bit-identical for every sharding.
#ifndef CBS_MCG_SEG
#define CBS_MCG_SEG 8
#endif
constexpr int MCG_SEG = CBS_MCG_SEG;           // source tiles per work item
template <int STABLE>
__global__ void __launch_bounds__(MCG_BLOCK, MCG_CTAS)
k_mc_gemm(const double* __restrict__ P, const double* __restrict__ Q, const double2* __restrict__ gp, McgShape sh,
          const int2* __restrict__ items, const int* __restrict__ srclist, const int* __restrict__ count,
          McAcc* __restrict__ part) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    double* stage_base = reinterpret_cast<double*>(smem_raw);
    double2* gp_s = reinterpret_cast<double2*>(smem_raw + MCG_STAGES * MCG_SMEM_STAGE);
    uint64_t* full = reinterpret_cast<uint64_t*>(smem_raw + MCG_STAGES * MCG_SMEM_STAGE + MCG_SMEM_GP);
    unsigned int* done = reinterpret_cast<unsigned int*>(full + MCG_STAGES);    // warps that have left each stage
    double* exp_tab = reinterpret_cast<double*>(smem_raw + MCG_STAGES * MCG_SMEM_STAGE + MCG_SMEM_GP + 2 * MCG_STAGES * sizeof(uint64_t) + 64);
    if (threadIdx.x < 64) exp_tab[threadIdx.x] = c_exp2_64[threadIdx.x];

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int wi = warp % MCG_WI, wj = warp / MCG_WI;
    const int2 item = items[blockIdx.x];
    const long itile = item.x;
    const long NJ = sh.rows / MCG_TJ;
    const int nsrc = min(MCG_SEG, count[itile] - item.y * MCG_SEG);       // source tiles of this work item
    __shared__ int src_s[MCG_SEG];
    if (threadIdx.x < MCG_SEG) src_s[threadIdx.x] = threadIdx.x < nsrc ? srclist[itile * NJ + (long)item.y * MCG_SEG + threadIdx.x] : 0;
    const int NCH = sh.KTP / MCG_KC;                              // k-chunks per source tile
    const int NKS = (NCH + MCG_SUB - 1) / MCG_SUB;                // hand-overs (stages) per source tile
    const int last_sub = NCH - (NKS - 1) * MCG_SUB;               // k-chunks in a tile's last hand-over (1..MCG_SUB)
    const long total = (long)nsrc * NKS;

    if (tid == 0) {
        for (int s = 0; s < MCG_STAGES; ++s) { mbar_init(&full[s], 1); done[s] = 0; }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    // copies of hand-over `c` (MCG_SUB P chunks and Q chunks and, with the first one of a source tile, its gamma/phi)
    auto issue = [&](long c) {
        const int ns = (int)(c % MCG_STAGES);
        const long njl = c / NKS; const int nks = (int)(c - njl * NKS);
        const uint32_t p_bytes = MCG_TI * MCG_KC * sizeof(double), q_bytes = MCG_TJ * MCG_KC * sizeof(double);
        double* sb = stage_base + (size_t)ns * MCG_STAGE_DBL;
        const int nsub = (nks == NKS - 1) ? last_sub : MCG_SUB;
        mbar_expect_tx(&full[ns], nsub * (p_bytes + q_bytes) + (nks == 0 ? (uint32_t)(MCG_TJ * sizeof(double2)) : 0u));
#pragma unroll
        for (int u = 0; u < MCG_SUB; ++u) {
            if (u >= nsub) break;
            const size_t kc = (size_t)nks * MCG_SUB + u;
            double* su = sb + u * (MCG_TI + MCG_TJ) * MCG_KC;
            bulk_g2s(su, P + (kc * sh.rows + itile * MCG_TI) * MCG_KC, p_bytes, &full[ns]);
            bulk_g2s(su + MCG_TI * MCG_KC, Q + (kc * sh.rows + (long)src_s[njl] * MCG_TJ) * MCG_KC, q_bytes, &full[ns]);
        }
        // gamma/phi of source tile njl go to buffer njl % MCG_GPBUF, last read in the epilogue of tile njl - MCG_GPBUF,
        // which every warp has finished before it leaves hand-over njl * NKS - MCG_STAGES
        // ((MCG_GPBUF - 1) * NKS >= MCG_STAGES + 1 is guaranteed by the host: KTP >= 2 stages).
        if (nks == 0) bulk_g2s(gp_s + (njl % MCG_GPBUF) * MCG_TJ, gp + (long)src_s[njl] * MCG_TJ, MCG_TJ * sizeof(double2), &full[ns]);
    };
    if (tid == 0)
        for (long c = 0; c < total && c < MCG_STAGES; ++c) issue(c);

    McAcc racc[MCG_MT];
#pragma unroll
    for (int mt = 0; mt < MCG_MT; ++mt) { racc[mt].m = -INFINITY; racc[mt].sw = racc[mt].sg = racc[mt].sp = 0.0; }
    const int frow = lane >> 2, fcol = lane & 3;

    {
        // ---- MMA warps ----
        double acc[MCG_MT][MCG_NT][2];
#pragma unroll
        for (int mt = 0; mt < MCG_MT; ++mt)
#pragma unroll
            for (int nt = 0; nt < MCG_NT; ++nt) acc[mt][nt][0] = acc[mt][nt][1] = 0.0;
        // fragment addressing: row = lane/4 within an 8-row MMA tile, column = lane%4 within a 4-wide k-step
        const int swz = ((frow >> 1) & 1) << 2;             // tile bases are multiples of 8, so bit 1 of the row is frow's
        const int a_off = (wi * (8 * MCG_MT) + frow) * MCG_KC, b_off = (wj * (8 * MCG_NT) + frow) * MCG_KC;
        long jl = 0; int kc = 0;
        int st = 0; uint32_t phase = 0;
        for (long it = 0; it < total; ++it) {
            mbar_wait(&full[st], phase);
            const int nsub = (kc == NKS - 1) ? last_sub : MCG_SUB;
#pragma unroll
            for (int u = 0; u < MCG_SUB; ++u) {
            if (u >= nsub) break;                            // warp-uniform: a tile's last hand-over may be short
            const double* ps = stage_base + (size_t)st * MCG_STAGE_DBL + u * (MCG_TI + MCG_TJ) * MCG_KC;
            const double* qs = ps + MCG_TI * MCG_KC;
#pragma unroll
            for (int ks = 0; ks < 2; ++ks) {
                const int col = (ks * 4 + fcol) ^ swz;
                double a[MCG_MT], b[MCG_NT];
#pragma unroll
                for (int mt = 0; mt < MCG_MT; ++mt) a[mt] = ps[a_off + mt * 8 * MCG_KC + col];
#pragma unroll
                for (int nt = 0; nt < MCG_NT; ++nt) b[nt] = qs[b_off + nt * 8 * MCG_KC + col];
#pragma unroll
                for (int mt = 0; mt < MCG_MT; ++mt)
#pragma unroll
                    for (int nt = 0; nt < MCG_NT; ++nt) dmma884(acc[mt][nt][0], acc[mt][nt][1], a[mt], b[nt]);
            }
            }
            __syncwarp();
            if (lane == 0) {                                 // this warp is done with stage st; the last one out refills it
                uint32_t old;
                asm volatile("atom.acq_rel.cta.shared::cta.add.u32 %0, [%1], 1;" : "=r"(old) : "r"(smem_u32(&done[st])) : "memory");
                if (old == MCG_CONSUMER_WARPS - 1) {
                    done[st] = 0;
                    if (it + MCG_STAGES < total) issue(it + MCG_STAGES);
                }
            }
            if (++st == MCG_STAGES) { st = 0; phase ^= 1u; }

            if (++kc == NKS) {                               // last k-chunk of source tile jl: fold the weights in
                const long jt = src_s[jl];
                const double2* g2 = gp_s + (jl % MCG_GPBUF) * MCG_TJ;
                if (jt == (itile * MCG_TI) / MCG_TJ) {     // mu[-i,]: the gene itself is not a source  helper_seq.R:11-14
                    const int ioff = (int)((itile * MCG_TI) % MCG_TJ);
#pragma unroll
                    for (int mt = 0; mt < MCG_MT; ++mt)
#pragma unroll
                        for (int nt = 0; nt < MCG_NT; ++nt)
#pragma unroll
                            for (int e = 0; e < 2; ++e)
                                if (wj * (8 * MCG_NT) + nt * 8 + 2 * fcol + e == ioff + wi * (8 * MCG_MT) + mt * 8 + frow) acc[mt][nt][e] = -INFINITY;
                }
                // running maximum of every target row (shared by the 4 lanes that hold the row) -> prune threshold
                double thr[MCG_MT];
#pragma unroll
                for (int mt = 0; mt < MCG_MT; ++mt) {
                    double mx = fmax(acc[mt][0][0], acc[mt][0][1]);
#pragma unroll
                    for (int nt = 1; nt < MCG_NT; ++nt) mx = fmax(mx, fmax(acc[mt][nt][0], acc[mt][nt][1]));
                    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
                    mx = fmax(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
                    if (STABLE) {
                        if (mx > racc[mt].m && mx > -1e29) {         // -1e30 stands for log(0): an impossible source
                            const double sc = exp(racc[mt].m - mx);  // exp(-inf) = 0 on the first hit
                            racc[mt].sw *= sc; racc[mt].sg *= sc; racc[mt].sp *= sc; racc[mt].m = mx;
                        }
                        thr[mt] = fmax(racc[mt].m - MCG_PRUNE, -1e29);
                    } else {
                        racc[mt].m = fmax(racc[mt].m, mx);
                        thr[mt] = fmax(racc[mt].m - MCG_PRUNE, -746.0);   // exp(L) == 0 below -745.14
                    }
                }
                // Fold the weights in.  A sizeable share of the pairs survives the pruning (about one in eight on
                // RNA-seq-like data), so every lane group has work: the 8 elements of a column group are evaluated
                // branch-free and independently (8 interleaved mcg_exp chains keep the FP64 pipe busy) and masked;
                // one warp vote skips a column group whose 256 pairs are all pruned (far-apart genes).
#pragma unroll
                for (int nt = 0; nt < MCG_NT; ++nt) {
                    bool need = false;
#pragma unroll
                    for (int mt = 0; mt < MCG_MT; ++mt) need |= (acc[mt][nt][0] > thr[mt]) | (acc[mt][nt][1] > thr[mt]);
                    if (__any_sync(0xffffffffu, need)) {
                        const double2 gp0 = g2[wj * (8 * MCG_NT) + nt * 8 + 2 * fcol], gp1 = g2[wj * (8 * MCG_NT) + nt * 8 + 2 * fcol + 1];
                        double w0[MCG_MT], w1[MCG_MT];
#pragma unroll
                        for (int mt = 0; mt < MCG_MT; ++mt) {
                            const double L0 = acc[mt][nt][0], L1 = acc[mt][nt][1];
                            const double e0 = mcg_exp(STABLE ? L0 - racc[mt].m : L0, exp_tab), e1 = mcg_exp(STABLE ? L1 - racc[mt].m : L1, exp_tab);
                            w0[mt] = (L0 > thr[mt]) ? e0 : 0.0;
                            w1[mt] = (L1 > thr[mt]) ? e1 : 0.0;
                        }
#pragma unroll
                        for (int mt = 0; mt < MCG_MT; ++mt) {
                            racc[mt].sw += w0[mt] + w1[mt];
                            racc[mt].sg = fma(w1[mt], gp1.x, fma(w0[mt], gp0.x, racc[mt].sg));
                            racc[mt].sp = fma(w1[mt], gp1.y, fma(w0[mt], gp0.y, racc[mt].sp));
                        }
                    }
#pragma unroll
                    for (int mt = 0; mt < MCG_MT; ++mt) acc[mt][nt][0] = acc[mt][nt][1] = 0.0;
                }
                kc = 0; ++jl;
            }
        }
    }
    // reduce over the 4 lanes that share a row, then over the two source-side warps
    __syncthreads();
    McAcc* red = reinterpret_cast<McAcc*>(smem_raw);        // [2][MCG_TI], aliases the (now idle) stage buffers
    {
#pragma unroll
        for (int mt = 0; mt < MCG_MT; ++mt) {
            McAcc a = racc[mt];
            for (int o = 1; o <= 2; o <<= 1) { const McAcc b = mc_shfl_xor(a, o); mc_merge(a, b, STABLE); }
            if (fcol == 0) red[wj * MCG_TI + wi * (8 * MCG_MT) + mt * 8 + frow] = a;
        }
    }
    __syncthreads();
    if (tid < MCG_TI) {
        McAcc a = red[tid];
        mc_merge(a, red[MCG_TI + tid], STABLE);
        part[(size_t)blockIdx.x * MCG_TI + tid] = a;
    }
}

// merge the work-item partials of every target row this rank owns, in segment order, and apply helper_seq.R:30-33.
// Row i of the sorted order is gene perm[i]; gamma/phi/gamma_star/phi_star are indexed by gene.
// grid: one block of MCG_TI threads per owned target tile (tiles[] lists them).
__global__ void __launch_bounds__(MCG_TI)
k_mcg_finalize(const McAcc* __restrict__ part, const int* __restrict__ tiles, const int* __restrict__ item_start,
               const int* __restrict__ perm, long G, int stable, const double* __restrict__ gamma,
               const double* __restrict__ phi, double* __restrict__ gamma_star, double* __restrict__ phi_star,
               unsigned long long* __restrict__ n_fallback) {
    const int t = tiles[blockIdx.x];
    const long i = (long)t * MCG_TI + threadIdx.x;
    if (i >= G) return;
    const int w0 = item_start[blockIdx.x], w1 = item_start[blockIdx.x + 1];
    McAcc a;
    a.m = -INFINITY; a.sw = a.sg = a.sp = 0.0;
    for (int w = w0; w < w1; ++w) mc_merge(a, part[(size_t)w * MCG_TI + threadIdx.x], stable);
    const long gi = perm[i];
    if (a.sw == 0 || !(a.sw == a.sw)) {
        gamma_star[gi] = gamma[gi]; phi_star[gi] = phi[gi];
        atomicAdd(n_fallback, 1ull);
    } else {
        gamma_star[gi] = a.sg / a.sw; phi_star[gi] = a.sp / a.sw;
    }
}

__global__ void k_mcg_pack_gp(const double* __restrict__ gamma, const double* __restrict__ phi, const int* __restrict__ perm,
                              long G, long rows, double2* __restrict__ gp) {
    const long j = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (j < rows) gp[j] = (j < G) ? make_double2(gamma[perm[j]], phi[perm[j]]) : make_double2(0.0, 0.0);
}

// ---- abundance order ------------------------------------------------------------------------------------------------
// key[g] = mean over the batch's samples of log(mu_hat[g, s]); the stable radix sort of the keys gives perm[].
__global__ void __launch_bounds__(128)
k_mcg_sort_key(const double* __restrict__ mu_hat, long G, int n, const int* __restrict__ cols, float* __restrict__ key,
               int* __restrict__ idx) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    // single precision and independent loads: the key only orders the genes (any fixed order gives the same sums up to
    // rounding, and the same bits for every GPU count); in double precision with one dependent log per load this kernel was
    // as long as the whole operand preparation of a shard at 8 GPUs
    float s = 0.f;
#pragma unroll 8
    for (int t = 0; t < n; ++t) s += __log2f(fmaxf((float)mu_hat[g + (long)cols[t] * G], 1e-37f));
    key[g] = s / (float)n;
    idx[g] = (int)g;
}
__global__ void k_mcg_iota(int* __restrict__ idx, long G) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g < G) idx[g] = (int)g;
}

// ---- source-tile lists -------------------------------------------------------------------------------------------------
// Without screening every target tile visits every source tile.  grid (ntiles64), block 128.
__global__ void k_mcg_list_all(int NJ, int* __restrict__ srclist, int* __restrict__ count) {
    const int t = blockIdx.x;
    for (int j = threadIdx.x; j < NJ; j += blockDim.x) srclist[(long)t * NJ + j] = j;
    if (threadIdx.x == 0) count[t] = NJ;
}

// With screening (tu_screen.cu): hi[J*rows + i] is an upper bound of max_{j in source tile J} log LH_ij and lo[J*rows + i] a
// lower bound of the same maximum over j != i.  Source tile J is kept for target tile t when some row of t can have a
// pair in J within MCG_PRUNE of the row's best pair:  hi[J][i] >= max_J' lo[J'][i] - MCG_PRUNE.  Whatever is dropped is
// below e^-40 of the row's largest weight, the bound the fold in k_mc_gemm applies anyway.  grid (ntiles64), block 256,
// NJ bytes of dynamic shared memory.  The 4 thread groups of a block split the source tiles.
__global__ void __launch_bounds__(256)
k_mcg_list_screened(const float* __restrict__ hi, const float* __restrict__ lo, long rows, int NJ, float prune, McgOwn own,
                    int* __restrict__ srclist, int* __restrict__ count) {
    extern __shared__ unsigned char flag_s[];                  // [NJ]
    __shared__ float mx[4][MCG_TI];
    const int t = blockIdx.x, r = threadIdx.x & (MCG_TI - 1), q = threadIdx.x / MCG_TI;
    if (!own.owns_tile64(t)) {                                 // another rank's targets: no bounds were computed here
        if (threadIdx.x == 0) count[t] = 0;
        return;
    }
    const long i = (long)t * MCG_TI + r;
    float m = -INFINITY;
#pragma unroll 4
    for (int J = q; J < NJ; J += 4) m = fmaxf(m, lo[(long)J * rows + i]);
    mx[q][r] = m;
    for (int J = threadIdx.x; J < NJ; J += 256) flag_s[J] = 0;
    __syncthreads();
    m = fmaxf(fmaxf(mx[0][r], mx[1][r]), fmaxf(mx[2][r], mx[3][r]));
    const float thr = m - prune;                               // -inf when the row has no finite pair: everything is kept
#pragma unroll 4
    for (int J = q; J < NJ; J += 4) if (!(hi[(long)J * rows + i] < thr)) flag_s[J] = 1;      // (NaN counts as kept)
    __syncthreads();
    if (threadIdx.x < 32) {
        const int lane = threadIdx.x;
        int n = 0;
        for (int J0 = 0; J0 < NJ; J0 += 32) {
            const int J = J0 + lane;
            const bool k = J < NJ && flag_s[J];
            const unsigned b = __ballot_sync(0xffffffffu, k);
            if (k) srclist[(long)t * NJ + n + __popc(b & ((1u << lane) - 1u))] = J;
            n += __popc(b);
        }
        if (lane == 0) count[t] = n;
    }
}

// Work items of the target tiles this rank owns, in tile order: tiles[k], item_start[k] .. item_start[k+1], items[w] =
// (tile, segment).  One block of 1024 threads, one tile per thread and round, block-wide exclusive scans of (owned,
// segments).  out[0] = number of items, out[1] = number of owned tiles, out[2] = sum of count[] over the owned tiles.
__global__ void __launch_bounds__(1024)
k_mcg_items(const int* __restrict__ count, int ntiles64, McgOwn own, int* __restrict__ tiles,
            int* __restrict__ item_start, int2* __restrict__ items, int* __restrict__ out) {
    __shared__ int s_k[1024], s_w[1024], s_c[1024];
    __shared__ int base_k, base_w, base_c;
    if (threadIdx.x == 0) { base_k = 0; base_w = 0; base_c = 0; }
    __syncthreads();
    for (int t0 = 0; t0 < ntiles64; t0 += 1024) {
        const int t = t0 + threadIdx.x;
        const bool mine = t < ntiles64 && own.owns_tile64(t);
        const int c = mine ? count[t] : 0, nseg = (c + MCG_SEG - 1) / MCG_SEG;
        s_k[threadIdx.x] = mine ? 1 : 0; s_w[threadIdx.x] = nseg; s_c[threadIdx.x] = c;
        __syncthreads();
        for (int o = 1; o < 1024; o <<= 1) {                  // inclusive Hillis-Steele scans
            int ak = 0, aw = 0, ac = 0;
            if ((int)threadIdx.x >= o) { ak = s_k[threadIdx.x - o]; aw = s_w[threadIdx.x - o]; ac = s_c[threadIdx.x - o]; }
            __syncthreads();
            s_k[threadIdx.x] += ak; s_w[threadIdx.x] += aw; s_c[threadIdx.x] += ac;
            __syncthreads();
        }
        if (mine) {
            const int k = base_k + s_k[threadIdx.x] - 1, w = base_w + s_w[threadIdx.x] - nseg;
            tiles[k] = t; item_start[k] = w;
            for (int sgm = 0; sgm < nseg; ++sgm) items[w + sgm] = make_int2(t, sgm);
        }
        __syncthreads();
        if (threadIdx.x == 1023) { base_k += s_k[1023]; base_w += s_w[1023]; base_c += s_c[1023]; }
        __syncthreads();
    }
    if (threadIdx.x == 0) { item_start[base_k] = base_w; out[0] = base_w; out[1] = base_k; out[2] = base_c; }
}

}  // namespace cbs
