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
#ifndef CBS_GLM_LANES_N
#define CBS_GLM_LANES_N 8
#endif
constexpr int GLM_LANES = CBS_GLM_LANES_N;
constexpr int GLM_LANES_PMAX = 16;
#ifndef CBS_GLM_MINBLOCKS
#define CBS_GLM_MINBLOCKS 6      // measured on C5: 4 -> 23.4, 5 -> 21.6, 6 -> 21.35 ms per step
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
#ifndef CBS_GLM_SWEEP_CAP
#define CBS_GLM_SWEEP_CAP 8      // measured on C5 (lanes kernel, ms per step): 12 -> 22.4, 9 -> 21.6, 8 -> 21.3, 7 -> 21.0, and 6 sends
                                 // the bulk of the genes to the second pass (65 ms): the cap must stay above the usual 5-7 sweeps
#endif
constexpr int GLM_SWEEP_CAP = CBS_GLM_SWEEP_CAP;
// One launch can carry two independent fits of the same genes (blockIdx.y picks the job): the plain fit and the predFC fit
// of a glmFit(prior.count > 0) call read the same counts and do not depend on each other, and once the genes are spread
// over several GPUs a launch is a fraction of a wave whose duration is the latency of one fit, not its throughput.
struct GlmJob {
    const double* yadd; double two_prior_over_meanlib; int mode;
    double* coef; double* fitted; double* alpha_out; int* strag; int* n_strag;
};
struct GlmJobs { GlmJob j[2]; };
__device__ __forceinline__ GlmJob glm_pick_job(const GlmJobs& jobs) {
    GlmJob a = jobs.j[0];
    if (blockIdx.y) a = jobs.j[1];
    return a;
}
template <int PM>                                            // PM >= design columns: sizes the per-thread packed X'WX (local memory)
__global__ void __launch_bounds__(128, CBS_GLM_MINBLOCKS)
k_glm_fit_lanes(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ logoff,
                const __grid_constant__ GlmJobs jobs, const double* __restrict__ phi,
                const int* __restrict__ phi_col, const DesignDev* __restrict__ design, const double* __restrict__ goff_in,
                const double* __restrict__ alpha_w, int n_alpha, int use_scr, int phi_by_class, int* __restrict__ iter_hist) {
    const GlmJob jb = glm_pick_job(jobs);
    const double* __restrict__ yadd = jb.yadd;
    const double two_prior_over_meanlib = jb.two_prior_over_meanlib;
    const int mode = jb.mode;
    double* __restrict__ coef = jb.coef; double* __restrict__ fitted = jb.fitted; double* __restrict__ alpha_out = jb.alpha_out;
    int* __restrict__ strag = jb.strag; int* __restrict__ n_strag = jb.n_strag;
    const long slot = ((long)blockIdx.x * blockDim.x + threadIdx.x) / GLM_LANES;
    const long g = g0 + slot;
    const int ln = threadIdx.x % GLM_LANES;
    if (g >= g1) return;                                    // whole groups leave together
    const unsigned gm = ((1u << GLM_LANES) - 1u) << ((threadIdx.x & 31) - ln);
    const DesignDev d = *design;
    double goff = goff_in ? goff_in[g] : 0.0;
    if (mode == 1) goff = log(exp(goff) + two_prior_over_meanlib);
    RowSrc src{cnt, G, logoff, mode == 1 ? yadd : nullptr, phi, phi_col, 0.0, nullptr, phi_by_class != 0};
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

// Second pass: one BLOCK (GLM_WIDE_WARPS warps) per abandoned gene.  The gene's counts are staged in shared memory (the sweeps then run
// at shared-memory instead of DRAM latency); a sweep is split over the design's classes (warp w takes classes w, w + GLM_WIDE_WARPS, ...:
// one exp per class, the class's samples over the 32 lanes, butterfly sums), the class sums are left in shared memory and
// warp 0 assembles deviance, X'WX and X'z from them in class order.  Warp 0 runs edgeR's Levenberg iteration exactly as
// fit_levenberg does (same lambda schedule, acceptance and stopping rules); the other warps wait for its sweep requests at
// a named barrier.  Everything that mutates the shared work arrays is done by lane 0 of warp 0 between __syncwarp()s; the
// other lanes of warp 0 follow the control flow on the same values.  The fit restarts from the null start; its sums
// associate differently than on 8 lanes, so the coefficients of these few genes differ from an 8-lane fit in the last
// places: what matters is that the route a gene takes does not depend on the sharding.  Needs the class form (d.ncls > 0).
// grid: any (blocks stride over the list).  Dynamic shared memory: glm_wide_smem_doubles<PM>() doubles + N ints.
template <int PM>
__host__ __device__ constexpr int glm_wide_smem_doubles() { return 3 * (PM * (PM + 1) / 2) + 5 * PM + 96 + 2 + 32 * PM; }

// chol_packed / chol_solve_packed (glm.cuh) with the rows of a column spread over the lanes of one warp.  Every element goes
// through the same operations as in the serial routines (the factorisation and the forward substitution in the same order, the
// back substitution subtracts its terms in the opposite order); a is a shared-memory array, p <= 32.  Returns false (on
// every lane) when a pivot is <= 0 or NaN.
__device__ __forceinline__ bool chol_packed_warp(double* a, int p, int lane) {
    for (int j = 0; j < p; ++j) {
        int ok = 1;
        if (lane == 0) {
            double ajj = a[tri(j, j)];
            for (int k = 0; k < j; ++k) ajj -= a[tri(j, k)] * a[tri(j, k)];
            ok = ajj > 0;
            if (ok) a[tri(j, j)] = sqrt(ajj);
        }
        ok = __shfl_sync(0xffffffffu, ok, 0);
        if (!ok) return false;
        __syncwarp();
        const int c = j + 1 + lane;
        if (c < p) {
            double v = a[tri(c, j)];
            for (int k = 0; k < j; ++k) v -= a[tri(j, k)] * a[tri(c, k)];
            a[tri(c, j)] = v / a[tri(j, j)];
        }
        __syncwarp();
    }
    return true;
}
__device__ __forceinline__ void chol_solve_packed_warp(const double* a, int p, double* b, int lane) {
    for (int i = 0; i < p; ++i) {                              // forward: b[c] -= a[c][i] b[i] in the order i = 0, 1, ... for every c
        if (lane == i) b[i] = b[i] / a[tri(i, i)];
        __syncwarp();
        if (lane > i && lane < p) b[lane] -= a[tri(lane, i)] * b[i];
        __syncwarp();
    }
    for (int i = p - 1; i >= 0; --i) {                         // backward: b[c] -= a[i][c] b[i] in the order i = p-1, p-2, ...
        if (lane == i) b[i] = b[i] / a[tri(i, i)];
        __syncwarp();
        if (lane < i) b[lane] -= a[tri(i, lane)] * b[i];
        __syncwarp();
    }
}
#ifndef CBS_GLM_WIDE_WARPS
#define CBS_GLM_WIDE_WARPS 16     // warps per abandoned gene: the classes of a sweep are dealt out over them (C5: 20 classes)
#endif
constexpr int GLM_WIDE_WARPS = CBS_GLM_WIDE_WARPS, GLM_WIDE_THREADS = 32 * GLM_WIDE_WARPS;
constexpr int GLM_WIDE_BLOCKS = 592;      // blocks stride over the list; the usual list is a few dozen genes, but data whose fits need more
                                          // sweeps than the cap must not queue behind 64 blocks (blocks beyond the list leave at once)
__device__ __forceinline__ void glm_wide_bar() { asm volatile("bar.sync 1, %0;" ::"n"(GLM_WIDE_THREADS) : "memory"); }

// class sums of one sweep at the coefficients bsrc for the classes of this warp: cls[c], cls[32+c], cls[64+c] = sum W, sum z, deviance
__device__ __forceinline__ void glm_wide_classes(const DesignDev& d, const Row& r, const double* bsrc, int warp, int lane,
                                                 double* cls) {
    for (int c = warp; c < d.ncls; c += GLM_WIDE_WARPS) {
        const int t0 = d.cell_start[c], t1 = d.cell_start[c + 1], k0 = d.rowptr[t0], k1 = d.rowptr[t0 + 1];
        double eta = r.goff;
        for (int k = k0; k < k1; ++k) eta += d.cval[k] * bsrc[d.cidx[k]];
        const double E = exp(eta);
        double sw = 0, sz = 0, dv = 0;
        const double phic = r.s.phi_by_class ? r.phi(d.col[t0]) : 0.0;
        for (int t = t0 + lane; t < t1; t += 32) {
            const int col = d.col[t];
            const double mu = E * r.lib(col), y = r.y(col), phi = r.s.phi_by_class ? phic : r.phi(col);
            dv += unit_nb_deviance(y, mu, phi);
            const double inv = rcp_pos(1 + mu * phi);
            sw += mu * inv; sz += (y - mu) * inv;
        }
        sw = lane_sum<32>(sw, 0xffffffffu); sz = lane_sum<32>(sz, 0xffffffffu); dv = lane_sum<32>(dv, 0xffffffffu);
        if (lane == 0) { cls[c] = sw; cls[32 + c] = sz; cls[64 + c] = dv; }
    }
}

template <int PM>
__global__ void __launch_bounds__(GLM_WIDE_THREADS)
k_glm_fit_wide(const int32_t* __restrict__ cnt, long G, long g0, int N, const double* __restrict__ logoff,
               const __grid_constant__ GlmJobs jobs, const double* __restrict__ phi,
               const int* __restrict__ phi_col, const DesignDev* __restrict__ design, const double* __restrict__ goff_in,
               const double* __restrict__ alpha_w, int n_alpha, int* __restrict__ iter_hist, int stage_counts, int phi_by_class) {
    const GlmJob jb = glm_pick_job(jobs);
    const double* __restrict__ yadd = jb.yadd;
    const double two_prior_over_meanlib = jb.two_prior_over_meanlib;
    const int mode = jb.mode;
    double* __restrict__ coef = jb.coef; double* __restrict__ fitted = jb.fitted; double* __restrict__ alpha_out = jb.alpha_out;
    const int* __restrict__ strag = jb.strag; const int* __restrict__ n_strag = jb.n_strag;
    extern __shared__ __align__(16) double glm_scr[];
    constexpr int PT = PM * (PM + 1) / 2;
    double* const xtwx = glm_scr;
    double* const xtwx_new = glm_scr + PT;
    double* const work = glm_scr + 2 * PT;
    double* const dl = glm_scr + 3 * PT;
    double* const dl_new = dl + PM;
    double* const dbeta = dl + 2 * PM;
    double* const beta_new = dl + 3 * PM;
    double* const beta_s = dl + 4 * PM;
    double* const cls = dl + 5 * PM;                          // [3][32] class sums of the current sweep
    int* const ctrl = reinterpret_cast<int*>(cls + 96);       // ctrl[0]: 1 = sweep requested, 2 = fit finished; ctrl[1]: 0 = at beta, 1 = at beta_new
    double* const Xc = cls + 98;                              // [32][PM] dense design row of every class
    int32_t* const xs = reinterpret_cast<int32_t*>(Xc + 32 * PM);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const DesignDev d = *design;
    const int p = d.p, pt = p * (p + 1) / 2;
    const int count = *n_strag;
    if ((int)blockIdx.x >= count) return;                     // (block-uniform) nothing for this block
    for (int k = threadIdx.x; k < 32 * PM; k += blockDim.x) Xc[k] = 0;
    __syncthreads();
    for (int c = threadIdx.x; c < d.ncls; c += blockDim.x) {
        const int t0 = d.cell_start[c];
        for (int k = d.rowptr[t0]; k < d.rowptr[t0 + 1]; ++k) Xc[c * PM + d.cidx[k]] = d.cval[k];
    }
    for (int i = blockIdx.x; i < count; i += gridDim.x) {
        const long g = g0 + strag[i];
        __syncthreads();                                      // the previous gene of this block is done with the shared arrays
        if (stage_counts) for (int c = threadIdx.x; c < N; c += blockDim.x) xs[c] = cnt[g + (long)c * G];
        __syncthreads();
        double goff = goff_in ? goff_in[g] : 0.0;
        if (mode == 1) goff = log(exp(goff) + two_prior_over_meanlib);
        RowSrc src{cnt, G, logoff, mode == 1 ? yadd : nullptr, phi, phi_col, 0.0, stage_counts ? xs : nullptr, phi_by_class != 0};
        Row r(src, g, goff);
        if (warp != 0) {                                      // workers: serve sweep requests until the fit is finished
            for (;;) {
                glm_wide_bar();
                if (ctrl[0] == 2) break;
                glm_wide_classes(d, r, ctrl[1] ? beta_new : beta_s, warp, lane, cls);
                glm_wide_bar();
            }
            continue;
        }
        // ---- warp 0: the Levenberg iteration (glm.cuh fit_levenberg, same rules) ----
        // one sweep at bsrc: deviance returned, X'WX -> xt, X'z -> dz
        auto sweep = [&](int at_new, double* xt, double* dz) -> double {
            if (lane == 0) { ctrl[0] = 1; ctrl[1] = at_new; }
            __syncwarp();
            glm_wide_bar();
            glm_wide_classes(d, r, at_new ? beta_new : beta_s, 0, lane, cls);
            glm_wide_bar();
            double dev = 0;
            for (int c = 0; c < d.ncls; ++c) dev += cls[64 + c];
            // X'WX and X'z from the class sums, one packed entry per lane and round, classes added in order (the zeros of
            // the dense design rows add exact zeros: the same sums as the sparse accumulation of lev_sweep)
            for (int e = lane; e < pt + p; e += 32) {
                double s = 0;
                if (e < pt) {
                    int i2 = 0;
                    while ((i2 + 1) * (i2 + 2) / 2 <= e) ++i2;
                    const int j2 = e - i2 * (i2 + 1) / 2;
                    for (int c = 0; c < d.ncls; ++c) s += Xc[c * PM + i2] * Xc[c * PM + j2] * cls[c];
                    xt[e] = s;
                } else {
                    const int i2 = e - pt;
                    for (int c = 0; c < d.ncls; ++c) s += Xc[c * PM + i2] * cls[32 + c];
                    dz[i2] = s;
                }
            }
            __syncwarp();
            return dev;
        };
        double ymax = 0, sum_exprs = 0, sum_weight = 0;
        const double eg0 = exp(r.goff);
        for (int t = lane; t < d.n; t += 32) {
            const int c = d.col[t];
            const double y = r.y(c);
            if (y > ymax) ymax = y;
            const double curN = r.lib(c) * eg0, w = curN / (1 + r.phi(c) * curN);
            sum_exprs += y * w / curN;
            sum_weight += w;
        }
        ymax = lane_max<32>(ymax, 0xffffffffu); sum_exprs = lane_sum<32>(sum_exprs, 0xffffffffu); sum_weight = lane_sum<32>(sum_weight, 0xffffffffu);
        int sweeps = -1;
        if (ymax < CBS_LOW_VALUE) {
            if (lane == 0) for (int k = 0; k < p; ++k) beta_s[k] = nan("");
            __syncwarp();
        } else {
            const double c0 = log(sum_exprs / sum_weight);
            if (lane == 0) for (int k = 0; k < p; ++k) beta_s[k] = d.start[k] * c0;
            __syncwarp();
            double dev = sweep(0, xtwx, dl);
            double max_info = -1, lambda = 0;
            int iter = 0;
            sweeps = 1;
            while ((++iter) <= 250) {
                for (int k = 0; k < p; ++k) if (xtwx[tri(k, k)] > max_info) max_info = xtwx[tri(k, k)];
                if (iter == 1) { lambda = max_info * 1e-6; if (lambda < 1e-13) lambda = 1e-13; }
                int lev = 0;
                bool low_dev = false, failed = false;
                while (++lev) {
                    for (;;) {
                        for (int k = lane; k < pt; k += 32) work[k] = xtwx[k];
                        __syncwarp();
                        if (lane < p) { work[tri(lane, lane)] += lambda; dbeta[lane] = dl[lane]; }
                        __syncwarp();
                        if (chol_packed_warp(work, p, lane)) break;
                        lambda *= 10;
                        if (lambda <= 0) lambda = CBS_RIDICULOUSLY_LOW;
                    }
                    chol_solve_packed_warp(work, p, dbeta, lane);
                    if (lane < p) beta_new[lane] = beta_s[lane] + dbeta[lane];
                    __syncwarp();
                    const double dev_new = sweep(1, xtwx_new, dl_new);
                    ++sweeps;
                    if (dev_new / ymax < CBS_SUPREMELY_LOW) low_dev = true;
                    if (dev_new <= dev || low_dev) {
                        if (lane < p) beta_s[lane] = beta_new[lane];
                        __syncwarp();
                        dev = dev_new;
                        break;
                    }
                    lambda *= 2;
                    if (lambda <= 0) lambda = CBS_RIDICULOUSLY_LOW;
                    if (lambda / max_info > 1 / CBS_SUPREMELY_LOW) { failed = true; break; }
                }
                if (failed || low_dev) break;
                double divergence = 0;
                for (int k = 0; k < p; ++k) divergence += dl[k] * dbeta[k];
                if (divergence < 1e-6) break;
                if (lev == 1) lambda /= 10;
                __syncwarp();
                for (int k = lane; k < pt; k += 32) xtwx[k] = xtwx_new[k];     // information at the accepted coefficients
                if (lane < p) dl[lane] = dl_new[lane];
                __syncwarp();
            }
        }
        if (lane == 0) ctrl[0] = 2;
        __syncwarp();
        glm_wide_bar();                                       // releases the workers
        if (iter_hist && lane == 0) atomicAdd(&iter_hist[max(0, min(sweeps, 255))], 1);
        double beta[PM], cellbeta[PM];
        for (int k = 0; k < p; ++k) { beta[k] = beta_s[k]; cellbeta[k] = 0; }
        glm_store<32, PM>(d, r, g, G, lane, beta, cellbeta, coef, fitted, alpha_w, n_alpha, alpha_out);
    }
}

// glmFit.default(prior.count) on genes [g0,g1): coef (G x p, may be null) from the prior-count refit when
// prior_count > 0 else from the plain fit; fitted (G x N, may be null) from the plain fit; alpha (G, may be null)
// = sum_b alpha_w[b] * coef[g,b].  d_lib is the per-sample library size on the device (N); h_lib the host copy.
void run_glm_fit(const int32_t* cnt, long G, long g0, long g1, int N, const double* d_logoff,
                        const std::vector<double>& h_lib, const double* d_phi, const int* d_phi_col,
                        const DesignDev* d_design, int h_p, const double* d_goff, double prior_count, double* d_coef,
                        double* d_fitted, const double* d_alpha_w, int n_alpha, double* d_alpha, cudaStream_t st, int class_form) {
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
    const bool two_pass = g_tune.glm_scr && g_tune.glm_two_pass && (class_form & 1);
    const int phi_by_class = (class_form & 2) ? 1 : 0;
    int* d_strag = g_dev.get<int>("glm_strag", 2 * (size_t)(g1 - g0));
    int* d_nstrag = g_dev.get<int>("glm_nstrag", 2);
    // the fits this call asks for: the plain one (fitted values, or everything when prior_count == 0) and the predFC one
    GlmJobs jobs{};
    int nj = 0;
    auto add_job = [&](const double* yadd, double tp, int mode, double* coef, double* fitted, double* alpha) {
        jobs.j[nj] = GlmJob{yadd, tp, mode, coef, fitted, alpha, two_pass ? d_strag + (size_t)nj * (g1 - g0) : (int*)nullptr,
                            d_nstrag + nj};
        ++nj;
    };
    if (d_fitted || (prior_count <= 0 && (d_coef || d_alpha)))
        add_job(nullptr, 0.0, 0, prior_count > 0 ? (double*)nullptr : d_coef, d_fitted, prior_count > 0 ? (double*)nullptr : d_alpha);
    if (prior_count > 0 && (d_coef || d_alpha))
        add_job(d_yadd, two_prior, 1, d_coef, (double*)nullptr, d_alpha);
    if (!nj) return;
    if (nj == 1) jobs.j[1] = jobs.j[0];
    if (!lanes) {
        for (int k = 0; k < nj; ++k) {
            const GlmJob& jb = jobs.j[k];
            CBS_LAUNCH(k_glm_fit, grid, 128, 0, st, cnt, G, g0, g1, d_logoff, jb.yadd, jb.two_prior_over_meanlib, d_phi, d_phi_col,
                       d_design, d_goff, jb.mode, jb.coef, jb.fitted, d_alpha_w, n_alpha, jb.alpha_out);
        }
        return;
    }
    // one shared-memory copy of the Levenberg work arrays per gene (16 genes per block)
#define CBS_GLM_LANES(PM_) do { \
        const size_t smem = g_tune.glm_scr ? (size_t)(128 / GLM_LANES) * lev_scratch_doubles<PM_>() * sizeof(double) : 0; \
        t_ctx->set_func_attr((const void*)k_glm_fit_lanes<PM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, \
                             (int)((128 / GLM_LANES) * lev_scratch_doubles<PM_>() * sizeof(double))); \
        CBS_CUDA(cudaMemsetAsync(d_nstrag, 0, 2 * sizeof(int), st)); \
        CBS_LAUNCH(k_glm_fit_lanes<PM_>, dim3(lgrid.x, (unsigned)nj), 128, smem, st, cnt, G, g0, g1, d_logoff, jobs, d_phi, \
                   d_phi_col, d_design, d_goff, d_alpha_w, n_alpha, g_tune.glm_scr ? 1 : 0, phi_by_class, d_hist); \
        if (two_pass) { \
            const int stage = N <= 32768; \
            const size_t wsmem = glm_wide_smem_doubles<PM_>() * sizeof(double) + (stage ? (size_t)N * sizeof(int32_t) : 0); \
            t_ctx->set_func_attr((const void*)k_glm_fit_wide<PM_>, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024); \
            CBS_LAUNCH(k_glm_fit_wide<PM_>, dim3(GLM_WIDE_BLOCKS, (unsigned)nj), GLM_WIDE_THREADS, wsmem, st, cnt, G, g0, N, d_logoff, jobs, d_phi, d_phi_col, \
                       d_design, d_goff, d_alpha_w, n_alpha, d_hist, stage, phi_by_class); \
        } } while (0)
    if (h_p <= 4) CBS_GLM_LANES(4); else if (h_p <= 8) CBS_GLM_LANES(8); else if (h_p <= 12) CBS_GLM_LANES(12); else CBS_GLM_LANES(16);
#undef CBS_GLM_LANES
}

}  // namespace cbs
