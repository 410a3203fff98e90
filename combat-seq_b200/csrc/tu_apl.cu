// tu_apl.cu — dispersion estimation (a5/a6): Cox-Reid adjusted profile likelihood kernels and their stage drivers.
//
// Replaces, per batch, estimateGLMCommonDisp (/root/reference/ComBat_seq.R:100-107) and estimateGLMTagwiseDisp
// (/root/reference/ComBat_seq.R:110-119) of edgeR 3.22: APL evaluation + Brent's optimize() on delta = phi^(1/4),
// then the 11-point grid + FMM spline maximiser.
#include "engine.hpp"
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>

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

// ---- a5/a6, one-way designs: lane groups with the batch's counts staged once per gene ---------------------------------
// One gene = LANES consecutive lanes (2 for cells of up to 128 samples: every lane redoes the exp, the division and the
// butterfly of a Newton step, so short cells want few lanes; 8 for longer ones); a block stages the batch's row of its
// 128 / LANES genes in shared memory as doubles
// (converted once), together with the batch's library sizes, and then evaluates EVERY dispersion of the group (11 grid
// points, or the batch's current Brent abscissa) on the staged row: the counts are read from memory once per launch
// instead of once per evaluation, the Newton sweeps of glm_one_group run on shared-memory operands with the samples of
// a cell split over the lanes (XOR-butterfly sums: every lane of a gene holds the same bits, so the data-dependent
// iteration stays uniform inside the group), and everything that does not depend on the dispersion (start values,
// sum of lgamma(y + 1)) is computed once per gene.  Used when the batch's design is one-way (every within-batch design of
// the BASELINE configs but C3b), p <= APLL_PM cells and n <= APLL_MAXN samples; anything else takes k_apl.
constexpr int APLL_PM = 4, APLL_MAXN = 512;
constexpr int APLL_WS = APLL_PM + 1;      // doubles per (design, gene) of the warm-start array: the cells' coefficients, sum lgamma(y + 1)
struct AplGroups { int n; int start[APL_MAX_EVALS + 1]; };      // evaluations [start[k], start[k+1]) share one design
// (Measured and dropped: staging the counts as int32 and converting on use, which halves the shared memory and lets 5-6 blocks
// share an SM instead of 3, is 1-16 % slower on C5: the kernel is bound by its dependent FP64 chains, not by occupancy.)
typedef double apl_stage_t;
__host__ __device__ inline int apll_row_stride(int n, int lanes) { return (n + 15) / 16 * 16 + lanes; }     // = lanes mod 16: the 16 / lanes genes of a half-warp hit disjoint banks
__device__ __forceinline__ double apl_yval(const apl_stage_t* y, int t) { return y[t]; }
__host__ __device__ inline size_t apll_smem_bytes(int n, int lanes) {
    const size_t rows = ((size_t)(128 / lanes) * apll_row_stride(n, lanes) * sizeof(apl_stage_t) + 15) / 16 * 16;
    return rows + 3 * (size_t)n * sizeof(double);
}

template <int APLL_LANES>
__global__ void __launch_bounds__(128)
k_apl_lanes(const int32_t* __restrict__ cnt, long G, long g0, long g1, const double* __restrict__ logoff,
            const DesignDev* __restrict__ designs, const unsigned char* __restrict__ sel, AplLaunch L, AplGroups GR,
            double* __restrict__ apl, const double* __restrict__ lgfact, const double* __restrict__ lgr,
            double* __restrict__ warm) {
    constexpr int APLL_GENES = 128 / APLL_LANES;
    extern __shared__ __align__(16) unsigned char apl_sm[];
    const int e0 = GR.start[blockIdx.y], e1 = GR.start[blockIdx.y + 1];
    const int di = L.design[e0];
    const DesignDev d = designs[di];
    const int n = d.n, p = d.p, ns = apll_row_stride(n, APLL_LANES);
    apl_stage_t* y_s = reinterpret_cast<apl_stage_t*>(apl_sm);                       // [genes][ns]
    double* lib_s = reinterpret_cast<double*>(apl_sm + ((size_t)APLL_GENES * ns * sizeof(apl_stage_t) + 15) / 16 * 16);   // [n] library sizes, [n] log library sizes, [n] columns (design order)
    double* off_s = lib_s + n;
    int* col_s = reinterpret_cast<int*>(off_s + n);
    const long gbase = g0 + (long)blockIdx.x * APLL_GENES;
    for (int t = threadIdx.x; t < n; t += blockDim.x) { const int c = d.col[t]; off_s[t] = logoff[2 * c]; lib_s[t] = logoff[2 * c + 1]; col_s[t] = c; }
    __syncthreads();
    // consecutive genes per sample: whole sectors.  Eight loads are in flight per thread before the first conversion (the
    // loop was the top stall of the kernel with one dependent load at a time; ncu, profiles/r02_ncu_k_apl_lanes_a.txt).
    for (int k0 = threadIdx.x; k0 < APLL_GENES * n; k0 += 8 * 128) {
        int v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int k = k0 + u * 128, gi = k & (APLL_GENES - 1), t = k / APLL_GENES;
            const long g = gbase + gi;
            v[u] = (k < APLL_GENES * n && g < g1) ? __ldg(cnt + g + (long)col_s[t] * G) : 0;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            const int k = k0 + u * 128, gi = k & (APLL_GENES - 1), t = k / APLL_GENES;
            if (k < APLL_GENES * n) y_s[gi * ns + t] = __hiloint2double(0x43300000, v[u]) - 4503599627370496.0;     // exact int -> double
        }
    }
    __syncthreads();
    const int gslot = threadIdx.x / APLL_LANES, ln = threadIdx.x % APLL_LANES;
    const long g = gbase + gslot;
    if (g >= g1) return;
    const unsigned gm = ((1u << APLL_LANES) - 1u) << ((threadIdx.x & 31) - ln);
    if (sel && !sel[(long)di * G + g]) {
        for (int e = e0 + ln; e < e1; e += APLL_LANES) apl[(long)e * G + g] = 0.0;
        return;
    }
    const apl_stage_t* y = y_s + gslot * ns;
    // Newton start of every cell: glm_one_group's own start value for the first evaluation of a gene, afterwards the
    // coefficient the previous evaluation converged to (the previous grid point, or the previous Brent abscissa of the batch
    // through `warm`, [design][gene][cells, sum lgamma(y + 1)]).  The iteration still runs to |step| < 1e-10, so it ends at the
    // same root (the likelihood is evaluated at a stationary point: a 1e-10 difference in the coefficient moves it by 1e-20),
    // in 2-3 steps instead of 6-9.  A cell without counts has no root: its coefficient is -1e8 in `warm` as in the result.
    double cbw[APLL_PM];
    double* wg = warm ? warm + ((long)di * G + g) * APLL_WS : nullptr;
#pragma unroll
    for (int l = 0; l < APLL_PM; ++l) cbw[l] = (wg && l < p) ? wg[l] : NAN;
    // dispersion-independent: glm_one_group's start value per cell and sum_s lgamma(y_s + 1); both are skipped from the second
    // Brent round on (ncu: a fifth of a round's stall samples), the sum comes back from `warm`
    double start0[APLL_PM];
    double sum_lgy1 = wg ? wg[APLL_PM] : NAN;
    bool cached = sum_lgy1 == sum_lgy1;
#pragma unroll
    for (int l = 0; l < APLL_PM; ++l) if (l < p && !(cbw[l] == cbw[l])) cached = false;
    if (cached) {
#pragma unroll
        for (int l = 0; l < APLL_PM; ++l) start0[l] = cbw[l] > -1e7 ? 0.0 : -INFINITY;     // (only "has counts" matters: cbw is the start)
    } else {
        sum_lgy1 = 0;
#pragma unroll 1
        for (int l = 0; l < p; ++l) {
            const int t0 = d.cell_start[l], t1 = d.cell_start[l + 1];
            double cur = 0, nz = 0;
            for (int t = t0 + ln; t < t1; t += APLL_LANES) {
                const double yy = apl_yval(y, t);
                if (yy > CBS_LOW_VALUE) { cur += yy / lib_s[t]; nz = 1; }
                sum_lgy1 += (lgfact && yy < (double)LGFACT_N) ? lgfact[(int)yy] : lgamma(yy + 1);
            }
            cur = lane_sum<APLL_LANES>(cur, gm); nz = lane_max<APLL_LANES>(nz, gm);
            start0[l] = nz > 0 ? log(cur / (double)(t1 - t0)) : -INFINITY;
        }
        sum_lgy1 = lane_sum<APLL_LANES>(sum_lgy1, gm);
    }

    for (int e = e0; e < e1; ++e) {
        const double phi = L.phi[e];
        const double* lgr_e = lgr ? lgr + (long)e * LGFACT_N : nullptr;
        const double rr = phi > 0 ? 1 / phi : 0.0, logrr = phi > 0 ? log(rr) : 0.0, lgrr = phi > 0 ? lgamma(rr) : 0.0;
        double Wc[APLL_PM];
        double ll = 0, nlive = 0, sumw = 0;
        // (not unrolled: four copies of the Newton and likelihood loops with their inlined exp / log / reciprocal made the kernel
        // 160 KB of code and a fifth of its stall samples were instruction fetches; ncu "no_instructions")
#pragma unroll 1
        for (int l = 0; l < p; ++l) {
            const int t0 = d.cell_start[l], t1 = d.cell_start[l + 1];
            // edgeR glm_one_group: Newton on the cell's coefficient, maxit 50, |step| < 1e-10
            double cur = start0[l];
            if (cur > -INFINITY) {
                if (cbw[l] == cbw[l] && cbw[l] > -1e7) cur = cbw[l];
                for (int it = 0; it < 50; ++it) {
                    const double ec = exp(cur);
                    double dl = 0, info = 0;
#pragma unroll 4
                    for (int t = t0 + ln; t < t1; t += APLL_LANES) {         // (independent reciprocal chains: the loop is latency-bound)
                        const double mu = ec * lib_s[t], inv = rcp_pos(1 + mu * phi);
                        dl += (apl_yval(y, t) - mu) * inv;
                        info += mu * inv;
                    }
                    dl = lane_sum<APLL_LANES>(dl, gm); info = lane_sum<APLL_LANES>(info, gm);
                    const double step = dl / info;                // (info may be 0 or tiny for a degenerate cell: the full division)
                    cur += step;
                    if (fabs(step) < 1e-10) break;
                }
            }
            const double cb = fmax(cur, -1e8);                   // mglmOneWay: pmax(beta, -1e8)
            cbw[l] = cb;
            const double ecell = exp(cb);
            double w = 0;
#pragma unroll 2
            for (int t = t0 + ln; t < t1; t += APLL_LANES) {
                const double mu = ecell * lib_s[t];
                if (mu == 0) continue;
                const double yy = apl_yval(y, t), logmu = cb + off_s[t];
                if (phi > 0) {
                    const double logmur = log_pos(mu + rr);
                    const double lgyr = (lgr_e && yy < (double)LGFACT_N) ? lgr_e[(int)yy] : lgamma(yy + rr);
                    ll += yy * logmu - yy * logmur - rr * logmur + lgyr;
                } else {
                    ll += yy * logmu - mu;
                }
                nlive += 1;
                w += mu * rcp_pos(1 + phi * mu);
            }
            w = lane_sum<APLL_LANES>(w, gm);
            Wc[l] = w;
            sumw += w;
        }
        ll = lane_sum<APLL_LANES>(ll, gm); nlive = lane_sum<APLL_LANES>(nlive, gm);
        const double sum_loglik = ll + nlive * (rr * logrr - lgrr) - sum_lgy1;      // (samples with mu == 0 have y == 0: lgamma(1) = 0)
        double adj;
        if (p == 1) adj = log(fabs(sumw)) / 2;
        else {
            double wm[APLL_PM * (APLL_PM + 1) / 2];
            for (int i = 0; i < p * (p + 1) / 2; ++i) wm[i] = 0;
#pragma unroll
            for (int l = 0; l < APLL_PM; ++l) {                    // X'WX from the cell sums of W (one design row per cell)
                if (l >= p) break;
                const int t0 = d.cell_start[l];
                for (int k = d.rowptr[t0]; k < d.rowptr[t0 + 1]; ++k)
                    for (int k2 = d.rowptr[t0]; k2 <= k; ++k2) wm[tri(d.cidx[k], d.cidx[k2])] += d.cval[k] * d.cval[k2] * Wc[l];
            }
            sytf2_packed(wm, p);
            const double log_low = log(CBS_LOW_VALUE);
            double sld = 0;
            for (int i = 0; i < p; ++i) {
                const double v = wm[tri(i, i)];
                sld += (v < CBS_LOW_VALUE || !isfinite(v)) ? log_low : log(v);
            }
            adj = sld * 0.5;
        }
        if (ln == 0) apl[(long)e * G + g] = sum_loglik - adj;
    }
    if (wg && ln == 0) {
#pragma unroll
        for (int l = 0; l < APLL_PM; ++l) if (l < p) wg[l] = cbw[l];
        wg[APLL_PM] = sum_lgy1;
    }
}

// APL of genes [g0,g1) for every job -> d_apl[job*G + g]
void launch_apl(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignDev* designs,
                const unsigned char* sel, const double* phi_gene, const std::vector<AplJob>& jobs, double* d_apl,
                cudaStream_t st, const std::vector<DesignDev>* h_designs, double* d_warm) {
    if (g1 <= g0) return;
    // lane-group kernel: scalar dispersions, every design of the launch one-way with few cells and a batch that fits shared memory
    bool lanes_ok = g_tune.apl_lanes && !phi_gene && h_designs != nullptr && !jobs.empty();
    int nmax = 0;
    if (lanes_ok)
        for (const AplJob& j : jobs) {
            const DesignDev& hd = (*h_designs)[j.design];
            lanes_ok = lanes_ok && hd.oneway && hd.p <= APLL_PM && hd.n <= APLL_MAXN && j.phi > 0;
            nmax = std::max(nmax, hd.n);
        }
    #ifndef CBS_APL_SHORT_LANES
#define CBS_APL_SHORT_LANES 2
#endif
    const int lanes = nmax <= 256 ? CBS_APL_SHORT_LANES : 8, genes = 128 / lanes;          // (cells are at most as long as the batch)
    lanes_ok = lanes_ok && apll_smem_bytes(nmax, lanes) <= 200 * 1024;
    if (lanes_ok) {
        double* lgfact = g_dev.get<double>("lgfact", (size_t)LGFACT_N);
        CBS_LAUNCH(k_lgfact_init, grid1(LGFACT_N, 256), 256, 0, st, lgfact);
        const size_t smem = apll_smem_bytes(nmax, lanes);
        t_ctx->set_func_attr((const void*)k_apl_lanes<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        t_ctx->set_func_attr((const void*)k_apl_lanes<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        t_ctx->set_func_attr((const void*)k_apl_lanes<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
        for (size_t e0 = 0; e0 < jobs.size();) {
            AplLaunch L; AplGroups GR;
            L.n_evals = 0; GR.n = 0;
            // whole groups (runs of equal design) per launch
            while (e0 < jobs.size()) {
                size_t e1 = e0;
                while (e1 < jobs.size() && jobs[e1].design == jobs[e0].design && e1 - e0 < 16) ++e1;
                if (L.n_evals + (int)(e1 - e0) > APL_MAX_EVALS) break;
                GR.start[GR.n++] = L.n_evals;
                for (size_t e = e0; e < e1; ++e) { L.design[L.n_evals] = jobs[e].design; L.phi[L.n_evals] = jobs[e].phi; ++L.n_evals; }
                e0 = e1;
            }
            GR.start[GR.n] = L.n_evals;
            const size_t first = e0 - (size_t)L.n_evals;
            double* t = g_dev.get<double>("lgr", (size_t)APL_MAX_EVALS * LGFACT_N);
            CBS_LAUNCH(k_lgr_init, dim3(LGFACT_N / 256, (unsigned)L.n_evals), 256, 0, st, L, t);
            dim3 grid((unsigned)((g1 - g0 + genes - 1) / genes), (unsigned)GR.n);
            if (lanes == 4)
                CBS_LAUNCH(k_apl_lanes<4>, grid, 128, smem, st, cnt, G, g0, g1, logoff, designs, sel, L, GR, d_apl + first * G,
                           (const double*)lgfact, (const double*)t, d_warm);
            else if (lanes == 2)
                CBS_LAUNCH(k_apl_lanes<2>, grid, 128, smem, st, cnt, G, g0, g1, logoff, designs, sel, L, GR, d_apl + first * G,
                           (const double*)lgfact, (const double*)t, d_warm);
            else
                CBS_LAUNCH(k_apl_lanes<8>, grid, 128, smem, st, cnt, G, g0, g1, logoff, designs, sel, L, GR, d_apl + first * G,
                           (const double*)lgfact, (const double*)t, d_warm);
        }
        return;
    }
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
    // Newton warm starts across the Brent rounds (k_apl_lanes): NaN = "no previous evaluation"
    double* d_warm = g_dev.get<double>("apl_warm", (size_t)nb * G * APLL_WS);
    CBS_CUDA(cudaMemsetAsync(d_warm, 0xff, (size_t)nb * G * APLL_WS * sizeof(double), st));
    int steps = 0;
    for (;;) {
        std::vector<AplJob> jobs;
        std::vector<int> who;
        for (int b = 0; b < nb; ++b)
            if (active[b] && !br[b].done) { const double x = br[b].query(); jobs.push_back({b, x * x * x * x}); who.push_back(b); }
        if (jobs.empty()) break;
        launch_apl(cnt, G, g0, g1, logoff, ds.d_array, sel, nullptr, jobs, d_apl, st, &ds.h_array, d_warm);
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
    if (!jobs.empty()) launch_apl(cnt, G, g0, g1, logoff, ds.d_array, sel, nullptr, jobs, d_apl, st, &ds.h_array);
    for (int b = 0; b < nb; ++b) {
        if (do_batch[b]) {
            CBS_LAUNCH(k_tagwise_max, grid1(g1 - g0, 128), 128, 0, st, d_apl + (size_t)job0[b] * G, G, g0, g1,
                       sel ? sel + (size_t)b * G : nullptr, common[b], d_phi_hat + (size_t)b * G);
        } else {
            CBS_LAUNCH(k_fill_apl, grid1(g1 - g0, 256), 256, 0, st, d_phi_hat + (size_t)b * G + g0, g1 - g0, common[b]);
        }
    }
}

// ---- prior.df > 0: weighted-likelihood empirical-Bayes moderation of the tagwise dispersions -----------------------------
// edgeR estimateGLMTagwiseDisp(prior.df) -> dispCoxReidInterpolateTagwise: the APL grid of every gene is blended with the
// trend of the grid over abundance, apl.smooth = (apl + prior.n * trend) / (1 + prior.n), prior.n = prior.df / (nlibs - ncoefs),
// trend = moving average (width floor(span * ntags), span = (10 / ntags)^0.23) of the APL rows ordered by AveLogCPM.
// ComBat-seq itself always passes prior.df = 0 (/root/reference/ComBat_seq.R:116), where this is the identity; the
// capability is exposed through cbs_tagwise_disp_prior.

// aveLogCPM(y, prior.count = 2, dispersion = 0.05): key of the abundance order; genes outside the selection sort last
__global__ void __launch_bounds__(128)
k_ave_log_cpm(const int32_t* __restrict__ cnt, long G, const double* __restrict__ logoff, const double* __restrict__ yadd,
              const DesignDev* __restrict__ onecell, const unsigned char* __restrict__ sel, double goff,
              double* __restrict__ key, int* __restrict__ idx) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    idx[g] = (int)g;
    if (sel && !sel[g]) { key[g] = INFINITY; return; }
    const DesignDev d = *onecell;
    RowSrc src{cnt, G, logoff, yadd, nullptr, nullptr, 0.05};
    Row r(src, g, goff);
    const double beta = fit_one_group(d, r, 0, d.n, 50, 1e-10);
    key[g] = (beta + log(1e6)) / log(2.0);
}

// colbuf[k * Gs + r] = apl[k * G + order[r]]: the grid columns in abundance order
__global__ void k_gather_sorted(const double* __restrict__ apl, long G, const int* __restrict__ order, long Gs,
                                double* __restrict__ colbuf) {
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (r < Gs) colbuf[(long)k * Gs + r] = apl[(long)k * G + order[r]];
}

// movingAverageByCol(full.length = TRUE) from the inclusive prefix sums cs of the sorted columns, blended into apl in place
__global__ void k_blend_trend(const double* __restrict__ cs, const int* __restrict__ order, long G, long Gs, int width,
                              double prior_n, double* __restrict__ apl) {
    const long r = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int k = blockIdx.y;
    if (r >= Gs) return;
    const long half1 = (width + 1) / 2, half2 = width / 2;
    const long lo = max(0L, r - half1 + 1), hi = min(Gs - 1, r + half2);
    const double* c = cs + (long)k * Gs;
    const double trend = (c[hi] - (lo > 0 ? c[lo - 1] : 0.0)) / (double)(hi - lo + 1);
    double* a = apl + (long)k * G + order[r];
    *a = (*a + prior_n * trend) / (1 + prior_n);
}

// estimateGLMTagwiseDisp(y, design, dispersion = common, prior.df) for ONE batch on one device (every gene of the batch
// takes part in the trend, so the stage does not shard by genes).  ds holds the batch's design, onecell an intercept-only
// design over the same columns.
void run_tagwise_disp_prior(const int32_t* cnt, long G, const double* logoff, const std::vector<double>& h_lib,
                            const DesignSet& ds, const DesignSet& onecell, const unsigned char* sel, double common,
                            double prior_df, cudaStream_t st, double* d_phi_hat) {
    std::vector<AplJob> jobs;
    for (int k = 0; k < 11; ++k) jobs.push_back({0, common * std::pow(2.0, -6.0 + k * (12.0 / 10))});
    double* d_apl = g_dev.get<double>("apl_grid", jobs.size() * (size_t)G);
    launch_apl(cnt, G, 0, G, logoff, ds.d_array, sel, nullptr, jobs, d_apl, st, &ds.h_array);
    const int n = ds.h_array[0].n, p = ds.h_array[0].p;
    if (prior_df > 0) {
        // abundance order of the selected genes
        double mean_lib = 0;
        for (int s = 0; s < n; ++s) mean_lib += h_lib[s];
        mean_lib /= n;
        std::vector<double> yadd(n);
        for (int s = 0; s < n; ++s) yadd[s] = 2.0 * h_lib[s] / mean_lib;                     // add_prior.cpp, prior.count = 2
        double* d_yadd = upload(g_dev, "alc_yadd", yadd.data(), (size_t)n, st);
        double* key = g_dev.get<double>("alc_key", (size_t)G);
        double* key2 = g_dev.get<double>("alc_key2", (size_t)G);
        int* idx = g_dev.get<int>("alc_idx", (size_t)G);
        int* order = g_dev.get<int>("alc_order", (size_t)G);
        CBS_LAUNCH(k_ave_log_cpm, grid1(G, 128), 128, 0, st, cnt, G, logoff, d_yadd, onecell.d_array, sel, std::log1p(4.0 / mean_lib), key, idx);
        size_t tmp_bytes = 0;
        CBS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, key, key2, idx, order, (int)G, 0, 64, st));
        void* tmp = g_dev.get("alc_sort_tmp", tmp_bytes);
        CBS_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, key, key2, idx, order, (int)G, 0, 64, st));
        // number of selected genes (they sort before the +inf keys)
        std::vector<unsigned char> h_sel(G, 1);
        if (sel) CBS_CUDA(cudaMemcpyAsync(h_sel.data(), sel, (size_t)G, cudaMemcpyDeviceToHost, st));
        CBS_CUDA(cudaStreamSynchronize(st));
        long Gs = 0;
        for (long g = 0; g < G; ++g) Gs += h_sel[g] ? 1 : 0;
        const double prior_n = prior_df / (double)(n - p);
        const double span = G > 10 ? std::pow(10.0 / (double)G, 0.23) : 1.0;
        int width = (int)std::floor(span * (double)Gs);
        if (width > Gs) width = (int)Gs;
        if (Gs > 0 && width > 1) {
            double* colbuf = g_dev.get<double>("alc_cols", (size_t)11 * Gs);
            double* cs = g_dev.get<double>("alc_cs", (size_t)11 * Gs);
            CBS_LAUNCH(k_gather_sorted, dim3((unsigned)((Gs + 255) / 256), 11), 256, 0, st, d_apl, G, order, Gs, colbuf);
            size_t scan_bytes = 0;
            CBS_CUDA(cub::DeviceScan::InclusiveSum(nullptr, scan_bytes, colbuf, cs, (int)Gs, st));
            void* stmp = g_dev.get("alc_scan_tmp", scan_bytes);
            for (int k = 0; k < 11; ++k)
                CBS_CUDA(cub::DeviceScan::InclusiveSum(stmp, scan_bytes, colbuf + (size_t)k * Gs, cs + (size_t)k * Gs, (int)Gs, st));
            CBS_LAUNCH(k_blend_trend, dim3((unsigned)((Gs + 255) / 256), 11), 256, 0, st, cs, order, G, Gs, width, prior_n, d_apl);
        } else if (Gs > 0) {
            // width <= 1: movingAverageByCol returns its input, the blend is the identity
        }
    }
    CBS_LAUNCH(k_tagwise_max, grid1(G, 128), 128, 0, st, d_apl, G, 0L, G, sel, common, d_phi_hat);
}

}  // namespace cbs
