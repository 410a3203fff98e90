// tu_mc.cu — the shrink=TRUE empirical-Bayes stage (a11): monte_carlo_int_NB for every batch.
//
// Replaces monte_carlo_int_NB (/root/reference/helper_seq.R:8-42) and its driver loop (/root/reference/ComBat_seq.R:142-165).
// Two paths: the FP64 tensor-core contraction over abundance-sorted genes with screened source-tile lists (mc_gemm.cuh,
// tu_screen.cu) and the exact pairwise kernel (mc.cuh: gene.subset.n, or a batch whose Chebyshev tail misses the tolerance).
#include "engine.hpp"
#include "mc.cuh"
#include "mc_gemm.cuh"
#include "screen.hpp"
#include <cub/device/device_radix_sort.cuh>
#include <mutex>

namespace cbs {

// number of Chebyshev terms for the range h = (uhi - ulo)/2 of u = log(r + c): the series of lgamma(x + e^u - c)
// converges like rho^-K with rho fixed by the singularities at Im u = +-pi above the end points of the interval.
static int mcg_terms(double h) {
    if (h < 1e-3) return 8;
    const double wr = 1.0, wi = M_PI / h;                     // w = 1 + i pi/h
    const double ar = wr * wr - wi * wi - 1.0, ai = 2 * wr * wi;   // w^2 - 1
    const double mod = std::sqrt(std::hypot(ar, ai)), arg = 0.5 * std::atan2(ai, ar);
    const double rho = std::hypot(wr + mod * std::cos(arg), wi + mod * std::sin(arg));
    int K = (int)std::ceil(std::log(1e14) / std::log(rho)) + 4;
    K = (K + 1) & ~1;
    return std::max(8, K);
}

static std::string sl(const char* name, int slot) { return std::string(name) + (slot ? "#1" : "#0"); }

// monte_carlo_int_NB for one batch by the direct pairwise kernel (every lgamma evaluated), targets [i0,i1)
static void run_mc_batch_direct(const int32_t* cnt, long G, int n, const int* d_cols, long i0, long i1, const double* d_mu_hat,
                                const double* d_gamma, const double* d_phi, int mc_mode, int subset_n, uint64_t seed,
                                double* d_gamma_star, double* d_phi_star, unsigned long long* d_nfb, cudaStream_t st) {
    double* A = g_dev.get<double>("mc_A", (size_t)n * G);
    double* cj = g_dev.get<double>("mc_cj", (size_t)G);
    double* rj = g_dev.get<double>("mc_rj", (size_t)G);
    double* lgr = g_dev.get<double>("mc_lgr", (size_t)G);
    double* di = g_dev.get<double>("mc_di", (size_t)G);
    CBS_LAUNCH(k_mc_prep_src, grid1(G, 128), 128, 0, st, d_mu_hat, G, n, d_cols, d_phi, A, cj, rj, lgr);
    CBS_LAUNCH(k_mc_prep_tgt, grid1(G, 128), 128, 0, st, cnt, G, n, d_cols, di);
    if (i1 <= i0) return;
    if (subset_n >= G - 1) subset_n = 0;      // sample(1:(G-1), G-1) is a permutation of all other genes
    const unsigned nblk = (unsigned)((i1 - i0 + MC_TI - 1) / MC_TI);
    // the targets' counts sit in shared memory (64 n bytes) when they fit; very wide batches read them through L1 instead
    size_t smem = (size_t)MC_TI * n * sizeof(double);
    const int xs_shared = smem <= 160 * 1024;
    if (!xs_shared) smem = 0;
    t_ctx->set_func_attr((const void*)k_mc_direct, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024);
    CBS_LAUNCH(k_mc_direct, dim3(nblk), 256, smem, st, cnt, G, n, d_cols, i0, i1, A, cj, rj, lgr, di, d_gamma, d_phi,
               mc_mode == CBS_MC_STABLE ? 1 : 0, subset_n, seed, xs_shared, d_gamma_star, d_phi_star, d_nfb);
}

// What the planning half of a contraction batch leaves for the execution half.
struct McPlan {
    bool ok = false;                 // false: the batch cannot use the contraction (range / too many terms)
    McgShape sh{};
    McgOwn own{0, 1};
    int slot = 0;
    double *P = nullptr, *Q = nullptr;
    double2* gp = nullptr;
    int *perm = nullptr, *srclist = nullptr, *count = nullptr, *tiles = nullptr, *item_start = nullptr;
    int2* items = nullptr;
    int* h_out = nullptr;            // pinned: n_items, n_owned_tiles, sum of counts, (tail bits: 2 ints)
    cudaEvent_t ready = nullptr;
};

// Planning half: sort, operands, screening, source-tile lists and work items of one batch, all enqueued on `st`
// without a host synchronisation; the counts the host needs are copied to pinned memory before the `ready` event.
static McPlan plan_mc_gemm(const int32_t* cnt, long G, int n, const int* d_cols, const double* d_mu_hat,
                           const double* d_gamma, const double* d_phi, double phi_lo, double phi_hi, int slot,
                           cudaStream_t st, cbs_comm* comm) {
    McPlan pl;
    pl.slot = slot;
    if (!(phi_lo > 0) || !std::isfinite(phi_hi)) return pl;
    McgShape& sh = pl.sh;
    sh.n = n;
    sh.ulo = std::log(1.0 / phi_hi + MCG_SHIFT);
    sh.uhi = std::log(1.0 / phi_lo + MCG_SHIFT);
    if (sh.uhi - sh.ulo < 0.02) { const double mid = 0.5 * (sh.ulo + sh.uhi); sh.ulo = mid - 0.01; sh.uhi = mid + 0.01; }
    sh.Kc = mcg_terms(0.5 * (sh.uhi - sh.ulo));
    if (sh.Kc > MCG_MAX_KC) return pl;
    sh.KT = n + sh.Kc + MCG_SHIFT + 1;
    // K is padded to whole 8-wide k-chunks only; at least two full hand-overs per source tile (gamma/phi buffering)
    sh.KTP = std::max((sh.KT + MCG_KC - 1) / MCG_KC * MCG_KC, 2 * MCG_KC * MCG_SUB);
    static_assert((MCG_GPBUF - 1) * 2 >= MCG_STAGES + 1, "gamma/phi buffers must outlive the stage ring");
    sh.rows = (G + MCG_TILE - 1) / MCG_TILE * MCG_TILE;
    if (!comm || comm->rank == 0) g_diag.mc_kc = sh.Kc;
    pl.own = McgOwn{comm ? comm->rank : 0, comm ? comm->world : 1};
    const McgOwn own = pl.own;
    const long NJ = sh.rows / MCG_TJ, ntiles64 = sh.rows / MCG_TI;
    const long own_rows = own.nblocks(sh.rows) * MCG_TILE;

    // abundance order of the batch
    float* key = g_dev.get<float>(sl("mcg_key", slot), (size_t)G);
    float* key2 = g_dev.get<float>(sl("mcg_key2", slot), (size_t)G);
    int* idx = g_dev.get<int>(sl("mcg_idx", slot), (size_t)G);
    pl.perm = g_dev.get<int>(sl("mcg_perm", slot), (size_t)sh.rows);
    CBS_LAUNCH(k_mcg_sort_key, grid1(G, 128), 128, 0, st, d_mu_hat, G, n, d_cols, key, idx);
    size_t tmp_bytes = 0;
    CBS_CUDA(cub::DeviceRadixSort::SortPairs(nullptr, tmp_bytes, key, key2, idx, pl.perm, (int)G, 0, 32, st));
    void* tmp = g_dev.get(sl("mcg_sort_tmp", slot), tmp_bytes);
    CBS_CUDA(cub::DeviceRadixSort::SortPairs(tmp, tmp_bytes, key, key2, idx, pl.perm, (int)G, 0, 32, st));
    g_launches.fetch_add(1, std::memory_order_relaxed);

    std::vector<double> dct((size_t)sh.Kc * sh.Kc);
    for (int k = 0; k < sh.Kc; ++k)
        for (int m = 0; m < sh.Kc; ++m) dct[(size_t)k * sh.Kc + m] = (k ? 2.0 : 1.0) / sh.Kc * std::cos(M_PI * k * (m + 0.5) / sh.Kc);
    double* h_dct = (double*)g_pin.get(sl("mcg_dct_h", slot), dct.size() * sizeof(double));
    memcpy(h_dct, dct.data(), dct.size() * sizeof(double));
    double* d_dct = upload(g_dev, sl("mcg_dct", slot), h_dct, dct.size(), st);
    pl.P = g_dev.get<double>(sl("mcg_P", slot), (size_t)sh.KTP * sh.rows);
    pl.Q = g_dev.get<double>(sl("mcg_Q", slot), (size_t)sh.KTP * sh.rows);
    double* Fn = g_dev.get<double>(sl("mcg_Fnode", slot), (size_t)sh.Kc * G);
    pl.gp = (double2*)g_dev.get<double>(sl("mcg_gp", slot), (size_t)2 * sh.rows);
    unsigned long long* d_tail = g_dev.get<unsigned long long>(sl("mcg_tail", slot), 1);
    CBS_CUDA(cudaMemsetAsync(d_tail, 0, sizeof(unsigned long long), st));
    double2* d_part = (double2*)g_dev.get<double>(sl("mcg_srcpart", slot), (size_t)2 * ((n + MCG_KC - 1) / MCG_KC) * sh.rows);
    float4* d_qsum = (float4*)g_dev.get<float>(sl("mcg_qsum", slot), (size_t)4 * sh.rows);
    CBS_LAUNCH(k_mcg_prep_src, dim3((unsigned)(sh.rows / 128), (unsigned)(sh.KTP / MCG_KC)), 128, 0, st, d_mu_hat, G, d_cols, d_phi, pl.perm, sh,
               pl.Q, d_part);
    CBS_LAUNCH(k_mcg_prep_src_fin, grid1(sh.rows, 128), 128, 0, st, G, d_phi, (const int*)pl.perm, sh, (const double2*)d_part, pl.Q, d_qsum);
    CBS_LAUNCH(k_mcg_pack_gp, grid1(sh.rows, 256), 256, 0, st, d_gamma, d_phi, pl.perm, G, sh.rows, pl.gp);
    if (own_rows > 0) {                     // P is only needed for this rank's target blocks (Q for all sources)
        int32_t* xs = g_dev.get<int32_t>(sl("mcg_xs", slot), (size_t)own_rows * n);
        CBS_LAUNCH(k_mcg_gather_counts, dim3((unsigned)((own_rows + 255) / 256), (unsigned)n), 256, 0, st, cnt, G, d_cols, pl.perm, sh, own, own_rows, xs);
        double* d_lgtab = g_dev.get<double>(sl("mcg_lgtab", slot), (size_t)sh.Kc * MCG_LGTAB);
        CBS_LAUNCH(k_mcg_node_table, dim3(MCG_LGTAB / 256, (unsigned)sh.Kc), 256, 0, st, sh, d_lgtab);
        CBS_LAUNCH(k_mcg_prep_nodes, dim3((unsigned)(own_rows / 128), (unsigned)sh.Kc), 128, 0, st, xs, G, sh, own, own_rows, d_lgtab, Fn);
        CBS_LAUNCH(k_mcg_prep_tgt, dim3((unsigned)((own_rows + 127) / 128), (unsigned)(sh.KTP / MCG_KC)), 128, 0, st, xs, G, sh, own, own_rows, Fn, d_dct, pl.P, d_tail);
    }
    // the series' tail is judged over ALL targets, so that every rank takes the decision a single GPU would
    if (comm && comm->world > 1) CBS_NCCL(nccl_api()->AllReduce(d_tail, d_tail, 1, NCCL_UINT64, NCCL_MAX, comm->nccl, st));

    // source-tile lists: screened (tcgen05 pass over all pairs in bf16x2) or all tiles
    pl.srclist = g_dev.get<int>(sl("mcg_srclist", slot), (size_t)ntiles64 * NJ);
    pl.count = g_dev.get<int>(sl("mcg_count", slot), (size_t)ntiles64);
    bool screened = false;
    if (g_tune.mc_band && own_rows > 0) {
        float* hi = g_dev.get<float>(sl("mcs_hi", slot), (size_t)NJ * sh.rows);
        float* lo = g_dev.get<float>(sl("mcs_lo", slot), (size_t)NJ * sh.rows);
        screened = mc_screen(pl.P, pl.Q, d_qsum, sh, own, slot, hi, lo, st);
        if (screened) {
            t_ctx->set_func_attr((const void*)k_mcg_list_screened, cudaFuncAttributeMaxDynamicSharedMemorySize, 64 * 1024);
            CBS_LAUNCH(k_mcg_list_screened, dim3((unsigned)ntiles64), 256, (size_t)NJ, st, hi, lo, sh.rows, (int)NJ,
                       (float)MCG_PRUNE, own, pl.srclist, pl.count);
        }
    }
    if (!screened) CBS_LAUNCH(k_mcg_list_all, dim3((unsigned)ntiles64), 128, 0, st, (int)NJ, pl.srclist, pl.count);
    pl.tiles = g_dev.get<int>(sl("mcg_tiles", slot), (size_t)ntiles64 + 1);
    pl.item_start = g_dev.get<int>(sl("mcg_item_start", slot), (size_t)ntiles64 + 2);
    pl.items = (int2*)g_dev.get<int>(sl("mcg_items", slot), (size_t)2 * ntiles64 * ((NJ + MCG_SEG - 1) / MCG_SEG));
    int* d_out = g_dev.get<int>(sl("mcg_out", slot), 8);
    CBS_LAUNCH(k_mcg_items, dim3(1), 1024, 0, st, pl.count, (int)ntiles64, own, pl.tiles, pl.item_start, pl.items, d_out);
    pl.h_out = (int*)g_pin.get(sl("mcg_out_h", slot), 8 * sizeof(int));
    CBS_CUDA(cudaMemcpyAsync(pl.h_out, d_out, 3 * sizeof(int), cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaMemcpyAsync(pl.h_out + 4, d_tail, sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    pl.ready = t_ctx->event(sl("mcg_ready", slot));
    CBS_CUDA(cudaEventRecord(pl.ready, st));
    pl.ok = true;
    return pl;
}

// Execution half: waits for the plan's counts, then launches the contraction over the work items and the finalize
// kernel.  Returns the Chebyshev tail estimate; nothing is launched when it exceeds `tol` (the caller redoes the batch
// exactly) unless force is set.
static double exec_mc_gemm(const McPlan& pl, long G, const double* d_gamma, const double* d_phi, int mc_mode, double tol,
                           bool force, double* d_gamma_star, double* d_phi_star, unsigned long long* d_nfb,
                           cudaStream_t st, bool* launched) {
    CBS_CUDA(cudaEventSynchronize(pl.ready));
    const int n_items = pl.h_out[0], n_tiles = pl.h_out[1];
    double tail;
    memcpy(&tail, pl.h_out + 4, sizeof tail);
    *launched = false;
    if (tail > tol && !force) return tail;
    *launched = true;
    if (pl.own.rank == 0) {
        g_diag.mc_tiles_done += pl.h_out[2];
        g_diag.mc_tiles_all += (double)n_tiles * (double)(pl.sh.rows / MCG_TJ);
    }
    if (n_tiles == 0) return tail;
    McAcc* part = (McAcc*)g_dev.get<double>("mcg_part", (size_t)std::max(n_items, 1) * MCG_TI * (sizeof(McAcc) / sizeof(double)));
    if (n_items > 0) {
        t_ctx->set_func_attr((const void*)k_mc_gemm<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MCG_SMEM_BYTES);
        t_ctx->set_func_attr((const void*)k_mc_gemm<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)MCG_SMEM_BYTES);
        t_ctx->set_func_attr((const void*)k_mc_gemm<0>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);   // two CTAs per SM
        t_ctx->set_func_attr((const void*)k_mc_gemm<1>, cudaFuncAttributePreferredSharedMemoryCarveout, 100);
        if (mc_mode == CBS_MC_STABLE)
            CBS_LAUNCH(k_mc_gemm<1>, dim3((unsigned)n_items), MCG_BLOCK, MCG_SMEM_BYTES, st, pl.P, pl.Q, pl.gp, pl.sh, pl.items, pl.srclist, pl.count, part);
        else
            CBS_LAUNCH(k_mc_gemm<0>, dim3((unsigned)n_items), MCG_BLOCK, MCG_SMEM_BYTES, st, pl.P, pl.Q, pl.gp, pl.sh, pl.items, pl.srclist, pl.count, part);
    }
    CBS_LAUNCH(k_mcg_finalize, dim3((unsigned)n_tiles), MCG_TI, 0, st, part, pl.tiles, pl.item_start, pl.perm, G,
               mc_mode == CBS_MC_STABLE ? 1 : 0, d_gamma, d_phi, d_gamma_star, d_phi_star, d_nfb);
    return tail;
}

// The whole stage.  bcols[b] = the columns of batch b.  gamma_hat / phi_hat / gamma_star / phi_star are G x B; every
// rank returns with the complete gamma_star / phi_star (own targets computed, the rest received through one all-reduce
// of disjointly filled arrays: x + 0 is exact, so the result does not depend on the number of ranks).
// [g0,g1): the rank's contiguous gene shard, used by the exact pairwise kernel; the contraction interleaves 128-row
// blocks of the abundance order over the ranks instead (the band is wider in the middle of the order).
void run_mc_stage(const int32_t* cnt, long G, int B, const std::vector<std::vector<int>>& bcols, long g0, long g1,
                  const double* d_mu_hat, const double* d_gamma_hat, const double* d_phi_hat, int mc_mode, int subset_n,
                  uint64_t seed, double* d_gamma_star, double* d_phi_star, unsigned long long* h_nfb, cudaStream_t st,
                  cbs_comm* comm) {
    const int world = comm ? comm->world : 1;
    unsigned long long* d_nfb = g_dev.get<unsigned long long>("mc_nfb", (size_t)B);
    CBS_CUDA(cudaMemsetAsync(d_nfb, 0, (size_t)B * sizeof(unsigned long long), st));
    if (world > 1) {
        CBS_CUDA(cudaMemsetAsync(d_gamma_star, 0, (size_t)G * B * sizeof(double), st));
        CBS_CUDA(cudaMemsetAsync(d_phi_star, 0, (size_t)G * B * sizeof(double), st));
    }
    if (subset_n >= G - 1) subset_n = 0;
    const bool want_gemm = subset_n <= 0 && g_tune.mc_algo != 1 && G >= 2;
    std::vector<int*> d_cols(B);
    {
        size_t tot = 0;
        for (int b = 0; b < B; ++b) tot += bcols[b].size();
        int* h = (int*)g_pin.get("mc_cols_h", tot * sizeof(int));
        int* d = g_dev.get<int>("mc_cols", tot);
        size_t o = 0;
        for (int b = 0; b < B; ++b) { memcpy(h + o, bcols[b].data(), bcols[b].size() * sizeof(int)); d_cols[b] = d + o; o += bcols[b].size(); }
        CBS_CUDA(cudaMemcpyAsync(d, h, tot * sizeof(int), cudaMemcpyHostToDevice, st));
    }
    std::vector<double> rng((size_t)2 * B, 0.0);
    if (want_gemm) {
        double* d_rng = g_dev.get<double>("mcg_range", (size_t)2 * B);
        CBS_LAUNCH(k_mcg_phi_range, dim3((unsigned)B), 256, 0, st, d_phi_hat, G, d_rng);
        CBS_CUDA(cudaMemcpyAsync(rng.data(), d_rng, rng.size() * sizeof(double), cudaMemcpyDeviceToHost, st));
        CBS_CUDA(cudaStreamSynchronize(st));
    }
    auto plan = [&](int b) {
        return plan_mc_gemm(cnt, G, (int)bcols[b].size(), d_cols[b], d_mu_hat, d_gamma_hat + (size_t)b * G,
                            d_phi_hat + (size_t)b * G, rng[2 * b], rng[2 * b + 1], b & 1, st, comm);
    };
    McPlan next;
    if (want_gemm) next = plan(0);
    for (int b = 0; b < B; ++b) {
        const int n = (int)bcols[b].size();
        const double *gam = d_gamma_hat + (size_t)b * G, *phi = d_phi_hat + (size_t)b * G;
        double *gs = d_gamma_star + (size_t)b * G, *ps = d_phi_star + (size_t)b * G;
        bool done = false;
        if (want_gemm) {
            const McPlan cur = next;
            if (b + 1 < B) next = plan(b + 1);                 // enqueued BEFORE this batch's contraction: no bubble at the hand-over
            if (cur.ok) {
                bool launched = false;
                const double tail = exec_mc_gemm(cur, G, gam, phi, mc_mode, g_tune.mc_tail_tol, g_tune.mc_algo == 2, gs, ps, d_nfb + b, st, &launched);
                if (!comm || comm->rank == 0) g_diag.mc_tail = std::max(g_diag.mc_tail, tail);
                done = launched;
            }
        }
        if (!done) run_mc_batch_direct(cnt, G, n, d_cols[b], g0, g1, d_mu_hat, gam, phi, mc_mode, subset_n, seed + (uint64_t)b, gs, ps, d_nfb + b, st);
    }
    if (world > 1) {
        CBS_NCCL(nccl_api()->AllReduce(d_gamma_star, d_gamma_star, (size_t)G * B, NCCL_FLOAT64, NCCL_SUM, comm->nccl, st));
        CBS_NCCL(nccl_api()->AllReduce(d_phi_star, d_phi_star, (size_t)G * B, NCCL_FLOAT64, NCCL_SUM, comm->nccl, st));
        CBS_NCCL(nccl_api()->AllReduce(d_nfb, d_nfb, (size_t)B, NCCL_UINT64, NCCL_SUM, comm->nccl, st));
    }
    CBS_CUDA(cudaMemcpyAsync(h_nfb, d_nfb, (size_t)B * sizeof(unsigned long long), cudaMemcpyDeviceToHost, st));
    CBS_CUDA(cudaStreamSynchronize(st));
}

// 16 independent DMMA.8x8x4 accumulator tiles per warp fed from shared memory fragments, 2048 rounds: the operand
// pattern of k_mc_gemm without its pipeline (the roofline denominator of the contraction; MEASURED_PEAKS.json has no
// FP64 figure).
__global__ void __launch_bounds__(256) k_fp64_mma_peak(double* out, int rounds) {
    __shared__ double frag[2][16][32];
    for (int i = threadIdx.x; i < 2 * 16 * 32; i += blockDim.x) (&frag[0][0][0])[i] = 1e-3 * (i % 7) + 0.5;
    __syncthreads();
    const int lane = threadIdx.x & 31;
    double c[16][2];
#pragma unroll
    for (int t = 0; t < 16; ++t) { c[t][0] = threadIdx.x; c[t][1] = t; }
#pragma unroll 1
    for (int i = 0; i < rounds; ++i) {
        double a[2], b[8];
#pragma unroll
        for (int t = 0; t < 2; ++t) a[t] = frag[0][(i + t) & 15][lane];
#pragma unroll
        for (int t = 0; t < 8; ++t) b[t] = frag[1][(i + t) & 15][lane];
#pragma unroll
        for (int t = 0; t < 16; ++t) dmma884(c[t][0], c[t][1], a[t >> 3], b[t & 7]);
    }
    double s = 0;
#pragma unroll
    for (int t = 0; t < 16; ++t) s += c[t][0] + c[t][1];
    if (s == 123.456) out[0] = s;
}

double fp64_mma_peak_tflops() {
    int dev = 0, sms = 0;
    CBS_CUDA(cudaGetDevice(&dev));
    CBS_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
    double* d = g_dev.get<double>("fp64_peak", 8);
    const int blocks = sms * 8, rounds = 2048;
    cudaEvent_t e0, e1;
    CBS_CUDA(cudaEventCreate(&e0)); CBS_CUDA(cudaEventCreate(&e1));
    double best = 0;
    for (int rep = 0; rep < 8; ++rep) {
        CBS_CUDA(cudaEventRecord(e0, 0));
        CBS_LAUNCH(k_fp64_mma_peak, dim3(blocks), 256, 0, 0, d, rounds);
        CBS_CUDA(cudaEventRecord(e1, 0));
        CBS_CUDA(cudaEventSynchronize(e1));
        float ms = 0;
        CBS_CUDA(cudaEventElapsedTime(&ms, e0, e1));
        const double flop = 2.0 * 256 /*fma per mma*/ * 16 * (double)rounds * 8 /*warps*/ * (double)blocks;
        if (rep > 1) best = std::max(best, flop / (ms * 1e-3) / 1e12);
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    return best;
}

}  // namespace cbs

// ---- monte_carlo_int_NB without any allocation (cbs_mc_int_static) ---------------------------------------------------
// No counterpart in the reference: added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The exact pairwise kernel's per-batch
// scratch (A, c_j, r_j, lgamma(r_j), d_i: (n + 4) G doubles, the column list and the fallback counter) lives in ONE file-scope
// device array of CBS_STATIC_SCRATCH_BYTES that every GPU which loads the library holds until exit.  Every call shares it
// and it keeps the last call's contents: what must start at zero (the fallback counter) is cleared by the call's first
// kernel, everything else is overwritten before it is read.  Two calls that overlap on one GPU race on it.
namespace cbs {

static __device__ __align__(256) unsigned char g_mc_static_scratch[CBS_STATIC_SCRATCH_BYTES];

// first kernel of a call: column list 0 .. n-1 and the cleared fallback counter
__global__ void k_mc_static_setup(int n, int* __restrict__ cols, unsigned long long* __restrict__ nfb) {
    const int t = blockIdx.x * blockDim.x + threadIdx.x;
    if (t < n) cols[t] = t;
    if (t == 0) *nfb = 0ull;
}
__global__ void k_mc_static_copy_nfb(const unsigned long long* __restrict__ src, unsigned long long* __restrict__ dst) { *dst = *src; }

// In host code the array's name is a host-side stand-in, not a device address: the address is looked up once per device,
// with that device current.
static std::mutex g_static_mutex;
static std::map<int, unsigned char*> g_static_ptr;
static int static_scratch_ptr(unsigned char** out) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return fail(CBS_ECUDA, "cudaGetDevice: %s", cudaGetErrorString(e));
    std::lock_guard<std::mutex> lk(g_static_mutex);
    auto it = g_static_ptr.find(dev);
    if (it == g_static_ptr.end()) {
        unsigned char* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_mc_static_scratch);
        if (e != cudaSuccess || !p) return fail(CBS_ECUDA, "cudaGetSymbolAddress(static scratch) on device %d: %s", dev, cudaGetErrorString(e));
        it = g_static_ptr.emplace(dev, p).first;
    }
    *out = it->second;
    return CBS_OK;
}

static size_t mc_static_bytes(int64_t G, int32_t n) {
    const size_t head = 256 + ((size_t)n * sizeof(int) + 255) / 256 * 256;       // counter, column list
    return head + ((size_t)n + 4) * (size_t)G * sizeof(double);
}

}  // namespace cbs

extern "C" int cbs_mc_int_static(int64_t G, int32_t n, const int32_t* counts_dev, const double* mu_dev, const double* gamma_dev,
                                 const double* phi_dev, int32_t mc_mode, double* gamma_star_dev, double* phi_star_dev,
                                 uint64_t* n_fallback_dev, void* stream) {
    using namespace cbs;
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || ndev < 1) { cudaGetLastError(); return fail(CBS_ENOGPU, "no CUDA device visible: libcombatseq has no CPU path"); }
    if (G < 2 || n < 1 || !counts_dev || !mu_dev || !gamma_dev || !phi_dev || !gamma_star_dev || !phi_star_dev)
        return fail(CBS_EINVAL, "null pointer / bad size");
    if (mc_mode != CBS_MC_REFERENCE && mc_mode != CBS_MC_STABLE) return fail(CBS_EINVAL, "mc_mode must be 0 or 1");
    if (mc_static_bytes(G, n) > (size_t)CBS_STATIC_SCRATCH_BYTES || n > 65536)
        return fail(CBS_EINVAL, "cbs_mc_int_static: 256 + 256*ceil(n/64) + 8 (n + 4) G = %zu bytes of scratch needed for G = %lld, n = %d; "
                                "the static array holds %d (use cbs_mc_int)", mc_static_bytes(G, n), (long long)G, (int)n, (int)CBS_STATIC_SCRATCH_BYTES);
    unsigned char* base = nullptr;
    if (int rc = static_scratch_ptr(&base)) return rc;
    cudaStream_t st = (cudaStream_t)stream;
    unsigned long long* nfb = reinterpret_cast<unsigned long long*>(base);
    int* cols = reinterpret_cast<int*>(base + 256);
    double* A = reinterpret_cast<double*>(base + 256 + ((size_t)n * sizeof(int) + 255) / 256 * 256);
    double *cj = A + (size_t)n * G, *rj = cj + G, *lgr = rj + G, *di = lgr + G;
    // the targets' counts sit in shared memory while the default 48 KB allow it (no function attribute is touched here)
    size_t smem = (size_t)MC_TI * n * sizeof(double);
    const int xs_shared = smem <= 48 * 1024;
    if (!xs_shared) smem = 0;
    const unsigned gblocks = (unsigned)((G + 127) / 128);
    k_mc_static_setup<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(n, cols, nfb);
    k_mc_prep_src<<<gblocks, 128, 0, st>>>(mu_dev, (long)G, n, cols, phi_dev, A, cj, rj, lgr);
    k_mc_prep_tgt<<<gblocks, 128, 0, st>>>(counts_dev, (long)G, n, cols, di);
    k_mc_direct<<<(unsigned)((G + MC_TI - 1) / MC_TI), 256, smem, st>>>(counts_dev, (long)G, n, cols, 0L, (long)G, A, cj, rj, lgr, di, gamma_dev, phi_dev,
                                                                      mc_mode == CBS_MC_STABLE ? 1 : 0, 0, 0ull, xs_shared, gamma_star_dev,
                                                                      phi_star_dev, nfb);
    if (n_fallback_dev) k_mc_static_copy_nfb<<<1, 1, 0, st>>>(nfb, reinterpret_cast<unsigned long long*>(n_fallback_dev));
    g_launches.fetch_add(n_fallback_dev ? 5 : 4, std::memory_order_relaxed);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(CBS_ECUDA, "cbs_mc_int_static launch: %s", cudaGetErrorString(e));
    return CBS_OK;
}
