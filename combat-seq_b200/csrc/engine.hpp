// engine.hpp — what the translation units of libcombatseq.so share: tuning knobs, diagnostics, the NCCL call macro,
// device copies of the design matrices and the stage drivers (one per group of kernels, each defined in the
// translation unit that owns the kernels: tu_apl.cu, tu_glm.cu, tu_mc.cu, tu_mq.cu; combatseq.cu orchestrates).
#pragma once
#include "common.cuh"
#include "host_design.hpp"
#include "glm.cuh"
#include <algorithm>
#include <numeric>

namespace cbs {

// ---- tuning knobs (cbs_set_tuning) and diagnostics (cbs_get_diag) ------------------------------------------------
struct Tuning {
    int mc_algo = 0;             // 0 = auto (tensor-core contraction when gene.subset.n is NULL), 1 = direct, 2 = contraction
    int mc_band = 1;             // 1 = abundance-sorted sources + screened source-tile lists for the contraction, 0 = all tiles
    int glm_scr = 1;             // 1 = Levenberg work arrays of the lane-group GLM kernel in shared memory (one copy per gene)
    int apl_fast = 1;            // 1 = lgamma tables + hoisted invariants in the APL kernel (same bits), 0 = per-sample calls
    int mq_bins = 1;             // 1 = match_quantiles over elements binned by count (uniform work per warp), 0 = lane = sample
    int apl_lanes = 1;           // 1 = lane-group APL kernel with the batch row staged in shared memory (one-way designs), 0 = k_apl
    int glm_two_pass = 1;        // 1 = fits that need more than 9 sweeps are redone by a warp per gene (k_glm_fit_wide)
    int glm_hist = 0;            // 1 = histogram of Levenberg iterations per gene (diagnostics; cbs_get_diag "glm_iter_*")
    double mc_tail_tol = 1e-8;   // largest accepted Chebyshev tail estimate (log-likelihood units)
};
struct Diag {
    double mc_tail = 0;          // largest tail estimate since the last reset
    int mc_kc = 0;               // Chebyshev terms used by the last contraction launch
    double mc_tiles_done = 0, mc_tiles_all = 0;   // source tiles visited / available, summed over the contraction launches
};
extern Tuning g_tune;
extern Diag g_diag;

#define CBS_NCCL(call)                                                                              \
    do {                                                                                            \
        int r_ = (call);                                                                            \
        if (r_ != 0) {                                                                              \
            cbs::NcclApi* a_ = cbs::nccl_api();                                                     \
            cbs::fail(CBS_ENCCL, "NCCL error %d (%s) at %s:%d", r_,                                 \
                      (a_ && a_->GetErrorString) ? a_->GetErrorString(r_) : "?", __FILE__, __LINE__); \
            throw cbs::CbsError{CBS_ENCCL};                                                         \
        }                                                                                           \
    } while (0)
constexpr int NCCL_INT8 = 0, NCCL_INT32 = 2, NCCL_UINT64 = 5, NCCL_FLOAT64 = 8, NCCL_SUM = 0, NCCL_MAX = 2;   // nccl.h enum values

// Device copies of HostDesigns: all arrays of all designs of a set packed into ONE cached device buffer (named, so that two
// sets alive at the same time do not share it) and uploaded with one copy: no cudaMalloc / cudaFree per call, which are
// device-synchronising and serialise the host threads of a single-process multi-GPU call.
struct DesignSet {
    explicit DesignSet(const char* name_) : name(name_) {}
    std::string name;
    std::vector<HostDesign> host;
    DesignDev* d_array = nullptr;      // device array of DesignDev
    std::vector<DesignDev> h_array;    // scalar fields valid on the host; pointers are device addresses
    void add(const HostDesign& h) { host.push_back(h); }
    void finish();
};

template <class T>
inline T* upload(DevCache& c, const std::string& name, const T* src, size_t count, cudaStream_t st = 0) {
    T* p = c.get<T>(name, count ? count : 1);
    if (count) CBS_CUDA(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, st));
    return p;
}
inline dim3 grid1(long n, int block) { return dim3((unsigned)((n + block - 1) / block)); }

// ---- gene sharding ------------------------------------------------------------------------------------------------
// The deterministic reduction unit of the common-dispersion objective is a chunk of APL_CHUNK genes (tree order inside a
// chunk, chunk partials added on the host in chunk order), and shards are whole chunks: disp_common and everything
// downstream are bit-identical for every GPU count.
constexpr int APL_CHUNK = 128;
void shard_range(long G, int rank, int world, long* g0, long* g1);

// ---- stage drivers ------------------------------------------------------------------------------------------------
struct AplJob { int design; double phi; };

// tu_apl.cu
void launch_apl(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignDev* designs,
                const unsigned char* sel, const double* phi_gene, const std::vector<AplJob>& jobs, double* d_apl, cudaStream_t st,
                const std::vector<DesignDev>* h_designs = nullptr,      // host copies of `designs` (enables the lane-group kernel)
                double* d_warm = nullptr);                               // [design][gene][4] Newton warm starts carried between launches
void run_common_disp(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignSet& ds,
                     const unsigned char* sel, cbs_comm* comm, cudaStream_t st, std::vector<double>& common, int* n_steps);
void run_tagwise_disp(const int32_t* cnt, long G, long g0, long g1, const double* logoff, const DesignSet& ds,
                      const unsigned char* sel, const std::vector<double>& common, const std::vector<char>& do_batch,
                      cudaStream_t st, double* d_phi_hat);

void run_tagwise_disp_prior(const int32_t* cnt, long G, const double* logoff, const std::vector<double>& h_lib,
                            const DesignSet& ds, const DesignSet& onecell, const unsigned char* sel, double common,
                            double prior_df, cudaStream_t st, double* d_phi_hat);

// tu_glm.cu
void run_glm_fit(const int32_t* cnt, long G, long g0, long g1, int N, const double* d_logoff,
                 const std::vector<double>& h_lib, const double* d_phi, const int* d_phi_col, const DesignDev* d_design,
                 int h_p, const double* d_goff, double prior_count, double* d_coef, double* d_fitted,
                 const double* d_alpha_w, int n_alpha, double* d_alpha, cudaStream_t st,
                 int class_form = 0);         // bit 0: the design is a Levenberg design in class form (HostDesign::ncls > 0): enables the
                                              // wide second pass; bit 1: every class lies inside one dispersion column (glm_class_flags)

// tu_mc.cu: monte_carlo_int_NB for every batch (ComBat_seq.R:142-165); see the definition for the sharding rules
void run_mc_stage(const int32_t* cnt, long G, int B, const std::vector<std::vector<int>>& bcols, long g0, long g1,
                  const double* d_mu_hat, const double* d_gamma_hat, const double* d_phi_hat, int mc_mode, int subset_n,
                  uint64_t seed, double* d_gamma_star, double* d_phi_star, unsigned long long* h_nfb, cudaStream_t st,
                  cbs_comm* comm);
double fp64_mma_peak_tflops();

// tu_mq.cu
void launch_match_quantiles(const int32_t* cnt, long G, long g0, long g1, const double* mu_hat, const int* batch_idx,
                            const int* order, int N, const double* phi_hat, const double* gamma_star,
                            const double* phi_star, const int* keep_idx, long G0, double* out, cudaStream_t st,
                            int si_begin = 0, int si_end = -1);      // samples [si_begin, si_end) of the batch order (-1: all)
void launch_match_quantiles_plain(const int32_t* cnt, long G, int n, const double* old_mu, const double* old_phi,
                                  const double* new_mu, const double* new_phi, double* out, cudaStream_t st);
void launch_simulate_counts(long G, int N, const double* base_mu, const double* col_factor, const int* level,
                            const double* fold, const double* phi_col, uint64_t seed, int32_t* out, cudaStream_t st);
void launch_nb_elementwise(int what, long n, const double* a, const double* size, const double* mu, double* out, cudaStream_t st);

}  // namespace cbs
