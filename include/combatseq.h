/*
 * combatseq.h — C ABI of the B200-native ComBat-seq engine (libcombatseq.so).
 *
 * This is the drop-in boundary for the reference's one public entry point,
 *     ComBat_seq(counts, batch, group, covar_mod, full_mod, shrink, shrink.disp, gene.subset.n)
 * (/root/reference/ComBat_seq.R:22-23).  The reference has no FFI of its own: an R wrapper
 * (combat-seq_b200/R/ComBat_seq.R) keeps as.matrix/as.factor/model.matrix and the confounding checks
 * (ComBat_seq.R:25-31,53-94) and hands plain arrays to cbs_combat_seq() through a .Call shim
 * (combat-seq_b200/R/r_shim.c).  In this repo the same entry points are bound with ctypes
 * (combat-seq_b200/__init__.py).  See INTEGRATION.md.
 *
 * Conventions
 *   - every matrix is COLUMN-MAJOR (R layout): element (g,s) of a G x N matrix is at g + s*G;
 *   - plain pointers and sizes only, caller owns every host buffer, the library owns all device memory;
 *   - every function returns CBS_OK (0) or a CBS_E* code and never throws; cbs_last_error() gives the text;
 *   - there is NO CPU fallback: without a CUDA device every compute entry point returns CBS_ENOGPU.
 */
#ifndef COMBATSEQ_H
#define COMBATSEQ_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CBS_VERSION 200

enum {
    CBS_OK = 0,
    CBS_EINVAL = 1,                 /* bad argument (message says which) */
    CBS_ESINGLE_SAMPLE_BATCH = 2,   /* "ComBat-seq doesn't support 1 sample per batch yet"   ComBat_seq.R:30 */
    CBS_ECONF_COVAR_BATCH = 3,      /* "The covariate is confounded with batch! ..."          ComBat_seq.R:86 */
    CBS_ECONF_COVARS = 4,           /* "The covariates are confounded! ..."                   ComBat_seq.R:88 */
    CBS_ECONF_COVAR_WITH_BATCH = 5, /* "At least one covariate is confounded with batch! ..." ComBat_seq.R:89 */
    CBS_ECUDA = 6,                  /* CUDA runtime error */
    CBS_ENCCL = 7,                  /* NCCL error / NCCL not loadable */
    CBS_ENOGPU = 8,                 /* no CUDA device visible: the library has no CPU path */
    CBS_ENOMEM = 9
};

/* Monte-Carlo weight mode (helper_seq.R:28-34) */
enum {
    CBS_MC_REFERENCE = 0,   /* weights are prod(dnbinom) in linear space: underflow to 0 and the
                               sum(LH)==0 fallback happen exactly where they do in R */
    CBS_MC_STABLE = 1       /* log-sum-exp weights (no underflow, no fallback); NOT reference behaviour */
};

/* stage indices of cbs_output.stage_ms */
enum {
    CBS_ST_H2D = 0, CBS_ST_PREP, CBS_ST_COMMON_DISP, CBS_ST_TAGWISE_DISP, CBS_ST_GLM, CBS_ST_MC,
    CBS_ST_MATCH_QUANTILES, CBS_ST_D2H, CBS_ST_TOTAL, CBS_N_STAGES
};

int         cbs_version(void);
int         cbs_device_count(void);           /* 0 when no CUDA device is usable */
const char* cbs_last_error(void);             /* thread-local text of the last failure */
int         cbs_shutdown(void);               /* release cached device buffers / communicators */
uint64_t    cbs_kernel_launches(void);        /* kernels launched by this library since load */

/* ---- measurement hooks (bench.py) ---------------------------------------------------------------------
 * cbs_profile_enable(1) clears the accumulators and brackets every kernel launch with CUDA events on the
 * launching stream; cbs_profile_dump() writes one "kernel_name total_ms launches" line per kernel.
 * cbs_fp64_peak() runs a dependent-chain-free DFMA microbenchmark and returns the achieved FP64 TFLOP/s
 * (2 flop per FMA): the roofline denominator of the FP64-pipe-bound kernels (MEASURED_PEAKS.json has none). */
int         cbs_profile_enable(int on);
int         cbs_profile_dump(char* buf, size_t len);
int         cbs_fp64_peak(double* tflops);
int         cbs_fp64_mma_peak(double* tflops);   /* same for the FP64 tensor pipe (mma.sync m8n8k4 f64, DMMA), operands from
                                                    shared memory as in the contraction kernel */

/* ---- tuning knobs and diagnostics ---------------------------------------------------------------------------
 * Defaults reproduce the reference; none of them changes what is computed beyond the stated tolerance.
 *   "mc_algo"      0 = auto, 1 = direct pairwise kernel (every lgamma evaluated), 2 = FP64 tensor-core
 *                  contraction with the Chebyshev expansion of the lgamma term
 *   "mc_band"      1 (default) = the contraction visits, per 64-gene target tile, only the 128-gene source tiles that
 *                  the tcgen05 screening pass cannot rule out (pairs more than 40 nats below the row's best pair carry
 *                  < 4e-18 of the weight and are pruned either way); 0 = every source tile
 *   "mc_tail_tol"  largest accepted Chebyshev tail estimate in log-likelihood units (default 1e-8); a batch that
 *                  exceeds it is redone with the direct kernel when mc_algo = 0
 *   "apl_fast"     1 (default) = lgamma tables, hoisted log/lgamma of 1/phi and product-form mu in the APL kernel
 *                  (last-place differences); 0 = one call per sample
 *   "mq_bins"      1 (default) = match_quantiles visits the elements binned by count, so that the lanes of a warp have the same
 *                  amount of work; 0 = one warp per gene, lane = sample (same bits)
 *   "glm_scr"      1 (default) = Levenberg work arrays of the lane-group GLM kernel in shared memory; 0 = local memory
 *                  (same bits)
 *   "glm_two_pass" 1 (default) = a Levenberg fit that needs more than 8 sweeps over the samples is abandoned by the lane-group
 *                  kernel and redone by a block per gene (its sums associate differently: last-place differences for those
 *                  genes); 0 = every fit finishes in the lane-group kernel
 *   "apl_lanes"    1 (default) = lane-group APL kernel with the batch's row staged in shared memory (one-way designs);
 *                  0 = one thread per (gene, evaluation)
 *   "glm_hist"     1 = collect the histogram of sweeps per fit (cbs_get_diag "glm_iter_<k>"); default 0
 * cbs_get_diag keys: "mc_tail" (largest tail estimate since the last reset), "mc_terms" (Chebyshev terms of the
 * last contraction launch), "mc_tiles_done" / "mc_tiles_all" (source tiles visited / available, summed over the target
 * tiles of rank 0 since the last reset), "glm_iter_<k>" (fits that took k sweeps, with "glm_hist"); cbs_get_diag("reset", &x)
 * clears them. */
int         cbs_set_tuning(const char* key, double value);
int         cbs_get_diag(const char* key, double* value);

/* ---- multi-process gene sharding (one process per GPU, e.g. under torchrun) ------------------------
 * Rank 0 calls cbs_comm_unique_id() and ships the 128 bytes to the other ranks by any means
 * (bench.py uses torch.distributed.broadcast); every rank then calls cbs_comm_create().
 * NCCL is dlopen'ed lazily, so single-GPU use needs no NCCL at all. */
typedef struct cbs_comm cbs_comm;
int cbs_comm_unique_id(void* id128);
int cbs_comm_create(const void* id128, int rank, int world, int device, cbs_comm** out);
int cbs_comm_destroy(cbs_comm* comm);
/* Kept-gene rows [g0, g1) owned by `rank` of `world`: contiguous blocks rounded up to the 128-gene reduction chunk, so
 * that the chunk partial sums of the common-dispersion objective (and everything downstream) do not depend on the
 * number of GPUs.  Host-only (needs no device). */
int cbs_shard_range(int64_t G, int32_t rank, int32_t world, int64_t* g0, int64_t* g1);

/* ---- the whole ComBat_seq() numeric body (ComBat_seq.R:33-203) -------------------------------------- */
typedef struct {
    int32_t G0, N, B, p_mod;        /* genes, samples, batches, ncol(mod) incl. the intercept column      */
    const double*  counts_f64;      /* G0 x N, non-negative integers below 2^31 (NA, negative or fractional values are
                                       rejected with CBS_EINVAL); exactly one of counts_f64/counts_i32 set        */
    const int32_t* counts_i32;      /* G0 x N                                                             */
    const int32_t* batch_idx;       /* N, 0-based level index of as.factor(batch)       ComBat_seq.R:28   */
    const double*  mod;             /* N x p_mod, column 0 is the intercept             ComBat_seq.R:57-73 */
    const uint8_t* use_fallback;    /* B, or NULL: library evaluates ComBat_seq.R:101 itself              */
    int32_t shrink, shrink_disp;    /* ComBat_seq.R:141,162                                               */
    int32_t gene_subset_n;          /* <= 0 means NULL (all genes)                      helper_seq.R:17    */
    uint64_t seed;                  /* seed of the gene_subset_n sampler (keyed Feistel permutation, not R's stream) */
    int32_t mc_mode;                /* CBS_MC_REFERENCE / CBS_MC_STABLE                                    */
    int32_t device;                 /* CUDA device ordinal for single-process use; <0 = current           */
    cbs_comm* comm;                 /* NULL = single GPU; else this process is one rank of a gene-sharded job */
    int32_t counts_on_device;       /* non-zero: counts_i32 is a DEVICE pointer (kernel-only benchmarking) */
    int32_t n_gpus;                 /* single-process multi-GPU (the R caller is one process): 1 = one device (`device`),
                                       k > 1 = genes sharded over devices 0..k-1 by one host thread each, 0 = all visible
                                       devices.  Needs comm == NULL and host buffers.  Results are bit-identical for every k. */
} cbs_input;

typedef struct {                    /* every pointer optional (NULL = not wanted); caller-allocated        */
    double*  adjusted;              /* G0 x N (R double holding integers)            ComBat_seq.R:199-203  */
    uint8_t* keep;                  /* G0, 1 = gene kept by the zero filter          ComBat_seq.R:33-37    */
    double*  lib_size;              /* N                                             ComBat_seq.R:43       */
    double*  disp_common;           /* B                                             ComBat_seq.R:100-107  */
    double*  phi_hat;               /* G0 x B, rows of dropped genes = NaN           ComBat_seq.R:138      */
    double*  gamma_hat;             /* G0 x B                                        ComBat_seq.R:136      */
    double*  gamma_star;            /* G0 x B                                        ComBat_seq.R:157-170  */
    double*  phi_star;              /* G0 x B  (phi_star_mat)                        ComBat_seq.R:159-170  */
    double*  mu_hat;                /* G0 x N                                        ComBat_seq.R:137      */
    double*  stage_ms;              /* CBS_N_STAGES host-side wall times per stage (ms)                    */
    int64_t* n_mc_fallback;         /* B: genes that took the sum(LH)==0 fallback    helper_seq.R:30-31    */
    int64_t  row_begin, row_end;    /* OUT: G0-row range this rank wrote (whole matrix when comm == NULL)  */
    int32_t  adjusted_on_device;    /* non-zero: `adjusted` is a DEVICE pointer                            */
} cbs_output;

int cbs_combat_seq(const cbs_input* in, cbs_output* out);

/* ---- per-subsystem entry points (host buffers in/out; used by the parity tests and kernel benches) --- */

/* NB distribution primitives, elementwise over n values; replaces R's dnbinom/pnbinom/qnbinom as called at
 * helper_seq.R:28,53,58.  logpmf = log dnbinom(x; mu, size), cdf = pnbinom(x; mu, size),
 * quantile = qnbinom(p; mu, size) (R 3.5 semantics: smallest y with pnbinom(y) >= p*(1-64*DBL_EPSILON)). */
int cbs_nb_logpmf(int64_t n, const double* x, const double* size, const double* mu, double* out);
int cbs_nb_cdf(int64_t n, const double* x, const double* size, const double* mu, double* out);
int cbs_nb_quantile(int64_t n, const double* p, const double* size, const double* mu, double* out);

/* Synthetic negative-binomial count matrix, generated on the device (generator half).  Replaces
 * the count generation of the reference's simulation study, simulations/sim_DEpipe.R:98-113
 * (polyester::simulate_experiment(reads_per_transcript, fold_changes, size) = rnbinom(mu = reads * fold change, size)):
 *   counts[g + s*G] ~ NB(mean = base_mu[g] * col_factor[s] * fold[g + level[s]*G], dispersion = phi_col[s])
 * G x N column-major int32 (clamped to 2^31 - 1); fold (G x n_levels, with level[N]) may be NULL; phi_col[s] <= 0 gives
 * Poisson.  Gamma-Poisson mixture with one Philox4x32-10 stream per element keyed by (seed, element index): the matrix
 * depends on the seed only, not on the launch shape.  counts (host) and / or counts_dev (device pointer) receive it. */
int cbs_simulate_counts(int64_t G, int32_t N, const double* base_mu, const double* col_factor, const int32_t* level,
                        int32_t n_levels, const double* fold, const double* phi_col, uint64_t seed, int32_t* counts,
                        void* counts_dev);

/* edgeR glmFit.default(y, design, dispersion, offset, prior.count) as called at ComBat_seq.R:130,134.
 * counts G x n int32; design n x p; disp G x n_disp_cols with disp_cols[s] selecting the column for sample s
 * (phi_matrix of ComBat_seq.R:122-125 without materialising it); offset = log(lib[s]) + gene_offset[g]
 * (gene_offset may be NULL).  coef (G x p) comes from the prior.count refit, fitted (G x n) from the plain fit. */
int cbs_glm_fit(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p,
                const double* disp, int32_t n_disp_cols, const int32_t* disp_cols,
                const double* lib, const double* gene_offset, double prior_count,
                double* coef, double* fitted);

/* edgeR adjustedProfileLik(dispersion, y, design, offset) for D dispersion values per gene:
 * apl is G x D; disp is G x D (per-gene) if disp_per_gene else D scalars.  Used by ComBat_seq.R:103-116. */
int cbs_apl(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p,
            const double* lib, const double* disp, int32_t D, int32_t disp_per_gene, double* apl);

/* estimateGLMCommonDisp(y, design, subset=nrow(y)) and estimateGLMTagwiseDisp(y, design, dispersion, prior.df=0)
 * for one batch: counts G x n (the batch's columns), design n x p or NULL (intercept only). */
int cbs_common_disp(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p, double* common);
int cbs_tagwise_disp(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p,
                     double common, double* tagwise);
/* The same with edgeR's weighted-likelihood empirical-Bayes moderation, estimateGLMTagwiseDisp(..., prior.df): the APL grid
 * of every gene is blended with the moving-average trend of the grid over abundance (aveLogCPM order) before the spline
 * maximisation.  ComBat_seq() itself passes prior.df = 0 (ComBat_seq.R:116), which cbs_tagwise_disp reproduces; this entry
 * exposes the capability for prior_df > 0 (one device). */
int cbs_tagwise_disp_prior(int64_t G, int32_t n, const int32_t* counts, const double* design, int32_t p,
                           double common, double prior_df, double* tagwise);

/* monte_carlo_int_NB(dat, mu, gamma, phi, gene.subset.n)  helper_seq.R:8-42 for one batch */
int cbs_mc_int(int64_t G, int32_t n, const int32_t* counts, const double* mu, const double* gamma,
               const double* phi, int32_t mc_mode, int32_t gene_subset_n, uint64_t seed,
               double* gamma_star, double* phi_star, int64_t* n_fallback);

/* monte_carlo_int_NB for one batch WITHOUT ANY ALLOCATION.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses
 * that forbid it).  Same result as cbs_mc_int(gene_subset_n = 0) by the exact pairwise kernel; every pointer is a DEVICE
 * pointer of the current device (counts G x n int32 and mu G x n column-major, gamma / phi / gamma_star / phi_star G,
 * n_fallback one uint64 or NULL); the kernels are enqueued on `stream` (a cudaStream_t, NULL = default stream) and the call
 * returns without synchronising: outputs are valid once the stream has been synchronised.  The kernels' scratch
 * (256 + 256*ceil(n/64) + 8 (n + 4) G bytes) lives in one static device array of CBS_STATIC_SCRATCH_BYTES that every GPU which
 * loads the library holds until exit; larger problems are rejected with CBS_EINVAL (use cbs_mc_int).  The call makes no
 * cudaMalloc / cudaFree and takes no lock around the GPU work: every call shares the array, which keeps the last call's
 * contents (the call clears what must start at zero), so calls that overlap on one GPU race. */
#define CBS_STATIC_SCRATCH_BYTES (8 << 20)
int cbs_mc_int_static(int64_t G, int32_t n, const int32_t* counts_dev, const double* mu_dev, const double* gamma_dev,
                      const double* phi_dev, int32_t mc_mode, double* gamma_star_dev, double* phi_star_dev,
                      uint64_t* n_fallback_dev, void* stream);

/* match_quantiles(counts_sub, old_mu, old_phi, new_mu, new_phi)  helper_seq.R:46-64 for one batch */
int cbs_match_quantiles(int64_t G, int32_t n, const int32_t* counts, const double* old_mu, const double* old_phi,
                        const double* new_mu, const double* new_phi, double* adjusted);

#ifdef __cplusplus
}
#endif
#endif /* COMBATSEQ_H */
