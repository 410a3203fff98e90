// ORACLE — TEST INFRASTRUCTURE ONLY (see ref_nmath.hpp for the rules).
//
// CPU restatement of the edgeR 3.22.x routines that /root/reference/ComBat_seq.R calls:
//   DGEList/getOffset (ComBat_seq.R:43,132), estimateGLMCommonDisp (:103,105), estimateGLMTagwiseDisp (:115),
//   glmFit (:130), glmFit.default (:134).
// edgeR is a Bioconductor dependency pinned only by the comment "bioconductor 3.7, edgeR 3.22.1"
// (ComBat_seq.R:42); its sources are NOT under /root/reference.  The algorithms are restated from the
// published edgeR 3.22 sources (R/glmfit.R, R/mglmOneWay.R, R/mglmOneGroup.R, R/mglmLevenberg.R,
// src/glm_one_group.cpp, src/glm_levenberg.cpp, src/nbdev.cpp, src/add_prior.cpp, R/adjustedProfileLik.R,
// src/R_compute_apl.cpp, src/adj_coxreid.cpp, R/dispCoxReid.R, R/dispCoxReidInterpolateTagwise.R,
// src/interpolator.cpp, src/fmm_spline.c) and from R's optimize()/Brent_fmin and LINPACK dqrdc2,
// LAPACK dsytf2/dpotrf.  PARITY UNPINNED against real edgeR (no R in this image).
//
// All matrices are column-major (R layout): element (g,s) of a G x N matrix is at g + s*G.
#pragma once
#include <vector>
#include <cstdint>

namespace ref {

// dispersion / offset broadcasting as edgeR's compressed matrices: value for (g,s)
struct Bcast {
    const double* p; long G; long N; int mode;   // 0 scalar, 1 per-gene vector (G), 2 per-sample vector (N), 3 full G x N col-major
    double at(long g, long s) const {
        switch (mode) { case 0: return p[0]; case 1: return p[g]; case 2: return p[s]; default: return p[g + s * G]; }
    }
};

int  qr_rank(const double* X, int n, int p, double tol = 1e-7);    // LINPACK dqrdc2 rank as R's qr()$rank
// designAsFactor: group id (0-based, by ascending hash value) of each design row; returns number of levels
int  design_as_factor(const double* design, int n, int p, std::vector<int>& group);

// src/glm_one_group.cpp; returns beta (may be -inf), start = NaN => default start
double glm_one_group(int n, const double* y, const double* offset, const double* disp, int maxit, double tol);

double unit_nb_deviance(double y, double mu, double phi);           // src/nbdev.cpp

// glmFit.default(y, design, dispersion, offset, prior.count): coefficients (G x p), fitted.values (G x n).
// prior_count > 0 => coefficients come from the predFC refit (R/predFC.R, src/add_prior.cpp), fitted values
// from the plain fit (R/glmfit.R).  Rows subset: only genes with rows[g] != 0 are fitted when rows != nullptr.
void glm_fit(const double* y, long G, int n, const double* design, int p, const Bcast& disp, const Bcast& offset,
             double prior_count, double* coef /*G x p*/, double* fitted /*G x n, may be null*/, int nthreads);

// adjustedProfileLik(dispersion, y, design, offset): apl[g]
void adjusted_profile_lik(const double* y, long G, int n, const double* design, int p, const Bcast& disp,
                          const Bcast& offset, double* apl, int nthreads);

double brent_fmin(double ax, double bx, double (*f)(double, void*), void* info, double tol);   // R optimize()

// estimateGLMCommonDisp.default(y, design, subset = nrow(y)) -> dispCoxReid
double estimate_common_disp(const double* y, long G, int n, const double* design, int p, int nthreads,
                            int* n_eval = nullptr);
// estimateGLMTagwiseDisp.default(y, design, dispersion = common, prior.df = 0)
void estimate_tagwise_disp(const double* y, long G, int n, const double* design, int p, double common,
                           double* tagwise, int nthreads, double prior_df = 0.0);

void   fmm_spline(int n, const double* x, const double* y, double* b, double* c, double* d);
double maximize_interpolant(int npts, const double* x, const double* y);   // src/interpolator.cpp find_max

}  // namespace ref
