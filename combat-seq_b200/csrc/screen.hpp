// screen.hpp — interface of the tcgen05 screening pass of the Monte-Carlo stage (tu_screen.cu).
#pragma once
#include "engine.hpp"
#include "mc_shape.hpp"

namespace cbs {

// For every target row i of this rank's 128-row blocks and every 128-wide source tile J (rows in the batch's abundance
// order, operands P / Q as built for k_mc_gemm):
//   hi[J*rows + i] >= max_{j in J}          log LH_ij - b_i        (an upper bound)
//   lo[J*rows + i] <= max_{j in J, j != i}  log LH_ij - b_i        (a lower bound)
// with one unknown offset b_i per row (the order-0 Chebyshev coefficient, common to all sources of a row, is left out).
// qsum[j] = (sqrt(sum_s A_js^2), max_v |lg_jv|, |c_j| or < 0 for an impossible source): per source row, from k_mcg_prep_src_fin.
// Returns false (nothing launched) when the shape is outside what the kernel supports; the caller then visits all tiles.
bool mc_screen(const double* P, const double* Q, const float4* qsum, const McgShape& sh, McgOwn own, int slot, float* hi, float* lo,
               cudaStream_t st);

}  // namespace cbs
