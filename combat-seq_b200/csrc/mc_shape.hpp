// mc_shape.hpp — shapes, layout and ownership rules shared by the Monte-Carlo contraction (mc_gemm.cuh) and its
// screening pass (tu_screen.cu).
#pragma once
#include <cuda_runtime.h>
#include <cstddef>

namespace cbs {

constexpr int MCG_TILE = 128;        // source tile (and the row block of the P/Q layout): 128 genes
constexpr int MCG_TI = 64;           // target rows per CTA of k_mc_gemm: CTA tile = 64 targets x 128 sources
constexpr int MCG_KC = 8;            // k-chunk (doubles): the unit of the global layout and of one bulk copy
constexpr int MCG_SHIFT = 4;         // c: counts below c use the one-hot table
// A source whose log-likelihood is more than MCG_PRUNE below the best one of the row carries a weight below e^-40 = 4e-18
// of the largest weight (see mc_gemm.cuh, "Pruning").
constexpr double MCG_PRUNE = 40.0;

struct McgShape {
    int n;          // samples in the batch
    int Kc;         // Chebyshev terms
    int KT;         // n + Kc + MCG_SHIFT + 1
    int KTP;        // KT rounded up to a multiple of MCG_KC (the last hand-over of a source tile may hold one k-chunk)
    long rows;      // matrix rows, G rounded up to a multiple of MCG_TILE
    double ulo, uhi;   // range of u = log(r + c)
};

// Storage of P and Q: chunk-major, [kc][row][8] doubles with an XOR swizzle of the in-chunk column so that the
// m8n8k4 fragment loads (8 rows x 4 columns per warp) are shared-memory bank-conflict free at a dense row stride
// of 8 doubles, and so that one (tile, chunk) block of 128 rows is 8 KB contiguous in global memory = one bulk copy.
__host__ __device__ __forceinline__ size_t mcg_index(long rows, long row, int k) {
    const int kc = k >> 3, kk = (k & 7) ^ ((int)((row >> 1) & 1) << 2);
    return ((size_t)kc * rows + row) * 8 + kk;
}

// Rows of P this rank builds: the 128-row blocks blk with blk % world == rank (its targets; see McgOwn).
struct McgOwn {
    int rank, world;
    __host__ __device__ long nblocks(long rows) const { const long nb = rows / MCG_TILE; return nb > rank ? (nb - rank + world - 1) / world : 0; }
    __host__ __device__ long row(long local) const {       // local row index (0 .. nblocks*128) -> sorted row
        const long lb = local / MCG_TILE;
        return (lb * world + rank) * MCG_TILE + (local - lb * MCG_TILE);
    }
    __host__ __device__ bool owns_tile64(long t64) const { return (int)((t64 / (MCG_TILE / MCG_TI)) % world) == rank; }
};
}  // namespace cbs
