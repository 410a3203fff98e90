// prep_kernels.cuh — the streaming kernels around the numeric stages (a1, a2, a12, a14): zero filter, int32 compaction
// with exact library sizes, per-batch row-sum masks, batch-free dispersion, output pre-fill.  HBM-bound, one pass each.
// Data layout: every per-gene matrix is column-major with the gene index fastest (R layout), so one thread per gene
// (consecutive threads = consecutive genes) gives fully coalesced loads and stores.  Included by combatseq.cu only.
#pragma once
#include <cuda_runtime.h>
#include <cstdint>

namespace cbs {

// ---- a1: zero filter (ComBat_seq.R:33-37): keep[g] = every batch has a non-zero count in row g ----------
// bad (one int, zero on entry) collects what the int32 engine cannot represent: bit 0 = NA/NaN/Inf, bit 1 = negative,
// bit 2 = not an integer, bit 3 = above INT32_MAX.  edgeR's DGEList stops on the first two as well ("NA counts not
// allowed", "Negative counts not allowed").
// grid (ceil(G0 / 128), segments of the samples): with one thread walking all N samples of a gene the kernel ran at a tenth of
// the HBM rate (469 blocks of serial loads); the segments' batch sets are combined by atomicOr (order-independent).
// seen: two 64-bit words per gene, zero on entry.
template <class T>
__global__ void k_keep_mask(const T* __restrict__ counts, long G0, int N, const int* __restrict__ batch_idx,
                            unsigned long long* __restrict__ seen, int* __restrict__ bad) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G0) return;
    const int s0 = (int)((long)N * blockIdx.y / gridDim.y), s1 = (int)((long)N * (blockIdx.y + 1) / gridDim.y);
    unsigned long long seen_lo = 0, seen_hi = 0;   // up to 128 batches as a bit set
    int flags = 0;
#pragma unroll 4
    for (int s = s0; s < s1; ++s) {
        const T v = counts[g + (long)s * G0];
        const double dv = (double)v;
        if (!(dv == dv) || dv - dv != 0.0) flags |= 1;
        else if (dv < 0) flags |= 2;
        else if (dv != rint(dv)) flags |= 4;
        else if (dv > 2147483647.0) flags |= 8;
        if (v != (T)0) { const int b = batch_idx[s]; if (b < 64) seen_lo |= 1ull << b; else seen_hi |= 1ull << (b - 64); }
    }
    if (seen_lo) atomicOr(&seen[2 * g], seen_lo);
    if (seen_hi) atomicOr(&seen[2 * g + 1], seen_hi);
    if (flags) atomicOr(bad, flags);
}
__global__ void k_keep_finish(const unsigned long long* __restrict__ seen, long G0, int B, unsigned char* __restrict__ keep) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g < G0) keep[g] = (__popcll(seen[2 * g]) + __popcll(seen[2 * g + 1]) == B) ? 1 : 0;
}

// ---- a1/a2: compaction to int32 + exact integer column sums (DGEList lib.size, ComBat_seq.R:43) -----------
// grid: (ceil(G/256), N).  cnt is G x N column-major over the kept genes.
template <class T>
__global__ void k_compact_colsum(const T* __restrict__ counts, long G0, const int* __restrict__ keep_idx, long G,
                                 int32_t* __restrict__ cnt, unsigned long long* __restrict__ colsum) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int s = blockIdx.y;
    long long v = 0;
    if (i < G) {
        v = (long long)counts[keep_idx[i] + (long)s * G0];
        cnt[i + (long)s * G] = (int32_t)v;
    }
    // integer sums are exact and order-independent, so atomics keep the result deterministic
    for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
    __shared__ long long part[32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    if (lane == 0) part[wid] = v;
    __syncthreads();
    if (wid == 0) {
        v = (lane < (int)(blockDim.x >> 5)) ? part[lane] : 0;
        for (int o = 16; o > 0; o >>= 1) v += __shfl_down_sync(0xffffffffu, v, o);
        if (lane == 0 && v != 0) atomicAdd(&colsum[s], (unsigned long long)v);
    }
}

// per-batch row sums of the kept genes: sel[b*G+g] = rowSums(counts[g, batch b]) >= 5  (edgeR min.row.sum)
__global__ void k_rowsum_sel(const int32_t* __restrict__ cnt, long G, int N, const int* __restrict__ batch_idx, int B,
                             unsigned char* __restrict__ sel) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int b = blockIdx.y;
    if (g >= G) return;
    long long rs = 0;
    for (int s = 0; s < N; ++s) if (batch_idx[s] == b) rs += cnt[g + (long)s * G];
    sel[(long)b * G + g] = (rs >= 5) ? 1 : 0;
}

__global__ void k_fill(double* __restrict__ p, long n, double v) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) p[i] = v;
}

// ---- a12: batch-free dispersion (ComBat_seq.R:178) ---------------------------------------------------------
__global__ void k_phi_star_mean(const double* __restrict__ phi_star_mat, long G, int B, double* __restrict__ phi_star) {
    const long g = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (g >= G) return;
    double s = 0;
    for (int b = 0; b < B; ++b) s += phi_star_mat[g + (long)b * G];
    phi_star[g] = s / B;                                    // rowMeans  ComBat_seq.R:178
}

// a14: output pre-fill with the original counts as doubles (rows of filtered genes pass through; ComBat_seq.R:202)
template <class T>
__global__ void k_counts_to_f64(const T* __restrict__ counts, long n, double* __restrict__ out) {
    const long i = (long)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = (double)counts[i];
}

}  // namespace cbs
