// tu_screen.cu — screening pass of the Monte-Carlo stage on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// monte_carlo_int_NB (/root/reference/helper_seq.R:8-42) weighs every source gene j by LH_ij for every target gene i, but
// only sources within e^-40 of a row's best weight contribute above 1e-17 of the result (mc_gemm.cuh, "Pruning"), and in
// the batch's abundance order those sources form a band around the diagonal.  This pass finds the band rigorously: it
// evaluates ALL pairs of a batch at reduced precision together with a bound on its own error,
//
//   S_ij = sum_k P_ik Q_jk              in bf16 x 2 (hi/lo split: Ph Qh + Ph Ql + Pl Qh), fp32 accumulation in TMEM
//   |S_ij - exact| <= SCR_EPS * sum_k |P_ik| |Q_jk|      (3 * 2^-18 from the dropped lo*lo and residual terms + the fp32
//                                                          accumulation), and, term by term,
//   sum_k |P_ik| |Q_jk| <= |x_i| |A_j|  (Cauchy-Schwarz over the samples; the two vectors are close to parallel)
//                        + sum_k |C_ik|  (the Chebyshev polynomials are bounded by 1)
//                        + (#counts of i below 4) * max_v |lgamma(v + r_j) - lgamma(v + 1)|
//                        + 2e-6 |c_j|   (c_j enters as three bf16 pieces, exact to 2^-24; the rest is its share of the
//                                        fp32 accumulation error),
//
// i.e. three rank-one terms; their column factors are replaced by their maxima over the 128 sources of a tile, so the bound
// is one number d per (row, tile), and every 128 x 128 tile reduces to the row maxima m of S itself: upper bound m + d, lower
// bound m - d.  The per-source constant c_j rides in the contraction as three bf16 columns (hi, mid, lo: 24 bits) against ones,
// so the epilogue is a pure row maximum of the accumulator.  k_mcg_list_screened (mc_gemm.cuh)
// keeps a source tile for a target tile iff an upper bound reaches a row's best lower bound - 40; the FP64 contraction then
// visits only those tiles.  tcgen05 has no FP64 kind, so the weights themselves stay on the DMMA path; what runs here is 30
// tcgen05.mma (128 x 256 x 16) per 128 x 256 pairs, a few per cent of the FP64 work it removes.
//
// Operand layout.  A row of P or Q is a sequence of 16-column chunks (7 chunks of samples, then 3 chunks holding the
// Chebyshev orders >= 1 and the one-hot columns at C5's shape); a 64-column k-tile holds two chunks as the four 16-column
// sub-blocks (hi, lo, hi, lo).  The cross terms of the split are MMAs between different sub-blocks of the SAME pair of
// k-tiles (A descriptor at hi, B descriptor at lo, ...), so no operand is stored or fetched twice.
//
// Kernel: one CTA = one 128-row target block x a range of 256-row source tile pairs.  The target block's k-tiles stay resident
// in shared memory; warp 0 streams the source rows' k-tiles (32 KB each) through a ring with bulk async copies (the global
// layout is the shared-memory image: 128 rows x 64 bf16 per 16 KB block, 128-byte swizzle); one lane of warp 1 issues
// tcgen05.mma (M = 128, N = 256, K = 16 per instruction: 12 KB of shared-memory operands per 128 tensor-core cycles, under
// the 128 B/clk of the SM, which N = 128 is not), A and B from shared memory descriptors, the accumulator in TMEM, double
// buffered (2 x 256 of the 512 columns); warps 2-9 read the accumulator back with tcgen05.ld and fold it into the row maxima
// while the next tile is being multiplied.
#include "screen.hpp"
#include <cuda_bf16.h>

namespace cbs {

#ifndef CBS_SCR_WAVES
#define CBS_SCR_WAVES 8
#endif
constexpr int SCR_BM = 128;                 // target rows per CTA (UMMA M)
constexpr int SCR_BN = 256;                 // source rows per MMA (UMMA N): two source tiles of the contraction
constexpr int SCR_KTILE = 64;               // bf16 per operand row and k-tile: 128 bytes, the swizzle span
constexpr int SCR_TILE_BYTES = SCR_BM * SCR_KTILE * 2;      // 16 KB: one k-tile of one 128-row block
constexpr int SCR_MAX_KT = 7;               // most k-tiles of an operand row the kernel supports (resident A: 112 KB)
constexpr int SCR_SMEM_BLOCKS = 13;         // 16 KB blocks that fit shared memory: resident A + ring of B (2 blocks per stage)
constexpr int SCR_EPI_WARPS = 8;            // two warps per TMEM lane quarter, one per source tile of the pair
constexpr int SCR_THREADS = 64 + 32 * SCR_EPI_WARPS;     // warp 0 producer, warp 1 MMA issuer, warps 2-9 epilogue
constexpr float SCR_EPS = 3.1e-5f;          // error of S relative to sum |P||Q| (see above), incl. the fp32 rounding of the bound
static_assert(SCR_BM == MCG_TILE && SCR_BN == 2 * MCG_TILE, "screening tiles are the contraction's source tiles");

struct ScrShape {
    int n, KS;       // samples; KS - n series columns (Chebyshev orders 1.., one-hot; no order-0 term, no c_j)
    int nch;         // 16-column chunks: the n sample columns, then the series columns and the three pieces of c_j, back to back
    int nk;          // k-tiles per operand row = ceil(nch / 2)
    int stages;      // ring of B stages (2 blocks each)
    long rows;       // rows of P / Q (multiple of 128)
    long nrb;        // 128-row blocks; the operands hold one more (zero) block so that the last pair can be read whole
    int NJ2;         // source tile pairs
};

// ---- operand conversion: FP64 P / Q (mc_gemm.cuh layout) -> bf16 tiles in the shared-memory image ------------------
// Element (row r, k-tile kt, column c of the k-tile) lives in block [kt][r / 128] (16 KB each) at byte (r % 128) * 128 +
// ((c / 8) ^ (r % 8)) * 16 + (c % 8) * 2: the 128-byte swizzle the UMMA descriptors below declare.  One thread writes one
// 16-byte chunk (8 columns = half a sub-block).  P is only built for the owned blocks, the other blocks are never read.
// Series part of a row: MC columns n+1 .. KT-2 (Chebyshev orders >= 1, one-hot), then three columns for c_j: ones in P, the
// bf16 pieces (hi, mid, lo) of c_j in Q, all in the hi sub-block (their lo sub-block entries are zero).
// pvec[r] (targets): SCR_EPS (|x_i|, sum |C_ik|, #small counts).  qtile[J] (source tiles): maxima over the tile's rows of
// |A_j|, max_v |lg_jv| and |c_j|, kept as the bit patterns of non-negative floats (atomicMax on ints; zero on entry).
__global__ void __launch_bounds__(256)
k_scr_convert(const double* __restrict__ X, McgShape sh, ScrShape ss, int is_q, McgOwn own, long n_rows,
              __nv_bfloat16* __restrict__ out, float4* __restrict__ pvec, int* __restrict__ qtile, const float4* __restrict__ qsum) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int chunks_per_row = ss.nk * 8;
    const long lr = t / chunks_per_row;
    const int ch = (int)(t - lr * chunks_per_row);
    if (lr >= n_rows) return;
    const long r = is_q ? lr : own.row(lr);
    const bool pad_row = r >= sh.rows;                            // the zero block behind Q (read by the last tile pair)
    const int kt = ch >> 3, c8 = ch & 7, sb = c8 >> 1, half = c8 & 1;
    const int c = 2 * kt + (sb >> 1), lo_part = sb & 1;           // chunk, hi / lo
    const int ser0 = ss.n + 1, ser1 = sh.KT - 1;                  // series part in MC columns: [ser0, ser1), then 3 columns of c_j
    __align__(16) __nv_bfloat16 v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        float f = 0.f;
        if (!pad_row) {
            int k = -1, cpart = -1;                               // MC column to convert, or piece of c_j
            const int q = c * 16 + half * 8 + e;                  // column of the packed screening row
            if (q < ss.n) k = q;
            else {
                const int kk = ser0 + (q - ss.n);
                if (kk < ser1) k = kk; else if (kk < ser1 + 3) cpart = kk - ser1;
            }
            if (k >= 0) {
                const double x = X[mcg_index(sh.rows, r, k)];
                const float h = __bfloat162float(__double2bfloat16(x));
                f = lo_part ? (float)(x - (double)h) : h;
            } else if (cpart >= 0 && !lo_part) {
                if (!is_q) f = 1.f;
                else {
                    const double x = X[mcg_index(sh.rows, r, sh.KT - 1)];
                    const float h = __bfloat162float(__double2bfloat16(x));
                    const float m = __bfloat162float(__double2bfloat16(x - (double)h));
                    f = cpart == 0 ? h : cpart == 1 ? m : (float)(x - (double)h - (double)m);
                }
            }
        } else if (is_q && !lo_part) {                            // padding sources: c_j = -1e30, never kept
            const int q = c * 16 + half * 8 + e;
            if (q >= ss.n && ser0 + (q - ss.n) == ser1) f = -1e30f;
        }
        v[e] = __float2bfloat16_rn(f);
    }
    const long blk = (long)kt * (ss.nrb + 1) + r / SCR_BM;
    const int rr = (int)(r % SCR_BM);
    char* dst = reinterpret_cast<char*>(out) + blk * SCR_TILE_BYTES + rr * 128 + ((c8 ^ (rr & 7)) << 4);
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(v);
    if (ch == 0 && !pad_row) {
        if (!is_q) {
            double s2 = 0, s1 = 0, hot = 0;
            for (int k = 0; k < ss.n; ++k) { const double x = X[mcg_index(sh.rows, r, k)]; s2 += x * x; }
            for (int k = ss.n + 1; k < ss.n + sh.Kc; ++k) s1 += fabs(X[mcg_index(sh.rows, r, k)]);
            for (int k = ss.n + sh.Kc; k < sh.KT - 1; ++k) hot += X[mcg_index(sh.rows, r, k)];
            pvec[r] = make_float4(SCR_EPS * (float)sqrt(s2), SCR_EPS * (float)s1, SCR_EPS * (float)hot, 0.f);
        } else {
            const float4 q = qsum[r];                              // row summary left by k_mcg_prep_src_fin
            int* q3 = qtile + 4 * (r / MCG_TILE);
            atomicMax(q3 + 0, __float_as_int(fminf(q.x * 1.000001f, 3e38f)));
            atomicMax(q3 + 1, __float_as_int(q.y * 1.000001f));
            if (q.z >= 0.f && q.z < 1e29f) atomicMax(q3 + 2, __float_as_int(q.z));     // (c_j = -1e30 marks an impossible source)
        }
    }
}

// ---- PTX wrappers ----------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t s_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void sbar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s_u32(bar)), "r"(count));
}
__device__ __forceinline__ void sbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void sbar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s_u32(bar)) : "memory");
}
__device__ __forceinline__ void sbar_wait(uint64_t* bar, uint32_t parity) {
    for (;;) {
        uint32_t ok;
        asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
                     : "=r"(ok) : "r"(s_u32(bar)), "r"(parity) : "memory");
        if (ok) return;
    }
}
__device__ __forceinline__ void sbulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 ::"r"(s_u32(dst)), "l"(src), "r"(bytes), "r"(s_u32(bar)) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// D[tmem] (+)= A[smem desc] * B[smem desc]^T, bf16 x bf16 -> fp32, issued by one thread for the CTA
__device__ __forceinline__ void tc_mma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
        ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// the mbarrier is arrived on (count 1) when every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(s_u32(bar)) : "memory");
}
// 32 lanes x 32 consecutive fp32 columns: thread l of the warp receives row (lane base + l), columns col .. col + 31
__device__ __forceinline__ void tc_ld32(uint32_t taddr, float* v) {
    uint32_t* u = reinterpret_cast<uint32_t*>(v);
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                 "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                 "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
                 : "=r"(u[0]), "=r"(u[1]), "=r"(u[2]), "=r"(u[3]), "=r"(u[4]), "=r"(u[5]), "=r"(u[6]), "=r"(u[7]),
                   "=r"(u[8]), "=r"(u[9]), "=r"(u[10]), "=r"(u[11]), "=r"(u[12]), "=r"(u[13]), "=r"(u[14]), "=r"(u[15]),
                   "=r"(u[16]), "=r"(u[17]), "=r"(u[18]), "=r"(u[19]), "=r"(u[20]), "=r"(u[21]), "=r"(u[22]), "=r"(u[23]),
                   "=r"(u[24]), "=r"(u[25]), "=r"(u[26]), "=r"(u[27]), "=r"(u[28]), "=r"(u[29]), "=r"(u[30]), "=r"(u[31])
                 : "r"(taddr));
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// Shared-memory matrix descriptor of a K-major operand tile stored with the 128-byte swizzle (cute::UMMA::SmemDescriptor):
// start address >> 4 in bits [0,14), leading byte offset (unused for swizzled K-major, 1) in [16,30), stride byte offset
// = 1024 B (8 rows x 128 B) >> 4 in [32,46), descriptor version 1 in [46,48), layout type SWIZZLE_128B = 2 in [61,64).
__device__ __forceinline__ uint64_t scr_desc(uint32_t smem_addr) {
    const uint32_t lo = ((smem_addr >> 4) & 0x3FFFu) | (1u << 16);
    const uint32_t hi = 64u | (1u << 14) | (2u << 29);
    return ((uint64_t)hi << 32) | lo;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D fp32 (bits 4-5 = 1), A and B bf16 (bits 7-9, 10-12 = 1), both
// K-major (bits 15, 16 = 0), N >> 3 in bits 17-22, M >> 4 in bits 24-28.
__host__ __device__ constexpr uint32_t scr_idesc() {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(SCR_BN >> 3) << 17) | ((uint32_t)(SCR_BM >> 4) << 24);
}

struct ScrSmem {
    // [SCR_SMEM_BLOCKS x 16 KB: the resident A k-tiles, then the B ring (32 KB stages)] [barriers] [tmem base]; the tiles
    // need 1024-byte alignment
    static constexpr size_t tile_bytes = (size_t)SCR_SMEM_BLOCKS * SCR_TILE_BYTES;
    static constexpr size_t bar_off = tile_bytes;
    static constexpr int max_stages = SCR_SMEM_BLOCKS / 2;
    static constexpr size_t total = bar_off + (5 + 2 * max_stages + 1) * sizeof(uint64_t) + 16 + 1024 /*alignment slack*/;
};
static_assert(ScrSmem::total <= 227 * 1024, "screening kernel: shared memory");

// items[w] = (target block, first source tile pair, number of pairs).  hi / lo are [NJ][rows], NJ = 128-row source tiles.
__global__ void __launch_bounds__(SCR_THREADS, 1)
k_scr_mma(const __nv_bfloat16* __restrict__ Pt, const __nv_bfloat16* __restrict__ Qt, const float4* __restrict__ pvec,
          const float4* __restrict__ qtile, ScrShape ss, const int3* __restrict__ items, float* __restrict__ hi,
          float* __restrict__ lo) {
    extern __shared__ unsigned char scr_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(scr_raw) + 1023) & ~(uintptr_t)1023);
    const int NST = ss.stages;
    unsigned char* a_s = smem;
    unsigned char* b_s = smem + (size_t)ss.nk * SCR_TILE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ScrSmem::bar_off);
    uint64_t* a_full = bars;                    // 1
    uint64_t* acc_full = bars + 1;              // 2
    uint64_t* acc_empty = bars + 3;             // 2
    uint64_t* b_full = bars + 5;                // NST
    uint64_t* b_empty = bars + 5 + ScrSmem::max_stages;     // NST
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5 + 2 * ScrSmem::max_stages);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int3 item = items[blockIdx.x];
    const long iblk = item.x;
    const int j0 = item.y, nj = item.z;
    const long nrb1 = ss.nrb + 1;
    constexpr uint32_t STAGE_BYTES = 2 * SCR_TILE_BYTES;

    if (threadIdx.x == 0) {
        sbar_init(a_full, 1);
        for (int s = 0; s < NST; ++s) { sbar_init(&b_full[s], 1); sbar_init(&b_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { sbar_init(&acc_full[s], 1); sbar_init(&acc_empty[s], SCR_EPI_WARPS); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {                            // TMEM: all 512 columns (two accumulator stages of 128 x 256 fp32)
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s_u32(tmem_slot)), "r"(512u) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = *tmem_slot;

    if (warp == 0) {
        // ===== producer =====
        if (lane == 0) {
            sbar_expect_tx(a_full, (uint32_t)ss.nk * SCR_TILE_BYTES);
            for (int kt = 0; kt < ss.nk; ++kt)
                sbulk_g2s(a_s + (size_t)kt * SCR_TILE_BYTES, reinterpret_cast<const char*>(Pt) + ((size_t)kt * nrb1 + iblk) * SCR_TILE_BYTES,
                          SCR_TILE_BYTES, a_full);
            int st = 0; uint32_t ph = 0;
            for (int jj = 0; jj < nj; ++jj) {
                for (int kt = 0; kt < ss.nk; ++kt) {
                    sbar_wait(&b_empty[st], ph ^ 1u);
                    sbar_expect_tx(&b_full[st], STAGE_BYTES);
                    sbulk_g2s(b_s + (size_t)st * STAGE_BYTES,               // two consecutive 128-row blocks of k-tile kt
                              reinterpret_cast<const char*>(Qt) + ((size_t)kt * nrb1 + 2 * (size_t)(j0 + jj)) * SCR_TILE_BYTES, STAGE_BYTES, &b_full[st]);
                    if (++st == NST) { st = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            sbar_wait(a_full, 0);
            tc_fence_after();
            const uint32_t a_addr = s_u32(a_s), b_addr = s_u32(b_s);
            const int nchunks = ss.nch;
            int st = 0; uint32_t ph = 0;
            for (int jj = 0; jj < nj; ++jj) {
                const int as = jj & 1;
                sbar_wait(&acc_empty[as], ((uint32_t)(jj >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d_s = tmem + (uint32_t)as * 256u;
                uint32_t started = 0u;
                for (int kt = 0; kt < ss.nk; ++kt) {
                    sbar_wait(&b_full[st], ph);
                    tc_fence_after();
                    const uint32_t at = a_addr + (uint32_t)kt * SCR_TILE_BYTES, bt = b_addr + (uint32_t)st * STAGE_BYTES;
#pragma unroll
                    for (int h = 0; h < 2; ++h) {                         // the k-tile's two chunks: sub-blocks (hi, lo) 32 bytes apart
                        if (2 * kt + h < nchunks) {
                            const uint32_t ah = at + (uint32_t)(2 * h) * 32u, bh = bt + (uint32_t)(2 * h) * 32u;
                            tc_mma_bf16(d_s, scr_desc(ah), scr_desc(bh), scr_idesc(), started); started = 1u;      // hi x hi
                            tc_mma_bf16(d_s, scr_desc(ah), scr_desc(bh + 32u), scr_idesc(), 1u);                   // hi x lo
                            tc_mma_bf16(d_s, scr_desc(ah + 32u), scr_desc(bh), scr_idesc(), 1u);                   // lo x hi
                        }
                    }
                    tc_commit(&b_empty[st]);                      // the stage is free once these MMAs have read it
                    if (++st == NST) { st = 0; ph ^= 1u; }
                }
                tc_commit(&acc_full[as]);                         // accumulator of pair jj complete
            }
        }
    } else {
        // ===== epilogue: warps 2..9; warp w may touch TMEM lanes 32 (w % 4) .. +31; the two warps of a lane quarter take one
        // source tile (128 columns) of the pair each: a row maximum of the accumulator (four independent chains), plus the
        // (row, tile) error bound.
        const int q = warp & 3, half = (warp - 2) >> 2;
        const int row = q * 32 + lane;                            // row of the target block = TMEM lane
        const long grow = iblk * SCR_BM + row;
        const float4 pv = pvec[grow];                             // SCR_EPS (|x_i|, sum |C_ik|, #small counts)
        constexpr int HC = SCR_BN / 2;                            // columns per warp = one source tile
        for (int jj = 0; jj < nj; ++jj) {
            const int as = jj & 1;
            const long J = 2 * (long)(j0 + jj) + half;            // source tile of this warp
            const float4 qt = qtile[min(J, ss.nrb)];              // max |A_j|, max |lg_jv|, max |c_j| over the tile
            sbar_wait(&acc_full[as], (uint32_t)(jj >> 1) & 1u);
            tc_fence_after();
            const uint32_t t_s = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)as * 256u + (uint32_t)(half * HC);
            const bool diag = J == iblk;
            float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll 1
            for (int c0 = 0; c0 < HC; c0 += 32) {
                float s[32];
                tc_ld32(t_s + (uint32_t)c0, s);
                tc_ld_wait();
                if (diag && (row >> 5) == (c0 >> 5)) {             // j == i is not a source (mu[-i,], helper_seq.R:11-14)
#pragma unroll
                    for (int c = 0; c < 32; ++c) m4[c & 3] = fmaxf(m4[c & 3], c == (row & 31) ? -INFINITY : s[c]);
                } else {
#pragma unroll
                    for (int c = 0; c < 32; ++c) m4[c & 3] = fmaxf(m4[c & 3], s[c]);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) sbar_arrive(&acc_empty[as]);
            if (J < ss.nrb) {
                const float m = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
                const float e = fmaf(pv.x, qt.x, fmaf(pv.z, qt.y, fmaf(2e-6f, qt.z, pv.y))) + 1e-6f * fabsf(m);
                __stcs(&hi[J * ss.rows + grow], m + e);          // streamed: read once by k_mcg_list_screened
                __stcs(&lo[J * ss.rows + grow], m - e);
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

// Work items of the screening kernel: every owned target block x ranges of NJ2 / nchunk source tile pairs, ordered
// range-major (the CTAs resident at the same time stream the same range of source rows: L2 reuse).
__global__ void k_scr_items(McgOwn own, long rows, int NJ2, int nchunk, int3* __restrict__ items) {
    const long lb = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long nb = own.nblocks(rows);
    if (lb >= nb) return;
    const int blk = (int)(lb * own.world + own.rank);
    for (int c = 0; c < nchunk; ++c) {
        const int a = (int)((long)NJ2 * c / nchunk), b = (int)((long)NJ2 * (c + 1) / nchunk);
        items[(long)c * nb + lb] = make_int3(blk, a, b - a);
    }
}

bool mc_screen(const double* P, const double* Q, const float4* qsum, const McgShape& sh, McgOwn own, int slot, float* hi, float* lo,
               cudaStream_t st) {
    ScrShape ss;
    ss.n = sh.n;
    ss.KS = sh.KT - 2;
    ss.nch = (ss.KS + 3 + 15) / 16;                              // KS columns + the three columns of c_j
    ss.nk = (ss.nch + 1) / 2;
    ss.rows = sh.rows;
    ss.nrb = sh.rows / SCR_BM;
    ss.NJ2 = (int)((ss.nrb + 1) / 2);
    if (ss.nk > SCR_MAX_KT) return false;                        // wider batches: the caller visits all tiles
    ss.stages = (SCR_SMEM_BLOCKS - ss.nk) / 2;
    int cc_major = 0, dev = 0;
    CBS_CUDA(cudaGetDevice(&dev));
    CBS_CUDA(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    if (cc_major != 10) return false;                            // tcgen05 / TMEM
    const std::string sfx = slot ? "#1" : "#0";
    const long own_blocks = own.nblocks(sh.rows);
    const size_t op_bytes = (size_t)ss.nk * (ss.nrb + 1) * SCR_TILE_BYTES;
    __nv_bfloat16* Pt = (__nv_bfloat16*)g_dev.get("mcs_Pt" + sfx, op_bytes);
    __nv_bfloat16* Qt = (__nv_bfloat16*)g_dev.get("mcs_Qt" + sfx, op_bytes);
    float4* pvec = (float4*)g_dev.get("mcs_pvec" + sfx, (size_t)sh.rows * sizeof(float4));
    int* qtile = g_dev.get<int>("mcs_qtile" + sfx, (size_t)4 * (ss.nrb + 1));
    CBS_CUDA(cudaMemsetAsync(qtile, 0, (size_t)4 * (ss.nrb + 1) * sizeof(int), st));
    const int cpr = ss.nk * 8;
    CBS_LAUNCH(k_scr_convert, grid1(own_blocks * SCR_BM * cpr, 256), 256, 0, st, P, sh, ss, 0, own, own_blocks * SCR_BM, Pt, pvec, (int*)nullptr, (const float4*)nullptr);
    CBS_LAUNCH(k_scr_convert, grid1((sh.rows + SCR_BM) * cpr, 256), 256, 0, st, Q, sh, ss, 1, own, sh.rows + SCR_BM, Qt, (float4*)nullptr, qtile, qsum);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // about 8 waves of CTAs and at least 4 ranges; every CTA re-loads its 128-row A operand once, so ranges should not be
    // too short either
    int nchunk = (int)std::min<long>(std::max<long>(4, ((long)CBS_SCR_WAVES * sms + own_blocks - 1) / std::max<long>(own_blocks, 1)), std::max(1, ss.NJ2 / 4));
    int3* items = (int3*)g_dev.get<int>("mcs_items" + sfx, (size_t)3 * own_blocks * nchunk + 3);
    CBS_LAUNCH(k_scr_items, grid1(own_blocks, 128), 128, 0, st, own, sh.rows, ss.NJ2, nchunk, items);
    t_ctx->set_func_attr((const void*)k_scr_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ScrSmem::total);
    CBS_LAUNCH(k_scr_mma, dim3((unsigned)(own_blocks * nchunk)), SCR_THREADS, ScrSmem::total, st, Pt, Qt, pvec, (const float4*)qtile, ss, items, hi, lo);
    return true;
}

}  // namespace cbs
