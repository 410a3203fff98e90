// tu_screen.cu — screening pass of the Monte-Carlo stage on the 5th-generation tensor cores (tcgen05 + TMEM).
//
// monte_carlo_int_NB (/root/reference/helper_seq.R:8-42) weighs every source gene j by LH_ij for every target gene i, but
// only sources within e^-40 of a row's best weight contribute above 1e-17 of the result (mc_gemm.cuh, "Pruning"), and in
// the batch's abundance order those sources form a band around the diagonal.  This pass finds the band rigorously: it
// evaluates ALL pairs of a batch at reduced precision together with a bound on its own error,
//
//   S_ij = sum_k P_ik Q_jk              in bf16 x 2 (hi/lo split: Ph Qh + Ph Ql + Pl Qh), fp32 accumulation in TMEM
//   D_ij = sum_k |P_ik| |Q_jk|          in bf16 (1 % is enough for a bound)
//   |S_ij - exact| <= SCR_EPS * D_ij    (3 * 2^-18 from the dropped lo*lo and residual terms + the fp32 accumulation)
//
// and reduces every 128 x 128 tile to per-row maxima of the upper bound S + SCR_EPS*D and of the lower bound S - SCR_EPS*D.
// k_mcg_list_screened (mc_gemm.cuh) keeps a source tile for a target tile iff an upper bound reaches a row's best lower
// bound - 40; the FP64 contraction then visits only those tiles.  tcgen05 has no FP64 kind, so the weights themselves stay on
// the DMMA path; what runs here is 2 * 576 * G^2 bf16 flop per batch, a few per cent of the FP64 work it removes.
//
// Kernel: one CTA = one 128-row target block x a range of 128-row source tiles.  Warp 0 streams the operand tiles with
// bulk async copies (the global layout is the shared-memory image: 128 rows x 64 bf16, 128-byte swizzle, one 16 KB block
// per copy); one lane of warp 1 issues tcgen05.mma (M = N = 128, K = 16 per instruction, A and B from shared memory
// descriptors, accumulators S and D in TMEM, double buffered: 2 x 256 of the 512 columns); warps 2-5 read the
// accumulators back with tcgen05.ld and fold them into the row maxima while the next tile is being multiplied.
#include "screen.hpp"
#include <cuda_bf16.h>

namespace cbs {

constexpr int SCR_BM = 128;                 // target rows per CTA (UMMA M)
constexpr int SCR_BN = 128;                 // source rows per tile (UMMA N) = MCG_TILE
constexpr int SCR_KTILE = 64;               // bf16 per operand row and k-tile: 128 bytes, the swizzle span
constexpr int SCR_TILE_BYTES = SCR_BM * SCR_KTILE * 2;      // 16 KB
constexpr int SCR_MAX_KT = 9;               // k-tiles of the resident A operand (4 segments of <= 144 columns)
constexpr int SCR_STAGES = 4;               // ring of B k-tiles
constexpr int SCR_THREADS = 192;            // warp 0 producer, warp 1 MMA issuer, warps 2-5 epilogue
constexpr float SCR_EPS = 3.1e-5f;          // error bound of S relative to D (see above), including 1 % for D's own rounding
constexpr float SCR_EPS_C = 2.5e-7f;        // fp32 rounding of the per-source constant c_j added in the epilogue
static_assert(SCR_BN == MCG_TILE, "screening tiles are the contraction's source tiles");

struct ScrShape {
    int K1;          // columns per segment: KS = KT - 2 screened columns (no order-0 Chebyshev term, no c_j), padded to 16
    int nsteps_s;    // UMMA steps (K = 16) that accumulate into S: 3 * K1 / 16
    int nsteps;      // all steps: 4 * K1 / 16
    int nk;          // k-tiles: ceil(4 * K1 / 64)
    long rows;       // rows of P / Q (multiple of 128)
    int NJ;          // source tiles
};

// ---- operand conversion: FP64 P / Q (mc_gemm.cuh layout) -> bf16 tiles in the shared-memory image ------------------
// Segment s of an operand row holds K1 columns: s = 0: hi, 1: hi (P) / lo (Q), 2: lo (P) / hi (Q), 3: |hi|.
// Element (row r, column c) lives in block [c / 64][r / 128] (16 KB each) at byte (r % 128) * 128 +
// (((c % 64) / 8) ^ (r % 8)) * 16 + (c % 8) * 2: the 128-byte swizzle the UMMA descriptors below declare.
// One thread writes one 16-byte chunk (8 columns).  is_q selects the source operand's segment rule; rows_used: P is only
// built for the owned blocks, the other blocks are never read.
__global__ void __launch_bounds__(256)
k_scr_convert(const double* __restrict__ X, McgShape sh, ScrShape ss, int is_q, McgOwn own, long n_chunks_rows,
              __nv_bfloat16* __restrict__ out, float* __restrict__ cq) {
    const long t = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const int chunks_per_row = ss.nk * 8;
    const long lr = t / chunks_per_row;
    const int ch = (int)(t - lr * chunks_per_row);
    if (lr >= n_chunks_rows) return;
    const long r = is_q ? lr : own.row(lr);
    const int KS = sh.KT - 2;
    __align__(16) __nv_bfloat16 v[8];
#pragma unroll
    for (int e = 0; e < 8; ++e) {
        const int c = ch * 8 + e;
        const int seg = c / ss.K1, k = c - seg * ss.K1;
        float f = 0.f;
        if (seg < 4 && k < KS) {
            const int km = k < sh.n ? k : k + 1;                           // skip the order-0 Chebyshev column
            const double x = X[mcg_index(sh.rows, r, km)];
            const __nv_bfloat16 h = __double2bfloat16(x);
            const bool want_lo = is_q ? (seg == 1) : (seg == 2);
            if (seg == 3) f = fabsf(__bfloat162float(__double2bfloat16(fabs(x))));
            else if (want_lo) f = (float)(x - (double)__bfloat162float(h));
            else f = __bfloat162float(h);
        }
        v[e] = __float2bfloat16_rn(f);
    }
    const long blk = (long)(ch / 8) * (sh.rows / SCR_BM) + r / SCR_BM;
    const int rr = (int)(r % SCR_BM), c8 = ch % 8;
    char* dst = reinterpret_cast<char*>(out) + blk * SCR_TILE_BYTES + rr * 128 + ((c8 ^ (rr & 7)) << 4);
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(v);
    if (is_q && ch == 0) cq[r] = (float)X[mcg_index(sh.rows, r, sh.KT - 1)];      // c_j, added in fp32 in the epilogue
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
// K-major (bits 15, 16 = 0), N >> 3 in bits 17-22, M >> 4 in bits 24-28; bit 14 negates B.
__host__ __device__ constexpr uint32_t scr_idesc(bool neg_b) {
    return (1u << 4) | (1u << 7) | (1u << 10) | ((neg_b ? 1u : 0u) << 14) | ((uint32_t)(SCR_BN >> 3) << 17) | ((uint32_t)(SCR_BM >> 4) << 24);
}

struct ScrSmem {
    // [A k-tiles (resident)] [B ring] [c_j double buffer] [barriers] [tmem base]; the tiles need 1024-byte alignment
    static constexpr size_t a_bytes = (size_t)SCR_MAX_KT * SCR_TILE_BYTES;
    static constexpr size_t b_bytes = (size_t)SCR_STAGES * SCR_TILE_BYTES;
    static constexpr size_t cq_off = a_bytes + b_bytes;
    static constexpr size_t bar_off = cq_off + 2 * SCR_BN * sizeof(float);
    static constexpr size_t total = bar_off + 16 * sizeof(uint64_t) + 16 + 1024 /*alignment slack*/;
};

// items[w] = (target block, first source tile, number of source tiles).  hi / lo are [NJ][rows].
__global__ void __launch_bounds__(SCR_THREADS, 1)
k_scr_mma(const __nv_bfloat16* __restrict__ Pt, const __nv_bfloat16* __restrict__ Qt, const float* __restrict__ cq,
          ScrShape ss, const int3* __restrict__ items, float* __restrict__ hi, float* __restrict__ lo) {
    extern __shared__ unsigned char scr_raw[];
    unsigned char* smem = reinterpret_cast<unsigned char*>((reinterpret_cast<uintptr_t>(scr_raw) + 1023) & ~(uintptr_t)1023);
    unsigned char* a_s = smem;
    unsigned char* b_s = smem + ScrSmem::a_bytes;
    float* cq_s = reinterpret_cast<float*>(smem + ScrSmem::cq_off);                  // [2][128]
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + ScrSmem::bar_off);
    uint64_t* a_full = bars;                    // 1
    uint64_t* b_full = bars + 1;                // SCR_STAGES
    uint64_t* b_empty = bars + 1 + SCR_STAGES;  // SCR_STAGES
    uint64_t* acc_full = bars + 1 + 2 * SCR_STAGES;     // 2
    uint64_t* acc_empty = bars + 3 + 2 * SCR_STAGES;    // 2
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int3 item = items[blockIdx.x];
    const long iblk = item.x;
    const int j0 = item.y, nj = item.z;
    const long nrb = ss.rows / SCR_BM;

    if (threadIdx.x == 0) {
        sbar_init(a_full, 1);
        for (int s = 0; s < SCR_STAGES; ++s) { sbar_init(&b_full[s], 1); sbar_init(&b_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { sbar_init(&acc_full[s], 1); sbar_init(&acc_empty[s], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 2) {                            // TMEM: all 512 columns (two accumulator stages of S and D)
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
                sbulk_g2s(a_s + (size_t)kt * SCR_TILE_BYTES, reinterpret_cast<const char*>(Pt) + ((size_t)kt * nrb + iblk) * SCR_TILE_BYTES,
                          SCR_TILE_BYTES, a_full);
            int st = 0; uint32_t ph = 0;
            for (int jj = 0; jj < nj; ++jj) {
                for (int kt = 0; kt < ss.nk; ++kt) {
                    sbar_wait(&b_empty[st], ph ^ 1u);
                    sbar_expect_tx(&b_full[st], SCR_TILE_BYTES);
                    sbulk_g2s(b_s + (size_t)st * SCR_TILE_BYTES,
                              reinterpret_cast<const char*>(Qt) + ((size_t)kt * nrb + (j0 + jj)) * SCR_TILE_BYTES, SCR_TILE_BYTES, &b_full[st]);
                    if (++st == SCR_STAGES) { st = 0; ph ^= 1u; }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            sbar_wait(a_full, 0);
            tc_fence_after();
            const uint32_t a_addr = s_u32(a_s), b_addr = s_u32(b_s);
            int st = 0; uint32_t ph = 0;
            for (int jj = 0; jj < nj; ++jj) {
                const int as = jj & 1;
                sbar_wait(&acc_empty[as], ((uint32_t)(jj >> 1) & 1u) ^ 1u);
                tc_fence_after();
                const uint32_t d_s = tmem + (uint32_t)as * 256u, d_d = d_s + 128u;
                for (int kt = 0; kt < ss.nk; ++kt) {
                    sbar_wait(&b_full[st], ph);
                    tc_fence_after();
#pragma unroll
                    for (int q = 0; q < SCR_KTILE / 16; ++q) {
                        const int step = kt * (SCR_KTILE / 16) + q;
                        if (step < ss.nsteps) {
                            const uint64_t da = scr_desc(a_addr + (uint32_t)kt * SCR_TILE_BYTES + (uint32_t)q * 32u);
                            const uint64_t db = scr_desc(b_addr + (uint32_t)st * SCR_TILE_BYTES + (uint32_t)q * 32u);
                            const bool to_d = step >= ss.nsteps_s;
                            const uint32_t acc = (step != 0 && step != ss.nsteps_s) ? 1u : 0u;
                            tc_mma_bf16(to_d ? d_d : d_s, da, db, scr_idesc(false), acc);
                        }
                    }
                    tc_commit(&b_empty[st]);                      // the stage is free once these MMAs have read it
                    if (++st == SCR_STAGES) { st = 0; ph ^= 1u; }
                }
                tc_commit(&acc_full[as]);                         // accumulators of tile jj complete
            }
        }
    } else {
        // ===== epilogue: warps 2..5; warp w may touch TMEM lanes 32 (w % 4) .. +31 =====
        const int q = warp & 3;
        const int row = q * 32 + lane;                            // row of the target block = TMEM lane
        const long grow = iblk * SCR_BM + row;
        const int et = threadIdx.x - 64;                          // 0..127
        for (int jj = 0; jj < nj; ++jj) {
            const int as = jj & 1;
            const int J = j0 + jj;
            cq_s[as * SCR_BN + et] = cq[(long)J * SCR_BN + et];
            asm volatile("bar.sync 1, 128;" ::: "memory");        // the 4 epilogue warps
            sbar_wait(&acc_full[as], (uint32_t)(jj >> 1) & 1u);
            tc_fence_after();
            const uint32_t t_s = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)as * 256u, t_d = t_s + 128u;
            const bool diag = (long)J == iblk;
            float mhi = -INFINITY, mlo = -INFINITY;
#pragma unroll 1
            for (int c0 = 0; c0 < SCR_BN; c0 += 32) {
                float s[32], d[32];
                tc_ld32(t_s + (uint32_t)c0, s);
                tc_ld32(t_d + (uint32_t)c0, d);
                tc_ld_wait();
#pragma unroll
                for (int c = 0; c < 32; ++c) {
                    const float cj = cq_s[as * SCR_BN + c0 + c];
                    const float v = s[c] + cj;
                    const float e = fmaf(SCR_EPS, d[c], SCR_EPS_C * fabsf(cj));
                    mhi = fmaxf(mhi, v + e);
                    const float l = v - e;
                    mlo = fmaxf(mlo, (diag && (c0 + c) == row) ? -INFINITY : l);
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) sbar_arrive(&acc_empty[as]);
            __stcs(&hi[(long)J * ss.rows + grow], mhi);            // streamed: read once by k_mcg_list_screened, keep L2 for the operands
            __stcs(&lo[(long)J * ss.rows + grow], mlo);
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512u) : "memory");
    }
}

// Work items of the screening kernel: every owned target block x source-tile ranges of NJ / nchunk tiles, ordered
// range-major: the CTAs that are resident at the same time stream the SAME range of source tiles (a few MB), so the B
// operand is read from HBM once per range and served from L2 to the other SMs; block-major order has every SM streaming a
// different part of the 70 MB operand and the pass runs at HBM speed.
__global__ void k_scr_items(McgOwn own, long nrb, int NJ, int nchunk, int3* __restrict__ items) {
    const long lb = (long)blockIdx.x * blockDim.x + threadIdx.x;
    const long nb = own.nblocks(nrb * SCR_BM);
    if (lb >= nb) return;
    const int blk = (int)(lb * own.world + own.rank);
    for (int c = 0; c < nchunk; ++c) {
        const int a = (int)((long)NJ * c / nchunk), b = (int)((long)NJ * (c + 1) / nchunk);
        items[(long)c * nb + lb] = make_int3(blk, a, b - a);
    }
}

bool mc_screen(const double* P, const double* Q, const McgShape& sh, McgOwn own, int slot, float* hi, float* lo,
               cudaStream_t st) {
    ScrShape ss;
    const int KS = sh.KT - 2;
    ss.K1 = (KS + 15) / 16 * 16;
    ss.nsteps_s = 3 * ss.K1 / 16;
    ss.nsteps = 4 * ss.K1 / 16;
    ss.nk = (4 * ss.K1 + SCR_KTILE - 1) / SCR_KTILE;
    ss.rows = sh.rows;
    ss.NJ = (int)(sh.rows / SCR_BN);
    if (ss.nk > SCR_MAX_KT) return false;                        // wider batches: the caller visits all tiles
    int cc_major = 0, dev = 0;
    CBS_CUDA(cudaGetDevice(&dev));
    CBS_CUDA(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, dev));
    if (cc_major != 10) return false;                            // tcgen05 / TMEM
    const std::string sfx = slot ? "#1" : "#0";
    const long nrb = sh.rows / SCR_BM, own_blocks = own.nblocks(sh.rows);
    const size_t op_elems = (size_t)ss.nk * nrb * SCR_BM * SCR_KTILE;
    __nv_bfloat16* Pt = (__nv_bfloat16*)g_dev.get("mcs_Pt" + sfx, op_elems * 2);
    __nv_bfloat16* Qt = (__nv_bfloat16*)g_dev.get("mcs_Qt" + sfx, op_elems * 2);
    float* cq = g_dev.get<float>("mcs_cq" + sfx, (size_t)sh.rows);
    const int cpr = ss.nk * 8;
    CBS_LAUNCH(k_scr_convert, grid1(own_blocks * SCR_BM * cpr, 256), 256, 0, st, P, sh, ss, 0, own, own_blocks * SCR_BM, Pt, (float*)nullptr);
    CBS_LAUNCH(k_scr_convert, grid1(sh.rows * cpr, 256), 256, 0, st, Q, sh, ss, 1, own, sh.rows, Qt, cq);
    int sms = 148;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // about 8 waves of CTAs and at least 4 ranges; every CTA re-loads its 128-row A operand once, so ranges should not be
    // too short either
    int nchunk = (int)std::min<long>(std::max<long>(4, (8L * sms + own_blocks - 1) / std::max<long>(own_blocks, 1)), std::max(1, ss.NJ / 8));
    int3* items = (int3*)g_dev.get<int>("mcs_items" + sfx, (size_t)3 * own_blocks * nchunk + 3);
    CBS_LAUNCH(k_scr_items, grid1(own_blocks, 128), 128, 0, st, own, nrb, ss.NJ, nchunk, items);
    t_ctx->set_func_attr((const void*)k_scr_mma, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ScrSmem::total);
    CBS_LAUNCH(k_scr_mma, dim3((unsigned)(own_blocks * nchunk)), SCR_THREADS, ScrSmem::total, st, Pt, Qt, cq, ss, items, hi, lo);
    return true;
}

}  // namespace cbs
