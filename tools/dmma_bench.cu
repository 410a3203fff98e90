// DMMA.8x8x4 throughput probes (development tool): how many warps per SM and which operand patterns reach the
// FP64 tensor peak.  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o dmma_bench dmma_bench.cu
#include <cuda_runtime.h>
#include <cstdio>
__device__ __forceinline__ void dmma(double& c0, double& c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};" : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
// MT x NT accumulator tiles per warp, operands from registers (distinct per tile row/col)
template <int MT, int NT>
__global__ void k_reg(double* out, int iters) {
    double acc[MT][NT][2], a[MT], b[NT];
    for (int m = 0; m < MT; ++m) a[m] = threadIdx.x * 1e-3 + m;
    for (int n = 0; n < NT; ++n) b[n] = threadIdx.x * 1e-4 + n;
    for (int m = 0; m < MT; ++m) for (int n = 0; n < NT; ++n) acc[m][n][0] = acc[m][n][1] = 0;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
            for (int n = 0; n < NT; ++n) dmma(acc[m][n][0], acc[m][n][1], a[m], b[n]);
    }
    double s = 0;
    for (int m = 0; m < MT; ++m) for (int n = 0; n < NT; ++n) s += acc[m][n][0] + acc[m][n][1];
    if (s == 123.456) out[0] = s;
}
// same, operands re-loaded from shared memory every k-step (the real kernel's pattern)
template <int MT, int NT>
__global__ void k_smem(double* out, int iters) {
    extern __shared__ double sm[];
    for (int i = threadIdx.x; i < 4096; i += blockDim.x) sm[i] = i * 1e-6;
    __syncthreads();
    double acc[MT][NT][2];
    for (int m = 0; m < MT; ++m) for (int n = 0; n < NT; ++n) acc[m][n][0] = acc[m][n][1] = 0;
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int off = ((lane >> 2) * 8 + ((lane & 3) ^ (((lane >> 3) & 1) << 2))) + w * 64;
    for (int i = 0; i < iters; ++i) {
        double a[MT], b[NT];
        const double* p = sm + ((i & 7) * 128) + off;
#pragma unroll
        for (int m = 0; m < MT; ++m) a[m] = p[m * 64];
#pragma unroll
        for (int n = 0; n < NT; ++n) b[n] = p[1024 + n * 64];
#pragma unroll
        for (int m = 0; m < MT; ++m)
#pragma unroll
            for (int n = 0; n < NT; ++n) dmma(acc[m][n][0], acc[m][n][1], a[m], b[n]);
    }
    double s = 0;
    for (int m = 0; m < MT; ++m) for (int n = 0; n < NT; ++n) s += acc[m][n][0] + acc[m][n][1];
    if (s == 123.456) out[0] = s;
}
template <class K>
static void run(const char* name, K kern, int tiles, int threads, int ctas_per_sm, size_t smem) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double* d; cudaMalloc(&d, 64);
    const int iters = 4096;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        cudaEventRecord(e0);
        kern<<<sms * ctas_per_sm, threads, smem>>>(d, iters);
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r && ms < best) best = ms;
    }
    const double flop = 512.0 * tiles * iters * (threads / 32) * (double)sms * ctas_per_sm;
    printf("%-28s warps/SM %2d : %.2f TFLOP/s  (%s)\n", name, threads / 32 * ctas_per_sm, flop / best / 1e9, cudaGetErrorString(cudaGetLastError()));
    cudaFree(d);
}
// Do DMMA (FP64 tensor) and DFMA (FP64 vector) share execution resources?  Even warps run DMMA 4x8 tiles, odd warps run
// 8 independent DFMA chains; mode 1 = DMMA warps only, 2 = DFMA warps only, 3 = both at once.
__global__ void k_mix(double* out, int iters, int mode) {
    const int w = threadIdx.x >> 5;
    double s = 0;
    if ((w & 4) == 0) {          // warps 0-3 (one per scheduler) DMMA, warps 4-7 DFMA
        if (!(mode & 1)) return;
        double acc[4][8][2], a[4], b[8];
        for (int m = 0; m < 4; ++m) a[m] = threadIdx.x * 1e-3 + m;
        for (int n = 0; n < 8; ++n) b[n] = threadIdx.x * 1e-4 + n;
        for (int m = 0; m < 4; ++m) for (int n = 0; n < 8; ++n) acc[m][n][0] = acc[m][n][1] = 0;
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int m = 0; m < 4; ++m)
#pragma unroll
                for (int n = 0; n < 8; ++n) dmma(acc[m][n][0], acc[m][n][1], a[m], b[n]);
        }
        for (int m = 0; m < 4; ++m) for (int n = 0; n < 8; ++n) s += acc[m][n][0] + acc[m][n][1];
    } else {
        if (!(mode & 2)) return;
        double c[8], x = 1.0 + threadIdx.x * 1e-9, y = 1e-9;
        for (int k = 0; k < 8; ++k) c[k] = k;
        for (int i = 0; i < iters; ++i) {
#pragma unroll
            for (int r = 0; r < 32; ++r)
#pragma unroll
                for (int k = 0; k < 8; ++k) c[k] = fma(c[k], x, y);
        }
        for (int k = 0; k < 8; ++k) s += c[k];
    }
    if (s == 123.456) out[0] = s;
}
static void run_mix(int mode) {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    double* d; cudaMalloc(&d, 64);
    const int iters = 4096;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int r = 0; r < 4; ++r) {
        cudaEventRecord(e0);
        k_mix<<<sms, 256>>>(d, iters, mode);        // 8 warps/SM: 4 DMMA + 4 DFMA, 1 + 1 per scheduler
        cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (r && ms < best) best = ms;
    }
    const double f_mma = (mode & 1) ? 512.0 * 32 * iters * 4 * (double)sms : 0;
    const double f_fma = (mode & 2) ? 2.0 * 32 * 8 * 32 * iters * 4 * (double)sms : 0;
    printf("mix mode %d: %.3f ms  DMMA %.2f TFLOP/s  DFMA %.2f TFLOP/s  (%s)\n", mode, best, f_mma / best / 1e9, f_fma / best / 1e9,
           cudaGetErrorString(cudaGetLastError()));
    cudaFree(d);
}
int main() {
    run("reg 4x8", k_reg<4, 8>, 32, 128, 1, 0);
    run("reg 4x8", k_reg<4, 8>, 32, 256, 1, 0);
    run("reg 4x8", k_reg<4, 8>, 32, 256, 2, 0);
    run("reg 2x4", k_reg<2, 4>, 8, 256, 1, 0);
    run("reg 2x4", k_reg<2, 4>, 8, 512, 1, 0);
    run("reg 2x4", k_reg<2, 4>, 8, 1024, 1, 0);
    run("reg 2x4", k_reg<2, 4>, 8, 1024, 2, 0);
    run("reg 4x4", k_reg<4, 4>, 16, 512, 1, 0);
    run("reg 2x8", k_reg<2, 8>, 16, 512, 1, 0);
    run("smem 4x8", k_smem<4, 8>, 32, 256, 1, 32768);
    run("smem 4x8", k_smem<4, 8>, 32, 256, 2, 32768);
    run_mix(1); run_mix(2); run_mix(3);
    return 0;
}
