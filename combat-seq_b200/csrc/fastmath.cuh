// fastmath.cuh — FP64 reciprocal and logarithm for POSITIVE NORMAL arguments, branch-free.
//
// The GLM / APL inner loops spend most of their instructions inside 1/x and log(x) (one reciprocal and up to two
// logarithms per sample and sweep).  CUDA's general routines carry range checks and slow paths for zeros, denormals,
// infinities and NaN (log: ~90 SASS instructions and 3 branches, 1/x: ~65 and 8, sm_100a); every argument in those loops
// is a positive normal double by construction (1 + mu*phi >= 1, mu + r, ratios of positive counts shifted by 1e-8), so
// the checks can go.  Results are within 1 ulp (reciprocal) / 2 ulp (logarithm) of the correctly rounded value, which is
// the accuracy of the library routines themselves (tests/test_hostmath.py checks the host compilation against libm).
#pragma once
#include <cuda_runtime.h>
#include <cmath>

namespace cbs {

// 1/a for a positive normal double whose reciprocal is normal as well
__host__ __device__ __forceinline__ double rcp_pos(double a) {
#ifdef __CUDA_ARCH__
    double r;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(r) : "d"(a));      // MUFU.RCP64H: about 20 good bits
    double e = fma(-a, r, 1.0);
    r = fma(r, e, r);
    e = fma(-a, r, 1.0);
    return fma(r, e, r);
#else
    return 1.0 / a;
#endif
}

// log(x) for a positive normal double: x = 2^e m, m in [sqrt(1/2), sqrt(2)), log m = 2 atanh(f), f = (m - 1)/(m + 1),
// |f| <= 0.1716, odd series to f^21 (truncation below 2e-18 relative)
__host__ __device__ __forceinline__ double log_pos(double x) {
#ifdef __CUDA_ARCH__
    int hi = __double2hiint(x);
    const int lo = __double2loint(x);
    int e = (hi >> 20) - 1023;
    hi = (hi & 0x000fffff) | 0x3ff00000;                        // m in [1, 2)
    if (hi >= 0x3ff6a09f) { hi -= 0x00100000; e += 1; }         // m >= sqrt(2) (to 20 bits): halve it
    const double m = __hiloint2double(hi, lo);
#else
    int e;
    double m = 2.0 * frexp(x, &e);                              // m in [1, 2)
    e -= 1;
    if (m >= 1.4142131805419922) { m *= 0.5; e += 1; }          // the same 20-bit threshold
#endif
    const double f = (m - 1.0) * rcp_pos(m + 1.0);
    const double f2 = f * f;
    double p = fma(f2, 1.0 / 21, 1.0 / 19);
    p = fma(p, f2, 1.0 / 17);
    p = fma(p, f2, 1.0 / 15);
    p = fma(p, f2, 1.0 / 13);
    p = fma(p, f2, 1.0 / 11);
    p = fma(p, f2, 1.0 / 9);
    p = fma(p, f2, 1.0 / 7);
    p = fma(p, f2, 1.0 / 5);
    p = fma(p, f2, 1.0 / 3);
    const double lm = fma(2.0 * f * f2, p, 2.0 * f);           // log m = 2 f + 2 f^3 (1/3 + f^2/5 + ...)
    const double ed = (double)e;
    return fma(ed, 6.93147180369123816490e-01, fma(ed, 1.90821492927058770002e-10, lm));     // e * (ln2_hi + ln2_lo) + log m
}

}  // namespace cbs
