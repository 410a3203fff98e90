// rng.cuh — counter-based negative-binomial sampler for the synthetic count generator.
//
// The reference's simulation study draws its count matrices with polyester::simulate_experiment
// (/root/reference/simulations/sim_DEpipe.R:98-113), whose core is rnbinom(mu = reads_per_transcript * fold change, size):
// a gamma-Poisson mixture.  Here every matrix element owns an independent Philox4x32-10 stream keyed by (seed, element
// index), so the matrix does not depend on the launch shape or on how it is sharded over GPUs:
//   lambda ~ Gamma(shape = 1/phi, scale = mu * phi)   (Marsaglia & Tsang 2000; shape < 1 by the u^(1/shape) boost)
//   x ~ Poisson(lambda)                               (product of uniforms below 10, Hoermann's PTRS above)
// Compiles for the host as well (tests/hostmath): the not-gpu tier checks the sampler's distribution without a device.
#pragma once
#include <cstdint>
#include <cmath>
#include <cuda_runtime.h>

#ifndef CBS_HD
#define CBS_HD __host__ __device__ inline
#endif

namespace cbs {

CBS_HD void philox4x32_10(uint32_t c[4], uint32_t k0, uint32_t k1) {
#pragma unroll
    for (int r = 0; r < 10; ++r) {
        const uint64_t p0 = (uint64_t)0xD2511F53u * c[0], p1 = (uint64_t)0xCD9E8D57u * c[2];
        const uint32_t n0 = (uint32_t)(p1 >> 32) ^ c[1] ^ k0, n2 = (uint32_t)(p0 >> 32) ^ c[3] ^ k1;
        c[1] = (uint32_t)p1; c[3] = (uint32_t)p0; c[0] = n0; c[2] = n2;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
}

// the stream of one matrix element: block b of the stream is Philox(counter = (index, b), key = seed)
struct ElemRng {
    uint64_t seed, index;
    uint32_t block = 0, buf[4];
    int have = 0;
    CBS_HD ElemRng(uint64_t seed_, uint64_t index_) : seed(seed_), index(index_) {}
    CBS_HD uint32_t next32() {
        if (have == 0) {
            buf[0] = (uint32_t)index; buf[1] = (uint32_t)(index >> 32); buf[2] = block++; buf[3] = 0x43425331u;
            philox4x32_10(buf, (uint32_t)seed, (uint32_t)(seed >> 32));
            have = 4;
        }
        return buf[--have];
    }
    CBS_HD double uniform() {                                   // 53 bits, in (0, 1)
        const uint32_t a = next32() >> 5, b = next32() >> 6;
        return ((double)a * 67108864.0 + (double)b + 0.5) * (1.0 / 9007199254740992.0);
    }
    CBS_HD double normal() {                                    // Box-Muller, one value per two uniforms
        const double u1 = uniform(), u2 = uniform();
        return sqrt(-2.0 * log(u1)) * cos(6.283185307179586 * u2);
    }
};

CBS_HD double sample_gamma(ElemRng& g, double shape) {          // scale 1
    const double a = shape < 1 ? shape + 1 : shape;
    const double d = a - 1.0 / 3.0, c = 1.0 / sqrt(9.0 * d);
    double v;
    for (;;) {
        const double x = g.normal();
        v = 1.0 + c * x;
        if (v <= 0) continue;
        v = v * v * v;
        const double u = g.uniform(), x2 = x * x;
        if (u < 1.0 - 0.0331 * x2 * x2) break;
        if (log(u) < 0.5 * x2 + d * (1.0 - v + log(v))) break;
    }
    double r = d * v;
    if (shape < 1) r *= pow(g.uniform(), 1.0 / shape);
    return r;
}

CBS_HD double sample_poisson(ElemRng& g, double lam) {
    if (!(lam > 0)) return 0;
    if (lam < 10) {                                             // Knuth: count uniforms until their product drops below e^-lam
        const double L = exp(-lam);
        double p = g.uniform();
        int k = 0;
        while (p > L) { ++k; p *= g.uniform(); }
        return k;
    }
    // PTRS (W. Hoermann, "The transformed rejection method for generating Poisson random variables", 1993)
    const double slam = sqrt(lam), loglam = log(lam), b = 0.931 + 2.53 * slam, a = -0.059 + 0.02483 * b;
    const double invalpha = 1.1239 + 1.1328 / (b - 3.4), vr = 0.9277 - 3.6224 / (b - 2);
    for (;;) {
        const double U = g.uniform() - 0.5, V = g.uniform(), us = 0.5 - fabs(U);
        const double k = floor((2 * a / us + b) * U + lam + 0.43);
        if (us >= 0.07 && V <= vr) return k;
        if (k < 0 || (us < 0.013 && V > us)) continue;
        if (log(V) + log(invalpha) - log(a / (us * us) + b) <= -lam + k * loglam - lgamma(k + 1)) return k;
    }
}

// one NB(mean mu, dispersion phi) draw; phi <= 0 is the Poisson limit.  Clamped to int32 (the engine's count type).
CBS_HD int32_t sample_nb(uint64_t seed, uint64_t index, double mu, double phi) {
    if (!(mu > 0) || !isfinite(mu)) return 0;
    ElemRng g(seed, index);
    double lam = mu;
    if (phi > 0 && isfinite(phi)) lam = sample_gamma(g, 1.0 / phi) * (mu * phi);
    const double x = sample_poisson(g, lam);
    return x < 2147483647.0 ? (int32_t)x : 2147483647;
}

}  // namespace cbs
