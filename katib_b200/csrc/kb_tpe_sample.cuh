// kb_tpe_sample.cuh -- the per-(candidate, parameter) draw of tpe.Sampler.Sample's first half, shared by the bulk
// sampling kernel (kb_tpe_sample.cu) and the fused small-problem kernel (kb_tpe_score.cu).  Not part of the ABI.
//
// goptuna draws from Go's math/rand (not reproducible outside Go), so only the distribution can be matched; the stream
// here is Philox4x32-10 keyed by the caller's seed with counter (candidate index, parameter, attempt).
#pragma once
#include <math.h>

#include "kb_internal.cuh"

struct Philox {
    uint32_t k0, k1;
    __device__ Philox(uint64_t seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
    __device__ uint4 operator()(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3) const {
        uint32_t a = k0, b = k1;
#pragma unroll
        for (int r = 0; r < 10; r++) {
            const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
            const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
            c0 = hi1 ^ c1 ^ a; c1 = lo1; c2 = hi0 ^ c3 ^ b; c3 = lo0;
            a += 0x9E3779B9u; b += 0xBB67AE85u;
        }
        return make_uint4(c0, c1, c2, c3);
    }
};
// 53-bit uniform in [0,1) from two 32-bit words
__device__ __forceinline__ double u01(uint32_t a, uint32_t b) {
    return (double)(((unsigned long long)a << 21) ^ (unsigned long long)(b >> 11)) * (1.0 / 9007199254740992.0);
}

#define SAMPLE_MAX_TRIES 4096

// Candidate `gi` of parameter d, in the internal (working-domain) representation:
//   numeric      Sampler.sampleFromGMM: pick a component of the BELOW mixture by its weight, draw N(mu, sigma), reject
//                until low <= x < high, then round to the quantum (Go math.Round: half away from zero)
//   categorical  Sampler.sampleFromCategoricalDist: inverse CDF over p_below
__device__ __forceinline__ double tpe_sample_one(const KbDim &dm, const Philox &ph, unsigned long long gi, int d) {
    if (dm.kind == KB_CATEGORICAL) {
        const uint4 r = ph((uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)d, 0u);
        const double u = u01(r.x, r.y);
        const double *p = dm.cat_p[0];
        double acc = 0.0;
        int c = dm.n_choices - 1;
        for (int e = 0; e < dm.n_choices; e++) { acc += p[e]; if (u < acc) { c = e; break; } }
        return (double)c;
    }
    const int K = dm.K[0];
    const double *w = dm.w[0], *mu = dm.mu[0], *sg = dm.sigma[0];
    double x = 0.5 * (dm.lo + dm.hi);
    // Rejection sampling, two attempts per round: one Box-Muller transform yields two independent normals (R cos, R sin),
    // each paired with its own component pick.  A lane whose first attempt falls outside [lo, hi) takes the second
    // without another logarithm, and a warp repeats the round only if some lane lost both -- the attempts are i.i.d.,
    // so the accepted draw has the distribution of attempt-after-attempt rejection (sampleFromGMM).
    for (uint32_t round = 0; round < SAMPLE_MAX_TRIES / 2; round++) {
        const uint4 r = ph((uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)d, round + 1u);
        const uint4 r2 = ph((uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)d | 0x80000000u, round + 1u);
        double u1 = u01(r2.x, r2.y);
        u1 = u1 < 1e-300 ? 1e-300 : u1;
        const double rad = sqrt(-2.0 * log(u1));
        double sn, cs;
        sincospi(2.0 * (((double)r.z + 0.5) * (1.0 / 4294967296.0)), &sn, &cs);
        bool done = false;
#pragma unroll
        for (int half = 0; half < 2 && !done; half++) {
            const double u = ((double)(half ? r.y : r.x) + 0.5) * (1.0 / 4294967296.0);
            double acc = 0.0;
            int a = K - 1;
            for (int j = 0; j < K; j++) { acc += w[j]; if (u < acc) { a = j; break; } }
            // _rn intrinsics: the same bits in every translation unit, with or without -fmad
            const double draw = __dadd_rn(mu[a], __dmul_rn(sg[a], __dmul_rn(rad, half ? sn : cs)));
            if (dm.lo <= draw && draw < dm.hi) { x = draw; done = true; }
        }
        if (done) break;
    }
    if (dm.q > 0.0) x = round(x / dm.q) * dm.q;
    return x;
}
