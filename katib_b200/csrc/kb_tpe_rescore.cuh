// kb_tpe_rescore.cuh -- the FP64 decision formula of the TPE path (gmmLogPDF below minus above, in the oracle's order:
// max, sum of exp, log; quantised: sum of w (Phi(ub) - Phi(lb))), one warp per candidate.  Shared by the staged
// pipeline's near-tie re-scoring (kb_tpe_score.cu) and the single-kernel GetSuggestions path (kb_tpe_small.cu).
#pragma once
#include <math.h>

#include "kb_internal.cuh"

#ifndef KB_EPS
#define KB_EPS 1e-12
#endif

// Every multiply that feeds an add is written with the _rn intrinsics: no FMA contraction, so the value is the same in
// every translation unit whatever its -fmad setting, and rounds like the oracle (built with -ffp-contract=off).
static __device__ double rescore_value(const KbDim &dm, double x, int lane) {
    double out[2];
    for (int side = 0; side < 2; side++) {
        const int K = dm.K[side];
        const double *mu = dm.mu[side], *sg = dm.sigma[side];
        if (dm.q > 0.0) {
            const double *w = dm.w[side];
            double ub = x + dm.q / 2.0; ub = ub > dm.hi ? dm.hi : ub;
            double lb = x - dm.q / 2.0; lb = lb < dm.lo ? dm.lo : lb;
            double p = 0.0;
            for (int j = lane; j < K; j += 32) {
                double den = sqrt(2.0) * sg[j];
                den = den < KB_EPS ? KB_EPS : den;
                const double cdf = 0.5 * (1.0 + erf((ub - mu[j]) / den)) - 0.5 * (1.0 + erf((lb - mu[j]) / den));
                p = __dadd_rn(p, __dmul_rn(w[j], cdf));
            }
            for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
            out[side] = log(p + KB_EPS) - log(dm.p_accept[side] + KB_EPS);
        } else {
            const double *logc = dm.logc[side];
            double m = -INFINITY;
            for (int j = lane; j < K; j += 32) {
                double s = sg[j] > KB_EPS ? sg[j] : KB_EPS;
                double dd = (x - mu[j]) / s;
                m = fmax(m, __dadd_rn(__dmul_rn(-0.5, __dmul_rn(dd, dd)), logc[j]));
            }
            for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
            double acc = 0.0;
            for (int j = lane; j < K; j += 32) {
                double s = sg[j] > KB_EPS ? sg[j] : KB_EPS;
                double dd = (x - mu[j]) / s;
                acc += exp(__dadd_rn(__dadd_rn(__dmul_rn(-0.5, __dmul_rn(dd, dd)), logc[j]), -m));
            }
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            out[side] = log(acc) + m;
        }
    }
    return out[0] - out[1];
}
