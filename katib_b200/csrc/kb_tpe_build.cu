// kb_tpe_build.cu -- history upload, below/above split and Parzen-estimator construction on the GPU.
//
// Replaces, per parameter, what goptuna's tpe.Sampler.Sample does before it draws candidates
// (module not in the reference tree, go.mod:8; reached from sample.go:48-90 through trial.Suggest*):
//   getObservationPairs      -> the dense values[n x D] / loss[n] arrays handed to kb_tpe_set_history
//   splitObservationPairs    -> split_kernel (stable order by loss, gamma(n) = min(ceil(0.1 n), 25))
//   NewParzenEstimator       -> parzen_kernel (sort, prior insert, neighbour sigmas, magic clip, weight ramp)
//   sampleCategorical tables -> categorical_kernel
// Everything here is FP64 with the operation order of the CPU oracle (oracle/tpe.c); the file is
// compiled with -fmad=false so that mu, sigma and the weights come out bit-identical to it.
// The FP32 "folded" tables consumed by the scoring kernels are derived at the end of parzen_kernel.
#include <math.h>

#include "kb_internal.cuh"

#define BUILD_THREADS 1024
#define KB_EPS 1e-12

static int next_pow2(int64_t v) {
    int p = 1;
    while (p < v) p <<= 1;
    return p;
}
static inline int pad_to(int v, int m) { return (v + m - 1) / m * m; }

// ---- bitonic sort of (key, idx) pairs by (key asc, idx asc), one CTA, arrays in global memory -------------
#define SORT_SMEM_MAX 1024
__device__ void bitonic_sort_pairs(double *keys, uint32_t *idx, int npad) {
    // small arrays (the GetSuggestions-latency shapes) are sorted in shared memory: same network, same result, but a
    // stage costs a barrier instead of a round trip to L2
    __shared__ double s_keys[SORT_SMEM_MAX];
    __shared__ uint32_t s_idx[SORT_SMEM_MAX];
    if (npad <= SORT_SMEM_MAX) {
        for (int i = threadIdx.x; i < npad; i += blockDim.x) { s_keys[i] = keys[i]; s_idx[i] = idx[i]; }
        __syncthreads();
        for (int k = 2; k <= npad; k <<= 1) {
            for (int j = k >> 1; j > 0; j >>= 1) {
                for (int i = threadIdx.x; i < npad; i += blockDim.x) {
                    int ixj = i ^ j;
                    if (ixj > i) {
                        double a = s_keys[i], b = s_keys[ixj];
                        uint32_t ia = s_idx[i], ib = s_idx[ixj];
                        bool gt = (a > b) || (a == b && ia > ib);
                        bool up = (i & k) == 0;
                        if (gt == up) { s_keys[i] = b; s_keys[ixj] = a; s_idx[i] = ib; s_idx[ixj] = ia; }
                    }
                }
                __syncthreads();
            }
        }
        for (int i = threadIdx.x; i < npad; i += blockDim.x) { keys[i] = s_keys[i]; idx[i] = s_idx[i]; }
        __syncthreads();
        return;
    }
    for (int k = 2; k <= npad; k <<= 1) {
        for (int j = k >> 1; j > 0; j >>= 1) {
            for (int i = threadIdx.x; i < npad; i += blockDim.x) {
                int ixj = i ^ j;
                if (ixj > i) {
                    double a = keys[i], b = keys[ixj];
                    uint32_t ia = idx[i], ib = idx[ixj];
                    bool gt = (a > b) || (a == b && ia > ib);
                    bool up = (i & k) == 0;
                    if (gt == up) { keys[i] = b; keys[ixj] = a; idx[i] = ib; idx[ixj] = ia; }
                }
            }
            __syncthreads();
        }
    }
}

__device__ __forceinline__ int tpe_gamma_dev(int n) {
    int a = (int)ceil(0.1 * (double)n);
    return a > 25 ? 25 : a;
}

// weight of the observation at position `pos` (trial order) among m observations: tpe.DefaultWeights
__device__ __forceinline__ double tpe_ramp(int pos, int m) {
    if (m < 25) return 1.0;
    int r = m - 25;
    if (pos >= r) return 1.0;
    double start = 1.0 / (double)m;
    if (r == 1) return start;
    double step = (1.0 - start) / (double)(r - 1);
    return start + (double)pos * step;  // -fmad=false: separately rounded, like Go
}

// ---- split: rank trials by loss; pos[0][i] / pos[1][i] = position of trial i inside below / above --------
__device__ void split_body(const double *__restrict__ loss, int n, int npad, double *keys, uint32_t *idx, int32_t *pos) {
    __shared__ int s_cnt[BUILD_THREADS];
    for (int i = threadIdx.x; i < npad; i += blockDim.x) {
        keys[i] = i < n ? loss[i] : INFINITY;
        idx[i] = i < n ? (uint32_t)i : 0xFFFFFFFFu;
    }
    __syncthreads();
    bitonic_sort_pairs(keys, idx, npad);
    const int nb = tpe_gamma_dev(n);
    // flag: pos[0][trial] = 1 if the trial is among the nb best
    for (int i = threadIdx.x; i < n; i += blockDim.x) pos[i] = 0;
    __syncthreads();
    for (int r = threadIdx.x; r < nb; r += blockDim.x) pos[idx[r]] = 1;
    __syncthreads();
    // positions in trial order: exclusive scan of the flags (segment per thread + scan of segment counts)
    const int seg = (n + blockDim.x - 1) / blockDim.x;
    const int b0 = threadIdx.x * seg, b1 = min(n, b0 + seg);
    int c = 0;
    for (int i = b0; i < b1; i++) c += pos[i];
    s_cnt[threadIdx.x] = c;
    __syncthreads();
    for (int off = 1; off < blockDim.x; off <<= 1) {
        int v = threadIdx.x >= off ? s_cnt[threadIdx.x - off] : 0;
        __syncthreads();
        s_cnt[threadIdx.x] += v;
        __syncthreads();
    }
    int below_before = s_cnt[threadIdx.x] - c;
    for (int i = b0; i < b1; i++) {
        int f = pos[i];
        int above_before = i - below_before;
        pos[i] = f ? below_before : -1;
        pos[n + i] = f ? -1 : above_before;
        below_before += f;
    }
}

__global__ void __launch_bounds__(BUILD_THREADS)
split_kernel(const double *__restrict__ loss, int n, int npad, double *keys, uint32_t *idx, int32_t *pos) {
    split_body(loss, n, npad, keys, idx, pos);
}

// ---- Parzen estimator of one (numeric parameter, side) ---------------------------------------------------
__device__ void parzen_body(KbDim &dm, int d, int side, const double *__restrict__ values, int n, int D, int nb,
                            const int32_t *__restrict__ pos, double *keys, uint32_t *idx, int merge) {
    __shared__ double s_red[BUILD_THREADS];
    __shared__ double s_scalar[2];
    __shared__ int s_int[BUILD_THREADS];
    const int m = side == 0 ? nb : n - nb;  // observations on this side
    const int K0 = m + 1;  // components before the optional merge
    int K = K0;
    int npad = 2;  // this side's own power of two (>= K): the below side holds at most 26 entries
    while (npad < K) npad <<= 1;
    const double low = dm.lo, high = dm.hi, shift = dm.shift;
    double *mu = dm.mu[side], *sg = dm.sigma[side], *w = dm.w[side], *logc = dm.logc[side];
    const int tid = threadIdx.x, nt = blockDim.x;

    for (int i = tid; i < npad; i += nt) { keys[i] = INFINITY; idx[i] = 0xFFFFFFFFu; }
    __syncthreads();
    for (int i = tid; i < n; i += nt) {
        int p = pos[side * n + i];
        if (p >= 0) { keys[p] = values[(size_t)i * D + d] - shift; idx[p] = (uint32_t)p; }
    }
    __syncthreads();
    bitonic_sort_pairs(keys, idx, npad);

    const double prior_mu = 0.5 * (low + high), prior_sigma = 1.0 * (high - low);
    // prior_pos = number of observations strictly below prior_mu (searchsorted, side='left')
    int cnt = 0;
    for (int i = tid; i < m; i += nt) cnt += keys[i] < prior_mu ? 1 : 0;
    s_int[tid] = cnt;
    __syncthreads();
    for (int off = nt >> 1; off > 0; off >>= 1) {
        if (tid < off) s_int[tid] += s_int[tid + off];
        __syncthreads();
    }
    const int prior_pos = s_int[0];
    __syncthreads();

    for (int j = tid; j < K; j += nt) {
        if (j == prior_pos) { mu[j] = prior_mu; w[j] = 1.0 /* prior_weight */; }
        else {
            int s = j < prior_pos ? j : j - 1;
            mu[j] = keys[s];
            w[j] = tpe_ramp((int)idx[s], m);
        }
    }
    __syncthreads();
    if (m == 0) {
        if (tid == 0) sg[0] = prior_sigma;
    } else {
        for (int j = tid; j < K; j += nt) {
            double left = j == 0 ? low : mu[j - 1];
            double right = j == K - 1 ? high : mu[j + 1];
            double a = mu[j] - left, b = right - mu[j];
            double s = a > b ? a : b;
            if (j == 0) s = mu[1] - mu[0];              // consider_endpoints = false
            if (j == K - 1) s = mu[K - 1] - mu[K - 2];
            sg[j] = s;
        }
    }
    __syncthreads();
    if (tid == 0) {  // sequential sum: the oracle's (and a Go for-loop's) rounding order
        double s = 0.0;
        for (int j = 0; j < K; j++) s += w[j];
        s_scalar[0] = s;
    }
    __syncthreads();
    {
        const double wsum = s_scalar[0];
        const double maxsigma = 1.0 * (high - low);
        double dd = 1.0 + (double)K;
        if (dd > 100.0) dd = 100.0;
        const double minsigma = 1.0 * (high - low) / dd;  // consider_magic_clip = true
        for (int j = tid; j < K; j += nt) {
            w[j] = w[j] / wsum;
            double s = sg[j];
            if (s < minsigma) s = minsigma;
            if (s > maxsigma) s = maxsigma;
            if (j == prior_pos) s = prior_sigma;
            sg[j] = s;
        }
    }
    __syncthreads();
    // Optional (kb_tpe_set_option KB_OPT_MERGE_DUPLICATES): components with identical (mu, sigma) are one component
    // with the summed weight -- the same mixture, fewer terms.  Histories of integer / stepped parameters consist
    // almost entirely of such duplicates (10 000 observations on 64 levels leave ~200 distinct components).  Equal
    // (mu, sigma) pairs are adjacent after the sort, so this is a run-length compaction; weights add up in order.
    if (merge && K > 1) {
        double *tmp_sg = reinterpret_cast<double *>(dm.tab[side]);  // the FP32 table area is not written yet
        for (int j = tid; j < K; j += nt) idx[j] = (j == 0 || mu[j] != mu[j - 1] || sg[j] != sg[j - 1]) ? 1u : 0u;
        __syncthreads();
        const int seg = (K + nt - 1) / nt, b0 = tid * seg, b1 = min(K, b0 + seg);
        int c = 0;
        for (int j = b0; j < b1; j++) c += (int)idx[j];
        s_int[tid] = c;
        __syncthreads();
        for (int off = 1; off < nt; off <<= 1) {
            int v = tid >= off ? s_int[tid - off] : 0;
            __syncthreads();
            s_int[tid] += v;
            __syncthreads();
        }
        const int Knew = s_int[nt - 1];
        int base = s_int[tid] - c;
        __syncthreads();
        for (int j = b0; j < b1; j++)
            if (idx[j]) { idx[j] = 0x80000000u | (uint32_t)base; base++; }
        __syncthreads();
        for (int j = tid; j < K; j += nt) {
            if (!(idx[j] & 0x80000000u)) continue;  // heads only: sum the run in order
            const int p = (int)(idx[j] & 0x7FFFFFFFu);
            double ws = w[j];
            for (int jj = j + 1; jj < K && !(idx[jj] & 0x80000000u); jj++) ws += w[jj];
            keys[p] = ws; logc[p] = mu[j]; tmp_sg[p] = sg[j];
        }
        __syncthreads();
        for (int p = tid; p < Knew; p += nt) { w[p] = keys[p]; mu[p] = logc[p]; sg[p] = tmp_sg[p]; }
        __syncthreads();
        K = Knew;
    }
    // p_accept = sum_j w_j (Phi(high) - Phi(low)); terms in parallel, summed in order by one thread
    for (int j = tid; j < K; j += nt) {
        double den = sqrt(2.0) * sg[j];
        if (den < KB_EPS) den = KB_EPS;
        double ch = 0.5 * (1.0 + erf((high - mu[j]) / den));
        double cl = 0.5 * (1.0 + erf((low - mu[j]) / den));
        keys[j] = w[j] * (ch - cl);  // keys[] is free again (npad >= K)
    }
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int j = 0; j < K; j++) s += keys[j];
        s_scalar[1] = s;
    }
    __syncthreads();
    const double p_accept = s_scalar[1];
    double lmax = -INFINITY;
    for (int j = tid; j < K; j += nt) {
        double Z = sqrt(2.0 * M_PI) * sg[j];
        double lc = log(w[j] / Z / p_accept);
        logc[j] = lc;
        lmax = fmax(lmax, lc);
    }
    s_red[tid] = lmax;
    __syncthreads();
    for (int off = nt >> 1; off > 0; off >>= 1) {
        if (tid < off) s_red[tid] = fmax(s_red[tid], s_red[tid + off]);
        __syncthreads();
    }
    const double L = s_red[0];
    const int Kpad = (K + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD;
    if (tid == 0) { dm.K[side] = K; dm.Kpad[side] = Kpad; dm.p_accept[side] = p_accept; dm.L[side] = L; }
    // FP32 folded tables in normalised coordinates x_n = (x - low) / (high - low)
    const double inv_range = 1.0 / (high - low);
    const double LOG2E = 1.4426950408889634;
    float4 *tab = dm.tab[side];
    const bool quant = dm.q > 0.0;
    double dd2 = 1.0 + (double)K0;
    if (dd2 > 100.0) dd2 = 100.0;
    const double minsigma2 = 1.0 * (high - low) / dd2;  // the magic-clip floor (same expression as above)
    if (quant) {
        // fit R(tau) = log2( erfcx(y-h) - e^{-4yh} erfcx(y+h) ) on y in [0, YMAX] by interpolation at M Chebyshev nodes.
        // Spread over the block (this sits on the latency path of every GetSuggestions: one thread doing all of it
        // took ~0.2 ms): node values by M threads, Chebyshev coefficients by N+1 threads, the basis change by one.
        const int N = KB_QF_DEG, M = 40;
        const double h = dm.q / (2.0 * sqrt(2.0) * minsigma2);
        __syncthreads();  // s_red is free again
        if (tid < M) {
            const double th = M_PI * ((double)tid + 0.5) / (double)M;
            const double y = (cos(th) + 1.0) * (KB_QF_YMAX / 2.0);
            s_red[tid] = log2(erfcx(y - h) - exp(-4.0 * y * h) * erfcx(y + h));
        }
        __syncthreads();
        if (tid <= N) {
            double c = 0.0;
            for (int k = 0; k < M; k++) c += s_red[k] * cos((double)tid * (M_PI * ((double)k + 0.5) / (double)M));
            s_red[64 + tid] = c * (tid == 0 ? 1.0 : 2.0) / (double)M;
        }
        __syncthreads();
        if (tid == 0) {
            // Chebyshev -> monomial: T0 = 1, T1 = x, T_{n+1} = 2 x T_n - T_{n-1}
            const double *cheb = s_red + 64;
            double mono[KB_QF_DEG + 1], tp[KB_QF_DEG + 1], tc[KB_QF_DEG + 1], tn[KB_QF_DEG + 1];
            for (int i = 0; i <= N; i++) { mono[i] = 0.0; tp[i] = 0.0; tc[i] = 0.0; }
            tp[0] = 1.0; tc[1] = 1.0;
            mono[0] += cheb[0];
            for (int i = 0; i <= N; i++) mono[i] += cheb[1] * tc[i];
            for (int nn = 2; nn <= N; nn++) {
                for (int i = 0; i <= N; i++) tn[i] = (i > 0 ? 2.0 * tc[i - 1] : 0.0) - tp[i];
                for (int i = 0; i <= N; i++) { mono[i] += cheb[nn] * tn[i]; tp[i] = tc[i]; tc[i] = tn[i]; }
            }
            for (int i = 0; i < 16; i++) dm.qf_coef[side][i] = i <= N ? (float)mono[i] : 0.f;
            dm.qf_h0[side] = (float)(sqrt(LOG2E) * h);
            dm.qf_tau_s[side] = (float)(2.0 / (sqrt(LOG2E) * KB_QF_YMAX));
            dm.qf_hw[side] = (float)(0.5 * dm.q * inv_range);
        }
        __syncthreads();
        // self-check of the fit: the fast path is switched off (h0 <= 0) when the interpolant misses the exact bin
        // integral by more than 1e-6 anywhere on a 3x finer grid (happens for bins much wider than sigma)
        double err = 0.0;
        if (tid <= 128) {
            const double y = KB_QF_YMAX * (double)tid / 128.0, tau = 2.0 * y / KB_QF_YMAX - 1.0;
            double pv = (double)dm.qf_coef[side][KB_QF_DEG];
            for (int i = KB_QF_DEG - 1; i >= 0; i--) pv = pv * tau + (double)dm.qf_coef[side][i];
            const double exact = log2(erfcx(y - h) - exp(-4.0 * y * h) * erfcx(y + h));
            err = fabs(exp2(pv - exact) - 1.0);
            if (!(err == err)) err = INFINITY;  // NaN fails the check
        }
        s_red[tid] = err;
        __syncthreads();
        for (int off = nt >> 1; off > 0; off >>= 1) {
            if (tid < off) s_red[tid] = fmax(s_red[tid], s_red[tid + off]);
            __syncthreads();
        }
        // ... and when the bin is so wide against sigma that a component further than YMAX still matters: the kernel
        // clamps tau there, which is only harmless while erfc(YMAX - h) is negligible (found by the randomised tests:
        // a step-3 integer on a range of 7 with 260 observations has h = 15)
        if (tid == 0 && (!(s_red[0] < 1e-6) || h > KB_QF_YMAX - 4.6)) dm.qf_h0[side] = -1.f;
        __syncthreads();
    }
    // look-up class (see KbDim::lut): bin integral of a floor-sigma component at every integer grid offset
    int G = 0;
    if (quant) {
        const double gcount = (high - low) / dm.q;
        const double gr = rint(gcount);
        if (fabs(gcount - gr) <= 1e-9 * gcount && gr >= 1.0 && 2.0 * gr - 1.0 <= (double)KB_LUT_MAX) G = (int)gr;
        const double rr = 1.0 / (sqrt(2.0) * minsigma2), h = 0.5 * dm.q * rr;
        for (int k = tid; k < 2 * G - 1; k += nt) {
            const double y = fabs((double)(k - (G - 1))) * dm.q * rr;
            dm.lut[side][k] = (float)(erfc(y - h) - erfc(y + h));
        }
        if (tid == 0) dm.lut_G[side] = G;
    }
    for (int j = tid; j < Kpad; j += nt) {
        float4 A, B;  // A = {-mu_hi, -mu_hi, s, s}  B = {-mu_lo*s, -mu_lo*s, t, t}
        if (j >= K) {
            A = make_float4(0.f, 0.f, 0.f, 0.f);
            B = quant ? make_float4(0.f, 0.f, 0.f, 0.f) : make_float4(0.f, 0.f, -INFINITY, -INFINITY);
        } else {
            double mun = (mu[j] - low) * inv_range;
            float mh = (float)mun;
            double ml = mun - (double)mh;
            if (quant) {
                double den = sqrt(2.0) * sg[j];
                if (den < KB_EPS) den = KB_EPS;
                float r = (float)(sqrt(LOG2E) * (high - low) / den);  // sqrt(log2 e) / (sqrt2 sigma_n)
                // a negative weight marks the sigma == magic-clip-floor class served by the fitted fast path
                const float wf = (float)w[j];
                const float ws = sg[j] == minsigma2 ? -wf : wf;
                const float mr = (float)(-ml * (double)r);
                int code = 0;  // grid index of mu, pre-scaled to the look-up table's row pitch (128 B), +64 so that 0 = none
                if (G > 0 && sg[j] == minsigma2) {
                    const double gi = (mu[j] - (low + 0.5 * dm.q)) / dm.q, gr = rint(gi);
                    if (fabs(gi - gr) <= 1e-9 && gr >= 0.0 && gr < (double)G) code = (int)gr * 128 + 64;
                }
                A = make_float4(-mh, -mh, r, r);
                B = make_float4(mr, mr, ws, __int_as_float(code));
            } else {
                double s = sg[j] > KB_EPS ? sg[j] : KB_EPS;
                float c = (float)(sqrt(0.5 * LOG2E) * (high - low) / s);
                const float mc = (float)(-ml * (double)c), kk = (float)((logc[j] - L) * LOG2E);
                A = make_float4(-mh, -mh, c, c);
                B = make_float4(mc, mc, kk, kk);
            }
        }
        tab[2 * j] = A;
        tab[2 * j + 1] = B;
        if (!quant) dm.tabc[side][j] = make_float4(A.x, A.z, B.x, B.z);
    }
    // Quantised parameters: split the entries, in order, into the look-up class (grid code != 0: dense weights + run
    // table, see KbDim::qw) and the rest (tabw).  A stable partition by a block scan over per-thread segments, then a
    // second scan over the look-up-class entries that numbers the runs of equal grid code.  Every run is padded with
    // zero weights to a multiple of four entries and starts on a multiple of four, so that the grid kernel reads it
    // as whole LDS.128s.
    if (quant) {
        __syncthreads();
        float *qw = dm.qw[side];
        float4 *to = dm.tabw[side];
        int4 *runs = dm.qruns[side];
        const int seg = (K + nt - 1) / nt, b0 = min(K, tid * seg), b1 = min(K, b0 + seg);
        int c = 0;
        for (int j = b0; j < b1; j++) c += __float_as_int(tab[2 * j + 1].w) != 0 ? 1 : 0;
        s_int[tid] = c;
        __syncthreads();
        for (int off = 1; off < nt; off <<= 1) {
            int v = tid >= off ? s_int[tid - off] : 0;
            __syncthreads();
            s_int[tid] += v;
            __syncthreads();
        }
        const int Kg = s_int[nt - 1], Ko = K - Kg;
        int bg = s_int[tid] - c, bo = b0 - bg;
        __syncthreads();
        // The other components that sit ON the grid too (wider sigma: the below side of a dense history, and the two
        // components at the ends of every run of equal values on the above side, whose neighbour distance is one step)
        // get a table row as well: for an on-grid candidate their bin integral depends only on the integer offset
        // too, with their own sigma.  Components with the same sigma share a row, so a handful of rows serves them
        // all; rows are numbered in order of first appearance while the budget lasts (KB_LUT2_FLOATS).  The entry's .w
        // then carries 0x40000000 | row << 16 | grid index of mu (bit 29: this entry is the row's first user).
        const int rowlen = 2 * G - 1;
        const int max_rows = G > 0 ? min(KB_LUT2_FLOATS / rowlen, 255) : 0;
        const int Kl = min(Ko, 2048);  // entries considered (the pairing below is quadratic)
        for (int j = b0; j < b1; j++) {
            const float4 Bj = tab[2 * j + 1];
            const int code = __float_as_int(Bj.w);
            if (code != 0) { keys[bg] = (double)fabsf(Bj.z); idx[bg] = (uint32_t)code; bg++; }
            else {
                uint32_t mark = 0u;
                if (max_rows > 0 && bo < Kl) {
                    const double gi = (mu[j] - (low + 0.5 * dm.q)) / dm.q, gr = rint(gi);
                    if (fabs(gi - gr) <= 1e-9 && gr >= 0.0 && gr < (double)G) mark = 0x80000000u | (uint32_t)gr;
                }
                keys[Kg + bo] = sg[j];  // keys[] and idx[] hold K <= npad entries: [0, Kg) look-up class, [Kg, K) the others
                idx[Kg + bo] = mark;
                to[2 * bo] = tab[2 * j]; to[2 * bo + 1] = Bj; bo++;
            }
        }
        const float4 Z = make_float4(0.f, 0.f, 0.f, 0.f);
        for (int j = Ko + tid; j < (Ko + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD; j += nt) { to[2 * j] = Z; to[2 * j + 1] = Z; }
        __syncthreads();
        for (int o = tid; o < Kl; o += nt) {  // pass 1: is this entry the first on-grid one with its sigma?
            if (!(idx[Kg + o] & 0x80000000u)) continue;
            const double so = keys[Kg + o];
            bool rep = true;
            for (int e = 0; e < o; e++)
                if ((idx[Kg + e] & 0x80000000u) && keys[Kg + e] == so) { rep = false; break; }
            float4 Bo = to[2 * o + 1];
            Bo.w = __int_as_float(0x40000000 | (rep ? 0x20000000 : 0));
            to[2 * o + 1] = Bo;
        }
        __syncthreads();
        for (int o = tid; o < Kl; o += nt) {  // pass 2: row = number of first-users before this sigma's first user
            if (!(idx[Kg + o] & 0x80000000u)) continue;
            const double so = keys[Kg + o];
            int f = o;
            for (int e = 0; e < o; e++)
                if ((idx[Kg + e] & 0x80000000u) && keys[Kg + e] == so) { f = e; break; }
            int row = 0;
            for (int e = 0; e < f; e++) row += (__float_as_int(to[2 * e + 1].w) & 0x20000000) ? 1 : 0;
            idx[Kg + o] = 0x80000000u | ((uint32_t)min(row, 0x7FFF) << 16) | (idx[Kg + o] & 0xFFFFu);
        }
        __syncthreads();
        {   // pass 3: final marks; the first users are listed by row, then the rows are computed entry by entry in parallel
            float *lut2 = dm.lut2[side];
            if (tid == 0) s_int[BUILD_THREADS - 1] = 0;  // number of rows (rows < 256 never reach this slot)
            __syncthreads();
            for (int o = tid; o < Kl; o += nt) {
                if (!(idx[Kg + o] & 0x80000000u)) continue;
                const int row = (int)((idx[Kg + o] >> 16) & 0x7FFFu), gmu = (int)(idx[Kg + o] & 0xFFFFu);
                float4 Bo = to[2 * o + 1];
                const bool rep = (__float_as_int(Bo.w) & 0x20000000) != 0;
                Bo.w = row < max_rows ? __int_as_float(0x40000000 | (rep ? 0x20000000 : 0) | (row << 16) | gmu) : 0.f;
                to[2 * o + 1] = Bo;
                if (rep && row < max_rows) { s_int[row] = o; atomicMax(&s_int[BUILD_THREADS - 1], row + 1); }
            }
            __syncthreads();
            const int n_rows = s_int[BUILD_THREADS - 1];
            for (int e = tid; e < n_rows * rowlen; e += nt) {
                const int row = e / rowlen, k = e - row * rowlen;
                double den = sqrt(2.0) * keys[Kg + s_int[row]];
                if (den < KB_EPS) den = KB_EPS;
                const double rr = 1.0 / den, h = 0.5 * dm.q * rr;
                const double y = fabs((double)(k - (G - 1))) * dm.q * rr;
                lut2[e] = (float)(erfc(y - h) - erfc(y + h));
            }
            __syncthreads();
        }
        const int seg2 = (Kg + nt - 1) / nt, g0 = min(Kg, tid * seg2), g1 = min(Kg, g0 + seg2);
        int h = 0;
        for (int p = g0; p < g1; p++) h += (p == 0 || idx[p] != idx[p - 1]) ? 1 : 0;
        s_int[tid] = h;
        __syncthreads();
        for (int off = 1; off < nt; off <<= 1) {
            int v = tid >= off ? s_int[tid - off] : 0;
            __syncthreads();
            s_int[tid] += v;
            __syncthreads();
        }
        const int n_runs = s_int[nt - 1];  // <= grid points <= KB_LUT_MAX / 2 (codes are non-decreasing in table order)
        const int r0 = s_int[tid] - h;     // runs that begin before this thread's segment
        {
            int r = r0;
            for (int p = g0; p < g1; p++)
                if (p == 0 || idx[p] != idx[p - 1]) { runs[r] = make_int4((int)idx[p], p, 0, 0); r++; }
        }
        __syncthreads();
        for (int q = tid; q < n_runs; q += nt) runs[q].z = (q + 1 < n_runs ? runs[q + 1].y : Kg) - runs[q].y;
        __syncthreads();
        if (tid == 0) {  // padded starts (at most 128 runs)
            int ps = 0;
            for (int q = 0; q < n_runs; q++) { runs[q].w = ps; ps += (runs[q].z + 3) & ~3; }
            s_int[0] = ps;
        }
        __syncthreads();
        const int Kgp = s_int[0];
        {   // scatter the weights to their padded places; the run of position p is r0 - 1 + (heads met so far)
            int r = r0 - 1;
            for (int p = g0; p < g1; p++) {
                if (p == 0 || idx[p] != idx[p - 1]) r++;
                qw[runs[r].w + (p - runs[r].y)] = (float)keys[p];
            }
        }
        for (int q = tid; q < n_runs; q += nt)
            for (int e = runs[q].w + runs[q].z; e < runs[q].w + ((runs[q].z + 3) & ~3); e++) qw[e] = 0.f;
        for (int e = Kgp + tid; e < (Kgp + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD; e += nt) qw[e] = 0.f;
        __syncthreads();
        for (int q = tid; q < n_runs; q += nt) { const int4 rq = runs[q]; runs[q] = make_int4(rq.x, rq.w, (rq.z + 3) & ~3, rq.z); }
        if (tid == 0) { dm.Kn[side] = Kgp; dm.Kw[side] = Ko; dm.n_runs[side] = n_runs; }
    }
    // KB_OPT_PRUNE_FAR (continuous parameters): the same entries split, in order, into floor-sigma ("narrow") and
    // other ("wide") components -- see KbDim::tabn.  A stable partition by a block scan over per-thread segments.
    if (!quant) {
        __syncthreads();
        const int seg = (K + nt - 1) / nt, b0 = min(K, tid * seg), b1 = min(K, b0 + seg);
        int c = 0;
        for (int j = b0; j < b1; j++) c += sg[j] == minsigma2 ? 1 : 0;
        s_int[tid] = c;
        __syncthreads();
        for (int off = 1; off < nt; off <<= 1) {
            int v = tid >= off ? s_int[tid - off] : 0;
            __syncthreads();
            s_int[tid] += v;
            __syncthreads();
        }
        const int Kn = s_int[nt - 1], Kw = K - Kn;
        int bn = s_int[tid] - c, bw = b0 - bn;
        float4 *tn = dm.tabn[side], *tw = dm.tabw[side];
        // ... and the same partition once more as the source of tpe_score_cont_local_kernel's per-block tables,
        // {normalised mu in FP64 (low word, high word), c', k'}: narrow entries first, wide ones from the padded Kn on
        const int Knp = (Kn + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD, Kwp = (Kw + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD;
        float4 *sn = dm.tabs[side], *sw = dm.tabs[side] + Knp;
        for (int j = b0; j < b1; j++) {
            const double mun = (mu[j] - low) * inv_range;
            const float4 S = make_float4(__int_as_float(__double2loint(mun)), __int_as_float(__double2hiint(mun)), tab[2 * j].z, tab[2 * j + 1].z);
            if (sg[j] == minsigma2) { tn[2 * bn] = tab[2 * j]; tn[2 * bn + 1] = tab[2 * j + 1]; sn[bn] = S; bn++; }
            else { tw[2 * bw] = tab[2 * j]; tw[2 * bw + 1] = tab[2 * j + 1]; sw[bw] = S; bw++; }
        }
        const float4 Z = make_float4(0.f, 0.f, 0.f, 0.f), NI = make_float4(0.f, 0.f, -INFINITY, -INFINITY);
        const float4 SP = make_float4(0.f, 0.f, 0.f, -INFINITY);  // mu = 0, c' = 0, k' = -inf: contributes 2^-inf
        for (int j = Kn + tid; j < Knp; j += nt) { tn[2 * j] = Z; tn[2 * j + 1] = NI; sn[j] = SP; }
        for (int j = Kw + tid; j < Kwp; j += nt) { tw[2 * j] = Z; tw[2 * j + 1] = NI; sw[j] = SP; }
        if (tid == 0) {
            dm.Kn[side] = Kn; dm.Kw[side] = Kw;
            dm.prune_w[side] = (float)(7.75 * minsigma2 * inv_range * (1.0 + 1e-6));
            dm.c0[side] = (float)(sqrt(0.5 * LOG2E) * (high - low) / minsigma2);  // c' of every narrow component
        }
    }
}

__global__ void __launch_bounds__(BUILD_THREADS)
parzen_kernel(KbDim *dims, const int *numeric_dims, const double *__restrict__ values, int n, int D, int nb,
              const int32_t *__restrict__ pos, int npad_stride, double *keys_all, uint32_t *idx_all, int merge) {
    const int slot = blockIdx.x, side = blockIdx.y;
    const int d = numeric_dims[slot];
    parzen_body(dims[d], d, side, values, n, D, nb, pos, keys_all + ((size_t)slot * 2 + side) * npad_stride,
                idx_all + ((size_t)slot * 2 + side) * npad_stride, merge);
}

// ---- categorical parameters: weighted bincount (trial order) + prior weight, log ratio, ranks ------------
#define CAT_CHUNK 2048
__device__ void categorical_body(KbDim &dm, int d, const double *__restrict__ values, int n, int D, int nb,
                                 const int32_t *__restrict__ pos) {
    __shared__ double s_sum[2];
    __shared__ int s_pos[CAT_CHUNK];
    __shared__ unsigned short s_cat[CAT_CHUNK];
    const int C = dm.n_choices;
    for (int side = 0; side < 2; side++) {
        const int m = side == 0 ? nb : n - nb;
        double *p = dm.cat_p[side];
        // every thread owns the categories c, c + blockDim, ...: their sums run over the trials IN ORDER (the
        // oracle's rounding); trials are staged through shared memory a chunk at a time so the scan is on-chip
        for (int c0 = 0; c0 < C; c0 += blockDim.x) {
            const int c = c0 + threadIdx.x;
            double acc = 0.0;
            for (int base = 0; base < n; base += CAT_CHUNK) {
                const int cnt = min(CAT_CHUNK, n - base);
                __syncthreads();
                for (int i = threadIdx.x; i < cnt; i += blockDim.x) {
                    s_pos[i] = pos[side * n + base + i];
                    s_cat[i] = (unsigned short)(int)values[(size_t)(base + i) * D + d];
                }
                __syncthreads();
                if (c < C)
                    for (int i = 0; i < cnt; i++)
                        if (s_pos[i] >= 0 && (int)s_cat[i] == c) acc += tpe_ramp(s_pos[i], m);
            }
            if (c < C) p[c] = acc + 1.0;  // prior_weight
        }
        __syncthreads();
        if (threadIdx.x == 0) {
            double sm = 0.0;
            for (int c = 0; c < C; c++) sm += p[c];
            s_sum[side] = sm;
        }
        __syncthreads();
        for (int c = threadIdx.x; c < C; c += blockDim.x) p[c] = p[c] / s_sum[side];
        __syncthreads();
    }
    for (int c = threadIdx.x; c < C; c += blockDim.x) dm.cat_score[c] = log(dm.cat_p[0][c]) - log(dm.cat_p[1][c]);
    __syncthreads();
    for (int c = threadIdx.x; c < C; c += blockDim.x) {  // rank = number of strictly smaller scores (ties share it)
        double sc = dm.cat_score[c];
        int r = 0;
        for (int e = 0; e < C; e++) r += dm.cat_score[e] < sc ? 1 : 0;
        dm.cat_rank[c] = r;
    }
    if (threadIdx.x == 0) { dm.K[0] = dm.K[1] = C; dm.Kpad[0] = dm.Kpad[1] = C; }
}
__global__ void categorical_kernel(KbDim *dims, const int *cat_dims, const double *__restrict__ values, int n, int D,
                                   int nb, const int32_t *__restrict__ pos) {
    const int d = cat_dims[blockIdx.x];
    categorical_body(dims[d], d, values, n, D, nb, pos);
}

// ---- small histories (the GetSuggestions-latency shapes): ONE launch builds everything.  CTA (d, side) ranks the
// trials by loss for itself -- the split is a 128-element sort there, cheaper to repeat than to wait for -- and goes
// straight on to its estimator (numeric: one side each; categorical: CTA (d, 0) does both sides).
__global__ void __launch_bounds__(BUILD_THREADS)
tpe_build_small_kernel(KbDim *dims, const double *__restrict__ values, const double *__restrict__ loss, int n, int D, int nb,
                       int npad, double *keys_all, uint32_t *idx_all, int32_t *pos_all, int merge) {
    const int d = blockIdx.x, side = blockIdx.y;
    KbDim &dm = dims[d];
    if (dm.kind == KB_CATEGORICAL && side == 1) return;
    const size_t me = (size_t)d * 2 + side;
    double *keys = keys_all + me * npad;
    uint32_t *idx = idx_all + me * npad;
    int32_t *pos = pos_all + me * 2 * (size_t)(n > 0 ? n : 1);
    split_body(loss, n, npad, keys, idx, pos);
    __syncthreads();
    if (dm.kind == KB_CATEGORICAL) categorical_body(dm, d, values, n, D, nb, pos);
    else parzen_body(dm, d, side, values, n, D, nb, pos, keys, idx, merge);
}
#define KB_BUILD_SMALL_MAX 1000  // trials up to which the single-launch build is used

// ---- host side -----------------------------------------------------------------------------------------------
int kb_tpe_free_history(kb_ctx *ctx) {
    cudaFree(ctx->d_values); cudaFree(ctx->d_loss); cudaFree(ctx->d_pos);
    cudaFree(ctx->d_sort_keys); cudaFree(ctx->d_sort_idx); cudaFree(ctx->d_tabmem);
    cudaFree(ctx->d_small_keys); cudaFree(ctx->d_small_idx); cudaFree(ctx->d_small_pos);
    ctx->d_small_keys = nullptr; ctx->d_small_idx = nullptr; ctx->d_small_pos = nullptr; ctx->small_cap = 0;
    ctx->d_values = ctx->d_loss = nullptr; ctx->d_pos = nullptr;
    ctx->d_sort_keys = nullptr; ctx->d_sort_idx = nullptr; ctx->d_tabmem = nullptr;
    ctx->hist_cap = 0; ctx->sort_cap = 0; ctx->tabmem_bytes = 0;
    ctx->history_ready = false; ctx->n_trials = 0;
    return KB_OK;
}

// Shared by kb_tpe_set_history (keep = 0) and kb_tpe_append_history (keep = the rows already on the device): rows
// [keep, keep + n_add) come from the host, then every table is rebuilt from the whole device-resident history.
static int tpe_load_history(kb_ctx *ctx, const char *who, int64_t keep64, int64_t n_add, const double *values,
                            const double *loss) {
    if (ctx->D == 0) return kb_fail(ctx, KB_ERR_STATE, "%s: call kb_space_set first", who);
    if (n_add < 0 || keep64 + n_add > (1 << 22))
        return kb_fail(ctx, KB_ERR_UNSUPPORTED, "%s: n_trials=%lld outside [0, 2^22]", who, (long long)(keep64 + n_add));
    if (n_add > 0 && (!values || !loss)) return kb_fail(ctx, KB_ERR_INVALID, "%s: NULL buffer", who);
    const int keep = (int)keep64, n = (int)(keep64 + n_add), D = ctx->D;
    cudaStream_t s = ctx->stream;
    ctx->history_ready = false;

    if (n > ctx->hist_cap) {
        double *old_values = ctx->d_values, *old_loss = ctx->d_loss;
        cudaFree(ctx->d_pos);
        ctx->d_values = ctx->d_loss = nullptr; ctx->d_pos = nullptr;
        int64_t cap = n + n / 4 + 64;
        ctx->hist_cap = 0;
        KB_CUDA(ctx, cudaMalloc(&ctx->d_values, sizeof(double) * (size_t)cap * D));
        KB_CUDA(ctx, cudaMalloc(&ctx->d_loss, sizeof(double) * (size_t)cap));
        KB_CUDA(ctx, cudaMalloc(&ctx->d_pos, sizeof(int32_t) * 2 * (size_t)cap));
        if (keep > 0) {  // growing under an append: the rows already uploaded move over on the device
            KB_CUDA(ctx, cudaStreamSynchronize(s));
            KB_CUDA(ctx, cudaMemcpy(ctx->d_values, old_values, sizeof(double) * (size_t)keep * D, cudaMemcpyDeviceToDevice));
            KB_CUDA(ctx, cudaMemcpy(ctx->d_loss, old_loss, sizeof(double) * (size_t)keep, cudaMemcpyDeviceToDevice));
        }
        cudaFree(old_values); cudaFree(old_loss);
        ctx->hist_cap = cap;
    }
    const int npad = next_pow2((int64_t)n + 1);
    const int Kcap = pad_to(n + 1, KB_TABLE_PAD);
    const int slots = ctx->n_numeric > 0 ? ctx->n_numeric : 1;
    if ((int64_t)npad * 2 * slots > ctx->sort_cap) {
        cudaFree(ctx->d_sort_keys); cudaFree(ctx->d_sort_idx);
        ctx->d_sort_keys = nullptr; ctx->d_sort_idx = nullptr;
        ctx->sort_cap = (int64_t)npad * 2 * slots;
        KB_CUDA(ctx, cudaMalloc(&ctx->d_sort_keys, sizeof(double) * (size_t)ctx->sort_cap));
        KB_CUDA(ctx, cudaMalloc(&ctx->d_sort_idx, sizeof(uint32_t) * (size_t)ctx->sort_cap));
    }
    // table arena: per numeric dim and side 4 f64 arrays + 1 float4 array of Kcap; per categorical dim 3 f64 + 1 i32 of C
    size_t bytes = 0;
    for (int d = 0; d < D; d++) {
        const KbDim &k = ctx->h_dims[(size_t)d];
        if (k.kind == KB_CATEGORICAL) bytes += (size_t)pad_to(k.n_choices, 32) * (3 * sizeof(double) + sizeof(double));
        else bytes += 2 * (((size_t)Kcap * (4 * sizeof(double) + 8 * sizeof(float4)) + 2 * KB_TABLE_PAD * sizeof(float4)) + KB_LUT_MAX * sizeof(float) +
                           (KB_LUT_MAX / 2) * sizeof(int4) + (k.q > 0 ? KB_LUT2_FLOATS * sizeof(float) : 0));
    }
    if (bytes > ctx->tabmem_bytes) {
        cudaFree(ctx->d_tabmem); ctx->d_tabmem = nullptr;
        KB_CUDA(ctx, cudaMalloc(&ctx->d_tabmem, bytes));
        ctx->tabmem_bytes = bytes;
    }
    char *p = (char *)ctx->d_tabmem;
    for (int d = 0; d < D; d++) {
        KbDim &k = ctx->h_dims[(size_t)d];
        if (k.kind == KB_CATEGORICAL) {
            size_t c = (size_t)pad_to(k.n_choices, 32);
            k.cat_p[0] = (double *)p; p += c * sizeof(double);
            k.cat_p[1] = (double *)p; p += c * sizeof(double);
            k.cat_score = (double *)p; p += c * sizeof(double);
            k.cat_rank = (int32_t *)p; p += c * sizeof(double);
        } else {
            for (int side = 0; side < 2; side++) {
                k.tab[side] = (float4 *)p; p += 2 * (size_t)Kcap * sizeof(float4);
                k.tabn[side] = (float4 *)p; p += 2 * (size_t)Kcap * sizeof(float4);
                k.tabw[side] = (float4 *)p; p += 2 * (size_t)Kcap * sizeof(float4);
                k.tabc[side] = (float4 *)p; p += (size_t)Kcap * sizeof(float4);
                k.tabs[side] = (float4 *)p; p += ((size_t)Kcap + 2 * KB_TABLE_PAD) * sizeof(float4);  // two padded parts
                k.mu[side] = (double *)p; p += (size_t)Kcap * sizeof(double);
                k.sigma[side] = (double *)p; p += (size_t)Kcap * sizeof(double);
                k.w[side] = (double *)p; p += (size_t)Kcap * sizeof(double);
                k.logc[side] = (double *)p; p += (size_t)Kcap * sizeof(double);
                k.lut[side] = (float *)p; p += KB_LUT_MAX * sizeof(float);
                k.qruns[side] = (int4 *)p; p += (KB_LUT_MAX / 2) * sizeof(int4);
                k.qw[side] = (float *)k.tabn[side];  // quantised parameters have no "narrow" table: the area holds qw
                k.lut2[side] = nullptr;
                if (k.q > 0) { k.lut2[side] = (float *)p; p += KB_LUT2_FLOATS * sizeof(float); }
            }
        }
    }
    // Uploads go through a pinned staging buffer of the context: a copy from pageable memory costs the driver a
    // staging pass and a synchronisation per call (three of them were a third of the GetSuggestions latency); from
    // pinned memory the three copies are plain asynchronous DMA.  The descriptor table is sent only when it changed
    // (the table pointers move when the history crosses a multiple of 32 rows).
    {
        const size_t dims_b = sizeof(KbDim) * (size_t)D, val_b = sizeof(double) * (size_t)n_add * D, loss_b = sizeof(double) * (size_t)n_add;
        // compared without K / Kpad: the host mirrors them after every build, the device copy is written by the kernels
        std::vector<KbDim> canon(ctx->h_dims);
        for (KbDim &k : canon) { k.K[0] = k.K[1] = 0; k.Kpad[0] = k.Kpad[1] = 0; }
        const bool dims_changed = ctx->h_dims_sent.size() != dims_b || memcmp(ctx->h_dims_sent.data(), canon.data(), dims_b) != 0;
        const size_t need = dims_b + val_b + loss_b;
        if (need <= ((size_t)4 << 20)) {
            if (need > ctx->up_bytes) {
                if (ctx->h_up) { cudaStreamSynchronize(s); cudaFreeHost(ctx->h_up); ctx->h_up = nullptr; ctx->up_bytes = 0; }
                const size_t want = std::max<size_t>(need * 2, (size_t)64 << 10);
                KB_CUDA(ctx, cudaMallocHost(&ctx->h_up, want));
                ctx->up_bytes = want;
                if (!ctx->ev_up) KB_CUDA(ctx, cudaEventCreateWithFlags(&ctx->ev_up, cudaEventDisableTiming));
            } else if (ctx->ev_up) {
                KB_CUDA(ctx, cudaEventSynchronize(ctx->ev_up));  // the previous call's copies have left the buffer
            }
            char *hp = (char *)ctx->h_up;
            if (dims_changed) {
                memcpy(hp, canon.data(), dims_b);
                KB_CUDA(ctx, cudaMemcpyAsync(ctx->d_dims, hp, dims_b, cudaMemcpyHostToDevice, s));
            }
            if (n_add > 0) {
                memcpy(hp + dims_b, values, val_b);
                memcpy(hp + dims_b + val_b, loss, loss_b);
                KB_CUDA(ctx, cudaMemcpyAsync(ctx->d_values + (size_t)keep * D, hp + dims_b, val_b, cudaMemcpyHostToDevice, s));
                KB_CUDA(ctx, cudaMemcpyAsync(ctx->d_loss + keep, hp + dims_b + val_b, loss_b, cudaMemcpyHostToDevice, s));
            }
            KB_CUDA(ctx, cudaEventRecord(ctx->ev_up, s));
        } else {  // a large history: straight from the caller's memory
            if (dims_changed) KB_CUDA(ctx, cudaMemcpy(ctx->d_dims, canon.data(), dims_b, cudaMemcpyHostToDevice));  // canon is a local
            KB_CUDA(ctx, cudaMemcpyAsync(ctx->d_values + (size_t)keep * D, values, val_b, cudaMemcpyHostToDevice, s));
            KB_CUDA(ctx, cudaMemcpyAsync(ctx->d_loss + keep, loss, loss_b, cudaMemcpyHostToDevice, s));
        }
        if (dims_changed) ctx->h_dims_sent.assign((const char *)canon.data(), (const char *)canon.data() + dims_b);
    }
    // The tables are built when something first needs them (kb_tpe_ensure_tables): a small history followed by a small
    // kb_tpe_suggest is answered by ONE kernel that builds what it needs in shared memory (kb_tpe_small.cu) and never
    // asks for them.  K / Kpad are known on the host (n_side + 1) either way.
    const int nb = [&] { int a = (int)ceil(0.1 * (double)n); return a > 25 ? 25 : a; }();
    for (int d = 0; d < D; d++) {
        KbDim &k = ctx->h_dims[(size_t)d];
        for (int side = 0; side < 2; side++) {
            k.K[side] = k.kind == KB_CATEGORICAL ? k.n_choices : (side == 0 ? nb : n - nb) + 1;
            k.Kpad[side] = k.kind == KB_CATEGORICAL ? k.n_choices : pad_to(k.K[side], KB_TABLE_PAD);
        }
    }
    ctx->n_trials = n;
    ctx->n_below = nb;
    ctx->tables_dirty = true;
    ctx->history_ready = true;
    return KB_OK;
}

// Build (once per history) the below / above split and every table of every parameter.  Called, under the context's
// lock and with its device current, by every entry point that reads the tables.
int kb_tpe_ensure_tables(kb_ctx *ctx) {
    if (!ctx->tables_dirty) return KB_OK;
    const int n = (int)ctx->n_trials, D = ctx->D, nb = ctx->n_below;
    const int npad = next_pow2((int64_t)n + 1);
    cudaStream_t s = ctx->stream;
    static const bool small_ok = [] { const char *e = getenv("KB_BUILD_SMALL"); return !(e && e[0] == '0'); }();
    if (n <= KB_BUILD_SMALL_MAX && small_ok) {
        const size_t blocks = (size_t)D * 2;
        if ((int64_t)(blocks * npad) > ctx->small_cap) {
            cudaFree(ctx->d_small_keys); cudaFree(ctx->d_small_idx); cudaFree(ctx->d_small_pos);
            ctx->d_small_keys = nullptr; ctx->d_small_idx = nullptr; ctx->d_small_pos = nullptr; ctx->small_cap = 0;
            KB_CUDA(ctx, cudaMalloc(&ctx->d_small_keys, sizeof(double) * blocks * npad));
            KB_CUDA(ctx, cudaMalloc(&ctx->d_small_idx, sizeof(uint32_t) * blocks * npad));
            KB_CUDA(ctx, cudaMalloc(&ctx->d_small_pos, sizeof(int32_t) * blocks * 2 * npad));  // npad > n
            ctx->small_cap = (int64_t)(blocks * npad);
        }
        KbProf pf(ctx, 5, s);
        // every phase of the build ends in a barrier; 256 threads make each of them a third as expensive as 1024, and a
        // few hundred trials give 1024 threads nothing to do anyway
        const int threads = n <= 384 ? 256 : BUILD_THREADS;
        tpe_build_small_kernel<<<dim3((unsigned)D, 2), threads, 0, s>>>(ctx->d_dims, ctx->d_values, ctx->d_loss, n, D, nb, npad,
                                                                             ctx->d_small_keys, ctx->d_small_idx, ctx->d_small_pos,
                                                                             ctx->tpe_merge ? 1 : 0);
        KB_LAUNCH_CHECK(ctx);
    } else {
    KbProf pf(ctx, 5, s);
    split_kernel<<<1, BUILD_THREADS, 0, s>>>(ctx->d_loss, n, npad, ctx->d_sort_keys, ctx->d_sort_idx, ctx->d_pos);
    KB_LAUNCH_CHECK(ctx);
    if (ctx->n_numeric > 0) {
        // numeric dims in slot order = cont and quant lists merged by numeric_slot
        std::vector<int> numeric((size_t)ctx->n_numeric);
        for (int d = 0; d < D; d++) if (ctx->h_dims[(size_t)d].numeric_slot >= 0) numeric[(size_t)ctx->h_dims[(size_t)d].numeric_slot] = d;
        if (!ctx->d_numeric_dims) {
            KB_CUDA(ctx, cudaMalloc(&ctx->d_numeric_dims, sizeof(int) * numeric.size()));
            KB_CUDA(ctx, cudaMemcpy(ctx->d_numeric_dims, numeric.data(), sizeof(int) * numeric.size(), cudaMemcpyHostToDevice));
        }
        parzen_kernel<<<dim3((unsigned)ctx->n_numeric, 2), BUILD_THREADS, 0, s>>>(
            ctx->d_dims, ctx->d_numeric_dims, ctx->d_values, n, D, nb, ctx->d_pos, npad, ctx->d_sort_keys, ctx->d_sort_idx,
            ctx->tpe_merge ? 1 : 0);
        KB_LAUNCH_CHECK(ctx);
    }
    if (ctx->n_cat > 0) {
        categorical_kernel<<<ctx->n_cat, 128, 0, s>>>(ctx->d_dims, ctx->d_cat_dims, ctx->d_values, n, D, nb, ctx->d_pos);
        KB_LAUNCH_CHECK(ctx);
    }
    }
    // p_accept and L stay on the device (kb_tpe_get_parzen fetches them), so nothing here synchronises: on the
    // context's stream the kernels above are simply ordered before any scoring launch; a *_device entry point given
    // another stream waits on ev_build (kb_order_after_build)
    KB_CUDA(ctx, cudaEventRecord(ctx->ev_build, s));
    ctx->tables_dirty = false;
    return KB_OK;
}

extern "C" int kb_tpe_set_history(kb_ctx *ctx, int64_t n_trials, const double *values, const double *loss) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    return tpe_load_history(ctx, "kb_tpe_set_history", 0, n_trials, values, loss);
}

extern "C" int kb_tpe_append_history(kb_ctx *ctx, int64_t n_new, const double *values, const double *loss) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (!ctx->history_ready) return kb_fail(ctx, KB_ERR_STATE, "kb_tpe_append_history: no history set (call kb_tpe_set_history first)");
    return tpe_load_history(ctx, "kb_tpe_append_history", ctx->n_trials, n_new, values, loss);
}

extern "C" int kb_tpe_get_parzen(kb_ctx *ctx, int param, int side, int *K, double *w, double *mu, double *sigma,
                                 int cap, double *p_accept) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (!ctx->history_ready) return kb_fail(ctx, KB_ERR_STATE, "kb_tpe_get_parzen: no history set");
    if (param < 0 || param >= ctx->D || (side != 0 && side != 1))
        return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_get_parzen: param/side out of range");
    { int rc = kb_tpe_ensure_tables(ctx); if (rc) return rc; }
    KbDim k = ctx->h_dims[(size_t)param];
    // kb_tpe_set_history returns without waiting; the context's stream is non-blocking, so plain cudaMemcpy does not
    // order behind the build kernels by itself
    KB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    if (ctx->tpe_merge && k.kind != KB_CATEGORICAL)  // merged tables: the component count is decided on the device
        KB_CUDA(ctx, cudaMemcpy(&k, ctx->d_dims + param, sizeof(KbDim), cudaMemcpyDeviceToHost));
    const int n = k.K[side];
    if (K) *K = n;
    if (cap < n) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_get_parzen: cap=%d < K=%d", cap, n);
    if (k.kind == KB_CATEGORICAL) {
        if (w) KB_CUDA(ctx, cudaMemcpy(w, k.cat_p[side], sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost));
        if (p_accept) *p_accept = 1.0;
        return KB_OK;
    }
    if (w) KB_CUDA(ctx, cudaMemcpy(w, k.w[side], sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost));
    if (mu) KB_CUDA(ctx, cudaMemcpy(mu, k.mu[side], sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost));
    if (sigma) KB_CUDA(ctx, cudaMemcpy(sigma, k.sigma[side], sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost));
    if (p_accept) {
        KbDim dk;
        KB_CUDA(ctx, cudaMemcpy(&dk, ctx->d_dims + param, sizeof(KbDim), cudaMemcpyDeviceToHost));
        *p_accept = dk.p_accept[side];
    }
    return KB_OK;
}

extern "C" int kb_tpe_set_option(kb_ctx *ctx, int option, int value) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    if (option == KB_OPT_MERGE_DUPLICATES) {
        if (ctx->tpe_merge == (value != 0)) return KB_OK;
        ctx->tpe_merge = value != 0;
        if (!ctx->history_ready) return KB_OK;
        ctx->tables_dirty = true;  // rebuilt under the new setting, from the history on the device, at the next use
        return KB_OK;
    }
    if (option == KB_OPT_PRUNE_FAR) {  // both table forms are always built: nothing to redo
        ctx->tpe_prune = value != 0;
        return KB_OK;
    }
    if (option == KB_OPT_LOCAL_ORIGIN) {
        if (value < 0 || value > 2) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_set_option: KB_OPT_LOCAL_ORIGIN takes 0, 1 or 2");
        ctx->tpe_local = value;
        return KB_OK;
    }
    return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_set_option: unknown option %d", option);
}
