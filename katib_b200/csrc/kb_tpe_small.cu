// kb_tpe_small.cu -- GetSuggestions at Katib's own size in ONE kernel: a few hundred trials, a few dozen EI candidates.
//
// Replaces, per parameter, the whole of goptuna's tpe.Sampler.Sample (module not in the reference tree, go.mod:8;
// reached from sample.go:48-90): getObservationPairs / splitObservationPairs / NewParzenEstimator for both sides,
// sampleFromGMM, gmmLogPDF and the argmax.  The staged path spends its time between kernels at this size (upload,
// split, Parzen tables in HBM, FP32 folding, sampling, scoring, selection: every phase a round trip to L2 or a
// launch).  Here one CTA per parameter keeps the trial losses, the parameter's values and both estimators in shared
// memory, builds only what the FP64 decision formula reads (mu, sigma, weights, log-coefficients, p_accept), draws
// the candidates from the same Philox stream as tpe_sample_kernel, scores them with rescore_value and picks the
// first maximum.  Nothing is written to the context's tables: they stay "not built" (kb_tpe_ensure_tables builds them
// if a later call needs them), so a service that alternates history upload and suggestion never builds them at all.
//
// The estimator arithmetic follows parzen_kernel / categorical_kernel (kb_tpe_build.cu) operation for operation --
// same sums in the same order -- and this file, like that one, is compiled with -fmad=false, so the tables equal the
// oracle's bit for bit; sorting is done by counting ranks (stable: key, then position), which gives the order of the
// bitonic network there.
#include <math.h>

#include "kb_internal.cuh"
#include "kb_tpe_sample.cuh"
#include "kb_tpe_rescore.cuh"

#define LN_MAXT 1024          // most threads per CTA (KB_LEAN_THREADS picks 256 / 512 / 1024 for measurements)
#define LN_THREADS ((int)blockDim.x)
#define LN_MAX_N 512          // trials
#define LN_MAX_K (LN_MAX_N + 1)
#define LN_MAX_CAND 256       // EI candidates
#define LN_MAX_CHOICES 256

// tpe.DefaultWeights: weight of the observation at position `pos` (trial order) among m observations
__device__ __forceinline__ double ln_ramp(int pos, int m) {
    if (m < 25) return 1.0;
    const int r = m - 25;
    if (pos >= r) return 1.0;
    const double start = 1.0 / (double)m;
    if (r == 1) return start;
    const double step = (1.0 - start) / (double)(r - 1);
    return start + (double)pos * step;  // -fmad=false: separately rounded, like Go
}

// block-wide sum of an int (all threads get it); s_red has LN_THREADS entries
__device__ __forceinline__ int ln_block_sum(int v, int *s_red) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = v;
    __syncthreads();
    int t = 0;
    for (int w = 0; w < (int)blockDim.x / 32; w++) t += s_red[w];
    return t;
}

// NewParzenEstimator for one side, entirely in shared memory.  obs_val / obs_pos: the side's m observations in any
// order with their position in trial order (the weight ramp and the tie rule of the sort need it).  Writes mu, sg, w,
// logc (K entries, K returned; fewer than m + 1 when `merge` folds duplicates) and p_accept.  Mirrors parzen_kernel.
__device__ int ln_parzen(const double *obs_val, const int *obs_pos, int m, double low, double high, int merge, double *mu,
                         double *sg, double *w, double *logc, double *tmp, int km, int *itmp, int *s_red, double *p_accept_out) {
    const int tid = threadIdx.x;
    int K = m + 1;
    const double prior_mu = 0.5 * (low + high), prior_sigma = 1.0 * (high - low);
    int cnt = 0;
    for (int i = tid; i < m; i += LN_THREADS) cnt += obs_val[i] < prior_mu ? 1 : 0;
    const int prior_pos = ln_block_sum(cnt, s_red);
    // sorted place of every observation: rank by (value, position), the order of the bitonic network of parzen_kernel
    for (int i = tid; i < m; i += LN_THREADS) {
        const double v = obs_val[i];
        const int p = obs_pos[i];
        int r = 0;
        for (int j = 0; j < m; j++) r += (obs_val[j] < v || (obs_val[j] == v && obs_pos[j] < p)) ? 1 : 0;
        const int at = r < prior_pos ? r : r + 1;
        mu[at] = v;
        w[at] = ln_ramp(p, m);
    }
    if (tid == 0) { mu[prior_pos] = prior_mu; w[prior_pos] = 1.0; }
    __syncthreads();
    if (m == 0) {
        if (tid == 0) sg[0] = prior_sigma;
    } else {
        for (int j = tid; j < K; j += LN_THREADS) {
            const double left = j == 0 ? low : mu[j - 1];
            const double right = j == K - 1 ? high : mu[j + 1];
            const double a = mu[j] - left, b = right - mu[j];
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
        tmp[0] = s;
    }
    __syncthreads();
    {
        const double wsum = tmp[0];
        const double maxsigma = 1.0 * (high - low);
        double dd = 1.0 + (double)K;
        if (dd > 100.0) dd = 100.0;
        const double minsigma = 1.0 * (high - low) / dd;  // consider_magic_clip = true
        __syncthreads();
        for (int j = tid; j < K; j += LN_THREADS) {
            w[j] = w[j] / wsum;
            double s = sg[j];
            if (s < minsigma) s = minsigma;
            if (s > maxsigma) s = maxsigma;
            if (j == prior_pos) s = prior_sigma;
            sg[j] = s;
        }
    }
    __syncthreads();
    if (merge && K > 1) {  // KB_OPT_MERGE_DUPLICATES: runs of identical (mu, sigma) become one component, weights summed in order
        int heads = 0;
        for (int j = tid; j < K; j += LN_THREADS) {
            const int h = (j == 0 || mu[j] != mu[j - 1] || sg[j] != sg[j - 1]) ? 1 : 0;
            itmp[j] = h;
            heads += h;
        }
        const int Knew = ln_block_sum(heads, s_red);
        for (int j = tid; j < K; j += LN_THREADS) {
            if (!itmp[j]) continue;
            int p = 0;
            for (int e = 0; e < j; e++) p += itmp[e];
            double ws = w[j];
            for (int jj = j + 1; jj < K && !itmp[jj]; jj++) ws += w[jj];
            tmp[p] = ws; logc[p] = mu[j]; tmp[km + p] = sg[j];
        }
        __syncthreads();
        for (int p = tid; p < Knew; p += LN_THREADS) { w[p] = tmp[p]; mu[p] = logc[p]; sg[p] = tmp[km + p]; }
        __syncthreads();
        K = Knew;
    }
    for (int j = tid; j < K; j += LN_THREADS) {
        double den = sqrt(2.0) * sg[j];
        if (den < KB_EPS) den = KB_EPS;
        const double ch = 0.5 * (1.0 + erf((high - mu[j]) / den));
        const double cl = 0.5 * (1.0 + erf((low - mu[j]) / den));
        tmp[j] = w[j] * (ch - cl);
    }
    __syncthreads();
    if (tid == 0) {
        double s = 0.0;
        for (int j = 0; j < K; j++) s += tmp[j];
        tmp[2 * km] = s;
    }
    __syncthreads();
    const double p_accept = tmp[2 * km];
    for (int j = tid; j < K; j += LN_THREADS) {
        const double Z = sqrt(2.0 * M_PI) * sg[j];
        logc[j] = log(w[j] / Z / p_accept);
    }
    __syncthreads();
    *p_accept_out = p_accept;
    return K;
}

// shared-memory plan of one CTA (doubles first, then ints); LA = max(n + 1, n_choices of the widest categorical)
struct LnPlan {
    int n, km, la, nc;
    __host__ __device__ size_t doubles() const { return (size_t)2 * n + (2 * km + 1) + 4 * (size_t)la + 4 * 32 + 2 * (size_t)nc + LN_MAXT; }
    __host__ __device__ size_t ints() const { return (size_t)3 * n + km + 2 * LN_MAXT; }
    __host__ __device__ size_t bytes() const { return doubles() * sizeof(double) + ints() * sizeof(int); }
};

__global__ void __launch_bounds__(LN_MAXT)
tpe_suggest_lean_kernel(const KbDim *__restrict__ dims, const double *__restrict__ values, const double *__restrict__ loss_g, int n,
                        int D, int nb, int merge, int n_cand, uint64_t seed, LnPlan plan, KbWinner *__restrict__ win) {
    extern __shared__ __align__(16) double ln_smem[];
    __shared__ KbDim s_dm;
    const int d = blockIdx.x, tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    double *s_loss = ln_smem, *s_val = s_loss + plan.n, *tmp = s_val + plan.n;
    double *a_mu = tmp + (2 * plan.km + 1), *a_sg = a_mu + plan.la, *a_w = a_sg + plan.la, *a_lc = a_w + plan.la;
    double *b_mu = a_lc + plan.la, *b_sg = b_mu + 32, *b_w = b_sg + 32, *b_lc = b_w + 32;
    double *s_x = b_lc + 32, *s_sc = s_x + plan.nc, *s_bs = s_sc + plan.nc;
    int *s_flag = reinterpret_cast<int *>(s_bs + LN_MAXT), *s_pos = s_flag + plan.n, *s_opos = s_pos + plan.n;
    int *itmp = s_opos + plan.n, *s_red = itmp + plan.km, *s_bi = s_red + LN_MAXT;
    if (tid == 0) s_dm = dims[d];
    for (int i = tid; i < n; i += LN_THREADS) s_loss[i] = loss_g[i];
    __syncthreads();
    const KbDim &g = s_dm;
    const bool cat = g.kind == KB_CATEGORICAL;
    for (int i = tid; i < n; i += LN_THREADS) s_val[i] = values[(size_t)i * D + d] - (cat ? 0.0 : g.shift);
    // splitObservationPairs: rank by (loss, trial index); the nb best are "below"
    for (int i = tid; i < n; i += LN_THREADS) {
        const double l = s_loss[i];
        int r = 0;
        for (int j = 0; j < n; j++) r += (s_loss[j] < l || (s_loss[j] == l && j < i)) ? 1 : 0;
        s_flag[i] = r < nb ? 0 : 1;  // side
    }
    __syncthreads();
    // position of every trial inside its side, in trial order
    for (int i = tid; i < n; i += LN_THREADS) {
        const int f = s_flag[i];
        int p = 0;
        for (int j = 0; j < i; j++) p += s_flag[j] == f ? 1 : 0;
        s_pos[i] = p;
    }
    __syncthreads();
    if (cat) {
        // sampleCategorical tables: weighted bincount in trial order + prior weight, normalised; score = log pl - log pg
        const int C = g.n_choices;
        double *p0 = a_mu, *p1 = a_sg, *sc = a_w;
        // every trial's ramp weight once, in parallel (the divisions inside ln_ramp are what a serial loop cannot afford);
        // the per-category sums below then run over the trials IN ORDER, the oracle's rounding
        for (int i = tid; i < n; i += LN_THREADS) {
            s_loss[i] = ln_ramp(s_pos[i], s_flag[i] == 0 ? nb : n - nb);
            itmp[i] = (int)s_val[i];
        }
        __syncthreads();
        for (int side = 0; side < 2; side++) {
            double *p = side == 0 ? p0 : p1;
            for (int c = tid; c < C; c += LN_THREADS) {
                double acc = 0.0;
                for (int i = 0; i < n; i++)
                    if (s_flag[i] == side && itmp[i] == c) acc += s_loss[i];
                p[c] = acc + 1.0;  // prior_weight
            }
            __syncthreads();
            if (tid == 0) { double sm = 0.0; for (int c = 0; c < C; c++) sm += p[c]; tmp[0] = sm; }
            __syncthreads();
            const double sm = tmp[0];
            __syncthreads();
            for (int c = tid; c < C; c += LN_THREADS) p[c] = p[c] / sm;
            __syncthreads();
        }
        for (int c = tid; c < C; c += LN_THREADS) sc[c] = log(p0[c]) - log(p1[c]);
        __syncthreads();
        if (tid == 0) { s_dm.cat_p[0] = p0; s_dm.cat_p[1] = p1; s_dm.cat_score = sc; s_dm.K[0] = s_dm.K[1] = C; }
    } else {
        double pa0 = 0.0, pa1 = 0.0;
        int K0 = 0, K1 = 0;
        // both sides' observations compacted in trial order into one array (below first, then above); the losses are not
        // needed any more, their array takes the values
        for (int i = tid; i < n; i += LN_THREADS) {
            const int at = s_flag[i] == 0 ? s_pos[i] : nb + s_pos[i];
            s_opos[at] = s_pos[i];
            itmp[at] = i;
        }
        __syncthreads();
        for (int e = tid; e < n; e += LN_THREADS) s_loss[e] = s_val[itmp[e]];
        __syncthreads();
        K0 = ln_parzen(s_loss, s_opos, nb, g.lo, g.hi, merge, b_mu, b_sg, b_w, b_lc, tmp, plan.km, itmp, s_red, &s_bs[0]);
        K1 = ln_parzen(s_loss + nb, s_opos + nb, n - nb, g.lo, g.hi, merge, a_mu, a_sg, a_w, a_lc, tmp, plan.km, itmp, s_red, &s_bs[1]);
        __syncthreads();
        pa0 = s_bs[0]; pa1 = s_bs[1];
        __syncthreads();
        if (tid == 0) {
            s_dm.mu[0] = b_mu; s_dm.sigma[0] = b_sg; s_dm.w[0] = b_w; s_dm.logc[0] = b_lc;
            s_dm.mu[1] = a_mu; s_dm.sigma[1] = a_sg; s_dm.w[1] = a_w; s_dm.logc[1] = a_lc;
            s_dm.K[0] = K0; s_dm.K[1] = K1; s_dm.p_accept[0] = pa0; s_dm.p_accept[1] = pa1;
        }
    }
    __syncthreads();
    // ---- tpe.Sampler.Sample's second half: draw, score in FP64 (one warp per candidate), first maximum
    const Philox ph(seed);
    for (int i = tid; i < n_cand; i += LN_THREADS) {
        const double x = tpe_sample_one(s_dm, ph, (unsigned long long)i, d);
        s_x[i] = x;
        if (cat) {
            int c = (int)x;
            c = c < 0 ? 0 : (c >= s_dm.n_choices ? s_dm.n_choices - 1 : c);
            s_sc[i] = s_dm.cat_score[c];
        }
    }
    __syncthreads();
    if (!cat) {
        for (int i = warp; i < n_cand; i += LN_THREADS / 32) {
            const double v = rescore_value(s_dm, s_x[i], lane);
            if (lane == 0) s_sc[i] = v;
        }
        __syncthreads();
    }
    double bs = -INFINITY;
    int bi = n_cand;
    for (int i = tid; i < n_cand; i += LN_THREADS) {  // ascending i per thread: ">" keeps the first maximum
        const double v = s_sc[i];
        if (bi == n_cand || v > bs) { bs = v; bi = i; }
    }
    s_bs[tid] = bs; s_bi[tid] = bi;
    __syncthreads();
    for (int off = LN_THREADS / 2; off > 0; off >>= 1) {
        if (tid < off) {
            const double os = s_bs[tid + off];
            const int oi = s_bi[tid + off];
            if (oi < n_cand && (s_bi[tid] == n_cand || os > s_bs[tid] || (os == s_bs[tid] && oi < s_bi[tid]))) { s_bs[tid] = os; s_bi[tid] = oi; }
        }
        __syncthreads();
    }
    if (tid == 0) {
        KbWinner w;
        const int b = s_bi[0];
        w.score = s_bs[0];
        w.index = b < n_cand ? (long long)b : -1;
        double v = b < n_cand ? s_x[b] : 0.0;
        if (s_dm.kind == KB_DISCRETE || s_dm.kind == KB_STEP_INT) {  // sampleDiscreteUniform: shift back and clamp
            v = v + s_dm.shift;
            v = fmin(fmax(v, s_dm.ext_lo), s_dm.ext_hi);
        }
        if (cat) {
            const int c = (int)v;
            v = (double)(c < 0 ? 0 : (c >= s_dm.n_choices ? s_dm.n_choices - 1 : c));
        }
        w.value = v;
        win[d] = w;
    }
}

// Can the single-kernel path answer kb_tpe_suggest(n_cand) for the history now on the device?
bool kb_tpe_lean_fits(const kb_ctx *ctx, int64_t n_cand) {
    static const bool on = [] { const char *e = getenv("KB_SUGGEST_LEAN"); return !(e && e[0] == '0'); }();
    if (!on || ctx->n_trials > LN_MAX_N || n_cand > LN_MAX_CAND || n_cand < 1) return false;
    for (const KbDim &k : ctx->h_dims)
        if (k.kind == KB_CATEGORICAL && k.n_choices > LN_MAX_CHOICES) return false;
    return true;
}

// history on the device (kb_tpe_set_history / kb_tpe_append_history), tables not needed: launch + one 24 B x D read-back
int kb_tpe_suggest_lean(kb_ctx *ctx, int64_t n_cand, uint64_t seed, double *out_value) {
    const int n = (int)ctx->n_trials, D = ctx->D;
    cudaStream_t s = ctx->stream;
    if (!ctx->d_win) KB_CUDA(ctx, cudaMalloc(&ctx->d_win, sizeof(KbWinner) * (size_t)(D + 1)));
    if (!ctx->h_win) KB_CUDA(ctx, cudaMallocHost(&ctx->h_win, sizeof(KbWinner) * (size_t)(D + 1)));
    int cmax = 0;
    for (const KbDim &k : ctx->h_dims) if (k.kind == KB_CATEGORICAL) cmax = k.n_choices > cmax ? k.n_choices : cmax;
    LnPlan plan;
    plan.n = n > 0 ? n : 1; plan.km = n + 1; plan.la = n + 1 > cmax ? n + 1 : cmax; plan.nc = (int)n_cand;
    const size_t bytes = plan.bytes();
    if (bytes > 40 * 1024 && !ctx->lean_attr_set) {
        KB_CUDA(ctx, cudaFuncSetAttribute(tpe_suggest_lean_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
        ctx->lean_attr_set = true;
    }
    {
        KbProf pf(ctx, 3, s);
        static const int threads = [] { const char *e = getenv("KB_LEAN_THREADS"); const int t = e ? atoi(e) : 0; return t == 256 || t == 512 || t == 1024 ? t : 512; }();
        tpe_suggest_lean_kernel<<<(unsigned)D, threads, bytes, s>>>(ctx->d_dims, ctx->d_values, ctx->d_loss, n, D, ctx->n_below,
                                                                       ctx->tpe_merge ? 1 : 0, (int)n_cand, seed, plan, ctx->h_win);
        KB_LAUNCH_CHECK(ctx);
    }
    // the D winners (24 B each) are written by the kernel straight into the pinned host mirror (mapped into the device's
    // address space under unified addressing): no copy to enqueue, the stream synchronisation is the only wait
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    for (int d = 0; d < D; d++) out_value[d] = ctx->h_win[d].value;
    return KB_OK;
}
