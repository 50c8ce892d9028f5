// kb_cmaes.cu -- CMA-ES ask / tell on the GPU in FP64.
//
// Replaces goptuna's cmaes.Optimizer (a Go port of the `cmaes` library's CMA class; module not in the reference tree,
// go.mod:8; created and driven by cmaes.Sampler.SampleRelative, reference converter.go:41-68 and sample.go:39):
//   NewOptimizer                      -> kb_cmaes_create   (strategy constants, Hansen's tutorial eq. 48-56)
//   Optimizer.eigenDecomposition      -> cma_ensure_eig    (kb_eigh_device; D = sqrt(max(lambda, eps)); C <- B D^2 B^T)
//   Optimizer.Ask / sampleSolution    -> kb_cmaes_ask      (X = m + sigma * Z diag(D) B^T : one DMMA GEMM per batch)
//   Optimizer.Tell                    -> kb_cmaes_tell     (y_w, p_sigma, sigma, h_sigma, p_c, active weights, rank-one
//                                                           + rank-mu update; three DMMA GEMMs + small kernels)
// All dense contractions go through kb_dgemm_nt (kb_linalg.cu).  State lives in HBM; sigma and the generation counter
// are mirrored on the host because tell() needs a handful of scalars back anyway.
#include <math.h>

#include <algorithm>
#include <numeric>

#include "kb_internal.cuh"
#include "kb_linalg.cuh"

#define CMA_EPS 1e-8
#define CMA_SIGMA_MAX 1e32
#define CMA_MAX_RESAMPLE 100

struct KbCma {
    int n = 0, lam = 0, mu = 0, g = 0;
    double mu_eff, c1, cmu, c_sigma, d_sigma, cc, chi_n, cm, sum_w;
    double sigma = 0;
    bool have_eig = false, bounded = false, have_bt = false;
    int last_sweeps = 0;
    double *d_Bt = nullptr;
    int cap = 0;  // rows the batch buffers hold (>= lam)
    std::vector<double> h_w;
    double *d_mean = nullptr, *d_C = nullptr, *d_ps = nullptr, *d_pc = nullptr, *d_B = nullptr, *d_D = nullptr,
           *d_Dinv = nullptr, *d_D2 = nullptr, *d_lam = nullptr, *d_C2 = nullptr, *d_w = nullptr, *d_wio = nullptr,
           *d_yw = nullptr, *d_c2yw = nullptr, *d_lo = nullptr, *d_hi = nullptr, *d_scal = nullptr;
    double *d_X = nullptr, *d_Xt = nullptr, *d_Y = nullptr, *d_Yt = nullptr, *d_V = nullptr, *d_Z = nullptr;
    int *d_perm = nullptr, *d_flag = nullptr, *d_count = nullptr;
    void *d_eigws = nullptr;
    double *d_fall = nullptr, *d_wrow = nullptr;  // sharded tell: all ranks' objective values; this rank's row weights
    double *d_rot[3] = {nullptr, nullptr, nullptr};  // n x n temporaries of the rotated eigensolve (cma_ensure_eig)
    double *d_lam2 = nullptr;
};

static void cma_free(KbCma *c) {
    if (!c) return;
    double *ptrs[] = {c->d_mean, c->d_C, c->d_ps, c->d_pc, c->d_B, c->d_D, c->d_Dinv, c->d_D2, c->d_lam, c->d_C2, c->d_w,
                      c->d_wio, c->d_yw, c->d_c2yw, c->d_lo, c->d_hi, c->d_scal, c->d_X, c->d_Xt, c->d_Y, c->d_Yt, c->d_V, c->d_Z};
    for (double *p : ptrs) cudaFree(p);
    cudaFree(c->d_perm); cudaFree(c->d_flag); cudaFree(c->d_count); cudaFree(c->d_eigws); cudaFree(c->d_Bt);
    cudaFree(c->d_fall); cudaFree(c->d_wrow);
    for (double *p : c->d_rot) cudaFree(p);
    cudaFree(c->d_lam2);
    delete c;
}
int kb_cmaes_free(kb_ctx *ctx) {
    cma_free(ctx->cma);
    ctx->cma = nullptr;
    return KB_OK;
}

// ---- small kernels ---------------------------------------------------------------------------------------------
__global__ void cma_eye_kernel(int n, double *C) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < (size_t)n * n) C[i] = (i / n == i % n) ? 1.0 : 0.0;
}
__global__ void cma_d_from_lambda_kernel(int n, const double *lam, double *D, double *Dinv, double *D2) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double l = lam[i] < 0.0 ? CMA_EPS : lam[i];
    const double d = sqrt(l);
    D[i] = d; Dinv[i] = 1.0 / d; D2[i] = d * d;
}
// X[k,i] = mean[i] + X[k,i] (X holds sigma*y); optional clip; only rows with flag != 0 when flag is given
__global__ void cma_finish_x_kernel(int lam, int n, const double *mean, const double *lo, const double *hi, const double *Xt,
                                    double *X, int *flag, int *count, int final_clip) {
    const int k = blockIdx.x;
    if (flag && flag[k] == 0) return;
    __shared__ int s_bad;
    if (threadIdx.x == 0) s_bad = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        double x = mean[i] + Xt[(size_t)k * n + i];
        if (lo) {
            if (x < lo[i] || x > hi[i]) {
                if (final_clip) x = fmin(fmax(x, lo[i]), hi[i]); else atomicExch(&s_bad, 1);
            }
        }
        X[(size_t)k * n + i] = x;
    }
    __syncthreads();
    if (threadIdx.x == 0 && flag) {
        if (s_bad && !final_clip) atomicAdd(count, 1); else flag[k] = 0;
    }
}
struct CmaPhilox {
    uint32_t k0, k1;
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
// z ~ N(0, I) for flagged rows: Philox keyed by seed, counter (generation, row, element, attempt)
__global__ void cma_draw_z_kernel(int lam, int n, uint64_t seed, uint32_t gen, uint32_t attempt, const int *flag, double *Z,
                                  uint32_t row0) {
    const int k = blockIdx.x;
    if (flag && flag[k] == 0) return;
    CmaPhilox ph{(uint32_t)seed, (uint32_t)(seed >> 32)};
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const uint4 r = ph(gen, row0 + (uint32_t)k, (uint32_t)i, attempt);  // row0: first population row of this rank
        double u1 = (double)(((unsigned long long)r.x << 21) ^ (unsigned long long)(r.y >> 11)) * (1.0 / 9007199254740992.0);
        const double u2 = (double)(((unsigned long long)r.z << 21) ^ (unsigned long long)(r.w >> 11)) * (1.0 / 9007199254740992.0);
        u1 = u1 < 1e-300 ? 1e-300 : u1;
        Z[(size_t)k * n + i] = sqrt(-2.0 * log(u1)) * cospi(2.0 * u2);
    }
}
// Y[k,:] = (X[perm[k],:] - mean) / sigma
__global__ void cma_sorted_y_kernel(int lam, int n, const double *X, const int *perm, const double *mean, double sigma, double *Y) {
    const int k = blockIdx.x;
    const int src = perm ? perm[k] : k;
    for (int i = threadIdx.x; i < n; i += blockDim.x) Y[(size_t)k * n + i] = (X[(size_t)src * n + i] - mean[i]) / sigma;
}
// population sharded over ranks: this rank's share of y_w, rows in caller order with their rank-based weights
__global__ void cma_yw_rows_kernel(int rows, int n, const double *Y, const double *wrow, double *yw) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double acc = 0.0;
    for (int k = 0; k < rows; k++) { const double w = wrow[k]; if (w > 0.0) acc += Y[(size_t)k * n + i] * w; }
    yw[i] = acc;
}
// y += a * x
__global__ void cma_axpy_kernel(size_t count, double *y, const double *x, double a) {
    const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i < count) y[i] += a * x[i];
}
// yw[i] = sum_{k<mu} w_k Y[k,i]; mean += cm * sigma * yw
__global__ void cma_yw_kernel(int mu, int n, const double *Y, const double *w, double *yw, double *mean, double cm_sigma) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    double acc = 0.0;
    for (int k = 0; k < mu; k++) acc += Y[(size_t)k * n + i] * w[k];
    yw[i] = acc;
    mean[i] += cm_sigma * acc;
}
// out[i] = sum_j M[i,j] v[j]; one warp per row
__global__ void cma_gemv_kernel(int n, const double *M, const double *v, double *out) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= n) return;
    double acc = 0.0;
    for (int j = lane; j < n; j += 32) acc += M[(size_t)row * n + j] * v[j];
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) out[row] = acc;
}
// p_sigma update and its squared norm (single CTA)
__global__ void cma_psigma_kernel(int n, double *ps, const double *c2yw, double c_sigma, double coef, double *scal) {
    __shared__ double s_red[256];
    double acc = 0.0;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const double v = (1.0 - c_sigma) * ps[i] + coef * c2yw[i];
        ps[i] = v;
        acc += v * v;
    }
    s_red[threadIdx.x] = acc;
    __syncthreads();
    for (int off = 128; off > 0; off >>= 1) { if (threadIdx.x < off) s_red[threadIdx.x] += s_red[threadIdx.x + off]; __syncthreads(); }
    if (threadIdx.x == 0) scal[0] = s_red[0];
}
__global__ void cma_pc_kernel(int n, double *pc, const double *yw, double cc, double cpc) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) pc[i] = (1.0 - cc) * pc[i] + cpc * yw[i];
}
// wio[k] = w_k (w_k >= 0) or w_k * n / (||C^{-1/2} y_k||^2 + eps); one warp per row of V
__global__ void cma_wio_kernel(int lam, int n, const double *V, const double *w, double *wio) {
    const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (row >= lam) return;
    const double wk = w[row];
    if (wk >= 0.0) { if (lane == 0) wio[row] = wk; return; }
    double acc = 0.0;
    for (int j = lane; j < n; j += 32) { const double v = V[(size_t)row * n + j]; acc += v * v; }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (lane == 0) wio[row] = wk * ((double)n / (acc + CMA_EPS));
}
// C = keep * C + c1 * pc pc^T
__global__ void cma_rank_one_kernel(int n, double *C, const double *pc, double keep, double c1) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)n * n) return;
    const int i = (int)(idx / n), j = (int)(idx % n);
    C[idx] = keep * C[idx] + c1 * pc[i] * pc[j];
}

// ---- host side ---------------------------------------------------------------------------------------------------
#define CMA_ALLOC(ptr, count) KB_CUDA(ctx, cudaMalloc(&(ptr), sizeof(*(ptr)) * (size_t)(count)))

static int cma_ensure_rows(kb_ctx *ctx, KbCma *c, int rows) {
    if (rows <= c->cap) return KB_OK;
    double **bufs[] = {&c->d_X, &c->d_Xt, &c->d_Y, &c->d_Yt, &c->d_V, &c->d_Z};
    for (double **b : bufs) { cudaFree(*b); *b = nullptr; }
    cudaFree(c->d_flag); c->d_flag = nullptr;
    cudaFree(c->d_perm); c->d_perm = nullptr;
    c->cap = 0;
    for (double **b : bufs) CMA_ALLOC(*b, (size_t)rows * c->n);
    CMA_ALLOC(c->d_flag, rows);
    CMA_ALLOC(c->d_perm, rows);
    c->cap = rows;
    return KB_OK;
}

extern "C" int kb_cmaes_create(kb_ctx *ctx, int n, const double *mean, double sigma, int popsize, const double *lo,
                               const double *hi) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (n < 1 || n > 4096) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_cmaes_create: n=%d outside [1,4096]", n);
    if (!mean || !(sigma > 0)) return kb_fail(ctx, KB_ERR_INVALID, "kb_cmaes_create: mean NULL or sigma <= 0");
    if ((lo == nullptr) != (hi == nullptr)) return kb_fail(ctx, KB_ERR_INVALID, "kb_cmaes_create: lo/hi must both be given");
    if (popsize < 0 || popsize > (1 << 20)) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_cmaes_create: popsize outside [0, 2^20]");
    kb_cmaes_free(ctx);
    KbCma *c = new KbCma();
    ctx->cma = c;
    c->n = n;
    if (popsize <= 0) popsize = 4 + (int)floor(3.0 * log((double)n));  // eq.48
    if (popsize < 2) popsize = 2;
    c->lam = popsize;
    c->mu = popsize / 2;
    const int lam = popsize, mu = c->mu;
    std::vector<double> wp((size_t)lam);
    double sb = 0, sb2 = 0, sa = 0, sa2 = 0;
    for (int i = 0; i < lam; i++) {  // eq.49
        wp[(size_t)i] = log(((double)lam + 1.0) / 2.0) - log((double)(i + 1));
        if (i < mu) { sb += wp[(size_t)i]; sb2 += wp[(size_t)i] * wp[(size_t)i]; } else { sa += wp[(size_t)i]; sa2 += wp[(size_t)i] * wp[(size_t)i]; }
    }
    const double dn = (double)n, mu_eff = sb * sb / sb2, mu_eff_minus = sa * sa / sa2, alpha_cov = 2.0;
    const double c1 = alpha_cov / ((dn + 1.3) * (dn + 1.3) + mu_eff);
    const double cmu_b = alpha_cov * (mu_eff - 2.0 + 1.0 / mu_eff) / ((dn + 2.0) * (dn + 2.0) + alpha_cov * mu_eff / 2.0);
    const double cmu = std::min(1.0 - c1 - 1e-8, cmu_b);
    const double min_alpha = std::min({1.0 + c1 / cmu, 1.0 + (2.0 * mu_eff_minus) / (mu_eff + 2.0), (1.0 - c1 - cmu) / (dn * cmu)});
    double pos = 0, neg = 0;
    for (double v : wp) { if (v > 0) pos += v; else if (v < 0) neg += fabs(v); }
    c->h_w.resize((size_t)lam);
    c->sum_w = 0;
    for (int i = 0; i < lam; i++) {
        c->h_w[(size_t)i] = wp[(size_t)i] >= 0 ? 1.0 / pos * wp[(size_t)i] : min_alpha / neg * wp[(size_t)i];
        c->sum_w += c->h_w[(size_t)i];
    }
    c->cm = 1.0; c->mu_eff = mu_eff; c->c1 = c1; c->cmu = cmu;
    c->c_sigma = (mu_eff + 2.0) / (dn + mu_eff + 5.0);  // eq.55
    c->d_sigma = 1.0 + 2.0 * std::max(0.0, sqrt((mu_eff - 1.0) / (dn + 1.0)) - 1.0) + c->c_sigma;
    c->cc = (4.0 + mu_eff / dn) / (dn + 4.0 + 2.0 * mu_eff / dn);  // eq.56
    c->chi_n = sqrt(dn) * (1.0 - (1.0 / (4.0 * dn)) + 1.0 / (21.0 * (dn * dn)));
    c->sigma = sigma;
    c->bounded = lo != nullptr;
    const size_t nn = (size_t)n * n;
    CMA_ALLOC(c->d_mean, n); CMA_ALLOC(c->d_C, nn); CMA_ALLOC(c->d_ps, n); CMA_ALLOC(c->d_pc, n); CMA_ALLOC(c->d_B, nn);
    CMA_ALLOC(c->d_D, n); CMA_ALLOC(c->d_Dinv, n); CMA_ALLOC(c->d_D2, n); CMA_ALLOC(c->d_lam, n); CMA_ALLOC(c->d_C2, nn);
    CMA_ALLOC(c->d_w, lam); CMA_ALLOC(c->d_wio, lam); CMA_ALLOC(c->d_yw, n); CMA_ALLOC(c->d_c2yw, n); CMA_ALLOC(c->d_scal, 8);
    CMA_ALLOC(c->d_count, 1); CMA_ALLOC(c->d_Bt, nn);
    KB_CUDA(ctx, cudaMalloc(&c->d_eigws, kb_eigh_workspace_bytes(n)));
    if (c->bounded) {
        CMA_ALLOC(c->d_lo, n); CMA_ALLOC(c->d_hi, n);
        KB_CUDA(ctx, cudaMemcpy(c->d_lo, lo, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
        KB_CUDA(ctx, cudaMemcpy(c->d_hi, hi, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
    }
    KB_CUDA(ctx, cudaMemcpy(c->d_mean, mean, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
    KB_CUDA(ctx, cudaMemcpy(c->d_w, c->h_w.data(), sizeof(double) * (size_t)lam, cudaMemcpyHostToDevice));
    KB_CUDA(ctx, cudaMemset(c->d_ps, 0, sizeof(double) * (size_t)n));
    KB_CUDA(ctx, cudaMemset(c->d_pc, 0, sizeof(double) * (size_t)n));
    cma_eye_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, ctx->stream>>>(n, c->d_C);
    KB_LAUNCH_CHECK(ctx);
    int rc = cma_ensure_rows(ctx, c, lam);
    if (rc) return rc;
    KB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return KB_OK;
}

static int cma_ensure_eig(kb_ctx *ctx, KbCma *c, cudaStream_t s) {
    if (c->have_eig) return KB_OK;
    const int n = c->n;
    // warm start from the previous generation's eigenvectors (d_Bt holds B^T once a factorisation exists)
    // no warm start: the Cholesky-preconditioned path (kb_linalg.cu) needs fewer sweeps than a warm start from B_prev,
    // whose benefit vanishes while the eigenvalues are still clustered (early generations)
    // What does pay once a factorisation exists: solve in the previous eigenbasis.  A' = B^T C B is nearly diagonal (C moves
    // by a few per cent per generation), its Cholesky factor is nearly diagonal too, and the one-sided Jacobi iteration on
    // that factor -- the same cold, preconditioned solve, no V0 warm start -- is done in a few sweeps instead of a dozen.
    // Eigenvectors of C are then B W.  Four extra n^3 DMMA products (~0.3 ms at n = 512) against ~4 ms of sweeps.  While
    // the spectrum is still clustered A' is no more diagonal than C and the sweep count is that of the direct solve.
    static const bool rotate_ok = [] { const char *e = getenv("KB_CMA_ROTATE"); return !(e && e[0] == '0'); }();
    int rc;
    if (rotate_ok && c->have_bt && n >= 64) {
        const size_t nn = (size_t)n * n;
        for (double *&p : c->d_rot) if (!p) KB_CUDA(ctx, cudaMalloc(&p, sizeof(double) * nn));
        if (!c->d_lam2) KB_CUDA(ctx, cudaMalloc(&c->d_lam2, sizeof(double) * (size_t)n));
        double *T = c->d_rot[0], *A2 = c->d_rot[1], *W = c->d_rot[2];
        if ((rc = kb_dgemm_nt(ctx, n, n, n, c->d_Bt, n, nullptr, c->d_C, n, T, n, 1.0, 0.0, s))) return rc;   // T  = B^T C   (C symmetric)
        if ((rc = kb_dgemm_nt(ctx, n, n, n, T, n, nullptr, c->d_Bt, n, A2, n, 1.0, 0.0, s))) return rc;       // A' = T B
        if ((rc = kb_eigh_device(ctx, n, A2, 0.0, c->d_lam2, W, c->d_eigws, s, nullptr, &c->last_sweeps))) return rc;
        if ((rc = kb_transpose(ctx, n, n, W, T, s))) return rc;                                               // T  = W^T
        if ((rc = kb_dgemm_nt(ctx, n, n, n, T, n, nullptr, c->d_B, n, A2, n, 1.0, 0.0, s))) return rc;        // A2 = W^T B^T: rows = new eigenvectors
        if ((rc = kb_eigh_canonical(ctx, n, A2, c->d_lam2, c->d_lam, c->d_B, s))) return rc;                  // order, signs -> d_B, d_lam
    } else if ((rc = kb_eigh_device(ctx, n, c->d_C, 0.0, c->d_lam, c->d_B, c->d_eigws, s, nullptr, &c->last_sweeps))) return rc;
    if ((rc = kb_transpose(ctx, n, n, c->d_B, c->d_Bt, s))) return rc;
    c->have_bt = true;
    cma_d_from_lambda_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, c->d_lam, c->d_D, c->d_Dinv, c->d_D2);
    KB_LAUNCH_CHECK(ctx);
    // C <- B D^2 B^T (the reference rebuilds C from its factors)
    if ((rc = kb_dgemm_nt(ctx, n, n, n, c->d_B, n, c->d_D2, c->d_B, n, c->d_C, n, 1.0, 0.0, s))) return rc;
    c->have_eig = true;
    return KB_OK;
}

extern "C" int kb_cmaes_ask(kb_ctx *ctx, int lam, const double *z, uint64_t seed, double *x) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    KbCma *c = ctx->cma;
    if (!c) return kb_fail(ctx, KB_ERR_STATE, "kb_cmaes_ask: call kb_cmaes_create first");
    if (lam < 1 || lam > (1 << 20) || !x) return kb_fail(ctx, KB_ERR_INVALID, "kb_cmaes_ask: lam outside [1, 2^20] or x NULL");
    cudaStream_t s = ctx->stream;
    const int n = c->n;
    int rc = cma_ensure_rows(ctx, c, lam);
    if (rc) return rc;
    if ((rc = cma_ensure_eig(ctx, c, s))) return rc;
    KbProf pf(ctx, 7, s);
    if (z) {
        KB_CUDA(ctx, cudaMemcpyAsync(c->d_Z, z, sizeof(double) * (size_t)lam * n, cudaMemcpyHostToDevice, s));
        if ((rc = kb_dgemm_nt(ctx, lam, n, n, c->d_Z, n, c->d_D, c->d_B, n, c->d_Xt, n, c->sigma, 0.0, s))) return rc;
        cma_finish_x_kernel<<<lam, 128, 0, s>>>(lam, n, c->d_mean, c->d_lo, c->d_hi, c->d_Xt, c->d_X, nullptr, nullptr, 1);
        KB_LAUNCH_CHECK(ctx);
    } else {
        // Optimizer.Ask: resample infeasible points up to 100 times, then repair (clip)
        std::vector<int> ones((size_t)lam, 1);
        KB_CUDA(ctx, cudaMemcpyAsync(c->d_flag, ones.data(), sizeof(int) * (size_t)lam, cudaMemcpyHostToDevice, s));
        for (int attempt = 0; attempt <= CMA_MAX_RESAMPLE; attempt++) {
            const int last = attempt == CMA_MAX_RESAMPLE || !c->bounded;
            KB_CUDA(ctx, cudaMemsetAsync(c->d_count, 0, sizeof(int), s));
            int world = 1, rank = 0;
            kb_comm_world(ctx, &world, &rank);  // sharded population: every rank draws its own rows of the global stream
            cma_draw_z_kernel<<<lam, 128, 0, s>>>(lam, n, seed, (uint32_t)c->g, (uint32_t)attempt, c->d_flag, c->d_Z,
                                                  (uint32_t)(world > 1 ? rank * lam : 0));
            KB_LAUNCH_CHECK(ctx);
            if ((rc = kb_dgemm_nt(ctx, lam, n, n, c->d_Z, n, c->d_D, c->d_B, n, c->d_Xt, n, c->sigma, 0.0, s))) return rc;
            cma_finish_x_kernel<<<lam, 128, 0, s>>>(lam, n, c->d_mean, c->d_lo, c->d_hi, c->d_Xt, c->d_X, c->d_flag, c->d_count, last);
            KB_LAUNCH_CHECK(ctx);
            if (last) break;
            int remaining = 0;
            KB_CUDA(ctx, cudaMemcpyAsync(&remaining, c->d_count, sizeof(int), cudaMemcpyDeviceToHost, s));
            KB_CUDA(ctx, cudaStreamSynchronize(s));
            if (remaining == 0) break;
        }
    }
    KB_CUDA(ctx, cudaMemcpyAsync(x, c->d_X, sizeof(double) * (size_t)lam * n, cudaMemcpyDeviceToHost, s));
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    return KB_OK;
}

// Population sharded by rows over the ranks of the communicator: state, eigensystem and every
// n-sized update are replicated; a rank holds lam / world rows of X.  Exchanges per generation: one all-gather of the
// lam objective values (the ranking is global), one all-reduce of the n-vector y_w and one of the n x n rank-mu partial.
// Sums meet in a different order than on one GPU, so states agree to rounding (1e-13 relative), not bit for bit.
static int cma_tell_sharded(kb_ctx *ctx, KbCma *c, const double *x, const double *f, int world, int rank, cudaStream_t s) {
    const int n = c->n, lam = c->lam;
    if (lam % world) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_cmaes_tell: popsize %d is not a multiple of the %d ranks", lam, world);
    const int L = lam / world, r0 = rank * L;
    if (!c->d_fall) { CMA_ALLOC(c->d_fall, lam); CMA_ALLOC(c->d_wrow, lam); }
    int rc;
    c->g += 1;
    KB_CUDA(ctx, cudaMemcpyAsync(c->d_fall + r0, f, sizeof(double) * (size_t)L, cudaMemcpyHostToDevice, s));
    if ((rc = kb_comm_allgather(ctx, c->d_fall + r0, c->d_fall, sizeof(double) * (size_t)L, s))) return rc;
    std::vector<double> fall((size_t)lam);
    KB_CUDA(ctx, cudaMemcpyAsync(fall.data(), c->d_fall, sizeof(double) * (size_t)lam, cudaMemcpyDeviceToHost, s));
    KB_CUDA(ctx, cudaMemcpyAsync(c->d_X, x, sizeof(double) * (size_t)L * n, cudaMemcpyHostToDevice, s));
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    std::vector<int> perm((size_t)lam);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return fall[(size_t)a] < fall[(size_t)b]; });
    std::vector<double> wrow((size_t)L, 0.0);
    for (int pos = 0; pos < lam; pos++) {
        const int gidx = perm[(size_t)pos];
        if (gidx >= r0 && gidx < r0 + L) wrow[(size_t)(gidx - r0)] = c->h_w[(size_t)pos];
    }
    KB_CUDA(ctx, cudaMemcpyAsync(c->d_wrow, wrow.data(), sizeof(double) * (size_t)L, cudaMemcpyHostToDevice, s));
    cma_sorted_y_kernel<<<L, 128, 0, s>>>(L, n, c->d_X, nullptr, c->d_mean, c->sigma, c->d_Y);
    KB_LAUNCH_CHECK(ctx);
    cma_yw_rows_kernel<<<(n + 127) / 128, 128, 0, s>>>(L, n, c->d_Y, c->d_wrow, c->d_yw);
    KB_LAUNCH_CHECK(ctx);
    if ((rc = kb_comm_allreduce_sum_f64(ctx, c->d_yw, (size_t)n, s))) return rc;
    cma_axpy_kernel<<<(n + 127) / 128, 128, 0, s>>>((size_t)n, c->d_mean, c->d_yw, c->cm * c->sigma);  // eq.41
    KB_LAUNCH_CHECK(ctx);
    if ((rc = kb_dgemm_nt(ctx, n, n, n, c->d_B, n, c->d_Dinv, c->d_B, n, c->d_C2, n, 1.0, 0.0, s))) return rc;
    cma_gemv_kernel<<<(n * 32 + 255) / 256, 256, 0, s>>>(n, c->d_C2, c->d_yw, c->d_c2yw);
    KB_LAUNCH_CHECK(ctx);
    cma_psigma_kernel<<<1, 256, 0, s>>>(n, c->d_ps, c->d_c2yw, c->c_sigma, sqrt(c->c_sigma * (2.0 - c->c_sigma) * c->mu_eff), c->d_scal);
    KB_LAUNCH_CHECK(ctx);
    double norm2 = 0.0;
    KB_CUDA(ctx, cudaMemcpyAsync(&norm2, c->d_scal, sizeof(double), cudaMemcpyDeviceToHost, s));
    if ((rc = kb_dgemm_nt(ctx, L, n, n, c->d_Y, n, nullptr, c->d_C2, n, c->d_V, n, 1.0, 0.0, s))) return rc;
    cma_wio_kernel<<<(L * 32 + 255) / 256, 256, 0, s>>>(L, n, c->d_V, c->d_wrow, c->d_wio);  // eq.46
    KB_LAUNCH_CHECK(ctx);
    if ((rc = kb_transpose(ctx, L, n, c->d_Y, c->d_Yt, s))) return rc;
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    const double norm_ps = sqrt(norm2);
    c->sigma *= exp((c->c_sigma / c->d_sigma) * (norm_ps / c->chi_n - 1.0));
    if (c->sigma > CMA_SIGMA_MAX) c->sigma = CMA_SIGMA_MAX;
    const double hl = norm_ps / sqrt(1.0 - pow(1.0 - c->c_sigma, 2.0 * (double)(c->g + 1)));
    const double hr = (1.4 + 2.0 / ((double)n + 1.0)) * c->chi_n;
    const double h_sigma = hl < hr ? 1.0 : 0.0;
    cma_pc_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, c->d_pc, c->d_yw, c->cc, h_sigma * sqrt(c->cc * (2.0 - c->cc) * c->mu_eff));  // eq.45
    KB_LAUNCH_CHECK(ctx);
    const double delta_h = (1.0 - h_sigma) * c->cc * (2.0 - c->cc);
    const double keep = 1.0 + c->c1 * delta_h - c->c1 - c->cmu * c->sum_w;
    const size_t nn = (size_t)n * n;
    cma_rank_one_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, s>>>(n, c->d_C, c->d_pc, keep, c->c1);  // eq.47
    KB_LAUNCH_CHECK(ctx);
    // this rank's share of sum_k w_k^o y_k y_k^T (C^-1/2 is no longer needed: its buffer takes the partial)
    if ((rc = kb_dgemm_nt(ctx, n, n, L, c->d_Yt, L, c->d_wio, c->d_Yt, L, c->d_C2, n, 1.0, 0.0, s))) return rc;
    if ((rc = kb_comm_allreduce_sum_f64(ctx, c->d_C2, nn, s))) return rc;
    cma_axpy_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, s>>>(nn, c->d_C, c->d_C2, c->cmu);
    KB_LAUNCH_CHECK(ctx);
    c->have_eig = false;
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    return KB_OK;
}

extern "C" int kb_cmaes_tell(kb_ctx *ctx, const double *x, const double *f) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    KbCma *c = ctx->cma;
    if (!c) return kb_fail(ctx, KB_ERR_STATE, "kb_cmaes_tell: call kb_cmaes_create first");
    if (!x || !f) return kb_fail(ctx, KB_ERR_INVALID, "kb_cmaes_tell: NULL buffer");
    cudaStream_t s = ctx->stream;
    const int n = c->n, lam = c->lam, mu = c->mu;
    int rc = cma_ensure_eig(ctx, c, s);  // Tell uses the factors of the C the population was drawn from
    if (rc) return rc;
    KbProf pf(ctx, 8, s);
    {
        int world = 1, rank = 0;
        kb_comm_world(ctx, &world, &rank);
        if (world > 1) return cma_tell_sharded(ctx, c, x, f, world, rank, s);  // x, f: this rank's lam / world rows
    }
    c->g += 1;
    std::vector<int> perm((size_t)lam);
    std::iota(perm.begin(), perm.end(), 0);
    std::stable_sort(perm.begin(), perm.end(), [&](int a, int b) { return f[a] < f[b]; });
    KB_CUDA(ctx, cudaMemcpyAsync(c->d_X, x, sizeof(double) * (size_t)lam * n, cudaMemcpyHostToDevice, s));
    KB_CUDA(ctx, cudaMemcpyAsync(c->d_perm, perm.data(), sizeof(int) * (size_t)lam, cudaMemcpyHostToDevice, s));
    cma_sorted_y_kernel<<<lam, 128, 0, s>>>(lam, n, c->d_X, c->d_perm, c->d_mean, c->sigma, c->d_Y);
    KB_LAUNCH_CHECK(ctx);
    cma_yw_kernel<<<(n + 127) / 128, 128, 0, s>>>(mu, n, c->d_Y, c->d_w, c->d_yw, c->d_mean, c->cm * c->sigma);  // eq.41
    KB_LAUNCH_CHECK(ctx);
    // C^(-1/2) = B D^-1 B^T
    if ((rc = kb_dgemm_nt(ctx, n, n, n, c->d_B, n, c->d_Dinv, c->d_B, n, c->d_C2, n, 1.0, 0.0, s))) return rc;
    cma_gemv_kernel<<<(n * 32 + 255) / 256, 256, 0, s>>>(n, c->d_C2, c->d_yw, c->d_c2yw);
    KB_LAUNCH_CHECK(ctx);
    cma_psigma_kernel<<<1, 256, 0, s>>>(n, c->d_ps, c->d_c2yw, c->c_sigma, sqrt(c->c_sigma * (2.0 - c->c_sigma) * c->mu_eff), c->d_scal);
    KB_LAUNCH_CHECK(ctx);
    double norm2 = 0.0;
    KB_CUDA(ctx, cudaMemcpyAsync(&norm2, c->d_scal, sizeof(double), cudaMemcpyDeviceToHost, s));
    // meanwhile: V = Y C^(-1/2) (rows = C^(-1/2) y_k) and the transposed Y for the rank-mu product
    if ((rc = kb_dgemm_nt(ctx, lam, n, n, c->d_Y, n, nullptr, c->d_C2, n, c->d_V, n, 1.0, 0.0, s))) return rc;
    cma_wio_kernel<<<(lam * 32 + 255) / 256, 256, 0, s>>>(lam, n, c->d_V, c->d_w, c->d_wio);  // eq.46
    KB_LAUNCH_CHECK(ctx);
    if ((rc = kb_transpose(ctx, lam, n, c->d_Y, c->d_Yt, s))) return rc;
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    const double norm_ps = sqrt(norm2);
    c->sigma *= exp((c->c_sigma / c->d_sigma) * (norm_ps / c->chi_n - 1.0));
    if (c->sigma > CMA_SIGMA_MAX) c->sigma = CMA_SIGMA_MAX;
    const double hl = norm_ps / sqrt(1.0 - pow(1.0 - c->c_sigma, 2.0 * (double)(c->g + 1)));
    const double hr = (1.4 + 2.0 / ((double)n + 1.0)) * c->chi_n;
    const double h_sigma = hl < hr ? 1.0 : 0.0;
    cma_pc_kernel<<<(n + 127) / 128, 128, 0, s>>>(n, c->d_pc, c->d_yw, c->cc, h_sigma * sqrt(c->cc * (2.0 - c->cc) * c->mu_eff));  // eq.45
    KB_LAUNCH_CHECK(ctx);
    const double delta_h = (1.0 - h_sigma) * c->cc * (2.0 - c->cc);
    const double keep = 1.0 + c->c1 * delta_h - c->c1 - c->cmu * c->sum_w;
    const size_t nn = (size_t)n * n;
    cma_rank_one_kernel<<<(unsigned)((nn + 255) / 256), 256, 0, s>>>(n, c->d_C, c->d_pc, keep, c->c1);  // eq.47
    KB_LAUNCH_CHECK(ctx);
    if ((rc = kb_dgemm_nt(ctx, n, n, lam, c->d_Yt, lam, c->d_wio, c->d_Yt, lam, c->d_C, n, c->cmu, 1.0, s))) return rc;
    c->have_eig = false;
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    return KB_OK;
}

extern "C" int kb_cmaes_get_state(kb_ctx *ctx, double *mean, double *sigma, double *C, double *p_sigma, double *p_c,
                                  double *B, double *D, int *generation) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    KbCma *c = ctx->cma;
    if (!c) return kb_fail(ctx, KB_ERR_STATE, "kb_cmaes_get_state: call kb_cmaes_create first");
    cudaStream_t s = ctx->stream;
    const int n = c->n;
    const size_t nn = (size_t)n * n;
    if (B || D) { int rc = cma_ensure_eig(ctx, c, s); if (rc) return rc; }
    if (mean) KB_CUDA(ctx, cudaMemcpyAsync(mean, c->d_mean, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, s));
    if (C) KB_CUDA(ctx, cudaMemcpyAsync(C, c->d_C, sizeof(double) * nn, cudaMemcpyDeviceToHost, s));
    if (p_sigma) KB_CUDA(ctx, cudaMemcpyAsync(p_sigma, c->d_ps, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, s));
    if (p_c) KB_CUDA(ctx, cudaMemcpyAsync(p_c, c->d_pc, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, s));
    if (B) KB_CUDA(ctx, cudaMemcpyAsync(B, c->d_B, sizeof(double) * nn, cudaMemcpyDeviceToHost, s));
    if (D) KB_CUDA(ctx, cudaMemcpyAsync(D, c->d_D, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, s));
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    if (sigma) *sigma = c->sigma;
    if (generation) *generation = c->g;
    return KB_OK;
}

extern "C" int kb_cmaes_set_eig(kb_ctx *ctx, const double *B, const double *D) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    KbCma *c = ctx->cma;
    if (!c) return kb_fail(ctx, KB_ERR_STATE, "kb_cmaes_set_eig: call kb_cmaes_create first");
    if (!B || !D) return kb_fail(ctx, KB_ERR_INVALID, "kb_cmaes_set_eig: NULL buffer");
    const int n = c->n;
    std::vector<double> lam((size_t)n);
    for (int i = 0; i < n; i++) lam[(size_t)i] = D[i] * D[i];
    KB_CUDA(ctx, cudaMemcpy(c->d_B, B, sizeof(double) * (size_t)n * n, cudaMemcpyHostToDevice));
    KB_CUDA(ctx, cudaMemcpy(c->d_lam, lam.data(), sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));
    cma_d_from_lambda_kernel<<<(n + 127) / 128, 128, 0, ctx->stream>>>(n, c->d_lam, c->d_D, c->d_Dinv, c->d_D2);
    KB_LAUNCH_CHECK(ctx);
    KB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    KB_CUDA(ctx, cudaMemcpy(c->d_D, D, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice));  // keep the caller's exact D
    {   // B^T follows B (the next factorisation is solved in this basis)
        int rc = kb_transpose(ctx, n, n, c->d_B, c->d_Bt, ctx->stream);
        if (rc) return rc;
        KB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
        c->have_bt = true;
    }
    c->have_eig = true;
    return KB_OK;
}
