// kb_noalloc.cu -- kb_tpe_scores_noalloc: TPE scoring of one continuous parameter for callers that cannot allocate
// GPU memory.
//
// Who needs this: a caller that is not allowed to allocate device memory at the moment it wants scores.  Typical
// cases: a plugin living inside a host application that owns all GPU memory through its own pool and forbids raw
// cudaMalloc; a process whose device memory was handed out in full at start-up; a latency-sensitive loop where
// cudaMalloc / cudaFree (which synchronise the device) must not appear.  Every other scoring entry point of this
// library keeps its scratch (staged candidates, folded tables) in buffers it cudaMalloc's through the kb_ctx; this
// one keeps its scratch in a fixed array that is part of the library image itself.
//
// The scratch: g_kb_noalloc_scratch below, a file-scope `static __device__` array of KB_NOALLOC_SCRATCH_BYTES
// (256 KiB + a 256-byte header).  Every GPU that loads the library carries that array from load until process exit.
// It bounds the problem: at most KB_NOALLOC_MAX_K mixture components per side; larger inputs are refused.
//
// Rules for callers
//   * The first call on a device resolves the array's device address (cudaGetSymbolAddress, with that device
//     current) and remembers it.  Make that first call before any stream capture begins.
//   * All calls on one device share the one array, and it keeps whatever the previous call left in it.  Nothing in it
//     needs to start at zero: each call rewrites the header and every table entry it will read before it reads them
//     (the fold kernel precedes the score kernel on the caller's stream).  For the same reason two calls that overlap
//     on the GPU (different streams, or two host threads) race on the tables: serialise calls per device.
//   * The entry point calls no cudaMalloc / cudaFree / cudaMallocAsync and no allocating library routine; inputs and
//     outputs are the caller's device pointers.
//
// Arithmetic: the same formulation as tpe_score_cont_kernel (kb_tpe_score.cu) -- normalised FP32 (hi, lo) inputs,
// 2^(k - t^2) per component with the fixed shift L = max_j log c_j, FP32 partial sums over 32 components collected
// into an (hi, lo) pair -- written separately here so that the allocating code path is untouched.
#include <math.h>

#include "kb_internal.cuh"

#define KB_NOALLOC_MAX_K 4096
#define KB_NOALLOC_HEADER_BYTES 256
#define KB_NOALLOC_TABLE_BYTES ((size_t)KB_NOALLOC_MAX_K * 2 * sizeof(float4))   /* per side: 128 KiB */
#define KB_NOALLOC_SCRATCH_BYTES (KB_NOALLOC_HEADER_BYTES + 2 * KB_NOALLOC_TABLE_BYTES)
#define KB_NOALLOC_MAX_DEVICES 64
#define NA_EPS 1e-12

// In host code this name is only a stand-in for the symbol; its device address comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_kb_noalloc_scratch[KB_NOALLOC_SCRATCH_BYTES];

struct NaHeader {   // first 256 bytes of the scratch
    double L[2];    // log-sum-exp shift per side
    int Kpad[2];    // components per side, padded to 32
};

#define NA_FOLD_THREADS 1024

// one CTA per side: p_accept, log-coefficients, their maximum, then the folded FP32 table
__global__ void __launch_bounds__(NA_FOLD_THREADS)
noalloc_fold_kernel(unsigned char *scratch, int K0, const double *w0, const double *mu0, const double *sg0, int K1,
                    const double *w1, const double *mu1, const double *sg1, double low, double high) {
    __shared__ double s_red[NA_FOLD_THREADS];
    const int side = blockIdx.x;
    const int K = side ? K1 : K0;
    const double *w = side ? w1 : w0, *mu = side ? mu1 : mu0, *sg = side ? sg1 : sg0;
    NaHeader *hdr = (NaHeader *)scratch;
    float4 *tab = (float4 *)(scratch + KB_NOALLOC_HEADER_BYTES + (size_t)side * KB_NOALLOC_TABLE_BYTES);
    const int tid = threadIdx.x;

    double acc = 0.0;
    for (int j = tid; j < K; j += NA_FOLD_THREADS) {
        double den = sqrt(2.0) * sg[j];
        den = den < NA_EPS ? NA_EPS : den;
        acc += w[j] * (0.5 * (1.0 + erf((high - mu[j]) / den)) - 0.5 * (1.0 + erf((low - mu[j]) / den)));
    }
    s_red[tid] = acc;
    __syncthreads();
    for (int off = NA_FOLD_THREADS / 2; off > 0; off >>= 1) {
        if (tid < off) s_red[tid] += s_red[tid + off];
        __syncthreads();
    }
    const double p_accept = s_red[0];
    __syncthreads();
    double lmax = -INFINITY;
    for (int j = tid; j < K; j += NA_FOLD_THREADS) {
        const double s = sg[j] > NA_EPS ? sg[j] : NA_EPS;  // the same floor the exponent uses below
        lmax = fmax(lmax, log(w[j] / (sqrt(2.0 * M_PI) * s) / p_accept));
    }
    s_red[tid] = lmax;
    __syncthreads();
    for (int off = NA_FOLD_THREADS / 2; off > 0; off >>= 1) {
        if (tid < off) s_red[tid] = fmax(s_red[tid], s_red[tid + off]);
        __syncthreads();
    }
    const double L = s_red[0];
    const int Kpad = (K + 31) / 32 * 32;
    if (tid == 0) { hdr->L[side] = L; hdr->Kpad[side] = Kpad; }
    const double inv_range = 1.0 / (high - low), LOG2E = 1.4426950408889634;
    for (int j = tid; j < Kpad; j += NA_FOLD_THREADS) {
        float4 A = make_float4(0.f, 0.f, 0.f, 0.f), B = make_float4(0.f, 0.f, -INFINITY, -INFINITY);
        if (j < K) {
            const double mun = (mu[j] - low) * inv_range;
            const float mh = (float)mun;
            const double ml = mun - (double)mh;
            const double s = sg[j] > NA_EPS ? sg[j] : NA_EPS;
            const float c = (float)(sqrt(0.5 * LOG2E) * (high - low) / s);
            const double lc = log(w[j] / (sqrt(2.0 * M_PI) * s) / p_accept);
            const float mc = (float)(-ml * (double)c), kk = (float)((lc - L) * LOG2E);
            A = make_float4(-mh, -mh, c, c);
            B = make_float4(mc, mc, kk, kk);
        }
        tab[2 * j] = A;
        tab[2 * j + 1] = B;
    }
}

#define NA_THREADS 256
#define NA_R 4
#define NA_CHUNK 256  // components per shared-memory stage (8 KB)

__global__ void __launch_bounds__(NA_THREADS)
noalloc_score_kernel(const unsigned char *scratch, const double *__restrict__ cand, long long n, double low, double high,
                     double *__restrict__ out) {
    __shared__ float4 s_tab[NA_CHUNK * 2];
    const NaHeader *hdr = (const NaHeader *)scratch;
    const long long base = (long long)blockIdx.x * (NA_THREADS * NA_R);
    const double inv = 1.0 / (high - low);
    float xh[NA_R], xl[NA_R];
#pragma unroll
    for (int r = 0; r < NA_R; r++) {
        const long long i = base + r * NA_THREADS + threadIdx.x;
        const double xn = i < n ? (cand[i] - low) * inv : 0.5;
        xh[r] = (float)xn;
        xl[r] = (float)(xn - (double)xh[r]);
    }
    double logS[2][NA_R];
    for (int side = 0; side < 2; side++) {
        const float4 *tab = (const float4 *)(scratch + KB_NOALLOC_HEADER_BYTES + (size_t)side * KB_NOALLOC_TABLE_BYTES);
        const int Kpad = hdr->Kpad[side];
        float Shi[NA_R], Slo[NA_R];
#pragma unroll
        for (int r = 0; r < NA_R; r++) { Shi[r] = 0.f; Slo[r] = 0.f; }
        for (int c0 = 0; c0 < Kpad; c0 += NA_CHUNK) {
            const int cnt = min(NA_CHUNK, Kpad - c0);
            __syncthreads();
            for (int k = threadIdx.x; k < cnt * 2; k += NA_THREADS) s_tab[k] = tab[(size_t)c0 * 2 + k];
            __syncthreads();
            for (int j0 = 0; j0 < cnt; j0 += 32) {
                float a32[NA_R];
#pragma unroll
                for (int r = 0; r < NA_R; r++) a32[r] = 0.f;
#pragma unroll 4
                for (int j = 0; j < 32; j++) {
                    const float4 A = s_tab[2 * (j0 + j)], B = s_tab[2 * (j0 + j) + 1];
#pragma unroll
                    for (int r = 0; r < NA_R; r++) {
                        const float t = fmaf((xh[r] + A.x) + xl[r], A.z, B.x);
                        a32[r] += kb_ex2(fmaf(-t, t, B.z));
                    }
                }
#pragma unroll
                for (int r = 0; r < NA_R; r++) {  // TwoSum into (hi, lo)
                    const float s = Shi[r] + a32[r], bb = s - Shi[r];
                    Slo[r] += (Shi[r] - (s - bb)) + (a32[r] - bb);
                    Shi[r] = s;
                }
            }
        }
#pragma unroll
        for (int r = 0; r < NA_R; r++) {
            // A candidate more than ~13 sigma from every component underflows the FP32 sum; the floor keeps the
            // logarithm finite (the score then reads L_below - L_above instead of NaN).  Tables that come from
            // kb_tpe_get_parzen never get here: their prior component spans the whole range.
            const double S = fmax((double)Shi[r] + (double)Slo[r], 1e-300);
            logS[side][r] = log(S) + hdr->L[side];
        }
    }
#pragma unroll
    for (int r = 0; r < NA_R; r++) {
        const long long i = base + r * NA_THREADS + threadIdx.x;
        if (i < n) out[i] = logS[0][r] - logS[1][r];
    }
}

// ---- set-up: once per device, with that device current ------------------------------------------------------------
static std::mutex g_na_mu;
static unsigned char *g_na_ptr[KB_NOALLOC_MAX_DEVICES] = {nullptr};

static int noalloc_scratch_address(int device, unsigned char **out) {
    std::lock_guard<std::mutex> lk(g_na_mu);
    if (!g_na_ptr[device]) {
        void *p = nullptr;
        cudaError_t e = cudaGetSymbolAddress(&p, g_kb_noalloc_scratch);  // `device` is current here (caller's guard)
        if (e != cudaSuccess || !p)
            return kb_fail(nullptr, KB_ERR_CUDA, "kb_tpe_scores_noalloc: cudaGetSymbolAddress on device %d: %s", device,
                           cudaGetErrorString(e));
        g_na_ptr[device] = (unsigned char *)p;
    }
    *out = g_na_ptr[device];
    return KB_OK;
}

extern "C" int kb_tpe_scores_noalloc(int device, int64_t n_cand, const double *cand_device, int K_below,
                                     const double *w_below_device, const double *mu_below_device,
                                     const double *sigma_below_device, int K_above, const double *w_above_device,
                                     const double *mu_above_device, const double *sigma_above_device, double low,
                                     double high, double *scores_device, void *stream) {
    if (device < 0 || device >= KB_NOALLOC_MAX_DEVICES)
        return kb_fail(nullptr, KB_ERR_INVALID, "kb_tpe_scores_noalloc: device %d outside [0,%d)", device, KB_NOALLOC_MAX_DEVICES);
    if (K_below < 1 || K_above < 1 || K_below > KB_NOALLOC_MAX_K || K_above > KB_NOALLOC_MAX_K)
        return kb_fail(nullptr, KB_ERR_UNSUPPORTED,
                       "kb_tpe_scores_noalloc: K_below=%d, K_above=%d; the static scratch holds at most %d components per side "
                       "(use kb_tpe_scores for larger histories)", K_below, K_above, KB_NOALLOC_MAX_K);
    if (n_cand < 0 || !(low < high)) return kb_fail(nullptr, KB_ERR_INVALID, "kb_tpe_scores_noalloc: n_cand < 0 or low >= high");
    if (!w_below_device || !mu_below_device || !sigma_below_device || !w_above_device || !mu_above_device ||
        !sigma_above_device || (n_cand > 0 && (!cand_device || !scores_device)))
        return kb_fail(nullptr, KB_ERR_INVALID, "kb_tpe_scores_noalloc: NULL device pointer");
    KbDeviceGuard g(device);
    unsigned char *scratch = nullptr;
    int rc = noalloc_scratch_address(device, &scratch);
    if (rc) return rc;
    if (n_cand == 0) return KB_OK;
    cudaStream_t s = (cudaStream_t)stream;
    noalloc_fold_kernel<<<2, NA_FOLD_THREADS, 0, s>>>(scratch, K_below, w_below_device, mu_below_device, sigma_below_device,
                                                      K_above, w_above_device, mu_above_device, sigma_above_device, low, high);
    const long long tile = NA_THREADS * NA_R;
    noalloc_score_kernel<<<(unsigned)((n_cand + tile - 1) / tile), NA_THREADS, 0, s>>>(scratch, cand_device, n_cand, low,
                                                                                      high, scores_device);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return kb_fail(nullptr, KB_ERR_CUDA, "kb_tpe_scores_noalloc: launch: %s", cudaGetErrorString(e));
    return KB_OK;
}
