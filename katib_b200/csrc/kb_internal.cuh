// kb_internal.cuh -- shared state and helpers of libkatib_b200.so (not part of the ABI).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <mutex>
#include <string>
#include <vector>

#include "../../include/katib_b200.h"

// ---- per-parameter descriptor, resident in HBM (kb_ctx::d_dims) -----------------------------------
// Working domain of goptuna's sampleNumerical for the parameter: [lo, hi], quantum q (0 = continuous),
// shift = amount subtracted from stored history values (DISCRETE only).
struct KbDim {
    int32_t kind;
    int32_t n_choices;
    double lo, hi, q, shift;  // sampling range of sampleNumerical
    double ext_lo, ext_hi;    // distribution bounds (for the final clamp of DISCRETE)
    // Parzen tables (FP64, parity + FP64 re-scoring); capacity kcap entries per side
    int32_t K[2];             // components per side (n_side + 1)
    int32_t Kpad[2];          // padded to KB_TABLE_PAD
    double *mu[2], *sigma[2], *w[2], *logc[2];  // logc = log(w / (sqrt(2pi) sigma p_accept))
    double p_accept[2];
    double L[2];              // max_j logc (fixed log-sum-exp shift)
    // FP32 folded tables for the scoring kernels: two float4 per component, each value duplicated so that the
    // packed FP32x2 instructions take it straight from an LDS.128; Kpad entries
    //   continuous: {-mu_hi x2, c' x2} {-mu_lo*c' x2, k' x2}   c' = sqrt(log2e/2)/sigma_n, k' = (logc-L)*log2e
    //   quantised : {-mu_hi x2, r' x2} {-mu_lo*r' x2, +-w x2}  r' = sqrt(log2e)/(sqrt2 sigma_n); w<0: sigma_min class
    float4 *tab[2];
    // continuous parameters, what tpe_score_cont_kernel streams: the same four numbers once, {-mu_hi, c', -mu_lo*c', k'}
    // in ONE float4 per component (FADD2 / FFMA2 take a 32-bit register as a both-halves operand, so nothing needs
    // duplicating: half the shared-memory loads and TMA bytes of `tab`); Kpad entries
    float4 *tabc[2];
    // ... and what tpe_score_cont_local_kernel streams: {normalised mu in FP64 (low word, high word), c', k'} in the
    // narrow / wide partition of tabn / tabw below (narrow entries first, padded to KB_TABLE_PAD, then the wide ones);
    // every thread block turns a stage of it into FP32 entries about its own origin (see that kernel).  c0 = the c' all
    // narrow components share.
    float4 *tabs[2];
    float c0[2];
    // categorical: probabilities and the log-ratio table
    double *cat_p[2];
    double *cat_score;        // log pl[c] - log pg[c]
    int32_t *cat_rank;        // dense rank of cat_score (ties share a rank)
    int32_t numeric_slot;     // index among numeric dims (-1 for categorical)
    // quantised parameters, components whose sigma was clipped to the magic-clip floor (nearly all of them once the
    // history is dense): Phi(b)-Phi(a) with the fixed half width h0 is one function of y = |x-mu|/(sqrt2 sigma):
    //   erfc(y-h0) - erfc(y+h0) = 2^( -(y'-h0')^2 + R(tau) ),  y' = sqrt(log2 e) y, tau = y' * qf_tau_s - 1,
    // R = degree-KB_QF_DEG polynomial fitted per (parameter, side) at build time (Chebyshev-node interpolation on
    // y in [0, KB_QF_YMAX]; beyond it the term is < e^-29 of a component's peak and tau is clamped).
    float qf_coef[2][16];
    float qf_h0[2];           // h0' = sqrt(log2 e) * q / (2 sqrt2 sigma_min)
    float qf_tau_s[2];        // 2 / (sqrt(log2 e) * KB_QF_YMAX)
    float qf_hw[2];           // normalised half bin width q_n / 2 the fit assumes
    // ... and, of those, the components whose mu lies ON the quantisation grid (every observation of an integer /
    // stepped parameter does): for a candidate on the grid the bin integral depends only on the integer offset
    // k = (x - mu) / q, so it is a look-up in lut[side][k + G - 1], G = grid points, 2G - 1 entries (<= KB_LUT_MAX).
    // The component's second float4 then carries (grid index * 128 + 64) as an integer in .w (0 = not this class).
    float *lut[2];
    int32_t lut_G[2];         // 0: class switched off for this (parameter, side)
    // The same components once more, laid out for the grid kernel (tpe_score_quant_grid_kernel): the look-up-class
    // components' weights as a dense float array in table (mu-sorted) order, qw[side][Kn], cut into runs of equal grid
    // point, qruns[side][n_runs] = {grid code, first index into qw, length, 0} -- a run's table value is fetched once
    // per candidate and every component of it costs half a packed FFMA.  All other components (the prior, wide sigmas,
    // anything off the grid) sit in tabw[side][Kw] in the usual two-float4 form.  Kn / Kw hold the two counts.
    // Runs are padded with zero weights to multiples of four entries (Kn counts the padded array); a run entry is
    // {grid code, first index, padded length, real length}.  Other components whose mu is on the grid as well carry
    // 0x40000000 | row << 16 | grid index in their .w and own row `row` of lut2[side] (2G - 1 floats per row): their
    // bin integral at every integer grid offset, with their own sigma.
    float *qw[2];
    int4 *qruns[2];
    int32_t n_runs[2];
    float *lut2[2];
    // continuous parameters, KB_OPT_PRUNE_FAR: the same table split into the components whose sigma sits on the
    // magic-clip floor ("narrow", still sorted by mu, Kn entries padded to KB_TABLE_PAD) and the rest ("wide", Kw
    // entries, padded).  A narrow component further than prune_w (normalised) from a candidate contributes less than
    // e^-30 of its own peak and is skipped; wide components (the prior, sparse regions) are always evaluated.
    float4 *tabn[2], *tabw[2];
    int32_t Kn[2], Kw[2];     // real counts; the tables are padded to KB_TABLE_PAD with zero-contribution entries
    float prune_w[2];         // 7.75 sigma_floor of the side, normalised
};
#define KB_LUT_MAX 256        // floats per (parameter, side): grids of up to 128 points
#define KB_LUT2_FLOATS 8192   // floats per (quantised parameter, side) for the per-component rows (32 rows of a 128-point grid)
#ifndef KB_QF_DEG
#define KB_QF_DEG 11
#endif
#ifndef KB_QF_YMAX
#define KB_QF_YMAX 6.0
#endif

#define KB_TABLE_PAD 32

// near-tie table (per numeric parameter): open-addressing set keyed by the candidate's f64 bits
struct KbTie {
    unsigned long long key;     // f64 bits of the candidate value, ~0 = empty
    unsigned long long minidx;  // lowest global index seen with that value
};
#define KB_TIE_SLOTS_MIN (1u << 12)
#define KB_TIE_SLOTS_MAX (1u << 20)
#define KB_TIE_EMPTY 0xFFFFFFFFFFFFFFFFull

struct KbSelect {                // per-parameter running selection state (device)
    unsigned int gmax_bits;      // orderable bits of the running FP32 maximum
    unsigned int overflow;
    unsigned int tie_count;
    unsigned int pad;
    unsigned long long cat_best; // categorical: (rank << 32) | ~idx32
};

struct KbWinner {                // per-parameter result (device + host mirror)
    double score;
    long long index;
    double value;
};

#define KB_PROF_N 16
struct KbComm;  // kb_comm.cu
struct KbCma;   // kb_cmaes.cu

struct kb_ctx {
    int device = 0;
    cudaDeviceProp prop{};
    cudaStream_t stream = nullptr;   // compute
    cudaStream_t copy_stream = nullptr;
    std::mutex mu;
    mutable std::string err;
    int64_t launches = 0;

    // search space
    int D = 0;
    std::vector<KbDim> h_dims;
    KbDim *d_dims = nullptr;
    int n_numeric = 0, n_cont = 0, n_quant = 0, n_cat = 0;
    std::vector<int> cont_dims, quant_dims, cat_dims;
    int *d_cont_dims = nullptr, *d_quant_dims = nullptr, *d_cat_dims = nullptr;

    // history
    int64_t n_trials = 0;
    double *d_values = nullptr, *d_loss = nullptr;
    int64_t hist_cap = 0;
    int n_below = 0;
    int32_t *d_pos = nullptr;        // [2][n]: position of trial i inside below / above, or -1
    double *d_sort_keys = nullptr;   // scratch for the bitonic sorts
    uint32_t *d_sort_idx = nullptr;
    int64_t sort_cap = 0;
    double *d_small_keys = nullptr;  // per-CTA scratch of the single-launch build (small histories)
    uint32_t *d_small_idx = nullptr;
    int32_t *d_small_pos = nullptr;
    int64_t small_cap = 0;
    double *d_tabmem = nullptr;      // arena behind the KbDim table pointers
    size_t tabmem_bytes = 0;
    bool history_ready = false;
    bool tables_dirty = false;       // history uploaded, tables not built yet (kb_tpe_ensure_tables)
    void *h_up = nullptr;            // pinned staging of the history upload (grow-only) and the event behind its copies
    size_t up_bytes = 0;
    cudaEvent_t ev_up = nullptr;
    std::vector<char> h_dims_sent;   // the descriptor table as last uploaded (d_dims is rewritten only when it differs)
    bool tpe_merge = false;          // KB_OPT_MERGE_DUPLICATES
    bool tpe_prune = false;          // KB_OPT_PRUNE_FAR
    int tpe_local = 1;               // KB_OPT_LOCAL_ORIGIN: 0 never, 1 when a chunk has >= KB_LOCAL_MIN_ROWS rows, 2 always
    bool lean_attr_set = false;      // dynamic shared-memory limit of the single-kernel suggest raised
    bool cont_attr_set = false;      // ... of tpe_score_cont_kernel
    bool local_attr_set = false;     // ... of tpe_score_cont_local_kernel
    bool quant_attr_set = false;     // dynamic shared-memory limit of the quantised score kernels raised
    float2 *d_stage_sorted = nullptr;  // [n_cont][chunk] stage_cont in bucket order (pruned scoring)
    int32_t *d_perm = nullptr;         // [n_cont][chunk] chunk-local index of the candidate at each sorted position
    int32_t *d_bucket = nullptr;       // [n_cont][KB_PRUNE_BUCKETS] histogram / running offsets

    // scoring scratch
    int64_t chunk_cap = 0;           // candidates per staged chunk
    double *d_cand_chunk[2] = {nullptr, nullptr};   // H2D landing buffers (row-major f64)
    float2 *d_stage_cont = nullptr;  // [n_cont][chunk]   (x_hi, x_lo) normalised
    float4 *d_stage_quant = nullptr; // [n_quant][chunk]  (lb_hi, lb_lo, ub_hi, ub_lo) normalised
    int *d_qflags = nullptr;         // [n_quant][tiles]  1 = tile left to the general quantised kernel by the grid kernel
    double *d_scores_chunk = nullptr; // [chunk x D] dense scores (kb_tpe_scores only)
    KbSelect *d_sel = nullptr;       // [D]
    KbTie *d_tie = nullptr;          // [n_numeric][tie_slots]
    KbWinner *d_win = nullptr;       // [D]
    KbWinner *d_win_all = nullptr;   // [world][D] (gather target)
    float *d_tie_score = nullptr;    // [n_numeric][tie_slots] FP32 score stored with each near-tie entry
    unsigned int tie_slots = 0;      // slots per numeric parameter in d_tie (power of two, sized from n_cand)
    unsigned int tie_slots_floor = 0;  // lower bound on tie_slots, raised after an overflow
    void *d_tie_list = nullptr;      // [n_numeric][tie_slots] KbCand: compacted near ties awaiting FP64 re-scoring
    unsigned int *d_tie_count = nullptr;  // [n_numeric]
    int *d_slot_of_dim = nullptr;    // [D] index of a parameter inside its cont / quant / cat list
    int *d_numeric_dims = nullptr;   // [n_numeric] parameter index per numeric slot
    KbWinner *h_win = nullptr;       // pinned host mirror of d_win
    cudaEvent_t ev_h2d[2] = {nullptr, nullptr}, ev_done[2] = {nullptr, nullptr};
    cudaEvent_t ev_build = nullptr;  // recorded on `stream` after every history build; caller streams wait on it

    uint32_t *d_sobol_v = nullptr;   // [32][KB_SOBOL_MAX_DIM] direction numbers, bit-major
    double *d_sobol_buf[2] = {nullptr, nullptr};   // kb_sobol_fill: alternating chunk buffers (grow-only)
    double *h_sobol_pin[2] = {nullptr, nullptr};   // ... and pinned staging for pageable destinations
    size_t sobol_buf_bytes = 0, sobol_pin_bytes = 0;

    KbComm *comm = nullptr;
    KbCma *cma = nullptr;
    int last_eigh_sweeps = 0;        // Jacobi sweeps of the most recent factorisation (kb_eigh_stats)

    // per-kernel-class CUDA-event timing (kb_profile / kb_profile_read); off by default
    bool profile = false;
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> prof_pending[KB_PROF_N];
    std::vector<cudaEvent_t> prof_pool;
    double prof_ms[KB_PROF_N] = {0};
    int64_t prof_launches[KB_PROF_N] = {0};
};

// records an event pair around one launch on stream s when profiling is on
struct KbProf {
    kb_ctx *ctx; int id; cudaStream_t s; cudaEvent_t e0 = nullptr, e1 = nullptr;
    KbProf(kb_ctx *c, int id_, cudaStream_t s_) : ctx(c), id(id_), s(s_) {
        if (!ctx->profile) return;
        auto take = [&]() { cudaEvent_t e; if (!ctx->prof_pool.empty()) { e = ctx->prof_pool.back(); ctx->prof_pool.pop_back(); } else cudaEventCreate(&e); return e; };
        e0 = take(); e1 = take();
        cudaEventRecord(e0, s);
    }
    ~KbProf() {
        if (!e0) return;
        cudaEventRecord(e1, s);
        ctx->prof_pending[id].push_back({e0, e1});
    }
};

// ---- error plumbing ---------------------------------------------------------------------------------
int kb_fail(const kb_ctx *ctx, int code, const char *fmt, ...);
#define KB_CUDA(ctx, call)                                                                         \
    do {                                                                                           \
        cudaError_t e__ = (call);                                                                  \
        if (e__ != cudaSuccess)                                                                    \
            return kb_fail(ctx, KB_ERR_CUDA, "%s:%d %s -> %s", __FILE__, __LINE__, #call,          \
                           cudaGetErrorString(e__));                                               \
    } while (0)
#define KB_LAUNCH_CHECK(ctx)                                                                       \
    do {                                                                                           \
        (ctx)->launches++;                                                                         \
        KB_CUDA(ctx, cudaGetLastError());                                                          \
    } while (0)

struct KbDeviceGuard {
    int prev = -1;
    explicit KbDeviceGuard(int dev) { cudaGetDevice(&prev); if (prev != dev) cudaSetDevice(dev); else prev = -1; }
    ~KbDeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

// A caller-supplied stream is not ordered after work queued on the context's own (non-blocking) stream: make it wait for
// the most recent history build before anything on it reads the tables.
static inline cudaError_t kb_order_after_build(kb_ctx *ctx, cudaStream_t s) {
    if (s == ctx->stream || !ctx->ev_build) return cudaSuccess;
    return cudaStreamWaitEvent(s, ctx->ev_build, 0);
}

// cross-file entry points
int kb_tpe_free(kb_ctx *ctx);
int kb_sobol_free(kb_ctx *ctx);
int kb_cmaes_free(kb_ctx *ctx);
int kb_comm_free(kb_ctx *ctx);
int kb_comm_allgather(kb_ctx *ctx, const void *send, void *recv, size_t bytes, cudaStream_t s);
int kb_comm_allreduce_sum_f64(kb_ctx *ctx, double *buf, size_t count, cudaStream_t s);
int kb_tpe_ensure_tables(kb_ctx *ctx);
bool kb_tpe_lean_fits(const kb_ctx *ctx, int64_t n_cand);
int kb_tpe_suggest_lean(kb_ctx *ctx, int64_t n_cand, uint64_t seed, double *out_value);
int kb_tpe_suggest_range(kb_ctx *ctx, uint64_t seed, int64_t first, int64_t count, int64_t *out_index, double *out_value,
                         double *out_score);

// ---- device helpers ------------------------------------------------------------------------------------
#ifdef __CUDACC__
__device__ __forceinline__ unsigned int kb_f32_orderable(float f) {
    unsigned int b = __float_as_uint(f);
    return (b & 0x80000000u) ? ~b : (b | 0x80000000u);
}
__device__ __forceinline__ float kb_f32_from_orderable(unsigned int u) {
    unsigned int b = (u & 0x80000000u) ? (u & 0x7FFFFFFFu) : ~u;
    return __uint_as_float(b);
}
__device__ __forceinline__ float kb_ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
__device__ __forceinline__ float kb_lg2(float x) {
    float y;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
#endif
